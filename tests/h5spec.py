"""Independent structural validator for the HDF5 subset the pipeline exchanges, written from the HDF5 File Format
Specification (version 0 superblock, version 1 group B-trees, local heaps, symbol nodes, version 1 object headers).

Shares no code with ``expecto_b200/hdf5.py``: it exists so that the writer is checked by something other than the
repo's own reader.  ``validate(path)`` raises AssertionError naming the violated rule and returns
``{name: {"shape", "dtype", "address", "size", "header"}}``.  The validator is itself pinned on a file libhdf5 wrote:
tests/test_hdf5_cpu.py runs it over the reference's example/example.vcf.shift_0.diff.h5 (h5py output).
"""
from __future__ import annotations

import os
import struct

UNDEF = 0xFFFFFFFFFFFFFFFF
SIGNATURE = bytes([0x89, 0x48, 0x44, 0x46, 0x0D, 0x0A, 0x1A, 0x0A])


def _u(buf, off, n):
    return int.from_bytes(buf[off:off + n], "little")


class _Extents:
    """Every structure's byte range: must lie inside the file and overlap no other."""

    def __init__(self, eof):
        self.eof, self.items = eof, []

    def add(self, what, lo, size):
        assert size >= 0 and lo >= 0 and lo + size <= self.eof, f"{what}: [{lo:#x}, {lo + size:#x}) outside the file ({self.eof:#x})"
        for w2, a, b in self.items:
            assert lo + size <= a or lo >= b or size == 0 or a == b, f"{what} [{lo:#x},{lo + size:#x}) overlaps {w2} [{a:#x},{b:#x})"
        self.items.append((what, lo, lo + size))


def _cstr(heap_data: bytes, off: int) -> str:
    assert 0 <= off < len(heap_data), f"heap offset {off} outside the data segment ({len(heap_data)})"
    end = heap_data.find(b"\x00", off)
    assert end >= 0, "heap string is not NUL-terminated inside the data segment"
    return heap_data[off:end].decode("ascii")


def _object_header_v1(buf, addr, ext, what):
    """III.C / IV.A.1: version 1 prefix = version, reserved, #messages (2), refcount (4), header size (4), then the
    messages start at the next 8-byte boundary; each message = type (2), size (2), flags (1), 3 reserved, data padded
    to 8.  Returns [(type, flags, data)] and checks that the messages fill the declared header size exactly."""
    assert buf[addr] == 1, f"{what}: object header version {buf[addr]} != 1"
    assert buf[addr + 1] == 0, f"{what}: reserved byte not zero"
    nmsg, refc, hsize = _u(buf, addr + 2, 2), _u(buf, addr + 4, 4), _u(buf, addr + 8, 4)
    assert refc == 1, f"{what}: object reference count {refc} != 1"
    assert addr % 8 == 0, f"{what}: object header at {addr:#x} is not 8-byte aligned"
    ext.add(what, addr, 16 + hsize)
    pos, end, msgs = addr + 16, addr + 16 + hsize, []
    for i in range(nmsg):
        assert pos + 8 <= end, f"{what}: message {i} starts beyond the header chunk"
        mtype, msize, flags = _u(buf, pos, 2), _u(buf, pos + 2, 2), buf[pos + 4]
        assert buf[pos + 5:pos + 8] == b"\x00\x00\x00", f"{what}: message {i} reserved bytes not zero"
        assert msize % 8 == 0, f"{what}: message {i} (type {mtype:#x}) size {msize} is not a multiple of 8"
        assert pos + 8 + msize <= end, f"{what}: message {i} (type {mtype:#x}) overruns the header chunk"
        assert mtype != 0x10, f"{what}: continuation messages are not expected in this subset"
        msgs.append((mtype, flags, bytes(buf[pos + 8:pos + 8 + msize])))
        pos += 8 + msize
    assert pos == end, f"{what}: messages end at {pos:#x}, header chunk ends at {end:#x} (gap must be a NIL message)"
    return msgs


def _dataspace(data, what):
    """IV.A.2.b version 1: version, rank, flags (bit 0: max dims present; bit 1: permutation), 5 reserved, dims, max."""
    assert data[0] == 1, f"{what}: dataspace version {data[0]}"
    rank, flags = data[1], data[2]
    assert flags in (0, 1), f"{what}: dataspace flags {flags}"
    need = 8 + 8 * rank * (2 if flags & 1 else 1)
    assert len(data) >= need, f"{what}: dataspace message too short"
    dims = [_u(data, 8 + 8 * i, 8) for i in range(rank)]
    if flags & 1:
        mx = [_u(data, 8 + 8 * rank + 8 * i, 8) for i in range(rank)]
        assert all(m == UNDEF or m >= d for m, d in zip(mx, dims)), f"{what}: max dims below dims"
    return tuple(dims)


def _datatype(data, what):
    """IV.A.2.d version 1: class+version byte, 3 bytes of class bit fields, size (4), properties."""
    cls, ver = data[0] & 0x0F, data[0] >> 4
    assert ver == 1, f"{what}: datatype version {ver}"
    size = _u(data, 4, 4)
    if cls == 1:  # floating point: bit field byte 0: bit 0 byte order, bits 1-3 padding, bits 4-5 mantissa norm
        b0, sign_loc = data[1], data[2]
        assert b0 & 1 == 0, f"{what}: big-endian floats are not what numpy '<f' writes"
        assert (b0 >> 4) & 3 == 2, f"{what}: mantissa normalisation {(b0 >> 4) & 3} != 2 (implied leading 1)"
        bit_off, prec = _u(data, 8, 2), _u(data, 10, 2)
        exp_loc, exp_size, man_loc, man_size, bias = data[12], data[13], data[14], data[15], _u(data, 16, 4)
        ieee = {4: (31, 32, 23, 8, 0, 23, 127), 8: (63, 64, 52, 11, 0, 52, 1023)}
        assert size in ieee, f"{what}: float size {size}"
        assert (sign_loc, prec, exp_loc, exp_size, man_loc, man_size, bias) == ieee[size] and bit_off == 0, \
            f"{what}: not IEEE binary{8 * size}"
        return f"<f{size}", size
    if cls == 3:  # string: bits 0-3 padding type, bits 4-7 character set
        assert data[1] & 0x0F in (0, 1, 2) and data[1] >> 4 in (0, 1), f"{what}: string padding / charset"
        assert size > 0, f"{what}: zero-size string"
        return f"S{size}", size
    raise AssertionError(f"{what}: datatype class {cls} not in this subset")


def validate(path: str) -> dict:
    buf = open(path, "rb").read()
    size = len(buf)
    # ---- II.A superblock version 0
    assert buf[:8] == SIGNATURE, "format signature"
    assert buf[8] == 0, "superblock version must be 0"
    assert buf[9] == 0 and buf[10] == 0 and buf[11] == 0 and buf[12] == 0, "free-space / root-entry / shared-header versions"
    so, sl = buf[13], buf[14]
    assert (so, sl) == (8, 8), f"size of offsets/lengths {so}/{sl}"
    assert buf[15] == 0, "reserved"
    leaf_k, internal_k = _u(buf, 16, 2), _u(buf, 18, 2)
    assert leaf_k >= 1 and internal_k >= 1, "group K values"
    assert _u(buf, 20, 4) == 0, "file consistency flags"
    base, free_addr, eof, drv = (_u(buf, 24 + 8 * i, 8) for i in range(4))
    assert base == 0, "base address"
    assert free_addr == UNDEF, "free-space info address must be undefined"
    assert eof == size, f"end-of-file address {eof:#x} != file size {size:#x}"
    assert drv == UNDEF, "driver information block must be undefined"
    ext = _Extents(size)
    ext.add("superblock", 0, 24 + 32 + 40)
    # root group symbol table entry (III.C): name offset, header address, cache type, reserved, 16-byte scratch
    ste = 56
    assert _u(buf, ste, 8) == 0, "root entry link name offset"
    root_hdr, cache = _u(buf, ste + 8, 8), _u(buf, ste + 16, 4)
    assert _u(buf, ste + 20, 4) == 0, "root entry reserved"
    assert cache == 1, "root entry cache type must be 1 (group: B-tree and heap addresses cached)"
    btree, heap = _u(buf, ste + 24, 8), _u(buf, ste + 32, 8)
    # ---- root object header: one symbol table message agreeing with the cached addresses
    msgs = _object_header_v1(buf, root_hdr, ext, "root object header")
    stm = [m for m in msgs if m[0] == 0x11]
    assert len(stm) == 1, "root group needs exactly one symbol table message"
    assert (_u(stm[0][2], 0, 8), _u(stm[0][2], 8, 8)) == (btree, heap), "symbol table message != cached scratch pad"
    # ---- III.D local heap
    assert buf[heap:heap + 4] == b"HEAP", "local heap signature"
    assert buf[heap + 4] == 0 and buf[heap + 5:heap + 8] == b"\x00\x00\x00", "local heap version / reserved"
    seg_size, free_head, seg_addr = _u(buf, heap + 8, 8), _u(buf, heap + 16, 8), _u(buf, heap + 24, 8)
    ext.add("local heap header", heap, 32)
    ext.add("local heap data", seg_addr, seg_size)
    assert seg_size % 8 == 0 and seg_addr % 8 == 0, "heap data segment alignment"
    heap_data = bytes(buf[seg_addr:seg_addr + seg_size])
    free_blocks, off, guard = [], free_head, 0
    while off != 1 and off != UNDEF:
        assert off % 8 == 0 and off + 16 <= seg_size, f"heap free block at {off} outside the segment"
        nxt, fsz = _u(heap_data, off, 8), _u(heap_data, off + 8, 8)
        assert fsz >= 16 and off + fsz <= seg_size, f"heap free block at {off}: size {fsz}"
        free_blocks.append((off, off + fsz))
        off, guard = nxt, guard + 1
        assert guard < 1024, "heap free list loops"
    assert heap_data[0:1] == b"\x00", "heap offset 0 must hold the empty string (root / key 0 name)"

    def in_free(o):
        return any(a <= o < b for a, b in free_blocks)

    # ---- III.A.1 version 1 B-tree, node type 0 (group nodes)
    names_seen = []

    def walk(addr, depth):
        assert depth < 8, "B-tree too deep"
        assert buf[addr:addr + 4] == b"TREE", f"B-tree node signature at {addr:#x}"
        ntype, level, used = buf[addr + 4], buf[addr + 5], _u(buf, addr + 6, 2)
        assert ntype == 0, "B-tree node type must be 0 (group)"
        assert 1 <= used <= 2 * internal_k, f"B-tree entries used {used} (K = {internal_k})"
        left, right = _u(buf, addr + 8, 8), _u(buf, addr + 16, 8)
        if depth == 0:
            assert left == UNDEF and right == UNDEF, "root B-tree node has no siblings"
        ext.add(f"B-tree node {addr:#x}", addr, 24 + 8 * (2 * internal_k + 1) + 8 * 2 * internal_k)
        keys = [_u(buf, addr + 24 + 16 * i, 8) for i in range(used + 1)]
        kids = [_u(buf, addr + 32 + 16 * i, 8) for i in range(used)]
        knames = [_cstr(heap_data, k) for k in keys]
        assert all(a < b for a, b in zip(knames, knames[1:])), f"B-tree keys not increasing: {knames}"
        sub = []
        for i, child in enumerate(kids):
            got = walk(child, depth + 1) if level > 0 else snod(child)
            sub.extend(got)
            assert got, "empty child"
            assert knames[i] < got[0] and got[-1] <= knames[i + 1], \
                f"child {i} names {got[0]!r}..{got[-1]!r} outside its key interval ({knames[i]!r}, {knames[i + 1]!r}]"
            if level == 0:
                assert got[-1] == knames[i + 1], "right key of a leaf must be its symbol node's largest name"
        return sub

    entries = {}

    def snod(addr):
        assert buf[addr:addr + 4] == b"SNOD", f"symbol node signature at {addr:#x}"
        assert buf[addr + 4] == 1 and buf[addr + 5] == 0, "symbol node version"
        n = _u(buf, addr + 6, 2)
        assert 1 <= n <= 2 * leaf_k, f"symbol node holds {n} symbols (leaf K = {leaf_k})"
        ext.add(f"symbol node {addr:#x}", addr, 8 + 2 * leaf_k * 40)
        got = []
        for i in range(n):
            e = addr + 8 + 40 * i
            noff, hdr, ctype = _u(buf, e, 8), _u(buf, e + 8, 8), _u(buf, e + 16, 4)
            assert _u(buf, e + 20, 4) == 0, "symbol entry reserved"
            assert not in_free(noff), f"name offset {noff} lies in a free heap block"
            name = _cstr(heap_data, noff)
            assert name, "empty link name"
            assert ctype == 0, f"{name}: cache type {ctype} (datasets cache nothing)"
            assert buf[e + 24:e + 40] == bytes(16), f"{name}: scratch pad not zero"
            got.append(name)
            entries[name] = hdr
        assert all(a < b for a, b in zip(got, got[1:])), f"symbol node names not strictly increasing: {got}"
        unused = buf[addr + 8 + 40 * n: addr + 8 + 40 * 2 * leaf_k]
        assert unused == bytes(len(unused)), "unused symbol node entries must be zero"
        names_seen.extend(got)
        return got

    walk(btree, 0)
    assert len(set(names_seen)) == len(names_seen), "duplicate link names"
    # ---- dataset object headers
    out = {}
    for name, hdr in entries.items():
        what = f"dataset {name!r}"
        msgs = _object_header_v1(buf, hdr, ext, what + " header")
        kinds = [m[0] for m in msgs]
        for need, label in ((1, "dataspace"), (3, "datatype"), (5, "fill value"), (8, "layout")):
            assert kinds.count(need) == 1, f"{what}: needs exactly one {label} message"
        shape = dtype = None
        for mtype, flags, data in msgs:
            if mtype == 1:
                shape = _dataspace(data, what)
            elif mtype == 3:
                dtype, elsize = _datatype(data, what)
                assert flags & 1, f"{what}: datatype message must be flagged constant"
            elif mtype == 5:
                assert data[0] in (1, 2, 3), f"{what}: fill value version {data[0]}"
                if data[0] == 2:
                    assert data[1] in (1, 2, 3) and data[2] in (0, 1, 2) and data[3] in (0, 1), f"{what}: fill value fields"
                    if data[3] == 1:
                        assert _u(data, 4, 4) + 8 <= len(data), f"{what}: fill value size"
            elif mtype == 0x12:
                assert data[0] == 1 and data[1:4] == b"\x00\x00\x00", f"{what}: modification time message version"
            elif mtype == 0:
                pass  # NIL: padding
            elif mtype != 8:
                raise AssertionError(f"{what}: unexpected message type {mtype:#x}")
        lay = [m for m in msgs if m[0] == 8][0][2]
        assert lay[0] == 3, f"{what}: layout version {lay[0]} != 3"
        assert lay[1] == 1, f"{what}: layout class {lay[1]} != 1 (contiguous)"
        daddr, dsize = _u(lay, 2, 8), _u(lay, 10, 8)
        n_el = 1
        for d in shape:
            n_el *= d
        assert dsize == n_el * elsize, f"{what}: layout size {dsize} != {n_el} elements x {elsize} bytes"
        if dsize == 0:
            assert daddr == UNDEF or daddr <= size, f"{what}: empty dataset address"
        else:
            assert daddr != UNDEF, f"{what}: data address undefined for a non-empty dataset"
            ext.add(what + " raw data", daddr, dsize)
        out[name] = {"shape": shape, "dtype": dtype, "address": daddr, "size": dsize, "header": hdr}
    return out


def read(path: str, name: str):
    """Raw row-major little-endian elements of one dataset, through the validator's own walk."""
    import numpy as np

    info = validate(path)[name]
    if info["size"] == 0:
        return np.zeros(info["shape"], dtype=info["dtype"])
    with open(path, "rb") as f:
        f.seek(info["address"])
        return np.frombuffer(f.read(info["size"]), dtype=info["dtype"]).reshape(info["shape"])


if __name__ == "__main__":
    import sys

    for p in sys.argv[1:]:
        print(p, validate(p), os.path.getsize(p))
