"""Dependency-free reader/writer for the small HDF5 subset the ExPecto pipeline exchanges.

chromatin.py:282-286 writes ``snps.shift_{s}.diff.h5`` with contiguous float32 datasets ``diff``, ``ref``,
``alt`` ([2N,2002], default h5py layout: no chunks, no compression, no attributes); upstream ExPecto writes a
single ``pred`` dataset (example/*.h5).  predict.py:173-190 reads them back.  h5py / libhdf5 are not in this
image, so this module emits exactly that on-disk form (superblock v0, one symbol-table group, v1 object
headers, contiguous layout — decoded byte-by-byte from example/example.vcf.shift_0.diff.h5) and
any h5py/libhdf5 >= 1.8 reads it.  The reader walks root -> B-tree -> SNOD -> object header -> layout.
"""
from __future__ import annotations

import os
import struct
import time

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF
_DATA_START = 0x800
_LEAF_K, _INTERNAL_K = 4, 16


def _pad8(b: bytes) -> bytes:
    return b + b"\x00" * ((-len(b)) % 8)


def _msg(mtype: int, data: bytes, flags: int = 0) -> bytes:
    data = _pad8(data)
    return struct.pack("<HHB3x", mtype, len(data), flags) + data


def _datatype_message(dtype: np.dtype) -> bytes:
    """Datatype message body for the three element types the pipeline stores: float32 (chromatin.py:282-286),
    float64 (``chromatin_preds``, geuvadis_predict_consensus.py:128) and fixed-length byte strings (numpy ``S<n>``,
    ``record_ids`` :129 — h5py maps them to null-padded ASCII strings)."""
    if dtype == np.dtype("<f4"):
        # IEEE little-endian float32: class 1 version 1; bit fields 0x20,0x1f,0x00; size 4;
        # bit offset 0, precision 32, exp loc 23, exp size 8, mantissa loc 0, mantissa size 23, bias 127
        return struct.pack("<BBBBI", 0x11, 0x20, 0x1F, 0x00, 4) + struct.pack("<HHBBBBI", 0, 32, 23, 8, 0, 23, 127)
    if dtype == np.dtype("<f8"):
        return struct.pack("<BBBBI", 0x11, 0x20, 0x3F, 0x00, 8) + struct.pack("<HHBBBBI", 0, 64, 52, 11, 0, 52, 1023)
    if dtype.kind == "S" and dtype.itemsize > 0:
        # class 3 (string) version 1; bit field 0: padding = 1 (null pad), character set = 0 (ASCII); no properties
        return struct.pack("<BBBBI", 0x13, 0x01, 0x00, 0x00, dtype.itemsize)
    raise ValueError(f"unsupported HDF5 element type {dtype}")


def _dataset_header(shape, data_addr: int, nbytes: int, mtime: int, dtype=np.dtype("<f4")) -> bytes:
    rank = len(shape)
    dataspace = struct.pack("<BBB5x", 1, rank, 1) + b"".join(struct.pack("<Q", d) for d in shape) * 2
    datatype = _datatype_message(np.dtype(dtype))
    fill = struct.pack("<BBBBI", 2, 2, 2, 1, 0)
    layout = struct.pack("<BBQQ", 3, 1, data_addr, nbytes)
    mt = struct.pack("<B3xI", 1, mtime & 0xFFFFFFFF)
    msgs = _msg(1, dataspace) + _msg(3, datatype, 1) + _msg(5, fill, 1) + _msg(8, layout) + _msg(0x12, mt)
    total = 0x100
    nil = total - len(msgs) - 8
    assert nil >= 0, "dataset rank too large for the fixed header"
    msgs += struct.pack("<HHB3x", 0, nil, 0) + b"\x00" * nil
    return struct.pack("<BBHII4x", 1, 0, 6, 1, total) + msgs


class H5Writer:
    """Write datasets (<= 8; float32 unless ``dtypes`` says float64 or ``S<n>``) into one v0-superblock file, laid out
    the way libhdf5 1.8/1.10 lays out ``h5py.File(path, 'w')`` + ``create_dataset(name, data=...)`` calls in the order
    of ``datasets`` (chromatin.py:283-285: diff, ref, alt): metadata is carved out of the first 2048-byte aggregator
    block in allocation order — root header, B-tree node, local heap, the first dataset's object header, the group's
    symbol node, then the remaining dataset headers — names enter the local heap in creation order, the symbol node
    lists them sorted, and each dataset's raw data follows the previous one's from 0x800.  For a single ``pred``
    dataset the first 2048 bytes are identical to example/example.vcf.shift_0.diff.h5 (written by h5py) except the
    4-byte modification time (tests/test_hdf5_cpu.py).

    ``write_rows(name, row0, array)`` pwrite()s row slabs, so several producers (GPU ranks, pipeline stages) can
    fill one contiguous dataset without an in-memory copy."""

    _HDR_SIZE = 0x110   # v1 object header: 16-byte prefix + 0x100 bytes of messages
    _SNOD_SIZE = 8 + 2 * _LEAF_K * 40

    def __init__(self, path: str, datasets: dict[str, tuple[int, ...]], create: bool = True, dtypes=None):
        """``create=False`` attaches to a file another process already created with the same datasets (a GPU
        rank writing its own row slab; the layout is a pure function of the dataset names, order and shapes)."""
        if not 1 <= len(datasets) <= 2 * _LEAF_K:
            raise ValueError("1..8 datasets per file")
        self.path = path
        self.order = list(datasets)            # creation order (heap offsets, header addresses, data addresses)
        self.names = sorted(datasets)          # symbol-node order
        self.shapes = {n: tuple(int(x) for x in datasets[n]) for n in self.order}
        self.dtypes = {n: np.dtype((dtypes or {}).get(n, "<f4")) for n in self.order}
        for n in self.order:
            _datatype_message(self.dtypes[n])
        # local heap data: "" at 0, then each name (creation order) NUL-terminated and padded to 8, then the free
        # block; libhdf5 starts a group's heap at 88 bytes (longer name lists get a larger segment here)
        heap = bytearray(8)
        self._name_off = {}
        for n in self.order:
            self._name_off[n] = len(heap)
            heap += _pad8(n.encode("ascii") + b"\x00")
        self._heap_free = len(heap)
        self._heap_size = max(88, (self._heap_free + 16 + 7) // 8 * 8)
        heap += struct.pack("<QQ", 1, self._heap_size - self._heap_free)
        self._heap = bytes(heap) + b"\x00" * (self._heap_size - len(heap))
        # metadata addresses: the first object header follows the heap data segment (0x320 for an 88-byte heap)
        self._hdr_addr, a = {}, 0x2A8 + 32 + self._heap_size
        for i, n in enumerate(self.order):
            self._hdr_addr[n] = a
            a += self._HDR_SIZE
            if i == 0:
                self._snod_addr = a
                a += self._SNOD_SIZE
        # three headers fit libhdf5's first 2048-byte metadata block; with more, libhdf5 would open a second block
        # after the raw data written so far — this writer keeps one (larger) leading block instead, equally valid
        self.data_start = _DATA_START if a <= _DATA_START else (a + 0x7FF) // 0x800 * 0x800
        self.addr, self.nbytes, off = {}, {}, self.data_start
        for n in self.order:
            nb = int(np.prod(self.shapes[n])) * self.dtypes[n].itemsize
            self.nbytes[n] = nb
            self.addr[n] = off            # empty datasets own no storage (layout address UNDEF, as libhdf5 leaves it)
            off += nb
            off += (-off) % 8
        self.eof = off
        if create:
            self._f = open(path, "wb+", buffering=0)
            self._f.write(self._metadata())
            self._f.truncate(self.eof)
        else:
            self._f = open(path, "rb+", buffering=0)
            if self._f.read(8) != _SIG or os.fstat(self._f.fileno()).st_size != self.eof:
                raise ValueError(f"{path} was not created with these datasets")

    def _metadata(self) -> bytes:
        btree_addr, heap_addr = 0x88, 0x2A8
        heap_data_addr = heap_addr + 32
        heap, name_off, free_off, heap_size = self._heap, self._name_off, self._heap_free, self._heap_size
        mtime = int(time.time())

        out = bytearray(self.data_start)
        sb = _SIG + struct.pack("<BBBBBBBBHHI", 0, 0, 0, 0, 0, 8, 8, 0, _LEAF_K, _INTERNAL_K, 0)
        sb += struct.pack("<QQQQ", 0, _UNDEF, self.eof, _UNDEF)
        sb += struct.pack("<QQII", 0, 0x60, 1, 0) + struct.pack("<QQ", btree_addr, heap_addr)  # root symbol entry
        out[0:len(sb)] = sb
        root = struct.pack("<BBHII4x", 1, 0, 1, 1, 0x18) + _msg(0x11, struct.pack("<QQ", btree_addr, heap_addr))
        out[0x60:0x60 + len(root)] = root
        bt = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, _UNDEF, _UNDEF)
        bt += struct.pack("<QQQ", 0, self._snod_addr, name_off[self.names[-1]])
        out[btree_addr:btree_addr + len(bt)] = bt
        hp = b"HEAP" + struct.pack("<B3xQQQ", 0, heap_size, free_off, heap_data_addr)
        out[heap_addr:heap_addr + len(hp)] = hp
        out[heap_data_addr:heap_data_addr + heap_size] = heap
        sn = b"SNOD" + struct.pack("<BBH", 1, 0, len(self.names))
        for n in self.names:
            sn += struct.pack("<QQII16x", name_off[n], self._hdr_addr[n], 0, 0)
        out[self._snod_addr:self._snod_addr + len(sn)] = sn
        for n in self.order:
            nb = self.nbytes[n]
            h = _dataset_header(self.shapes[n], self.addr[n] if nb else _UNDEF, nb, mtime, self.dtypes[n])
            out[self._hdr_addr[n]:self._hdr_addr[n] + len(h)] = h
        return bytes(out)

    def write_rows(self, name: str, row0: int, arr: np.ndarray) -> None:
        shape = self.shapes[name]
        dt = self.dtypes[name]
        arr = np.ascontiguousarray(arr, dtype=dt)
        row_elems = int(np.prod(shape[1:])) if len(shape) > 1 else 1
        if arr.size % row_elems or row0 * row_elems + arr.size > int(np.prod(shape)):
            raise ValueError("slab does not fit the dataset")
        if arr.size == 0:
            return
        # positional writes: no shared file offset, so slabs can be written from several threads at once
        mv = memoryview(arr.reshape(-1).view(np.uint8))
        off, fd = self.addr[name] + row0 * row_elems * dt.itemsize, self._f.fileno()
        while len(mv):
            k = os.pwrite(fd, mv[:1 << 30], off)
            mv, off = mv[k:], off + k

    def write(self, name: str, arr: np.ndarray) -> None:
        if tuple(arr.shape) != self.shapes[name]:
            raise ValueError(f"{name}: shape {arr.shape} != {self.shapes[name]}")
        self.write_rows(name, 0, arr)

    def close(self) -> None:
        if self._f:
            self._f.close()
            self._f = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def write_datasets(path: str, arrays: dict[str, np.ndarray]) -> None:
    """``h5py.File(path,'w'); create_dataset(name, data=...)`` (chromatin.py:282-286; geuvadis_predict_consensus.py:
    127-133).  float64 and ``S<n>`` arrays keep their type; everything else is stored as float32."""
    arrays = {k: np.asarray(v) for k, v in arrays.items()}
    dtypes = {k: (v.dtype if v.dtype == np.float64 or (v.dtype.kind == "S" and v.dtype.itemsize) else np.dtype("<f4"))
              for k, v in arrays.items()}
    with H5Writer(path, {k: v.shape for k, v in arrays.items()}, dtypes=dtypes) as w:
        for k, v in arrays.items():
            w.write(k, v)


class H5Dataset:
    """Lazy view of one contiguous dataset: ``.shape`` and row slicing, as predict.py:174-190 uses h5py."""

    def __init__(self, path: str, shape, dtype, addr: int):
        self.path, self.shape, self.dtype, self.addr = path, tuple(shape), np.dtype(dtype), addr
        self._mm = None

    def _map(self):
        if self._mm is None and (self.addr is None or 0 in self.shape):
            self._mm = np.zeros(self.shape, dtype=self.dtype)   # HDF5 fill value of unallocated storage
        if self._mm is None:
            self._mm = np.memmap(self.path, dtype=self.dtype, mode="r", offset=self.addr, shape=self.shape)
        return self._mm

    def read_rows_into(self, dst: np.ndarray, row0: int, row1: int) -> None:
        """Rows [row0, row1) straight into ``dst`` (C-contiguous, same dtype; e.g. a pinned buffer) with positional
        reads — no intermediate copy, no shared file offset, the GIL is released while the kernel copies."""
        row_bytes = int(np.prod(self.shape[1:])) * np.dtype(self.dtype).itemsize if len(self.shape) > 1 else \
            np.dtype(self.dtype).itemsize
        mv = memoryview(dst.reshape(-1)).cast("B")[:(row1 - row0) * row_bytes]
        if self.addr is None:          # unallocated storage reads as the fill value
            mv[:] = bytes(len(mv))
            return
        off = self.addr + row0 * row_bytes
        fd = os.open(self.path, os.O_RDONLY)
        try:
            while len(mv):
                k = os.preadv(fd, [mv[:1 << 30]], off)
                if k <= 0:
                    raise OSError(f"{self.path}: short read at byte {off}")
                mv, off = mv[k:], off + k
        finally:
            os.close(fd)

    def __getitem__(self, idx):
        return np.asarray(self._map()[idx])

    def __array__(self, dtype=None, copy=None):
        a = np.asarray(self._map())
        if dtype is not None and np.dtype(dtype) != a.dtype:
            return a.astype(dtype)
        return a.copy() if copy else a


class H5File:
    """Minimal ``h5py.File(path,'r')``: ``f['diff']`` -> H5Dataset, ``list(f)`` -> dataset names."""

    def __init__(self, path: str, mode: str = "r"):
        if mode != "r":
            raise ValueError("H5File is read-only; use H5Writer / write_datasets to write")
        self.path = path
        with open(path, "rb") as f:
            self._meta = f.read(1 << 20)
        b = self._meta
        if b[:8] != _SIG or b[8] != 0:
            raise ValueError("not an HDF5 file with a version-0 superblock")
        if b[13] != 8 or b[14] != 8:
            raise ValueError("only 8-byte offsets/lengths are supported")
        self._base = struct.unpack_from("<Q", b, 24)[0]
        btree, heap = struct.unpack_from("<QQ", b, 24 + 32 + 24)  # scratch pad of the root symbol table entry
        if struct.unpack_from("<I", b, 24 + 32 + 16)[0] != 1:
            btree, heap = self._symbol_table_message(struct.unpack_from("<Q", b, 24 + 32 + 8)[0])
        self._datasets = {}
        heap_data = struct.unpack_from("<Q", b, heap + 24)[0]
        self._walk_btree(btree, heap_data)

    def _bytes(self, addr: int, n: int) -> bytes:
        if addr + n <= len(self._meta):
            return self._meta[addr:addr + n]
        with open(self.path, "rb") as f:
            f.seek(addr)
            return f.read(n)

    def _messages(self, addr: int):
        ver, _, nmsg, _, hsize = struct.unpack_from("<BBHII", self._bytes(addr, 16))
        if ver != 1:
            raise ValueError("only version-1 object headers are supported")
        blocks = [(addr + 16, hsize)]
        out = []
        while blocks and len(out) < nmsg:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end and len(out) < nmsg:
                mtype, msize, _flags = struct.unpack_from("<HHB", self._bytes(pos, 8))
                data = self._bytes(pos + 8, msize)
                if mtype == 0x10:  # continuation
                    blocks.append(struct.unpack_from("<QQ", data))
                out.append((mtype, data))
                pos += 8 + msize
        return out

    def _symbol_table_message(self, addr: int):
        for mtype, data in self._messages(addr):
            if mtype == 0x11:
                return struct.unpack_from("<QQ", data)
        raise ValueError("root group has no symbol table")

    def _walk_btree(self, addr: int, heap_data: int) -> None:
        b = self._bytes(addr, 24)
        if b[:4] != b"TREE":
            raise ValueError("bad B-tree node")
        level, used = b[5], struct.unpack_from("<H", b, 6)[0]
        body = self._bytes(addr + 24, (2 * used + 1) * 8)
        for i in range(used):
            child = struct.unpack_from("<Q", body, 8 + 16 * i)[0]
            if level > 0:
                self._walk_btree(child, heap_data)
            else:
                self._read_snod(child, heap_data)

    def _read_snod(self, addr: int, heap_data: int) -> None:
        b = self._bytes(addr, 8)
        if b[:4] != b"SNOD":
            raise ValueError("bad symbol node")
        n = struct.unpack_from("<H", b, 6)[0]
        ents = self._bytes(addr + 8, 40 * n)
        for i in range(n):
            name_off, hdr = struct.unpack_from("<QQ", ents, 40 * i)
            raw = self._bytes(heap_data + name_off, 256)
            name = raw[:raw.index(b"\x00")].decode("ascii")
            ds = self._read_dataset(hdr)
            if ds is not None:
                self._datasets[name] = ds

    def _read_dataset(self, addr: int):
        shape = dtype = data_addr = None
        for mtype, data in self._messages(addr):
            if mtype == 1:
                rank = data[1]
                off = 8 if data[0] == 1 else 4
                shape = struct.unpack_from("<%dQ" % rank, data, off)
            elif mtype == 3:
                cls, size = data[0] & 0x0F, struct.unpack_from("<I", data, 4)[0]
                big = data[1] & 1
                if cls == 1:
                    dtype = (">f%d" if big else "<f%d") % size
                elif cls == 0:
                    signed = (data[1] >> 3) & 1
                    dtype = (">" if big else "<") + ("i" if signed else "u") + str(size)
                elif cls == 3:
                    dtype = "S%d" % size
            elif mtype == 8:
                if data[0] != 3 or data[1] != 1:
                    raise ValueError("only contiguous (layout v3 class 1) datasets are supported")
                data_addr = struct.unpack_from("<Q", data, 2)[0]
        if shape is None or dtype is None or data_addr is None:
            return None
        if data_addr == _UNDEF:  # no storage allocated (empty dataset, or never written): reads as zeros / empty
            data_addr = None
        return H5Dataset(self.path, shape, dtype, None if data_addr is None else self._base + data_addr)

    def __getitem__(self, name: str) -> H5Dataset:
        return self._datasets[name]

    def __contains__(self, name: str) -> bool:
        return name in self._datasets

    def __iter__(self):
        return iter(self._datasets)

    def keys(self):
        return self._datasets.keys()

    def close(self) -> None:
        pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def File(path: str, mode: str = "r"):
    """``h5py.File`` look-alike for reading; writing goes through :func:`write_datasets`."""
    return H5File(path, mode)
