"""FASTA -> packed genome on the device (include/beluga_b200.h: bx_fasta_find_headers / bx_fasta_count /
bx_fasta_pack; replaces pyfasta's index + slicing, chromatin.py:44,205, and the genome side of encodeSeqs' mydict,
chromatin.py:155-160).  Integer / byte work: every word must equal the host packer's, which tests/test_windows_cpu.py
pins base by base against the reference's mydict semantics."""
import os

import numpy as np
import pytest

from expecto_b200 import synth
from expecto_b200.genome import PackedGenome

pytestmark = pytest.mark.gpu


def _same(a: PackedGenome, b: PackedGenome):
    assert a.names == b.names and a.offsets == b.offsets and a.lengths == b.lengths and a.n_bases == b.n_bases
    assert np.array_equal(a.bases2, b.bases2)
    assert np.array_equal(a.nmask, b.nmask)
    assert np.array_equal(a.invalid, b.invalid)


def _write_ragged(path, seqs, eol=b"\r\n", widths=(7, 60, 1, 33, 4096, 9000)):
    with open(path, "wb") as f:
        for name, s in seqs.items():
            f.write(f">{name} description text\n".encode())
            i, k = 0, 0
            while i < len(s):
                w = widths[k % len(widths)]
                f.write(s[i:i + w].encode() + eol)
                i, k = i + w, k + 1


def test_device_packer_equals_host_packer_on_ragged_crlf_fasta(tmp_path):
    """Line widths from 1 to 9000 characters: 32-base groups shared by two 4096-byte blocks (atomic merge), blocks
    without any line end whose first base is not group-aligned (129 groups in one block), CRLF line ends."""
    rng = np.random.default_rng(5)
    alphabet = np.frombuffer(b"ACGTacgtNnH-RYhx", dtype=np.uint8)   # incl. characters outside mydict
    chr1 = alphabet[rng.integers(0, len(alphabet), size=70_001)].tobytes().decode()
    chr1 = chr1[:100] + ">" + chr1[101:]        # a '>' inside a line is an invalid base, not a header
    seqs = {"chr1": chr1,
            "chrEmpty": "",
            "chr2": "ACGT" * 8,                                         # exactly one mask word
            "chr3": alphabet[rng.integers(0, 8, size=4096 * 3 + 17)].tobytes().decode(),
            "chr4": "N"}
    fa = str(tmp_path / "ragged.fa")
    _write_ragged(fa, seqs)
    host = PackedGenome.from_fasta(fa, cache=False)
    assert host.lengths == {k: len(v) for k, v in seqs.items()} and host.invalid.size > 0
    dev = PackedGenome.from_fasta(fa, cache=False, device=0)
    _same(dev, host)


def test_device_packer_plain_fasta_no_trailing_newline_and_cache(tmp_path, small_genome):
    fa = str(tmp_path / "g.fa")
    synth.write_fasta(fa, small_genome, width=60)
    with open(fa, "rb+") as f:          # drop the final newline
        f.seek(-1, os.SEEK_END)
        f.truncate()
    dev = PackedGenome.from_fasta(fa, device=0)
    _same(dev, PackedGenome.from_sequences(small_genome))
    assert os.path.exists(fa + ".bx2.npz")
    _same(PackedGenome.from_fasta(fa, device=0), dev)        # served from the cache


def test_engine_uses_the_device_planes(engine, tmp_path, small_genome):
    """A genome packed on the engine's device is used in place (no host round trip) and encodes the same windows."""
    from expecto_b200.windows import tss_descs
    fa = str(tmp_path / "g.fa")
    synth.write_fasta(fa, small_genome, width=50)
    dev = PackedGenome.from_fasta(fa, cache=False, device=engine.device)
    host = PackedGenome.from_sequences(small_genome)
    name = next(iter(small_genome))
    cols = ([name, name], np.array([5000, 9000]), np.array([1, -1]))
    out = []
    for g in (dev, host):
        engine.set_genome(g)
        descs = tss_descs(g, *cols, [-400, 0, 200])
        out.append(engine.encode_windows(descs, np.zeros(1, dtype=np.uint8)).cpu().numpy())
    assert engine._bases2 is not None and np.array_equal(out[0], out[1])
    engine.set_genome(dev)
    assert engine._bases2.data_ptr() == dev._device_planes[1].data_ptr()


def test_fasta_kernels_stay_inside_their_buffers(tmp_path):
    """compute-sanitizer is closed on this GPU pool, so the write side is checked with guard words: every output buffer
    of bx_fasta_count / bx_fasta_pack / bx_fasta_find_headers sits between sentinel regions that must come back intact
    (contig placed in the middle of the planes, body starting at an odd file offset, invalid list exactly full)."""
    import ctypes as C

    import torch

    from expecto_b200._lib import check, lib
    L = lib()
    rng = np.random.default_rng(3)
    body = np.frombuffer(b"ACGTacgtNXY", dtype=np.uint8)[rng.integers(0, 11, size=50_003)]
    lines = [body[i:i + 61].tobytes() for i in range(0, body.size, 61)]
    raw = b">c1 x\n" + b"\n".join(lines) + b"\n>c2\nACGT\n"
    n_text = len(raw)
    G, SENT = 64, 0x5A
    text = torch.full((G * 16 + (n_text + 15) // 16 * 16 + 16,), 10, dtype=torch.uint8, device="cuda")
    text[G * 16: G * 16 + n_text] = torch.from_numpy(np.frombuffer(raw, dtype=np.uint8).copy()).cuda()
    tptr = text.data_ptr() + G * 16
    st = torch.cuda.current_stream().cuda_stream

    def guarded(n, dtype):
        t = torch.full((n + 2 * G,), SENT, dtype=torch.uint8, device="cuda") if dtype == torch.uint8 else None
        if t is None:
            t = torch.empty(n + 2 * G, dtype=dtype, device="cuda")
            t.view(torch.uint8).fill_(SENT)
        return t

    def intact(t, n):
        b = t.view(torch.uint8).cpu().numpy()
        w = t.element_size()
        return bool((b[:G * w] == SENT).all() and (b[(G + n) * w:] == SENT).all())

    # headers
    pos = guarded(2, torch.int64)
    cnt = guarded(1, torch.int64)
    check(L.bx_fasta_find_headers(tptr, n_text, pos.data_ptr() + G * 8, 2, cnt.data_ptr() + G * 8, st))
    torch.cuda.synchronize()
    assert intact(pos, 2) and intact(cnt, 1) and int(cnt[G]) == 2
    starts = sorted(int(x) for x in pos[G:G + 2].cpu())
    assert starts == [0, raw.index(b">c2")]
    # count + scan of contig 1 (body starts at file offset 6: not 16-byte aligned)
    b0, b1 = 6, starts[1]
    nb = int(L.bx_fasta_blocks(b0, b1))
    off = guarded(nb + 1, torch.int64)
    n_out = C.c_int64(0)
    check(L.bx_fasta_count(tptr, b0, b1, off.data_ptr() + G * 8, C.byref(n_out), st))
    assert n_out.value == body.size and intact(off, nb + 1)
    # pack in the middle of guarded planes; the invalid list is exactly as long as the number of X / Y characters
    n_bad = int(np.isin(body, np.frombuffer(b"XY", dtype=np.uint8)).sum())
    base_off = 4096
    total = base_off + (body.size + 31) // 32 * 32 + 4096
    bases2 = guarded(total // 16, torch.int32)
    nmask = guarded(total // 32, torch.int32)
    bases2[G:G + total // 16] = 0
    nmask[G:G + total // 32] = -1
    inv = guarded(n_bad, torch.int64)
    icnt = guarded(1, torch.int64)
    icnt[G] = 0
    check(L.bx_fasta_pack(tptr, b0, b1, off.data_ptr() + G * 8, base_off, bases2.data_ptr() + G * 4, nmask.data_ptr() + G * 4,
                          inv.data_ptr() + G * 8, n_bad, icnt.data_ptr() + G * 8, st))
    torch.cuda.synchronize()
    assert intact(bases2, total // 16) and intact(nmask, total // 32) and intact(inv, n_bad) and intact(icnt, 1)
    assert int(icnt[G]) == n_bad
    want = PackedGenome.from_sequences({"c1": body.tobytes()})
    w0 = want.offsets["c1"]
    n32 = (body.size + 31) // 32 * 32
    got2 = bases2[G + base_off // 16: G + (base_off + n32) // 16].cpu().numpy().view(np.uint32)
    gotm = nmask[G + base_off // 32: G + (base_off + n32) // 32].cpu().numpy().view(np.uint32)
    assert np.array_equal(got2, want.bases2[w0 // 16:(w0 + n32) // 16])
    assert np.array_equal(gotm, want.nmask[w0 // 32:(w0 + n32) // 32])
    assert sorted(int(x) - base_off for x in inv[G:G + n_bad].cpu()) == sorted(int(x) - w0 for x in want.invalid)
    # a list shorter than the number of invalid characters: the count still reports all, nothing is written past it
    inv2 = guarded(3, torch.int64)
    icnt[G] = 0
    check(L.bx_fasta_pack(tptr, b0, b1, off.data_ptr() + G * 8, base_off, bases2.data_ptr() + G * 4, nmask.data_ptr() + G * 4,
                          inv2.data_ptr() + G * 8, 3, icnt.data_ptr() + G * 8, st))
    torch.cuda.synchronize()
    assert intact(inv2, 3) and int(icnt[G]) == n_bad
