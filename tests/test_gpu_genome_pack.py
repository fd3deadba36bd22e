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
