"""Segment mode (shared activations across the shifted windows of a variant / gene, bx_forward_segments) must be
bit-identical to the per-window path on the same windows."""
import numpy as np
import pytest
import torch

from expecto_b200.genome import PackedGenome
from expecto_b200.windows import (chromatin_segment_descs, chromatin_shift_list, chromatin_window_descs,
                                  segment_eligible, tss_descs, tss_segment_descs)
from oracle import golden_inputs as GI

pytestmark = pytest.mark.gpu


def test_chromatin_segments_equal_windows_bitwise(engine, small_genome):
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = GI.edge_variants(small_genome)
    ok = segment_eligible([len(r[3]) for r in rows], [len(r[4]) for r in rows])
    rows = [r for r, e in zip(rows, ok) if e]
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    n = len(rows)
    sd, codes, k_of = chromatin_segment_descs(pg, chroms, pos, refs, alts, 800)
    seg = engine.forward_segments(sd, codes, n, 4, 3600, k_of, shift_major=True)
    assert seg.shape == (9, n, 4, 2002)
    for si, shift in enumerate(chromatin_shift_list(800)):
        wd, wcodes = chromatin_window_descs(pg, chroms, pos, refs, alts, shift)
        win = engine.forward_windows(wd, wcodes).reshape(n, 4, 2002)
        assert torch.equal(seg[si], win), shift


def test_tss_segments_equal_windows_bitwise(engine, small_genome):
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    shifts = np.arange(-20000, 20000, 200)
    chroms, tss, strand = ["chr1", "chr1"], [30000, 25000], [1, -1]
    sd, k_of, seg_len = tss_segment_descs(pg, chroms, tss, strand, shifts)
    assert seg_len == 41800 and engine.segment_capacity(seg_len) >= 4
    seg = engine.forward_segments(sd, np.zeros(1, np.uint8), 2, 2, seg_len, k_of, shift_major=False)
    wd = tss_descs(pg, chroms, tss, strand, shifts)
    win = engine.forward_windows(wd, np.zeros(1, np.uint8)).reshape(2, 200, 2, 2002)
    assert torch.equal(seg, win)


def test_segment_chunking_and_argument_errors(engine, small_genome):
    """More groups than one scratch chunk holds (group-major order chunks transparently)."""
    from expecto_b200._lib import BelugaError
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    shifts = np.arange(-400, 600, 200)  # 5 shifts -> 2800-bp segments
    cap = engine.segment_capacity(2800) // 2
    n = cap + 3
    rng = np.random.default_rng(0)
    tss = rng.integers(5000, 50000, size=n)
    strand = np.where(rng.random(n) < 0.5, 1, -1)
    sd, k_of, seg_len = tss_segment_descs(pg, ["chr1"] * n, tss, strand, shifts)
    seg = engine.forward_segments(sd, np.zeros(1, np.uint8), n, 2, seg_len, k_of, shift_major=False)
    wd = tss_descs(pg, ["chr1"] * n, tss, strand, shifts)
    win = engine.forward_windows(wd, np.zeros(1, np.uint8)).reshape(n, 5, 2, 2002)
    assert torch.equal(seg, win)
    with pytest.raises(BelugaError):   # shift-major output cannot be chunked inside the library
        engine.forward_segments(sd, np.zeros(1, np.uint8), n, 2, seg_len, k_of, shift_major=True)
    with pytest.raises(BelugaError):   # seg_len must match the shift count
        engine.forward_segments(sd, np.zeros(1, np.uint8), n, 2, seg_len + 200, k_of, shift_major=False)
