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


def test_chromatin_cli_segments_and_windows_modes_write_identical_files(engine, cfg1_genome, golden, tmp_path):
    """Edge VCF (one variant — the 120-bp deletion — falls back to the per-window path inside segment mode)."""
    from expecto_b200 import chromatin, hdf5
    g = golden("chromatin_edge_golden.npz")
    vcf = tmp_path / "edge.vcf"
    vcf.write_text("".join("\t".join(str(x) for x in r) + "\n" for r in g["rows"]))
    pg = PackedGenome.from_sequences(cfg1_genome)
    infos = {}
    for mode in ("segments", "windows"):
        args = chromatin.build_parser().parse_args([str(vcf), "--output_dir", str(tmp_path / mode), "--mode", mode,
                                                    "--upstream_layout"])
        infos[mode] = chromatin.run(args, engine=engine, genome=pg)
        # --upstream_layout files land next to the input; keep each mode's copy
        for s in chromatin.shift_list(800):
            (tmp_path / f"edge.vcf.shift_{s}.diff.h5").rename(tmp_path / mode / f"pred_{s}.h5")
    assert infos["segments"]["mode"] == "segments" and infos["segments"]["window_variants"] == 1
    assert infos["segments"]["segment_variants"] == 5
    assert list(infos["segments"]["counters"]) == list(infos["windows"]["counters"]) == list(g["counters"])
    for s in chromatin.shift_list(800):
        with hdf5.File(str(tmp_path / "segments" / f"snps.shift_{s}.diff.h5")) as a, \
                hdf5.File(str(tmp_path / "windows" / f"snps.shift_{s}.diff.h5")) as b:
            for k in ("diff", "ref", "alt"):
                assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), (s, k)
            d = np.asarray(a["diff"])
        with hdf5.File(str(tmp_path / "segments" / f"pred_{s}.h5")) as p:
            assert np.array_equal(np.asarray(p["pred"]), d)


def test_features_cli_segments_equal_windows(engine, cfg1_genome):
    from expecto_b200.compute_expecto_features import gene_features
    from expecto_b200.transform import TSS_SHIFTS, fixed_pos_weights
    pg = PackedGenome.from_sequences(cfg1_genome)
    engine.set_genome(pg)
    genes = [("g1", "chr1", 5_000_000, 1), ("g2", "chr1", 7_654_321, -1), ("g3", "chr1", 30_000, 1)]
    a = gene_features(engine, pg, genes, TSS_SHIFTS, fixed_pos_weights(), 2, mode="segments")
    b = gene_features(engine, pg, genes, TSS_SHIFTS, fixed_pos_weights(), 2, mode="windows")
    assert np.array_equal(a, b)
