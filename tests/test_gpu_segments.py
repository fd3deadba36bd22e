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


def _substitutions(small_genome, n_random=40):
    """SNVs / MNVs incl. ref-mismatch, lower case, N alleles and the longest handled MNV, on both ends of contigs."""
    from expecto_b200.windows import delta_eligible
    rows = [r for r in GI.edge_variants(small_genome) if delta_eligible([len(r[3])], [len(r[4])])[0]]
    rng = np.random.default_rng(11)
    names = sorted(small_genome)
    for i in range(n_random):
        c = names[i % len(names)]
        p = int(rng.integers(3000, len(small_genome[c]) - 3000))
        ln = int(rng.choice([1, 1, 1, 2, 3, 8, 16]))
        ref = small_genome[c][p - 1:p - 1 + ln].upper()
        if i % 7 == 0:
            ref = "".join("ACGT"[(("ACGT".find(b) + 1) % 4)] if b in "ACGT" else "A" for b in ref)  # ref mismatch
        alt = "".join("ACGTN"[int(rng.integers(0, 5))] for _ in range(ln))
        rows.append((c, p, f"sub{i}", ref, alt.lower() if i % 5 == 0 else alt))
    return rows


def test_delta_patches_equal_full_segments_and_windows_bitwise(engine, small_genome):
    """Substitutions: ref segments in full + alt patches (bx_forward_segments_delta) == four full segments ==
    per-window path, bit for bit, for every shift, strand and allele."""
    from expecto_b200.windows import chromatin_delta_descs
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = _substitutions(small_genome)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    n = len(rows)
    rd, pd, codes, k_of, row0, plen = chromatin_delta_descs(pg, chroms, pos, refs, alts, 800)
    delta = engine.forward_segments_delta(rd, pd, codes, n, 3600, plen, row0, k_of, shift_major=True)
    sd, scodes, k2 = chromatin_segment_descs(pg, chroms, pos, refs, alts, 800)
    full = engine.forward_segments(sd, scodes, n, 4, 3600, k2, shift_major=True)
    assert delta.shape == full.shape == (9, n, 4, 2002)
    assert torch.equal(delta, full)
    assert not torch.equal(full[:, :, 0], full[:, :, 2])          # the alleles do change the predictions
    for si, shift in enumerate(chromatin_shift_list(800)):
        wd, wcodes = chromatin_window_descs(pg, chroms, pos, refs, alts, shift)
        assert torch.equal(delta[si], engine.forward_windows(wd, wcodes).reshape(n, 4, 2002)), shift
    # group-major order and a single shift (maxshift 0: one window per segment, phase A only)
    gm = engine.forward_segments_delta(rd, pd, codes, n, 3600, plen, row0, k_of, shift_major=False)
    assert torch.equal(gm, full.permute(1, 0, 2, 3))
    rd0, pd0, codes0, k0, row00, plen0 = chromatin_delta_descs(pg, chroms, pos, refs, alts, 0)
    one = engine.forward_segments_delta(rd0, pd0, codes0, n, 2000, plen0, row00, k0, shift_major=True)
    wd, wcodes = chromatin_window_descs(pg, chroms, pos, refs, alts, 0)
    assert torch.equal(one[0], engine.forward_windows(wd, wcodes).reshape(n, 4, 2002))
    from expecto_b200._lib import BelugaError
    with pytest.raises(BelugaError):
        engine.forward_segments_delta(rd, pd, codes, n, 3600, plen + 8, row0, k_of, shift_major=True)
    with pytest.raises(BelugaError):
        engine.forward_segments_delta(rd, pd, codes, n, 3600, plen, 3600 // 16, k_of, shift_major=True)


@pytest.mark.parametrize("dtype,cta_group,maxshift", [("f16", 2, 400), ("f16", 1, 1000), ("bf16", 2, 800)])
def test_delta_mode_other_shift_ranges_and_engines(small_genome, dtype, cta_group, maxshift):
    """Delta mode (conv patches + FC1 slice reuse) for 5 / 11 shifts, single-CTA tiles and bf16 operands."""
    from expecto_b200 import synth
    from expecto_b200.engine import BelugaEngine
    from expecto_b200.windows import chromatin_delta_descs
    eng = BelugaEngine(0, dtype=dtype, cta_group=cta_group, capacity=512)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype=dtype))
    pg = PackedGenome.from_sequences(small_genome)
    eng.set_genome(pg)
    rows = _substitutions(small_genome, 30)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    n = len(rows)
    seg_len = 2000 + 2 * maxshift
    rd, pd, codes, k_of, row0, plen = chromatin_delta_descs(pg, chroms, pos, refs, alts, maxshift)
    cap = eng.segment_capacity(seg_len) // 2
    assert cap >= 8
    m = min(n, cap)
    delta = eng.forward_segments_delta(rd[:2 * m], pd[:2 * m], codes, m, seg_len, plen, row0, k_of, shift_major=True)
    for si, shift in enumerate(chromatin_shift_list(maxshift)):
        wd, wcodes = chromatin_window_descs(pg, chroms[:m], pos[:m], refs[:m], alts[:m], shift)
        assert torch.equal(delta[si], eng.forward_windows(wd, wcodes).reshape(m, 4, 2002)), shift
    eng.close()


def test_segment_loop_with_and_without_delta_is_identical(engine, small_genome):
    """The pipelined chromatin loop: a mix of substitutions, indels and window-path variants gives the same slabs
    whether substitutions take the patch path or not."""
    from expecto_b200.chromatin import predict_all_shifts_segments
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = GI.edge_variants(small_genome) + _substitutions(small_genome, 12)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    n = len(rows)
    got = {}
    for delta in (True, False):
        store = {s: {k: np.full((2 * n, 2002), np.nan, np.float32) for k in ("diff", "ref", "alt")}
                 for s in chromatin_shift_list(800)}

        def make(s):
            def sink(name, row0, arr):
                store[s][name][row0:row0 + len(arr)] = arr
            return sink
        info = predict_all_shifts_segments(engine, pg, chroms, pos, refs, alts, 800,
                                           {s: make(s) for s in store}, delta=delta)
        assert (info["delta_variants"] > 0) == delta
        got[delta] = store
    for s in got[True]:
        for k in ("diff", "ref", "alt"):
            assert not np.isnan(got[True][s][k]).any()
            assert np.array_equal(got[True][s][k], got[False][s][k]), (s, k)


def test_averaged_device_effects_equal_strand_average_of_the_file_rows(engine, small_genome):
    """predict --vcf's source (predict_all_shifts_averaged_device) == (fwd + rc) / 2 of the rows the chromatin step
    writes, for substitutions, indels and window-path variants alike (predict.py:186-190)."""
    from expecto_b200.chromatin import predict_all_shifts_averaged_device, predict_all_shifts_segments
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    rows = GI.edge_variants(small_genome) + _substitutions(small_genome, 10)
    chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    n = len(rows)
    shifts = chromatin_shift_list(800)
    store = {s: {k: np.zeros((2 * n, 2002), np.float32) for k in ("diff", "ref", "alt")} for s in shifts}

    def make(s):
        def sink(name, row0, arr):
            store[s][name][row0:row0 + len(arr)] = arr
        return sink
    info = predict_all_shifts_segments(engine, pg, chroms, pos, refs, alts, 800, {s: make(s) for s in shifts})
    assert info["window_variants"] > 0 and info["delta_variants"] > 0
    dev = predict_all_shifts_averaged_device(engine, pg, chroms, pos, refs, alts, 800)
    for si, s in enumerate(shifts):
        for k in ("ref", "alt", "diff"):
            want = (torch.from_numpy(store[s][k][:n]) + torch.from_numpy(store[s][k][n:])) / 2.0
            assert torch.equal(dev[k][si].cpu(), want), (s, k)


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
