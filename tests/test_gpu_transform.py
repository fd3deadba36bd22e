"""Kernel 4 (decay-weighted shift reductions), the feature pipeline, and size-independent properties of the
forward at BASELINE batch sizes — all through the C ABI on a B200."""
import ctypes as C
import io

import numpy as np
import pytest
import torch

from expecto_b200 import _lib
from expecto_b200.genome import PackedGenome
from expecto_b200.windows import chromatin_window_descs
from oracle import beluga_oracle as O

pytestmark = pytest.mark.gpu
FEAT_RTOL = 1e-4


def _rel_err(ours, want):
    """max |ours - want| relative to the feature scale (floor 1e-3: the masked half is exactly 0)."""
    want = np.asarray(want, dtype=np.float64)
    return float(np.abs(np.asarray(ours, dtype=np.float64) - want).max() / max(np.abs(want).max(), 1e-3))


def test_fixed_transform_kernel_vs_oracle_f64():
    from expecto_b200.transform import decay_transform_fixed
    rng = np.random.default_rng(0)
    probs = rng.random((5, 200, 2, 2002), dtype=np.float32)
    pw = O.fixed_pos_weights()
    out = decay_transform_fixed(torch.from_numpy(probs).cuda(), pw).cpu().numpy()
    assert out.shape == (5, 20020) and out.dtype == np.float32
    for g in range(5):
        want = O.fixed_transform(probs[g, :, 0], probs[g, :, 1], pw)
        assert _rel_err(out[g], want) <= FEAT_RTOL
    # exact zeros where the basis is masked: upstream bases x shifts > 0 do not exist, and vice versa
    assert np.all(out.reshape(5, 10, 2002)[:, :5] >= 0)


def test_fixed_transform_ragged_shift_counts():
    from expecto_b200.transform import decay_transform_fixed
    rng = np.random.default_rng(1)
    for S in (1, 3, 9, 201):
        probs = rng.random((2, S, 2, 2002), dtype=np.float32)
        w = rng.standard_normal((10, S))
        out = decay_transform_fixed(torch.from_numpy(probs).cuda(), w).cpu().numpy()
        want = np.stack([O.fixed_transform(probs[g, :, 0], probs[g, :, 1], w) for g in range(2)])
        assert _rel_err(out, want) <= FEAT_RTOL, S


def test_variant_transform_kernel_vs_oracle_f64():
    from expecto_b200.transform import decay_transform_variant
    rng = np.random.default_rng(2)
    n = 257
    x = (rng.standard_normal((9, n, 2002)) * 0.05).astype(np.float32)
    dist = rng.integers(-40000, 40000, size=n).astype(np.float64)
    dist[:8] = [0, 1, -1, 199, 200, -200, 800, -800]  # the floor / sign boundaries of predict.py:95-108
    strand = np.where(rng.random(n) < 0.5, "+", "-")
    out = decay_transform_variant(torch.from_numpy(x).cuda(), dist, strand).cpu().numpy()
    want = O.variant_transform(list(x), O.variant_weights(dist, strand))
    assert out.shape == (n, 20020)
    assert _rel_err(out, want) <= FEAT_RTOL


def test_feature_pipeline_matches_compute_expecto_features_script(engine, cfg1_genome, golden, tmp_path):
    """compute_expecto_features.py drop-in (cfg4 shape: 200 shifts x fwd/rc per gene) vs the reference subprocess."""
    from expecto_b200 import compute_expecto_features as cef
    g = golden("features_golden.npz")
    anno = tmp_path / "geneanno.csv"
    with open(anno, "w") as f:
        f.write("id,symbol,seqnames,strand,TSS,CAGE_representative_TSS,type\n")
        for gid, chrom, strand, tss in g["genes"]:
            f.write(f"{gid},SYM,{chrom},{strand},{tss},{tss},protein_coding\n")
    tssf = tmp_path / "tss.tsv"
    tssf.write_text("\tens_id\tchrom\ttss\tstrand\tx\tis_default_tss\n0\tENSG_NOMATCH\tchr1\t1000000\t+\t0\tTrue\n")
    args = cef.build_parser().parse_args([str(anno), str(tssf), "-o", str(tmp_path / "out")])
    cef.run(args, engine=engine, genome=PackedGenome.from_sequences(cfg1_genome))
    ours = np.load(tmp_path / "out" / "Xreducedall.2002.representative_tss_top.npy")
    want = g["features"]
    assert ours.shape == want.shape == (2, 20020) and ours.dtype == np.float64
    # features are sums of up to 100 probabilities each carrying <= 2e-3 of 16-bit operand error; the 1e-4 bar is
    # on the transform (tests above, same inputs); end to end the bar is the probability tolerance summed
    w_abs = np.abs(O.fixed_pos_weights()).sum(axis=1)
    bound = np.repeat(w_abs, 2002) * 2e-3
    assert np.all(np.abs(ours - want) <= bound[None, :])
    assert _rel_err(ours, want) <= 2e-3


def test_feature_pipeline_tss_override_matches_reference_script(engine, cfg1_genome, golden, tmp_path):
    """The TSS-override rows of compute_expecto_features.py:54-69 through the drop-in (--tss_is_hg19 = the identity
    liftover the reference ran behind): per-gene feature sums of the reference's .npy, within the summed tolerance."""
    from expecto_b200 import compute_expecto_features as cef
    g = golden("features_override_golden.npz")
    (tmp_path / "anno.csv").write_text(str(g["anno"]))
    (tmp_path / "tss.tsv").write_text(str(g["tss"]))
    args = cef.build_parser().parse_args([str(tmp_path / "anno.csv"), str(tmp_path / "tss.tsv"), "-o",
                                          str(tmp_path / "out"), "--tss_is_hg19"])
    cef.run(args, engine=engine, genome=PackedGenome.from_sequences(cfg1_genome))
    ours = np.load(tmp_path / "out" / "Xreducedall.2002.representative_tss_top.npy")
    assert ours.shape == (2, 20020)
    bound = float(np.abs(O.fixed_pos_weights()).sum() * 2002 * 2e-3)     # every probability within 2e-3
    assert np.all(np.abs(ours.sum(axis=1) - g["feature_sums"]) <= bound)
    assert np.all(np.abs(ours.sum(axis=1) - g["feature_sums"]) <= 2e-3 * np.abs(g["feature_sums"]))


def _check_sed(ref, alt, sed, want, tag, pearson_min=0.999, rel_max=0.05):
    """absolute bars of 2e-2 / 5e-3 cannot fail on this fixture (median |SED| 4.1e-3, max 1.4e-2: an
    all-zero SED column would pass).  SED is the scientific output, so it is held to a correlation and a relative
    bar; the measured figures are printed (pytest -s) and quoted in DESIGN.md section 3."""
    from helpers import pearson
    w_ref, w_alt, w_sed = want["REF"].values, want["ALT"].values, want["SED"].values
    scale = float(np.abs(w_sed).max())
    err = float(np.abs(sed - w_sed).max())
    r = pearson(sed, w_sed)
    print(f"[sed {tag}] max|SED| {scale:.3e}  max err {err:.3e} ({err / scale:.2%})  pearson {r:.6f}  "
          f"REF err {np.abs(ref - w_ref).max():.2e}  ALT err {np.abs(alt - w_alt).max():.2e}")
    assert np.abs(ref - w_ref).max() <= 2e-3 * np.abs(w_ref).max() + 1e-3
    assert np.abs(alt - w_alt).max() <= 2e-3 * np.abs(w_alt).max() + 1e-3
    assert r >= pearson_min, r
    assert err <= rel_max * scale, (err, scale)
    assert np.all(np.sign(sed[np.abs(w_sed) > 0.2 * scale]) == np.sign(w_sed[np.abs(w_sed) > 0.2 * scale]))


def test_predict_pipeline_matches_predict_script(engine, cfg1_genome, golden, tmp_path):
    """chromatin -> strand average -> per-variant transform -> gblinear scores, vs predict.py's sed.tsv."""
    import pandas as pd
    from expecto_b200.transform import decay_transform_variant
    g = golden("predict_golden.npz")
    sed = pd.read_csv(io.StringIO(str(g["sed_tsv"])), sep="\t")
    rows = [l.split("\t") for l in str(golden("chromatin_cfg1_golden.npz")["snps_hg19_vcf"]).splitlines()[2:]]
    pg = PackedGenome.from_sequences(cfg1_genome)
    engine.set_genome(pg)
    chroms, pos = [r[0] for r in rows], np.array([int(r[1]) for r in rows])
    refs, alts = [r[3] for r in rows], [r[4] for r in rows]
    stacks = {"ref": [], "alt": [], "diff": []}
    for s in O.SHIFTS_CHROMATIN(800):
        descs, codes = chromatin_window_descs(pg, chroms, pos, refs, alts, s)
        r, a, d = engine.chromatin_shift_host(descs, codes, len(rows), average=True)
        for k, v in (("ref", r), ("alt", a), ("diff", d)):
            stacks[k].append(v.clone())
    dist, strand = sed["dist"].values.astype(np.float64), sed["strand"].values
    feats = {k: decay_transform_variant(torch.stack(v).cuda(), dist, strand).cpu().numpy().astype(np.float64)
             for k, v in stacks.items()}
    assert np.abs(feats["diff"][0] - g["diff_features_row0"]).max() <= 9 * 4e-3
    w = g["model_w"].astype(np.float64)
    s_ref, s_alt = (O.gblinear_predict(feats[key], w, 0.25, 0.5) for key in ("ref", "alt"))
    _check_sed(s_ref, s_alt, s_alt - s_ref, sed, "transform kernels + float64 scoring, f16 forward")


# ---------------------------------------------------------------- properties at BASELINE sizes
def test_batch_invariance_and_chunking_at_4096(engine):
    """Every window's output is independent of its batch position and of the capacity chunking (engine
    capacity is 512, so 4096 windows cross 8 chunks); also covers ragged tails (4096+37)."""
    from expecto_b200 import synth
    base = torch.from_numpy(synth.random_onehot(64, seed=9)).cuda()
    n = 4096 + 37
    idx = torch.arange(n, device="cuda") % 64
    big = base[idx].contiguous()
    out = engine.forward_onehot(big)
    first = engine.forward_onehot(base)
    torch.cuda.synchronize()
    assert torch.equal(out, first[idx]), "results depend on batch position"
    assert bool(((out > 0) & (out < 1)).all())


def test_reverse_complement_flag_equals_explicit_flip(engine, small_genome):
    """forward(window with RC flag) == forward(onehot[:, ::-1, ::-1]) bit for bit (chromatin.py:170)."""
    from expecto_b200.windows import tss_descs
    pg = PackedGenome.from_sequences(small_genome)
    engine.set_genome(pg)
    descs = tss_descs(pg, ["chr1"] * 3, [20000, 30000, 40000], [1, 1, 1], [0, 200])
    oh = engine.encode_windows(descs, np.zeros(1, np.uint8)).reshape(3, 2, 2, 4, 2000)
    assert torch.equal(oh[:, :, 1], torch.flip(oh[:, :, 0], dims=(-2, -1)))
    p = engine.forward_windows(descs, np.zeros(1, np.uint8)).reshape(3, 2, 2, 2002)
    q = engine.forward_onehot(torch.flip(oh[:, :, 0], dims=(-2, -1)).contiguous().reshape(-1, 4, 2000))
    assert torch.equal(p[:, :, 1].reshape(-1, 2002), q)


def test_bf16_operands_run_and_are_looser(weights_f16, golden):
    """bf16 is what BASELINE names; fp16 operands are the default because bf16 misses the 2e-3 bar."""
    from expecto_b200 import synth
    from expecto_b200.engine import BelugaEngine
    from oracle import golden_inputs as GI
    eng = BelugaEngine(0, dtype="bf16", cta_group=2, capacity=16)
    eng.load_state_dict(weights_f16)
    x = synth.random_onehot(GI.BELUGA_N, seed=GI.BELUGA_SEED)
    x[0, :, 100:110] = 0
    err = np.abs(eng.forward_onehot(torch.from_numpy(x).cuda()).cpu().numpy() - golden("beluga_golden.npz")["probs"]).max()
    eng.close()
    assert err <= 3e-2


def test_cta_group_1_matches_cta_group_2(engine, weights_f16):
    from expecto_b200 import synth
    from expecto_b200.engine import BelugaEngine
    eng = BelugaEngine(0, dtype="f16", cta_group=1, capacity=32)
    eng.load_state_dict(weights_f16)
    x = torch.from_numpy(synth.random_onehot(20, seed=4)).cuda()
    a, b = eng.forward_onehot(x).cpu().numpy(), engine.forward_onehot(x).cpu().numpy()
    eng.close()
    assert np.abs(a - b).max() <= 2e-5  # same operands; only the fp32 split-K atomic order differs


def test_state_errors(small_genome):
    from expecto_b200.engine import BelugaEngine
    eng = BelugaEngine(0, capacity=8)
    with pytest.raises(_lib.BelugaError):
        eng.forward_onehot(torch.zeros((1, 4, 2000), dtype=torch.uint8, device="cuda"))
    eng.close()


def test_fused_transform_scoring_kernel_vs_oracle():
    """decay_score_variant == gblinear(variant_transform(x)) for several models incl. a keep-mask-style zero block."""
    from expecto_b200.transform import decay_score_variant
    rng = np.random.default_rng(3)
    n, m = 131, 11
    x = (rng.standard_normal((9, n, 2002)) * 0.05).astype(np.float32)
    dist = rng.integers(-30000, 30000, size=n).astype(np.float64)
    strand = np.where(rng.random(n) < 0.5, "+", "-")
    W = (rng.standard_normal((m, 20020)) * 0.01).astype(np.float32)
    W[3].reshape(10, 2002)[:, 100:700] = 0
    off = rng.standard_normal(m).astype(np.float32)
    ours = decay_score_variant(torch.from_numpy(x).cuda(), dist, strand, W, off).cpu().numpy()
    feats = O.variant_transform(list(x), O.variant_weights(dist, strand))
    want = feats @ W.astype(np.float64).T + off[None, :]
    assert ours.shape == (n, m)
    assert np.abs(ours - want).max() <= 1e-4 * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("n,m", [(131, 11), (700, 218), (256, 1), (1, 300)])
def test_tensor_core_scoring_kernel_vs_float64_oracle(n, m):
    """bx_decay_score_variant_tc (features generated in shared memory, one tcgen05 GEMM with split fp16 operands) vs
    gblinear(variant_transform(x)) in float64 (predict.py:87-160): partial tiles, one model, more than 256 models
    (two launches), variants at distance 0 / on shift boundaries, a keep-mask-style zero block; and vs the SIMT
    kernel.  The bar is fp32-level agreement: SED = ALT - REF is a difference of scores that agree in 3-4 digits."""
    from expecto_b200.transform import ScoreModels, decay_score_variant
    rng = np.random.default_rng(3 + n)
    x = rng.random((9, n, 2002)).astype(np.float32)           # probabilities, as ref / alt rows are
    dist = rng.integers(-30000, 30000, size=n).astype(np.float64)
    dist[:min(n, 6)] = [0, 200, -200, 199, -800, 800][:min(n, 6)]
    strand = np.where(rng.random(n) < 0.5, "+", "-")
    W = (rng.standard_normal((m, 20020)) * 0.01).astype(np.float32)
    W[m // 3].reshape(10, 2002)[:, 100:700] = 0
    off = rng.standard_normal(m).astype(np.float32)
    models = ScoreModels(W, off, "cuda:0")
    xd = torch.from_numpy(x).cuda()
    tc = decay_score_variant(xd, dist, strand, models).cpu().numpy()
    simt = decay_score_variant(xd, dist, strand, models, tensor_cores=False).cpu().numpy()
    feats = O.variant_transform(list(x), O.variant_weights(dist, strand))
    want = feats @ W.astype(np.float64).T + off[None, :]
    scale = max(1.0, np.abs(want).max())
    assert tc.shape == (n, m)
    # what remains is the tensor core's truncating accumulation over the 1260-MMA main chain (score_tc.cuh)
    assert np.abs(tc - want).max() <= 5e-5 * scale, np.abs(tc - want).max() / scale
    assert np.abs(simt - want).max() <= 1e-4 * scale
    # a second call reuses the prepared weights and is deterministic
    assert np.array_equal(decay_score_variant(xd, dist, strand, models).cpu().numpy(), tc)


def test_tensor_core_scoring_padded_rows_equal_dense_rows_bitwise():
    """x with rows X_PITCH floats apart (transform.padded_rows: what FC2's epilogue leaves for predict) against the same
    values in dense 8008-byte rows: identical scores, the padding columns (NaN here) are never read; select_variants
    keeps the padding through a gene-row gather; odd pitches and pitches below 2002 are refused by the library."""
    import ctypes as C

    from expecto_b200 import _lib
    from expecto_b200.transform import X_PITCH, ScoreModels, decay_score_variant, padded_rows, select_variants, _row_pitch
    rng = np.random.default_rng(17)
    n, m = 300, 37
    buf = torch.full((9, n, X_PITCH), float("nan"), device="cuda")
    x = buf[..., :2002]
    x.copy_(torch.from_numpy(rng.random((9, n, 2002)).astype(np.float32)))
    assert _row_pitch(x) == X_PITCH and _row_pitch(padded_rows((9, 5), "cuda")) == X_PITCH
    dist = rng.integers(-3000, 3000, size=n).astype(np.float64)
    strand = np.where(rng.random(n) < 0.5, "+", "-")
    models = ScoreModels((rng.standard_normal((m, 20020)) * 0.01).astype(np.float32), rng.standard_normal(m).astype(np.float32), "cuda:0")
    padded = decay_score_variant(x, dist, strand, models)
    dense = decay_score_variant(x.contiguous(), dist, strand, models)
    assert torch.isfinite(padded).all() and torch.equal(padded, dense)
    sel = torch.from_numpy(rng.integers(0, n, size=211)).cuda()
    xs = select_variants(x, sel)
    assert _row_pitch(xs) == X_PITCH and torch.equal(xs, x.index_select(1, sel))
    assert torch.equal(decay_score_variant(xs, dist[sel.cpu().numpy()], strand[sel.cpu().numpy()], models), dense[sel])
    # edge tiles of the TMA-staged kernel: one row, a tile less / more than one row, an empty call
    for n1 in (1, 255, 257):
        xe = padded_rows((9, n1), "cuda")
        xe.copy_(x[:, :n1])
        assert _row_pitch(xe) == X_PITCH
        assert torch.equal(decay_score_variant(xe, dist[:n1], strand[:n1], models), dense[:n1])
    assert decay_score_variant(padded_rows((9, 0), "cuda"), dist[:0], strand[:0], models).shape == (0, m)
    # several tiles per CTA pair and 218 models: the kernel that stages x by TMA (aligned rows) against the one that
    # loads it through registers (dense rows) — same operand layout and MMA order, so every bit agrees
    n2 = 256 * 2 * 74 + 77
    big = padded_rows((9, n2), "cuda")
    big.uniform_()
    d2 = torch.from_numpy(rng.integers(-20000, 20000, size=n2).astype(np.float64)).cuda()
    s2 = torch.from_numpy(np.where(rng.random(n2) < 0.5, 1, -1).astype(np.int8)).cuda()
    m2 = ScoreModels((rng.standard_normal((218, 20020)) * 0.01).astype(np.float32), rng.standard_normal(218).astype(np.float32), "cuda:0")
    a = decay_score_variant(big, d2, s2, m2)
    assert torch.equal(a, decay_score_variant(big.contiguous(), d2, s2, m2))
    assert torch.equal(a, decay_score_variant(big, d2, s2, m2))
    del big
    # rows grouped by side of the TSS: whole K steps are zero for every row of most tiles and are skipped (tile classes
    # 1 and 2, score_tile_class_kernel) — the scores are those of the unsorted run, bit for bit
    from expecto_b200.transform import side_order
    d2h, s2h = d2.cpu().numpy(), s2.cpu().numpy()
    o = side_order(d2h, s2h)
    pos = (d2h[o, None] + np.array([0, -200, -400, -600, -800, 200, 400, 600, 800])[None, :]) * np.where(s2h[o] > 0, 1, -1)[:, None]
    cls = np.where((pos < 0).all(1), 0, np.where((pos > 0).all(1), 2, 1))
    assert (np.diff(cls) >= 0).all() and {0, 1, 2} <= set(cls.tolist())
    ot = torch.from_numpy(o).cuda()
    big = padded_rows((9, n2), "cuda")
    torch.manual_seed(5)
    big.uniform_()
    base = decay_score_variant(big, d2, s2, m2)
    srt = select_variants(big, ot)
    assert _row_pitch(srt) == X_PITCH
    assert torch.equal(decay_score_variant(srt, d2[ot], s2[ot], m2), base[ot])
    del big, srt
    L = _lib.lib()
    out = torch.empty((n, m), device="cuda")
    for bad in (2001, 2003, 0):
        assert L.bx_decay_score_variant_tc(x.data_ptr(), bad, None, None, None, None, None, n, 9, m, out.data_ptr(), None) == _lib.BX_E_ARG
        assert b"x_row_pitch" in L.bx_last_error()


def test_predict_cli_matches_predict_script(engine, cfg1_genome, golden, tmp_path):
    """chromatin CLI -> predict CLI (fused transform + scoring) vs the reference predict.py's sed.tsv."""
    import pandas as pd
    from expecto_b200 import chromatin, predict
    g = golden("predict_golden.npz")
    want = pd.read_csv(io.StringIO(str(g["sed_tsv"])), sep="\t")
    rows = [l.split("\t") for l in str(golden("chromatin_cfg1_golden.npz")["snps_hg19_vcf"]).splitlines()[2:]]
    vcf = tmp_path / "example.vcf"
    vcf.write_text("".join("\t".join(r) + "\n" for r in rows))
    out = tmp_path / "chromatin_out"
    chromatin.run(chromatin.build_parser().parse_args([str(vcf), "--output_dir", str(out)]), engine=engine,
                  genome=PackedGenome.from_sequences(cfg1_genome))
    genefile = tmp_path / "closestgene"
    with open(genefile, "w") as f:
        for r, (_, w) in zip(rows, want.iterrows()):
            f.write("\t".join([r[0], str(int(r[1]) - 1), r[1], r[3], r[4], w["strand"], w["gene"], str(-int(w["dist"]))]) + "\n")
    np.savez(tmp_path / "model.npz", weights=g["model_w"], bias=np.float32(0.25), base_score=np.float32(0.5))
    feats_tsv = tmp_path / "features.tsv"
    with open(feats_tsv, "w") as f:
        f.write("\tCell type\tAssay\tTreatment\tAssay type\tSource\n")
        for i in range(2002):
            f.write(f"{i}\tcell{i}\tassay{i}\tNone\t{'DNase' if i % 3 == 0 else 'TF'}\tsrc\n")
    args = predict.build_parser().parse_args([
        "--model_save_file", str(tmp_path / "model.npz"), "--belugaFeatures", str(feats_tsv), "--coorFile_chromatin",
        str(out / "snps_hg19.vcf"), "--geneFile", str(genefile), "--snpEffectFilePattern",
        str(out / "snps.shift_SHIFT.diff.h5"), "-o", str(tmp_path / "pred")])
    predict.run(args)
    ours = pd.read_csv(tmp_path / "pred" / "sed.tsv", sep="\t")
    assert list(ours.columns) == list(want.columns)
    for c in ("index", "0", "1", "2", "3", "4", "dist", "gene", "strand"):
        assert [str(v) for v in ours[c]] == [str(v) for v in want[c]], c
    _check_sed(ours["REF"].values, ours["ALT"].values, ours["SED"].values, want, "predict CLI, f16 forward")
    # SED is the float64 difference of the float32 scores (predict.py:264), printed like the reference prints it
    assert np.abs(ours["SED"].values - (ours["ALT"].values - ours["REF"].values)).max() <= 1e-12
    assert len(repr(float(ours["REF"].values[0]))) > 12          # widened float32, e.g. 3.9250710010528564
    for name in ("sed_sorted_by_magnitude.tsv", "sed_sorted_by_proportion.tsv"):
        assert (tmp_path / "pred" / name).exists()
    # one pass, no .diff.h5 in between (--vcf): byte-identical tables, with the variants in two HBM blocks
    args = predict.build_parser().parse_args([
        "--model_save_file", str(tmp_path / "model.npz"), "--belugaFeatures", str(feats_tsv), "--geneFile",
        str(genefile), "--vcf", str(vcf), "--vcf_block", "6", "-o", str(tmp_path / "pred_fused")])
    info = predict.run(args, engine=engine, genome=PackedGenome.from_sequences(cfg1_genome))
    assert info["n_rows"] == len(ours)
    assert (tmp_path / "pred_fused" / "snps_hg19.vcf").read_bytes() == (out / "snps_hg19.vcf").read_bytes()
    for name in ("sed.tsv", "sed_sorted_by_magnitude.tsv", "sed_sorted_by_proportion.tsv"):
        assert (tmp_path / "pred_fused" / name).read_bytes() == (tmp_path / "pred" / name).read_bytes(), name


def test_predict_one_pass_with_split_operands_reproduces_sed(weights_f16, cfg1_genome, golden, tmp_path):
    """The same predict.py fixture with --dtype f16x3 (one pass, --vcf): the fp32 reference's SED to within 1 %% of its
    largest value and Pearson >= 0.9999 — what the split-operand forward buys for the pipeline's scientific output."""
    import pandas as pd
    from expecto_b200 import predict
    from expecto_b200.engine import BelugaEngine
    g = golden("predict_golden.npz")
    want = pd.read_csv(io.StringIO(str(g["sed_tsv"])), sep="\t")
    rows = [l.split("\t") for l in str(golden("chromatin_cfg1_golden.npz")["snps_hg19_vcf"]).splitlines()[2:]]
    vcf = tmp_path / "example.vcf"
    vcf.write_text("".join("\t".join(r) + "\n" for r in rows))
    genefile = tmp_path / "closestgene"
    with open(genefile, "w") as f:
        for r, (_, w) in zip(rows, want.iterrows()):
            f.write("\t".join([r[0], str(int(r[1]) - 1), r[1], r[3], r[4], w["strand"], w["gene"], str(-int(w["dist"]))]) + "\n")
    np.savez(tmp_path / "model.npz", weights=g["model_w"], bias=np.float32(0.25), base_score=np.float32(0.5))
    feats_tsv = tmp_path / "features.tsv"
    with open(feats_tsv, "w") as f:
        f.write("\tCell type\tAssay\tTreatment\tAssay type\tSource\n")
        for i in range(2002):
            f.write(f"{i}\tcell{i}\tassay{i}\tNone\t{'DNase' if i % 3 == 0 else 'TF'}\tsrc\n")
    eng = BelugaEngine(0, dtype="f16x3", capacity=256)
    eng.load_state_dict(weights_f16)
    args = predict.build_parser().parse_args([
        "--model_save_file", str(tmp_path / "model.npz"), "--belugaFeatures", str(feats_tsv), "--geneFile",
        str(genefile), "--vcf", str(vcf), "-o", str(tmp_path / "pred")])
    predict.run(args, engine=eng, genome=PackedGenome.from_sequences(cfg1_genome))
    eng.close()
    ours = pd.read_csv(tmp_path / "pred" / "sed.tsv", sep="\t")
    _check_sed(ours["REF"].values, ours["ALT"].values, ours["SED"].values, want, "predict CLI, f16x3 forward",
               pearson_min=0.9999, rel_max=0.01)


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_two_devices_in_one_process(weights_f16):
    """Contexts on two GPUs of ONE process (the C ABI allows one context per GPU): kernel attributes are set per
    device, context-free entry points follow the tensors' device; both GPUs give bit-identical results."""
    from expecto_b200 import synth
    from expecto_b200.engine import BelugaEngine, chromatin_outputs
    from expecto_b200.transform import decay_transform_fixed, fixed_pos_weights
    x = torch.from_numpy(synth.random_onehot(8, seed=9))
    outs = []
    for d in (0, 1):
        eng = BelugaEngine(d, dtype="f16", cta_group=2, capacity=64)
        eng.load_state_dict(weights_f16)
        probs = eng.forward_onehot(x.to(f"cuda:{d}"))
        ref, alt, diff = chromatin_outputs(probs, 2)
        feats = decay_transform_fixed(torch.rand((3, 200, 2, 2002), generator=torch.Generator().manual_seed(1)).to(f"cuda:{d}"),
                                      fixed_pos_weights())
        torch.cuda.synchronize(d)
        outs.append((probs.cpu(), diff.cpu(), feats.cpu()))
        eng.close()
    for a, b in zip(*outs):
        assert torch.equal(a, b)


def test_output_buffers_are_written_inside_their_bounds(engine):
    """Guard rows around the caller's output buffers (compute-sanitizer is closed on this pool): the FC2 epilogue
    (float2 stores, 2002-column rows), the per-window / segment forwards and both decay transforms must leave the rows
    before and after their output untouched, for batch sizes that do not fill a tile."""
    from expecto_b200 import synth
    from expecto_b200.transform import decay_transform_fixed, decay_transform_variant, fixed_pos_weights
    G, SENT = 3, -123.25

    def guarded(rows, cols):
        buf = torch.full((rows + 2 * G, cols), SENT, dtype=torch.float32, device="cuda")
        return buf, buf[G:G + rows]

    def intact(buf, rows):
        return bool((buf[:G] == SENT).all() and (buf[G + rows:] == SENT).all())

    for n in (1, 37, 129):
        x = torch.from_numpy(synth.random_onehot(n, seed=n)).cuda()
        buf, out = guarded(n, 2002)
        engine.forward_onehot(x, out)
        torch.cuda.synchronize()
        assert intact(buf, n) and bool(((out > 0) & (out < 1)).all())
        assert torch.equal(out, engine.forward_onehot(x))
    probs = torch.rand((5, 200, 2, 2002), dtype=torch.float32, device="cuda")
    buf, out = guarded(5, 20020)
    decay_transform_fixed(probs, fixed_pos_weights(), out=out)
    torch.cuda.synchronize()
    assert intact(buf, 5) and torch.equal(out, decay_transform_fixed(probs, fixed_pos_weights()))
    xs = torch.rand((9, 7, 2002), dtype=torch.float32, device="cuda")
    dist = np.array([-900.0, -1.0, 0.0, 1.0, 350.0, 12345.0, -40000.0])
    strand = np.array(["+", "-", "+", "-", "+", "-", "+"])
    buf, out = guarded(7, 20020)
    decay_transform_variant(xs, dist, strand, out=out)
    torch.cuda.synchronize()
    assert intact(buf, 7) and torch.equal(out, decay_transform_variant(xs, dist, strand))
