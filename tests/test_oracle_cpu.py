"""The CPU oracle (oracle/beluga_oracle.py) against the fixtures the UNMODIFIED reference produced
(oracle/make_golden.py, run in the build container where /root/reference exists)."""
import io

import numpy as np
import pytest

from oracle import beluga_oracle as O
from oracle import golden_inputs as GI


def _vcf_rows(text):
    return [l.split("\t") for l in str(text).splitlines() if l and not l.startswith("#")]


def test_encode_matches_reference_fixture(small_genome, golden):
    g = golden("encode_golden.npz")
    rows = [[r[0], int(r[1]), r[2], r[3], r[4]] for r in g["rows"]]
    dg = O.DictGenome(small_genome)
    for k, (shift, which) in enumerate(zip(g["shifts"], g["which"])):
        seqs = [O.fetch_seqs(dg, r[0], r[1], r[3], r[4], shift=int(shift))[0 if which == "ref" else 1] for r in rows]
        enc = O.encode_seqs(seqs)
        want = np.unpackbits(g["packed"][k], axis=1)[:, :8000].reshape(-1, 4, 2000).astype(bool)
        assert np.array_equal(enc, want), (shift, which)
    odd = [str(s) for s in g["odd"]]
    want = np.unpackbits(g["odd_packed"], axis=1)[:, :8000].reshape(-1, 4, 2000).astype(bool)
    assert np.array_equal(O.encode_seqs(odd), want)


def test_encode_keyerror_on_unknown_base():
    with pytest.raises(KeyError):
        O.encode_seqs(["ACGR" * 500])


def test_beluga_forward_matches_reference_fixture(weights_f16, golden):
    from expecto_b200 import synth
    x = synth.random_onehot(GI.BELUGA_N, seed=GI.BELUGA_SEED)
    x[0, :, 100:110] = 0
    ours = O.beluga_forward(x.astype(np.float32), weights_f16)
    assert np.abs(ours - golden("beluga_golden.npz")["probs"]).max() < 2e-6


def test_chromatin_shift_matches_reference_script(cfg1_genome, weights_f16, golden):
    """chromatin.py run as a subprocess on example.vcf (cfg1): shift 0 ref/alt + the printed counters."""
    g = golden("chromatin_cfg1_golden.npz")
    rows = O.normalise_vcf_rows(_vcf_rows(g["snps_hg19_vcf"]))
    assert len(rows) == 10
    d, r, a, cnt = O.chromatin_shift(O.DictGenome(cfg1_genome), rows, weights_f16, 0)
    assert list(cnt) == list(g["counters"])
    assert np.abs(r - g["ref_0"]).max() < 2e-6 and np.abs(a - g["alt_0"]).max() < 2e-6
    assert np.allclose(d.astype(np.float64).sum(axis=1), g["row_sums"][0, 0], atol=1e-4)


def test_filter_and_prefix_rules(golden):
    g = golden("chromatin_edge_golden.npz")
    rows = [list(r) for r in g["rows"]]
    kept = O.normalise_vcf_rows(rows)
    assert [r[0] for r in kept] == ["chr1", "chr2", "chrX", "chr1", "chr1", "chr2"]
    assert g["ref_0"].shape == (2 * len(kept), 2002)


def test_fixed_transform_matches_reference_script(cfg1_genome, weights_f16, golden):
    """compute_expecto_features.py subprocess output for gene 0 (400 windows on the CPU oracle)."""
    g = golden("features_golden.npz")
    gid, chrom, strand, tss = g["genes"][0]
    dg = O.DictGenome(cfg1_genome)
    enc = O.encode_seqs(O.tss_windows(dg, str(chrom), int(tss), 1 if strand == "+" else -1), concat_rc=False)
    fwd = np.vstack([O.beluga_forward(enc[i:i + 50].astype(np.float32), weights_f16) for i in range(0, 200, 50)])
    rc_in = np.ascontiguousarray(enc[:, ::-1, ::-1]).astype(np.float32)
    rc = np.vstack([O.beluga_forward(rc_in[i:i + 50], weights_f16) for i in range(0, 200, 50)])
    feats = O.fixed_transform(fwd, rc, O.fixed_pos_weights())
    ref = g["features"][0]
    assert feats.dtype == np.float64 and feats.shape == (20020,)
    assert np.abs(feats - ref).max() / np.abs(ref).max() < 1e-6


def test_variant_transform_matches_predict_script(cfg1_genome, weights_f16, golden):
    """predict.py subprocess: the diff feature row of variant 0 and its REF/ALT scores in sed.tsv."""
    import pandas as pd
    g = golden("predict_golden.npz")
    sed = pd.read_csv(io.StringIO(str(g["sed_tsv"])), sep="\t")
    rows = O.normalise_vcf_rows(_vcf_rows(golden("chromatin_cfg1_golden.npz")["snps_hg19_vcf"]))[:1]
    dg = O.DictGenome(cfg1_genome)
    data = {"diff": [], "ref": [], "alt": []}
    for s in O.SHIFTS_CHROMATIN(800):
        d, r, a, _ = O.chromatin_shift(dg, rows, weights_f16, s)
        for k, v in (("diff", d), ("ref", r), ("alt", a)):
            data[k].append(O.strand_average(v))
    w = O.variant_weights(np.array([sed["dist"][0]]), np.array([sed["strand"][0]]))
    feats = O.variant_transform(data["diff"], w)
    ref = g["diff_features_row0"]
    # the reference ran all 10 variants in one oneDNN batch, this test runs variant 0 alone: fp32 summation order
    # differs at the 1e-6 level per probability, 9 shifts are summed
    assert np.abs(feats[0] - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max())
    for key, name in (("ref", "REF"), ("alt", "ALT")):
        score = O.gblinear_predict(O.variant_transform(data[key], w), g["model_w"], 0.25, 0.5)
        assert abs(float(score[0]) - float(sed[name][0])) < 1e-4


def test_variant_weights_edge_cases():
    """predict.py:87-109: at p == 0 both halves carry weight 1; floor(|p|/200) steps."""
    w = O.variant_weights(np.array([0, 199, 200, -1]), np.array(["+", "+", "-", "-"]))
    w0 = w[0]  # shift 0
    assert np.allclose(w0[0], 1.0)
    assert np.allclose(w0[1, :5], 0) and np.allclose(w0[1, 5:], 1.0)           # p = 199 -> floor 0, upstream half off
    assert np.allclose(w0[2, :5], np.exp(-np.array([0.01, 0.02, 0.05, 0.1, 0.2])))  # '-' strand: p = -200
    assert np.allclose(w0[3, 5:], 1.0) and np.allclose(w0[3, :5], 0)           # '-' strand: p = +1


def test_tss_override_rules_match_reference_script(golden, tmp_path):
    """compute_expecto_features.py:54-69 with a TSS table that matches both genes (oracle/make_golden.py
    stage_features_override ran the reference behind the identity liftover shim and checked which coordinates it
    used): a non-default row replaces chrom, TSS and strand; a default row still replaces the strand."""
    from expecto_b200 import compute_expecto_features as cef
    g = golden("features_override_golden.npz")
    (tmp_path / "anno.csv").write_text(str(g["anno"]))
    (tmp_path / "tss.tsv").write_text(str(g["tss"]))
    got = cef.read_genes(str(tmp_path / "anno.csv"), str(tmp_path / "tss.tsv"), tss_is_hg19=True)
    want = [(a, b, int(c), int(d)) for a, b, c, d in g["expected"]]
    assert got == want
    with pytest.raises(NotImplementedError):      # without --tss_is_hg19 the override would need the liftover package
        cef.read_genes(str(tmp_path / "anno.csv"), str(tmp_path / "tss.tsv"))
