"""Host logic of the predict.py drop-in: model loaders, gene-row run lengths, keep-mask expansion."""
import os

import numpy as np
import pandas as pd
import pytest

from expecto_b200 import gblinear, predict
from expecto_b200.predict import build_parser, expand_weights, num_repeats
from expecto_b200.transform import chromatin_shifts, fixed_pos_weights
from oracle import beluga_oracle as O


def test_gblinear_formats_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    m = gblinear.GBLinearModel(rng.standard_normal(20020).astype(np.float32), bias=0.25, base_score=0.5)
    gblinear.save_binary(str(tmp_path / "m.save"), m)
    b = gblinear.load_model(str(tmp_path / "m.save"))
    assert np.array_equal(b.weights, m.weights) and b.bias == 0.25 and b.base_score == 0.5
    np.savez(tmp_path / "m.npz", weights=m.weights, bias=np.float32(0.25), base_score=np.float32(0.5))
    z = gblinear.load_model(str(tmp_path / "m.npz"))
    assert np.array_equal(z.weights, m.weights) and z.offset == 0.75
    with open(tmp_path / "m.dump", "w") as f:
        f.write("booster[0]:\nbias:\n0.25\nweight:\n" + "\n".join(repr(float(v)) for v in m.weights[:50]) + "\n")
    t = gblinear.load_model(str(tmp_path / "m.dump"))
    assert np.array_equal(t.weights, m.weights[:50]) and t.bias == 0.25
    x = rng.standard_normal((3, 20020))
    assert np.allclose(O.gblinear_predict(x, b.weights, b.bias, b.base_score), 0.75 + x @ m.weights.astype(np.float64))


def test_binary_loader_rejects_other_boosters(tmp_path):
    p = tmp_path / "tree.save"
    import struct
    blob = struct.pack("<fIiii", 0.5, 10, 0, 0, 0) + b"\x00" * 116
    for name in ("reg:linear", "gbtree"):
        blob += struct.pack("<Q", len(name)) + name.encode()
    p.write_bytes(blob + b"\x00" * 200)
    with pytest.raises(ValueError):
        gblinear.load_binary(str(p))


def test_num_repeats_matches_reference_loop():
    rows = [["chr1", 1, 2, "A", "C", "+", "g1", 5], ["chr1", 1, 2, "A", "C", "-", "g2", -7],
            ["chr1", 9, 10, "G", "T", "+", "g3", 1], ["chr2", 9, 10, "G", "T", "+", "g3", 1],
            ["chr2", 9, 10, "G", "T", "-", "g4", 2], ["chr2", 9, 10, "G", "T", "-", "g5", 3]]
    assert num_repeats(pd.DataFrame(rows)) == [2, 1, 3]
    assert num_repeats(pd.DataFrame(rows[:1])) == [1]


def test_expand_weights_zero_fills_dropped_columns():
    mask = np.ones(2002, dtype=bool)
    mask[[0, 5, 2001]] = False
    m = gblinear.GBLinearModel(np.arange(10 * 1999, dtype=np.float32), 0.1, 0.5)
    W, off = expand_weights([m], mask)
    assert W.shape == (1, 20020) and np.isclose(off[0], 0.6)
    W3 = W.reshape(10, 2002)
    assert np.all(W3[:, ~mask] == 0) and np.array_equal(W3[:, mask].ravel(), m.weights)
    with pytest.raises(ValueError):
        expand_weights([gblinear.GBLinearModel(np.zeros(20020, np.float32), 0, 0)], mask)


def test_cli_defaults_match_reference():
    a = build_parser().parse_args(["-o", "x"])
    assert (a.nfeatures, a.fixeddist, a.maxshift, a.batchSize, a.threads) == (2002, 0, 800, 500, 16)
    assert chromatin_shifts(800) == O.SHIFTS_CHROMATIN(800)
    assert np.array_equal(fixed_pos_weights(), O.fixed_pos_weights())


@pytest.mark.parametrize("dt", [np.float32, np.float64])
def test_sed_tables_byte_identical_to_three_pandas_passes(tmp_path, dt):
    """write_sed_tables formats every row once; the three files must equal predict.py:255-280's three to_csv calls
    byte for byte, including NaN / inf proportions, zero effects and repeated magnitudes."""
    import pandas as pd
    from expecto_b200.predict import write_sed_tables
    rng = np.random.default_rng(0)
    n = 400
    ref = rng.standard_normal(n).astype(dt)
    alt = ref + rng.standard_normal(n).astype(dt) * dt(1e-3)
    alt[5] = ref[5]; ref[7] = 0; alt[7] = 0; alt[9] = -ref[9]; alt[20:30] = ref[20:30] + dt(0.5); ref[40] = np.nan
    df = pd.DataFrame({"index": np.arange(n), 0: ["chr1"] * n, 1: rng.integers(1, 10 ** 6, n),
                       2: [f"rs{i}" for i in range(n)], "dist": rng.integers(-20000, 20000, n),
                       "gene": [f"G{i % 7}" for i in range(n)], "strand": ["+"] * n})
    df = pd.concat([df] + [pd.DataFrame(v, columns=[k]) for k, v in
                           (("REF", ref), ("ALT", alt), ("SED", alt - ref))], axis=1)   # the reference's columns
    ours, theirs = tmp_path / "ours", tmp_path / "theirs"
    ours.mkdir(); theirs.mkdir()
    write_sed_tables(df, str(ours))
    df.to_csv(f"{theirs}/sed.tsv", header=True, sep="\t", index=False)
    s = df.copy()
    s["SED_MAGNITUDES"] = np.abs(s["SED"])
    s.sort_values(by="SED_MAGNITUDES", axis=0, ascending=False).to_csv(
        f"{theirs}/sed_sorted_by_magnitude.tsv", header=True, sep="\t", index=False)
    s = df.copy()
    s["SED_PROPORTION"] = np.abs(s["SED"] / ((s["REF"] + s["ALT"]) / 2))
    s.sort_values(by="SED_PROPORTION", axis=0, ascending=False).to_csv(
        f"{theirs}/sed_sorted_by_proportion.tsv", header=True, sep="\t", index=False)
    for name in ("sed.tsv", "sed_sorted_by_magnitude.tsv", "sed_sorted_by_proportion.tsv"):
        assert (ours / name).read_bytes() == (theirs / name).read_bytes(), name


def test_model_list_columns_round_trip(tmp_path):
    """--modelList extension: the reference's columns keep pandas' text, the extra SED_<model> columns are written by
    the fast path and must read back to the same float32 values (incl. NaN as empty)."""
    import pandas as pd
    from expecto_b200.predict import write_sed_tables
    rng = np.random.default_rng(1)
    n = 300
    ref = rng.standard_normal(n).astype(np.float32)
    alt = ref + (rng.standard_normal(n) * 1e-3).astype(np.float32)
    extra = (rng.standard_normal((n, 5)) * np.float32(10.0) ** rng.integers(-8, 8, (n, 5))).astype(np.float32)
    extra[3, 2] = np.nan; extra[4, 1] = 0.0; extra[5, 0] = 1e-5; extra[6, 0] = 123456789.0
    df = pd.DataFrame({"index": np.arange(n), 0: ["chr1"] * n, "gene": [f"G{i}" for i in range(n)]})
    cols = [("REF", ref), ("ALT", alt), ("SED", alt - ref)] + [(f"SED_model{i}.save", extra[:, i]) for i in range(5)]
    df = pd.concat([df] + [pd.DataFrame(v, columns=[k]) for k, v in cols], axis=1)
    write_sed_tables(df, str(tmp_path))
    base = df[list(df.columns)[:6]]
    got_text = (tmp_path / "sed.tsv").read_text().split("\n")
    want_text = base.to_csv(None, header=True, sep="\t", index=False).split("\n")
    assert all(g.startswith(w + "\t") for g, w in zip(got_text[:-1], want_text[:-1]))
    for name in ("sed.tsv", "sed_sorted_by_magnitude.tsv", "sed_sorted_by_proportion.tsv"):
        back = pd.read_csv(tmp_path / name, sep="\t")
        assert list(back.columns)[:11] == [str(c) for c in df.columns]
        back = back.sort_values("index").reset_index(drop=True)
        for i in range(5):
            assert np.array_equal(back[f"SED_model{i}.save"].values.astype(np.float32), extra[:, i], equal_nan=True)


def test_merge_predict_batched_natural_order_and_index(tmp_path):
    """merge_predict_batched.py:22-26: chunk directories in natural order (chunk_2 before chunk_10), the chunk count
    asserted, pandas' default index (restarting in every chunk) written as the first column."""
    import pandas as pd
    from expecto_b200 import merge_predict_batched as mpb
    frames = {}
    for k in (10, 2, 0, 1):
        d = tmp_path / "batches" / f"chunk_{k}"
        d.mkdir(parents=True)
        frames[k] = pd.DataFrame({"index": [0, 1], 0: ["chr1", "chr2"], 1: [100 + k, 200 + k], "SED": [0.125 * k, -1.5e-7]})
        frames[k].to_csv(d / "sed.tsv", sep="\t", index=False)
    args = mpb.build_parser().parse_args(["--batch_dir", str(tmp_path / "batches"), "--n_chunks", "4", "-o", str(tmp_path / "out")])
    out = mpb.run(args)
    want = pd.concat([pd.read_csv(tmp_path / "batches" / f"chunk_{k}" / "sed.tsv", sep="\t") for k in (0, 1, 2, 10)])
    import io
    buf = io.StringIO()
    want.to_csv(buf, sep="\t")
    assert open(out).read() == buf.getvalue()
    assert open(out).read().split("\n")[1].startswith("0\t0\tchr1\t100")
    args.n_chunks = 5
    with pytest.raises(AssertionError):
        mpb.run(args)


def test_gblinear_binary_layout_built_from_the_xgboost_0_7_structs(tmp_path):
    """A ``.save`` file assembled here byte by byte from xgboost 0.7's serialisation code (no shared code with
    gblinear.save_binary), with the optional pieces a real file can carry: the "binf" prefix LearnerImpl::Load skips,
    contain_extra_attrs = 1 with an attribute vector after the booster, and contain_eval_metrics = 1."""
    import struct
    rng = np.random.default_rng(8)
    w = rng.standard_normal(20020).astype("<f4")
    bias, base_score = np.float32(-0.125), np.float32(0.5)

    def dmlc_string(b):                       # dmlc::Stream::Write(std::string): uint64 length + bytes
        return struct.pack("<Q", len(b)) + b

    def build(prefix, extra_attrs, eval_metrics):
        # src/learner.cc: struct LearnerModelParam { float base_score; unsigned num_feature; int num_class;
        #                 int contain_extra_attrs; int contain_eval_metrics; int reserved[29]; }  -> 20 + 116 bytes
        out = prefix + struct.pack("<fIiii", base_score, w.size, 0, extra_attrs, eval_metrics) + bytes(4 * 29)
        out += dmlc_string(b"reg:linear") + dmlc_string(b"gblinear")      # name_obj_, name_gbm_
        # src/gbm/gblinear.cc: struct GBLinearModelParam { unsigned num_feature; int num_output_group;
        #                      int reserved[32]; } -> 8 + 128 bytes; then std::vector<bst_float> weight with the
        #                      bias of output group 0 after the num_feature feature weights
        out += struct.pack("<Ii", w.size, 1) + bytes(4 * 32)
        out += struct.pack("<Q", w.size + 1) + w.tobytes() + struct.pack("<f", bias)
        if extra_attrs:                       # std::vector<std::pair<std::string, std::string>>
            out += struct.pack("<Q", 1) + dmlc_string(b"best_iteration") + dmlc_string(b"99")
        if eval_metrics:                      # std::vector<std::string>
            out += struct.pack("<Q", 1) + dmlc_string(b"rmse")
        return out

    assert len(build(b"", 0, 0)) == 136 + (8 + 10) + (8 + 8) + 136 + 8 + 4 * 20021
    for i, (prefix, ea, em) in enumerate([(b"", 0, 0), (b"binf", 0, 0), (b"", 1, 1), (b"binf", 1, 0)]):
        p = tmp_path / f"m{i}.save"
        p.write_bytes(build(prefix, ea, em))
        m = gblinear.load_model(str(p))
        assert np.array_equal(m.weights, w) and m.bias == float(bias) and m.base_score == float(base_score)
    # the repo's own writer produces the minimal form of the same layout
    gblinear.save_binary(str(tmp_path / "own.save"), gblinear.GBLinearModel(w, float(bias), float(base_score)))
    assert (tmp_path / "own.save").read_bytes() == build(b"", 0, 0)
    bad = bytearray(build(b"", 0, 0))
    bad[136 + 8 + 10 + 8:136 + 8 + 10 + 8 + 8] = b"gbtree\x00\x00"
    (tmp_path / "tree.save").write_bytes(bytes(bad))
    with pytest.raises(ValueError):
        gblinear.load_model(str(tmp_path / "tree.save"))


def test_gblinear_text_dump_matches_the_reference_parser(tmp_path):
    """The text dump (train.py:158 ``bst.dump_model``) against the reference's own reading of it:
    predict_by_cluster.py:73-75 takes ``get_dump()[0]``, i.e. the file without its ``booster[0]:`` line, and reads
    the bias from line 1 and the weights from line 3 on.  The three statements are lifted from the reference with
    ``ast`` and executed (skipped where /root/reference does not exist)."""
    import ast
    ref = "/root/reference/predict_by_cluster.py"
    if not os.path.exists(ref):
        pytest.skip("reference checkout only exists in the build container")
    rng = np.random.default_rng(9)
    m = gblinear.GBLinearModel(rng.standard_normal(2002 * 10).astype(np.float32) * 1e-3, bias=-0.0625, base_score=0.5)
    path = tmp_path / "m.dump"
    gblinear.save_text_dump(str(path), m)
    back = gblinear.load_model(str(path))
    assert np.array_equal(back.weights, m.weights) and back.bias == m.bias
    fn = [n for n in ast.parse(open(ref).read()).body if isinstance(n, ast.FunctionDef) and n.name == "interpret_model"][0]
    fn.body = fn.body[:3] + [ast.parse("return bias, weights").body[0]]     # dump = ...; bias = ...; weights = ...

    class Booster:
        def get_dump(self):
            return [path.read_text().split("\n", 1)[1]]                     # xgboost: one string per booster

    ns = {"np": np}
    exec(compile(ast.fix_missing_locations(ast.Module(body=[fn], type_ignores=[])), ref, "exec"), ns)
    bias, weights = ns["interpret_model"](Booster(), None, None)
    assert bias == m.bias and np.array_equal(weights.astype(np.float32), m.weights)


def test_keep_mask_matches_reference_cluster_utils(golden, tmp_path):
    """predict.keep_mask against cluster_utils.get_keep_mask run by oracle/make_golden.py on the reference's own
    2002-feature table: all 16 combinations of the four feature-class flags (cluster_utils.py:8-49)."""
    import pandas as pd
    g = golden("keep_mask_golden.npz")
    tsv = tmp_path / "features.tsv"
    pd.DataFrame({"Cell type": "x", "Assay": g["assay"], "Treatment": "", "Assay type": g["assay_type"]},
                 index=np.arange(1, len(g["assay"]) + 1)).to_csv(tsv, sep="\t")

    class Flags:
        intersect_with_lambert = False

    for (tf, dn, hi, po), packed in zip(g["combos"], g["masks"]):
        a = Flags()
        a.no_tf_features, a.no_dnase_features, a.no_histone_features, a.no_pol2 = bool(tf), bool(dn), bool(hi), bool(po)
        want = np.unpackbits(packed)[:len(g["assay"])].astype(bool)
        assert np.array_equal(predict.keep_mask(str(tsv), a), want), (tf, dn, hi, po)


def test_lambert_intersection_matches_reference_cluster_utils(golden, tmp_path):
    """--intersect_with_lambert against cluster_utils.get_keep_mask(…, intersect_with_lambert=True) run by
    oracle/make_golden.py on seeded stand-ins for the two resource tables (cluster_utils.py:26-42): duplicated mapping
    rows, rows without an approved symbol, case changes, alone and combined with --no_dnase_features / --no_pol2."""
    import pandas as pd
    g = golden("keep_mask_lambert_golden.npz")
    tsv = tmp_path / "features.tsv"
    pd.DataFrame({"Cell type": "x", "Assay": g["assay"], "Treatment": "", "Assay type": g["assay_type"]},
                 index=np.arange(1, len(g["assay"]) + 1)).to_csv(tsv, sep="\t")
    (tmp_path / "lambert.csv").write_text(str(g["lambert_csv"]))
    (tmp_path / "mapping.csv").write_text(str(g["mapping_csv"]))

    class Flags:
        no_tf_features = no_histone_features = False
        intersect_with_lambert = True
        lambert_csv, hgnc_mapping_csv = str(tmp_path / "lambert.csv"), str(tmp_path / "mapping.csv")

    kept = []
    for (dn, po), packed in zip(g["combos"], g["masks"]):
        a = Flags()
        a.no_dnase_features, a.no_pol2 = bool(dn), bool(po)
        want = np.unpackbits(packed)[:len(g["assay"])].astype(bool)
        assert np.array_equal(predict.keep_mask(str(tsv), a), want), (dn, po)
        kept.append(int(want.sum()))
    assert 50 < kept[0] < len(g["assay"]) - 50 and kept[3] < kept[0]      # the fixture exercises both outcomes


def test_num_repeats_matches_reference_function(golden):
    """predict.num_repeats against predict.py's own get_num_repeats (:202-213, exec'd by oracle/make_golden.py) on
    seeded gene tables: singleton groups, a variant that reappears later, float-typed position columns."""
    import io
    g = golden("num_repeats_golden.npz")
    for text, want in zip(g["tables"], g["expected"]):
        df = pd.read_csv(io.StringIO(str(text)), sep="\t", header=None, comment="#")
        assert num_repeats(df) == [int(x) for x in str(want).split(",")]


def test_padded_rows_and_variant_selection_keep_the_row_pitch():
    """transform.padded_rows / _row_pitch / select_variants (the row layout FC2's epilogue writes and the scoring kernel
    reads: rows X_PITCH floats apart): views over padded buffers report their pitch, dense and oddly strided tensors
    do not, and a gene-row gather keeps the padding while returning what index_select returns."""
    import torch

    from expecto_b200.transform import N_FEATURES, X_PITCH, _row_pitch, padded_rows, select_variants
    assert X_PITCH % 32 == 0 and X_PITCH >= N_FEATURES          # rows start on 128-byte boundaries
    for n in (0, 1, 2, 77):
        x = padded_rows((9, n), "cpu")
        assert x.shape == (9, n, N_FEATURES) and _row_pitch(x) == X_PITCH
    x = padded_rows((9, 40), "cpu")
    x.copy_(torch.arange(9 * 40 * N_FEATURES, dtype=torch.float32).reshape(9, 40, N_FEATURES))
    assert _row_pitch(x.contiguous()) == N_FEATURES
    assert _row_pitch(x[:, ::2]) == 2 * X_PITCH                   # every other row: still one pitch, twice as wide
    assert _row_pitch(x.transpose(0, 1)) == 0                     # shift-minor layouts are not row-padded tensors
    assert _row_pitch(x[:, 3:17]) == 0                            # a row slice breaks the shift stride
    sel = torch.tensor([5, 5, 0, 39, 17])
    got = select_variants(x, sel)
    assert _row_pitch(got) == X_PITCH and torch.equal(got, x.index_select(1, sel))
    dense = x.contiguous()
    assert torch.equal(select_variants(dense, sel), dense.index_select(1, sel))


def test_side_order_groups_rows_by_side_of_the_tss():
    """transform.side_order: a stable sort by (all nine shifted positions upstream, straddling, all downstream) with
    predict.py:87-109's sign convention — the '+' / '-' strings of the gene table or +-1 integers."""
    from expecto_b200.transform import side_order
    dist = np.array([5000.0, -5000.0, 0.0, 801.0, -801.0, 800.0, -800.0, 300.0, 12.0, -9000.0])
    strand = np.array(["+", "+", "-", "+", "+", "-", "-", "+", "-", "-"])
    sh = np.array(chromatin_shifts(800), dtype=np.float64)
    sgn = np.where(strand == "+", 1.0, -1.0)
    pos = (dist[:, None] + sh[None, :]) * sgn[:, None]
    key = np.where((pos < 0).all(1), 0, np.where((pos > 0).all(1), 2, 1))
    assert key.tolist() == [2, 0, 1, 2, 0, 1, 1, 1, 1, 2]           # +-800 touch position 0: straddling
    o = side_order(dist, strand)
    assert sorted(o.tolist()) == list(range(10)) and (np.diff(key[o]) >= 0).all()
    for k in (0, 1, 2):                                              # stable within a group
        assert (np.diff(o[key[o] == k]) > 0).all()
    assert np.array_equal(side_order(dist, np.where(strand == "+", 1, -1).astype(np.int8)), o)
    assert side_order(np.zeros(0), np.zeros(0, np.int8)).size == 0
