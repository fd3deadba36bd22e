"""Host side of the consensus-sequence front end (geuvadis_predict_consensus.py) and of the per-gene prediction dumps
(replicate_expecto_features.py), and the oracle restatements against the reference-generated fixtures."""
import os

import numpy as np
import pytest

from expecto_b200 import geuvadis_predict_consensus as gpc
from expecto_b200 import gblinear, hdf5
from expecto_b200.genome import PackedGenome
from expecto_b200.windows import tss_descs, tss_segment_descs
from oracle import beluga_oracle as O
from oracle import golden_inputs as GI

from helpers import decode_descs

SHIFTS = list(range(-20000, 20000, 200))


@pytest.fixture(scope="module")
def consensus_dir(tmp_path_factory):
    root = str(tmp_path_factory.mktemp("consensus"))
    return GI.write_consensus_dir(root)


def test_natural_order_and_chunks(tmp_path):
    names = ["a10", "a2", "7sk", "a1bg", "b", "a2m", "ensg00000000002", "ensg00000000011"]
    for n in names:
        os.makedirs(tmp_path / n)
    want = ["7sk", "a1bg", "a2", "a2m", "a10", "b", "ensg00000000002", "ensg00000000011"]
    assert gpc.select_genes(str(tmp_path), None, None) == want
    chunks = [gpc.select_genes(str(tmp_path), 3, i) for i in range(3)]
    assert chunks == [list(x) for x in np.array_split(want, 3)]          # geuvadis_predict_consensus.py:66-69
    ranks = [gpc.select_genes(str(tmp_path), None, None, r, 3) for r in range(3)]
    assert sum(ranks, []) == want
    assert sum((gpc.select_genes(str(tmp_path), 3, 0, r, 2) for r in range(2)), []) == chunks[0]
    with pytest.raises(AssertionError):
        gpc.select_genes(str(tmp_path), 20, 19)


def test_records_padded_and_named_like_the_reference(consensus_dir, golden):
    cdir, genes_file, names = consensus_dir
    g = golden("consensus_golden.npz")
    assert list(g["genes"]) == names
    strands = gpc.read_gene_strands(genes_file)
    assert [strands.loc[n, "strand"] for n in names] == ["+", "-"]
    for gi, gene in enumerate(names):
        files = sorted(os.path.join(cdir, gene, "samples", f) for f in os.listdir(os.path.join(cdir, gene, "samples")))
        pairs = list(gpc.sample_seqs_and_ids(files))
        assert sorted(p[1] for p in pairs) == list(g[f"ids_{gi}"])        # ids the reference itself wrote
        recs = {f"{rid}|{stem}": (rid, seq) for stem, rs in GI.consensus_records(gi).items() for rid, seq in rs}
        for seq, rec in pairs:
            assert seq == O.consensus_pad(*recs[rec]) and len(seq) == gpc.ENFORMER_SEQ_LENGTH


def test_bad_interval_is_refused(tmp_path):
    p = tmp_path / "x.fa"
    p.write_text(">chr1:100-200\nACGT\n")
    with pytest.raises(AssertionError):
        list(gpc.sample_seqs_and_ids([str(p)]))


@pytest.mark.parametrize("strand", ["+", "-"])
def test_window_descriptors_cut_the_reference_windows(strand):
    """The packed +-21 kb slice + tss_descs / tss_segment_descs reproduce get_seq_shifts_for_sample_seq (:210-243)."""
    rid, raw = GI.consensus_records(0)["NA003"][0]                         # missing its beginning: N-padded in front
    seq = O.consensus_pad(rid, raw)
    tss_i = len(seq) // 2
    genome = PackedGenome.from_sequences({"r0": seq[tss_i - gpc.HALF_SPAN: tss_i + gpc.HALF_SPAN + 2]})
    sgn = gpc.strand_sign(strand)
    want = O.encode_seqs(O.consensus_seq_shifts(seq, strand, SHIFTS))      # forward windows, then reverse complements
    descs = tss_descs(genome, ["r0"], [gpc.HALF_SPAN + 1], [sgn], SHIFTS).reshape(200, 2)
    some = [0, 1, 66, 99, 100, 101, 198, 199]
    got = decode_descs(genome, descs[some].reshape(-1), np.zeros(1, np.uint8)).reshape(len(some), 2, 4, 2000)
    assert np.array_equal(got[:, 0], want[:200][some]) and np.array_equal(got[:, 1], want[200:][some])
    assert want[0].sum() < 2000 or strand == "-"                           # the padding reaches the first '+' window
    seg, k_of, seg_len = tss_segment_descs(genome, ["r0"], [gpc.HALF_SPAN + 1], [sgn], SHIFTS)
    assert seg_len == 41800 and len(seg) == 2
    full = decode_descs(genome, seg[:1], np.zeros(1, np.uint8), length=seg_len)[0]
    for k in some:                                                         # window k of the forward segment
        off = 200 * k if sgn > 0 else 200 * (199 - k)
        assert np.array_equal(full[:, off:off + 2000], want[k])


def test_oracle_chromatin_matches_reference_fixture(weights_f16, golden):
    """Six windows of the front-padded record through the CPU oracle vs geuvadis_predict_consensus.py's own output."""
    g = golden("consensus_golden.npz")
    ids = list(g["ids_0"])
    rec = [i for i in ids if i.endswith("|NA003") and ":-" in i][0]
    rid, raw = GI.consensus_records(0)["NA003"][0]
    keep = [1, 4, 8]                                                        # positions in CONSENSUS_KEEP_SHIFTS
    shifts = [SHIFTS[GI.CONSENSUS_KEEP_SHIFTS[k]] for k in keep]
    ours = O.consensus_chromatin(O.consensus_pad(rid, raw), "+", weights_f16, shifts)
    want = g["preds_0"][ids.index(rec)][keep]
    assert ours.dtype == np.float64 and np.abs(ours[:, ::GI.CONSENSUS_FEATURE_STEP] - want).max() < 2e-6


def test_replicate_oracle_matches_reference_fixture(cfg1_genome, weights_f16, golden):
    g = golden("replicate_golden.npz")
    gid, chrom, strand, tss = g["genes"][1]
    keep = [0, 4, 8]
    shifts = [SHIFTS[GI.CONSENSUS_KEEP_SHIFTS[k]] for k in keep]
    ours = O.replicate_prediction(O.DictGenome(cfg1_genome), str(chrom), int(tss), 1 if strand == "+" else -1,
                                  weights_f16, shifts)
    assert ours.dtype == np.float32 and np.abs(ours - g["pred_1"][keep]).max() < 2e-6


def test_legacy_features_and_scores_vs_oracle():
    rng = np.random.default_rng(4)
    preds = rng.random((3, 200, 2002))
    feats = gpc.legacy_features(preds, np.asarray(SHIFTS))
    want = O.consensus_features(preds)
    assert feats.shape == (3, 20030) and feats.dtype == np.float64
    assert np.all(feats.reshape(3, 10, 2003)[:, :, 0] == 0)
    assert np.abs(feats - want).max() <= 1e-12 * np.abs(want).max()
    w, bias, base = GI.consensus_model()
    model = gblinear.GBLinearModel(w, float(bias), float(base))
    got = gpc.gblinear_scores(feats, model)
    assert got.dtype == np.float32
    assert np.abs(got - O.gblinear_predict(want, w, float(bias), float(base))).max() < 1e-5
    with pytest.raises(ValueError):
        gpc.gblinear_scores(feats[:, :20020], model)


def test_exp_only_rescoring_and_skip(tmp_path, consensus_dir, capsys):
    """--exp_only (:86-88): stored chromatin predictions -> the two .h5 files, no GPU involved; existing outputs are
    skipped unless --overwrite (:79-82)."""
    cdir, genes_file, names = consensus_dir
    w, bias, base = GI.consensus_model()
    model_path = str(tmp_path / "model.npz")
    np.savez(model_path, weights=w, bias=bias, base_score=base)
    out = tmp_path / "out"
    rng = np.random.default_rng(8)
    stored = {}
    for gene in names:
        os.makedirs(out / gene)
        preds = rng.random((2, 200, 2002))
        ids = np.array([f"chr1:1-393216|{gene}_a", "chr1:-5-393210|b"], "S")
        hdf5.write_datasets(str(out / gene / f"{gene}_chromatin.h5"), {"chromatin_preds": preds, "record_ids": ids})
        stored[gene] = (preds, ids)
    args = gpc.build_parser().parse_args([model_path + " ", cdir, genes_file, "-o", str(out), "--exp_only"])
    assert gpc.run(args) == names
    for gene in names:
        preds, ids = stored[gene]
        with hdf5.File(str(out / gene / f"{gene}.h5")) as f:
            got, got_ids = np.asarray(f["expecto_preds"]), np.asarray(f["record_ids"])
        assert got.dtype == np.float32 and got_ids.dtype == ids.dtype and list(got_ids) == list(ids)
        want = O.gblinear_predict(O.consensus_features(preds), w, float(bias), float(base))
        assert np.abs(got - want).max() < 1e-5
        with hdf5.File(str(out / gene / f"{gene}_chromatin.h5")) as f:
            assert np.array_equal(np.asarray(f["chromatin_preds"]), preds)
    capsys.readouterr()
    assert gpc.run(args) == []
    assert capsys.readouterr().out.count("Skipping gene") == len(names)
    args.overwrite = True
    assert gpc.run(args) == names


def test_replicate_reads_seven_columns(tmp_path):
    from expecto_b200 import replicate_expecto_features as ref
    p = tmp_path / "anno.csv"
    p.write_text("id,symbol,seqnames,strand,TSS,CAGE_representative_TSS,type\nG1,S,chr1,+,10,12,pc\nG2,S,chrX,-,20,25,pc\n")
    assert ref.read_genes(str(p)) == [("G1", "chr1", 12, 1), ("G2", "chrX", 25, -1)]
    p.write_text("id,symbol\nG1,S\n")
    with pytest.raises(ValueError):
        ref.read_genes(str(p))


# ---- geuvadis_predict_ref_all_genes.py -----------------------------------------------------------------------------
def _golden_ref_csv(golden):
    import io
    import pandas as pd
    return pd.read_csv(io.StringIO(str(golden("ref_all_genes_golden.npz")["csv"])))


def test_ref_all_genes_host_side(tmp_path, consensus_dir):
    from expecto_b200 import geuvadis_predict_ref_all_genes as rag
    cdir, genes_file, names = consensus_dir
    GI.write_ref_fastas(cdir)
    df = rag.read_genes(genes_file)
    # the symbol falls back to the ens_id and keeps its case (geuvadis_predict_ref_all_genes.py:57-59)
    assert list(df.index) == ["Abc1", "ENSG00000000002"] and list(df["strand"]) == ["+", "-"]
    assert [g.lower() for g in df.index] == names
    for gi, gene in enumerate(names):
        seq = rag.read_ref_sequence(f"{cdir}/{gene}/ref.fa")
        assert seq == O.consensus_pad(*GI.consensus_ref_record(gi)) and len(seq) == O.ENFORMER_SEQ_LENGTH
        assert seq == seq.upper()
    two = tmp_path / "two.fa"
    two.write_text(">chr1:1-393216\nACGT\n>chr1:1-393216\nACGT\n")
    with pytest.raises(AssertionError):
        rag.read_ref_sequence(str(two))                                 # :117-118 exactly one record


def test_ref_all_genes_oracle_matches_reference_csv(golden):
    """One gene (400 windows through the CPU Beluga): the restatement reproduces the reference script's score."""
    from expecto_b200 import synth
    want = _golden_ref_csv(golden)
    assert list(want.columns) == ["genes", "ref_preds"] and list(want["genes"]) == ["Abc1", "ENSG00000000002"]
    sd = synth.make_beluga_weights(GI.WEIGHT_SEED, round_dtype="f16")
    w, bias, base = GI.consensus_model()
    got = O.ref_all_genes_scores([GI.consensus_ref_record(0)], [GI.CONSENSUS_GENES[0][4]], sd, w, float(bias), float(base))
    assert got.shape == () and abs(float(got) - float(want["ref_preds"][0])) < 1e-5
