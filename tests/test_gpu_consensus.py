"""SURVEY §8f rank 4 on a B200: the consensus-sequence front end (geuvadis_predict_consensus.py) and the per-gene
prediction dumps (replicate_expecto_features.py), through the drop-in CLIs, against the reference's own outputs."""
import os

import numpy as np
import pytest

from expecto_b200 import geuvadis_predict_consensus as gpc
from expecto_b200 import hdf5
from expecto_b200 import replicate_expecto_features as rep
from expecto_b200.genome import PackedGenome
from oracle import beluga_oracle as O
from oracle import golden_inputs as GI

pytestmark = pytest.mark.gpu
PROB_TOL = 2e-3      # fp16 operands / fp32 accumulate vs the fp32 reference (same bar as test_gpu_parity.py)


def _read(path):
    with hdf5.File(path) as f:
        return {k: np.array(f[k][...], copy=True) for k in f}


def _run_consensus(engine, tmp_path, out_name, extra=()):
    cdir, genes_file, names = GI.write_consensus_dir(str(tmp_path / "in"))
    w, bias, base = GI.consensus_model()
    model = str(tmp_path / "model.npz")
    np.savez(model, weights=w, bias=bias, base_score=base)
    args = gpc.build_parser().parse_args([model, cdir, genes_file, "-o", str(tmp_path / out_name), *extra])
    assert gpc.run(args, engine=engine) == names
    return names, args


def test_consensus_cli_matches_reference_script(engine, golden, tmp_path):
    g = golden("consensus_golden.npz")
    names, args = _run_consensus(engine, tmp_path, "seg")
    w, bias, base = GI.consensus_model()
    w_abs = np.abs(O.fixed_pos_weights()).sum(axis=1)
    # a score is offset + sum_j w_j x_j with |x error| <= PROB_TOL * sum of that basis function's position weights
    score_bound = float((np.abs(w.reshape(10, 2003)[:, 1:]) * w_abs[:, None]).sum()) * PROB_TOL
    for gi, gene in enumerate(names):
        chrom = _read(str(tmp_path / "seg" / gene / f"{gene}_chromatin.h5"))
        exp = _read(str(tmp_path / "seg" / gene / f"{gene}.h5"))
        ids = [x.decode() for x in chrom["record_ids"]]
        assert sorted(ids) == list(g[f"ids_{gi}"]) and list(exp["record_ids"]) == list(chrom["record_ids"])
        order = np.argsort(ids)
        preds = chrom["chromatin_preds"]
        assert preds.dtype == np.float64 and preds.shape == (len(ids), 200, 2002)
        sub = preds[order][:, GI.CONSENSUS_KEEP_SHIFTS][:, :, ::GI.CONSENSUS_FEATURE_STEP]
        assert np.abs(sub - g[f"preds_{gi}"]).max() <= PROB_TOL
        assert exp["expecto_preds"].dtype == np.float32
        assert np.abs(exp["expecto_preds"][order] - g[f"expecto_{gi}"]).max() <= score_bound
        # the stored scores are exactly the reference's formula applied to the stored chromatin predictions
        want = O.gblinear_predict(O.consensus_features(preds), w, float(bias), float(base))
        assert np.abs(exp["expecto_preds"] - want).max() < 1e-5


def test_consensus_segment_mode_is_bit_identical_to_windows_and_exp_only_rescoring(engine, tmp_path):
    names, _ = _run_consensus(engine, tmp_path, "seg", ["--records_per_batch", "3"])
    _, wargs = _run_consensus(engine, tmp_path, "win", ["--mode", "windows"])
    first = {}
    for gene in names:
        a = _read(str(tmp_path / "seg" / gene / f"{gene}_chromatin.h5"))
        b = _read(str(tmp_path / "win" / gene / f"{gene}_chromatin.h5"))
        ia, ib = np.argsort(a["record_ids"]), np.argsort(b["record_ids"])
        assert np.array_equal(a["record_ids"][ia], b["record_ids"][ib])
        assert np.array_equal(a["chromatin_preds"][ia], b["chromatin_preds"][ib])
        first[gene] = _read(str(tmp_path / "win" / gene / f"{gene}.h5"))
        os.remove(str(tmp_path / "win" / gene / f"{gene}.h5"))
    wargs.exp_only = True
    assert gpc.run(wargs) == names                                          # no engine: scoring only
    for gene in names:
        again = _read(str(tmp_path / "win" / gene / f"{gene}.h5"))
        assert np.array_equal(again["expecto_preds"], first[gene]["expecto_preds"])
        assert np.array_equal(again["record_ids"], first[gene]["record_ids"])


def test_replicate_cli_matches_reference_script(engine, cfg1_genome, golden, tmp_path):
    g = golden("replicate_golden.npz")
    anno = tmp_path / "geneanno.csv"
    with open(anno, "w") as f:
        f.write("id,symbol,seqnames,strand,TSS,CAGE_representative_TSS,type\n")
        for gid, chrom, strand, tss in g["genes"]:
            f.write(f"{gid},SYM,{chrom},{strand},{tss},{tss},protein_coding\n")
    genome = PackedGenome.from_sequences(cfg1_genome)
    outs = {}
    for mode in ("segments", "windows"):
        args = rep.build_parser().parse_args([str(anno), "-o", str(tmp_path / mode), "--mode", mode,
                                              "--genes_per_batch", "1" if mode == "segments" else "32"])
        assert rep.run(args, engine=engine, genome=genome) == len(g["genes"])
        outs[mode] = [np.load(tmp_path / mode / f"{gid}.npy") for gid, *_ in g["genes"]]
    for i in range(len(g["genes"])):
        a = outs["segments"][i]
        assert a.dtype == np.float32 and a.shape == (200, 2002)
        assert np.array_equal(a, outs["windows"][i])
        assert np.abs(a[GI.CONSENSUS_KEEP_SHIFTS] - g[f"pred_{i}"]).max() <= PROB_TOL
    # compute_expecto_features' file is the position-weight reduction of these dumps (same engine, same windows)
    from expecto_b200.compute_expecto_features import gene_features
    from expecto_b200.transform import TSS_SHIFTS, fixed_pos_weights
    genes = [(gid, str(chrom), int(tss), 1 if strand == "+" else -1) for gid, chrom, strand, tss in g["genes"]]
    feats = gene_features(engine, genome, genes, TSS_SHIFTS, fixed_pos_weights(TSS_SHIFTS))
    for i in range(len(genes)):
        red = np.sum(O.fixed_pos_weights()[:, :, None] * outs["segments"][i][None].astype(np.float64), axis=1).flatten()
        assert np.abs(red - feats[i]).max() <= 1e-4 * np.abs(red).max()


def test_ref_all_genes_cli_matches_reference_script(engine, golden, tmp_path):
    """geuvadis_predict_ref_all_genes.py drop-in: same ref_preds.csv (genes, order, case), scores within the bound the
    2e-3 probability tolerance implies; segment mode == per-window mode bit for bit."""
    import io
    import pandas as pd
    from expecto_b200 import geuvadis_predict_ref_all_genes as rag
    want = pd.read_csv(io.StringIO(str(golden("ref_all_genes_golden.npz")["csv"])))
    cdir, genes_file, names = GI.write_consensus_dir(str(tmp_path / "in"))
    GI.write_ref_fastas(cdir)
    w, bias, base = GI.consensus_model()
    model = str(tmp_path / "model.npz")
    np.savez(model, weights=w, bias=bias, base_score=base)
    outs = {}
    for mode in ("segments", "windows"):
        args = rag.build_parser().parse_args([model, cdir, genes_file, "-o", str(tmp_path / mode), "--mode", mode,
                                              "--genes_per_batch", "1" if mode == "windows" else "64"])
        df = rag.run(args, engine=engine)
        text = open(tmp_path / mode / "ref_preds.csv").read()
        got = pd.read_csv(io.StringIO(text))
        assert list(got.columns) == ["genes", "ref_preds"] and list(got["genes"]) == list(want["genes"])
        assert df["ref_preds"].dtype == np.float32
        outs[mode] = text
        w_abs = np.abs(O.fixed_pos_weights()).sum(axis=1)
        bound = float((np.abs(w.reshape(10, 2003)[:, 1:]) * w_abs[:, None]).sum()) * PROB_TOL
        assert np.abs(got["ref_preds"].to_numpy() - want["ref_preds"].to_numpy()).max() <= bound
    assert outs["segments"] == outs["windows"]
