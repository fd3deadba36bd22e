#!/usr/bin/env python
"""Pin the oracle against the reference itself and write the golden fixtures under tests/golden/.

Runs ONLY in the build container (needs /root/reference; the GPU box does not have it).  Everything the
reference computes here is computed by the reference's own, unmodified code:

* ``expecto_utils.encodeSeqs`` and ``Beluga.Beluga`` are imported from /root/reference;
* ``fetchSeqs`` is lifted out of chromatin.py with ``ast`` and exec'd (chromatin.py is a module-level script
  and cannot be imported) — executed, not copied: no reference source enters this repository;
* ``chromatin.py``, ``compute_expecto_features.py`` and ``predict.py`` run as subprocesses in a scratch
  directory holding synthetic ``resources/hg19.fa`` + ``resources/deepsea.beluga.pth``, behind the import shims in
  oracle/shims (pyfasta, h5py, liftover, xgboost, matplotlib, natsort, Bio, seaborn — third-party packages absent
  from this image); ``geuvadis_predict_consensus.py``, ``geuvadis_predict_ref_all_genes.py`` and
  ``replicate_expecto_features.py`` run the same way.

Every stage asserts that oracle/beluga_oracle.py reproduces the reference output, then stores a small slice
as a fixture.  Usage:  python oracle/make_golden.py [--workdir /tmp/bx_golden] [--skip-scripts]
"""
from __future__ import annotations

import argparse
import ast
import os
import shutil
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)

from expecto_b200 import synth  # noqa: E402
from oracle import beluga_oracle as O  # noqa: E402
from oracle import golden_inputs as GI  # noqa: E402


def load_reference_fetchseqs(genome):
    """exec the reference's own fetchSeqs (chromatin.py:175-209) against ``genome``."""
    src = open(os.path.join(REF, "chromatin.py")).read()
    tree = ast.parse(src)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "fetchSeqs"][0]
    ns = {"genome": genome}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "chromatin.py", "exec"), ns)
    return ns["fetchSeqs"]


def stage_encode():
    sys.path.insert(0, REF)
    from expecto_utils import encodeSeqs as ref_encode  # the reference's canonical encoder

    genome = synth.make_genome(GI.SMALL_CONTIGS, seed=GI.SMALL_GENOME_SEED)
    g = O.DictGenome(genome)
    ref_fetch = load_reference_fetchseqs(g)
    rows = GI.edge_variants(genome)
    packed, meta = [], []
    for shift in GI.ENCODE_SHIFTS:
        refseqs, altseqs = [], []
        for chrom, pos, _id, ref, alt in rows:
            a = ref_fetch(chrom, pos, ref, alt, shift=shift)
            b = O.fetch_seqs(g, chrom, pos, ref, alt, shift=shift)
            assert a == b, ("fetchSeqs restatement differs", chrom, pos, ref, alt, shift)
            refseqs.append(a[0]), altseqs.append(a[1])
        for name, seqs in (("ref", refseqs), ("alt", altseqs)):
            r = ref_encode(seqs)
            o = O.encode_seqs(seqs)
            assert r.dtype == np.bool_ and np.array_equal(r, o), "encodeSeqs restatement differs"
            packed.append(np.packbits(r.reshape(r.shape[0], -1), axis=1))
            meta.append((shift, name))
    # shifts at and beyond +-1049, where the variant leaves the 2100-bp fetch and fetchSeqs' splice index goes negative
    # or past the end (Python slice semantics): the restatement must follow the reference there too (no fixture;
    # tests/test_windows_cpu.py checks the descriptors against the oracle at the same shifts)
    for shift in GI.EXTREME_SHIFTS:
        for chrom, pos, _id, ref, alt in rows:
            a = ref_fetch(chrom, pos, ref, alt, shift=shift)
            assert a == O.fetch_seqs(g, chrom, pos, ref, alt, shift=shift), ("fetchSeqs restatement differs", shift)
            assert np.array_equal(ref_encode([a[0], a[1]]), O.encode_seqs([a[0], a[1]]))
    # random variants of every kind at random shifts (the generator of tools/fuzz_segments.py and of
    # tests/test_windows_cpu.py::test_random_variants_descriptors_match_oracle): oracle == the reference's own functions
    import importlib.util
    spec = importlib.util.spec_from_file_location("fuzz_segments", os.path.join(ROOT, "tools", "fuzz_segments.py"))
    fuzz = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(fuzz)
    n_random = 0
    for seed in (0, 1, 2):
        rng = np.random.default_rng(100 + seed)
        rrows = fuzz.random_variants(rng, genome, 40, margin=3200)
        for shift in [0] + [int(x) for x in rng.choice(np.arange(-2200, 2201, 50), size=4, replace=False)]:
            for chrom, pos, _id, ref, alt in rrows:
                a = ref_fetch(chrom, pos, ref, alt, shift=shift)
                assert a == O.fetch_seqs(g, chrom, pos, ref, alt, shift=shift), ("fetchSeqs restatement differs", ref, alt, shift)
                assert np.array_equal(ref_encode([a[0], a[1]]), O.encode_seqs([a[0], a[1]]))
                n_random += 1
    print(f"[encode] {n_random} random variant-shifts: oracle == reference fetchSeqs / encodeSeqs")
    # odd-length / short strings straight into the encoder (crop arithmetic, chromatin.py:164)
    rng = np.random.default_rng(5)
    odd = [synth.random_sequence(rng, n) for n in (2000, 2001, 2099, 2100, 2101, 2103, 2110, 1990, 1999)]
    assert np.array_equal(ref_encode(odd), O.encode_seqs(odd))
    np.savez_compressed(os.path.join(GOLD, "encode_golden.npz"),
                        rows=np.array([[str(x) for x in r] for r in rows]),
                        shifts=np.array([m[0] for m in meta]), which=np.array([m[1] for m in meta]),
                        packed=np.stack(packed), odd=np.array(odd),
                        odd_packed=np.packbits(ref_encode(odd).reshape(2 * len(odd), -1), axis=1))
    print(f"[encode] {len(rows)} variants x {len(GI.ENCODE_SHIFTS)} shifts: oracle == reference, fixture written")


def stage_beluga(sd):
    import torch
    sys.path.insert(0, REF)
    from Beluga import Beluga

    model = Beluga()
    model.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    model.eval()
    x = synth.random_onehot(GI.BELUGA_N, seed=GI.BELUGA_SEED)
    x[0, :, 100:110] = 0
    with torch.no_grad():
        ref = model.forward(torch.from_numpy(x.astype(np.float32)).unsqueeze(2)).numpy()
    ours = O.beluga_forward(x.astype(np.float32), sd)
    err = np.abs(ref - ours).max()
    assert err < 2e-6, f"Beluga restatement differs by {err}"
    np.savez_compressed(os.path.join(GOLD, "beluga_golden.npz"), probs=ref.astype(np.float32))
    print(f"[beluga] {GI.BELUGA_N} windows: |oracle - reference| max {err:.2e}; p mean {ref.mean():.3f}, "
          f"frac>0.5 {np.mean(ref > 0.5):.3f}")


def run_ref_script(script, args, cwd):
    env = dict(os.environ)
    env["PYTHONPATH"] = os.path.join(ROOT, "oracle", "shims")
    env["BX_REPO_ROOT"] = ROOT
    t0 = time.time()
    r = subprocess.run([sys.executable, os.path.join(REF, script)] + args, cwd=cwd, env=env, capture_output=True,
                       text=True)
    if r.returncode != 0:
        raise RuntimeError(f"{script} failed:\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}")
    print(f"[{script}] ok in {time.time() - t0:.1f}s")
    return r.stdout


def prepare_workdir(work, sd):
    import torch
    os.makedirs(os.path.join(work, "resources"), exist_ok=True)
    genome = GI.cfg1_genome()
    fa = os.path.join(work, "resources", "hg19.fa")
    if not os.path.exists(fa):
        synth.write_fasta(fa, genome)
    pth = os.path.join(work, "resources", "deepsea.beluga.pth")
    if not os.path.exists(pth):
        torch.save({k: torch.from_numpy(v) for k, v in sd.items()}, pth)
    return genome


def stage_chromatin(work, sd, genome):
    shutil.copy(os.path.join(REF, "example", "example.vcf"), os.path.join(work, "example.vcf"))
    out = run_ref_script("chromatin.py", ["example.vcf", "--output_dir", "out_example"], work)
    counters = [int(l.rsplit(":", 1)[1]) for l in out.splitlines() if l.startswith("Number of")]
    g = O.DictGenome(genome)
    rows = O.normalise_vcf_rows([l.rstrip("\n").split("\t") for l in open(os.path.join(work, "example.vcf"))])
    shifts = O.SHIFTS_CHROMATIN(800)
    sums = np.zeros((len(shifts), 3, 20), dtype=np.float64)
    keep = {}
    for si, s in enumerate(shifts):
        z = np.load(os.path.join(work, "out_example", f"snps.shift_{s}.diff.h5.npz"))
        assert np.array_equal(z["diff"], z["alt"] - z["ref"])
        for di, k in enumerate(("diff", "ref", "alt")):
            sums[si, di] = z[k].astype(np.float64).sum(axis=1)
        if s in GI.CFG1_KEEP_SHIFTS:
            d, r, a, cnt = O.chromatin_shift(g, rows, sd, s)
            err = max(np.abs(r - z["ref"]).max(), np.abs(a - z["alt"]).max())
            assert err < 2e-6, f"chromatin restatement differs at shift {s}: {err}"
            if s == 0:
                assert list(cnt) == counters, (cnt, counters)
            keep[f"ref_{s}"], keep[f"alt_{s}"] = z["ref"], z["alt"]
    vcf_text = open(os.path.join(work, "out_example", "snps_hg19.vcf")).read()
    np.savez_compressed(os.path.join(GOLD, "chromatin_cfg1_golden.npz"), row_sums=sums, counters=np.array(counters),
                        snps_hg19_vcf=np.array(vcf_text), **keep)
    print(f"[chromatin cfg1] counters {counters}; oracle == reference on shifts {GI.CFG1_KEEP_SHIFTS}")

    # edge-case VCF (indels, ref mismatch, lower-case alleles, un-prefixed and non-CHRS contigs), 3 shifts
    rows_e = GI.cfg1_edge_variants(genome)
    synth.write_vcf(os.path.join(work, "edge.vcf"), rows_e, header=True)
    out = run_ref_script("chromatin.py", ["edge.vcf", "--output_dir", "out_edge", "--maxshift", "200"], work)
    counters_e = [int(l.rsplit(":", 1)[1]) for l in out.splitlines() if l.startswith("Number of")]
    rows_n = O.normalise_vcf_rows(rows_e)
    keep = {}
    for s in O.SHIFTS_CHROMATIN(200):
        z = np.load(os.path.join(work, "out_edge", f"snps.shift_{s}.diff.h5.npz"))
        d, r, a, cnt = O.chromatin_shift(g, rows_n, sd, s)
        err = max(np.abs(r - z["ref"]).max(), np.abs(a - z["alt"]).max())
        assert err < 2e-6 and z["ref"].shape[0] == 2 * len(rows_n), (s, err)
        if s == 0:
            assert list(cnt) == counters_e
        keep[f"ref_{s}"], keep[f"alt_{s}"] = z["ref"], z["alt"]
    np.savez_compressed(os.path.join(GOLD, "chromatin_edge_golden.npz"),
                        rows=np.array([[str(x) for x in r] for r in rows_e]), counters=np.array(counters_e),
                        snps_hg19_vcf=np.array(open(os.path.join(work, "out_edge", "snps_hg19.vcf")).read()), **keep)
    print(f"[chromatin edge] {len(rows_e)} rows -> {len(rows_n)} kept; counters {counters_e}")


def stage_features(work, sd, genome):
    genes = GI.cfg4_genes()
    with open(os.path.join(work, "geneanno_small.csv"), "w") as f:
        f.write("id,symbol,seqnames,strand,TSS,CAGE_representative_TSS,type\n")
        for gid, chrom, strand, tss in genes:
            f.write(f"{gid},SYM,{chrom},{strand},{tss},{tss},protein_coding\n")
    with open(os.path.join(work, "tss_none.tsv"), "w") as f:
        f.write("\tens_id\tchrom\ttss\tstrand\tx\tis_default_tss\n0\tENSG_NOMATCH\tchr1\t1000000\t+\t0\tTrue\n")
    run_ref_script("compute_expecto_features.py", ["geneanno_small.csv", "tss_none.tsv", "-o", "out_features"], work)
    ref = np.load(os.path.join(work, "out_features", "Xreducedall.2002.representative_tss_top.npy"))
    g = O.DictGenome(genome)
    pw = O.fixed_pos_weights()
    for i, (gid, chrom, strand, tss) in enumerate(genes):
        seqs = O.tss_windows(g, chrom, tss, 1 if strand == "+" else -1)
        enc = O.encode_seqs(seqs, concat_rc=False)
        fwd = O.beluga_forward(enc.astype(np.float32), sd)
        rc = O.beluga_forward(np.ascontiguousarray(enc[:, ::-1, ::-1]).astype(np.float32), sd)
        ours = O.fixed_transform(fwd, rc, pw)
        rel = np.abs(ours - ref[i]).max() / np.abs(ref[i]).max()
        assert rel < 1e-6, f"feature restatement differs for gene {i}: {rel}"
    np.savez_compressed(os.path.join(GOLD, "features_golden.npz"), features=ref,
                        genes=np.array([[str(x) for x in gg] for gg in genes]))
    print(f"[features] {len(genes)} genes: oracle == reference, shape {ref.shape} {ref.dtype}")


def stage_features_override(work, sd, genome):
    """compute_expecto_features.py with a TSS table that matches both genes (:54-69; liftover = identity shim): gene 1
    has a non-default TSS -> chrom, TSS and strand come from the table; gene 2 is marked default -> only its strand is
    taken from the table (the reference overwrites ``strand`` at :56 before it looks at ``is_default_tss``).  The
    output must be what the oracle computes at exactly those coordinates; they are stored for the host-logic test."""
    genes = GI.cfg4_genes()
    (g1, c1, s1, t1), (g2, c2, s2, t2) = genes
    with open(os.path.join(work, "tss_override.tsv"), "w") as f:
        f.write("\tens_id\tchrom\ttss\tstrand\tx\tis_default_tss\n")
        f.write(f"0\t{g1}\t{c1}\t{t1 + 400}\t-\t0\tFalse\n")
        f.write(f"1\t{g2}\tchr2\t123\t+\t0\tTrue\n")
    run_ref_script("compute_expecto_features.py", ["geneanno_small.csv", "tss_override.tsv", "-o", "out_features_override"], work)
    ref = np.load(os.path.join(work, "out_features_override", "Xreducedall.2002.representative_tss_top.npy"))
    expected = [(g1, c1, t1 + 400, -1), (g2, c2, t2, 1)]
    g = O.DictGenome(genome)
    pw = O.fixed_pos_weights()
    for i, (gid, chrom, tss, sgn) in enumerate(expected):
        enc = O.encode_seqs(O.tss_windows(g, chrom, tss, sgn), concat_rc=False)
        fwd = O.beluga_forward(enc.astype(np.float32), sd)
        rc = O.beluga_forward(np.ascontiguousarray(enc[:, ::-1, ::-1]).astype(np.float32), sd)
        rel = np.abs(O.fixed_transform(fwd, rc, pw) - ref[i]).max() / np.abs(ref[i]).max()
        assert rel < 1e-6, f"TSS override: the reference did not use {expected[i]} (rel {rel})"
    np.savez_compressed(os.path.join(GOLD, "features_override_golden.npz"),
                        anno=np.array(open(os.path.join(work, "geneanno_small.csv")).read()),
                        tss=np.array(open(os.path.join(work, "tss_override.tsv")).read()),
                        expected=np.array([[str(x) for x in e] for e in expected]),
                        feature_sums=ref.sum(axis=1))
    print(f"[features override] reference used {expected}: fixture written")


def stage_predict(work):
    rng = np.random.default_rng(GI.GBLINEAR_SEED)
    w = (rng.standard_normal(20020) * 0.01).astype(np.float32)
    np.savez(os.path.join(work, "model.npz"), weights=w, bias=np.float32(0.25), base_score=np.float32(0.5))
    shutil.copy(os.path.join(REF, "example", "example.vcf.bed.sorted.bed.closestgene"), work)
    run_ref_script("predict.py", [
        "--model_save_file", "model.npz", "--belugaFeatures",
        os.path.join(REF, "resources", "deepsea_beluga_2002_features.tsv"), "--coorFile_chromatin",
        "out_example/snps_hg19.vcf", "--geneFile", "example.vcf.bed.sorted.bed.closestgene",
        "--snpEffectFilePattern", "out_example/snps.shift_SHIFT.diff.h5", "-o", "out_predict"], work)
    import pandas as pd
    sed = pd.read_csv(os.path.join(work, "out_predict", "sed.tsv"), sep="\t")
    # restatement: strand average -> weights -> transform -> gblinear
    shifts = O.SHIFTS_CHROMATIN(800)
    data = {k: [] for k in ("diff", "ref", "alt")}
    for s in shifts:
        z = np.load(os.path.join(work, "out_example", f"snps.shift_{s}.diff.h5.npz"))
        for k in data:
            data[k].append(O.strand_average(z[k]))
    gene = pd.read_csv(os.path.join(work, "example.vcf.bed.sorted.bed.closestgene"), sep="\t", header=None)
    dist = -np.asarray(gene.iloc[:, -1])
    strand = np.asarray(gene.iloc[:, -3])
    wts = O.variant_weights(dist, strand)
    ref_s = O.gblinear_predict(O.variant_transform(data["ref"], wts), w, 0.25, 0.5)
    alt_s = O.gblinear_predict(O.variant_transform(data["alt"], wts), w, 0.25, 0.5)
    assert np.abs(ref_s - sed["REF"].values).max() < 1e-4 and np.abs(alt_s - sed["ALT"].values).max() < 1e-4
    feats = O.variant_transform(data["diff"], wts)
    np.savez_compressed(os.path.join(GOLD, "predict_golden.npz"),
                        sed_tsv=np.array(open(os.path.join(work, "out_predict", "sed.tsv")).read()),
                        diff_features_row0=feats[0], model_w=w)
    print(f"[predict] sed.tsv rows {len(sed)}; oracle == reference (REF/ALT within 1e-4)")


def closest_gene_inputs():
    """A 700-gene slice of the reference's own resources/geneanno.csv (chr1, chr13 incl. two genes sharing one TSS,
    chrX) and a VCF that exercises ties, exact-TSS hits, dense and gene-free neighbourhoods."""
    import pandas as pd
    g = pd.read_csv(os.path.join(REF, "resources", "geneanno.csv"), index_col=0)
    sub = pd.concat([g[g["seqnames"] == "chr1"].iloc[:400], g[g["seqnames"] == "chr13"], g[g["seqnames"] == "chrX"].iloc[:60]])
    rng = np.random.default_rng(12)
    rows = [l.rstrip("\n").split("\t") for l in open(os.path.join(REF, "example", "example.vcf"))]
    rows = [[r[0], int(r[1]), r[2], r[3], r[4]] for r in rows]
    for chrom in ("chr1", "chr13", "chrX"):
        t = np.sort(sub[sub["seqnames"] == chrom]["CAGE_representative_TSS"].values)
        for k in range(12):
            i = int(rng.integers(0, len(t) - 1))
            rows.append([chrom, int((t[i] + t[i + 1]) // 2), f"mid{k}", "A", "C"])       # midpoints: ties when the gap is even
            rows.append([chrom, int(t[i] + rng.integers(-30000, 30000)), f"rnd{k}", "G", "T"])
        rows.append([chrom, int(t[5]), "on_tss", "A", "G"])
        rows.append([chrom, int(t[-1] + 5_000_000), "far_right", "C", "T"])
        rows.append([chrom, max(1, int(t[0] - 900_000)), "far_left", "C", "A"])
    rows.append(["chr13", 57715052, "shared_tss", "A", "T"])      # PRR20A / PRR20C share this TSS: first annotation row wins
    rows.append(["chr13", 57715052 + 20300, "rf_edge_plus", "A", "T"])
    rows.append(["chr13", 57715052 - 20499, "rf_edge_minus", "A", "T"])
    return sub, rows


def stage_closest_genes(work):
    sub, rows = closest_gene_inputs()
    anno = os.path.join(work, "geneanno_sub.csv")
    sub.to_csv(anno)
    vcf = os.path.join(work, "closest_in.vcf")
    with open(vcf, "w") as f:
        for r in rows:
            f.write("\t".join(str(x) for x in r) + "\n")
    out = {}
    for mode, flag in (("closest", []), ("rf", ["--all_in_receptive_field"])):
        run_ref_script("make_closest_genes_file.py", ["closest_in.vcf", "--geneanno_file", "geneanno_sub.csv", "-o",
                                                      f"out_closest_{mode}"] + flag, work)
        out[f"{mode}_tsv"] = np.array(open(os.path.join(work, f"out_closest_{mode}", "closest_genes.tsv")).read())
        out[f"{mode}_vcf"] = np.array(open(os.path.join(work, f"out_closest_{mode}", "snps_hg19.vcf")).read())
    np.savez_compressed(os.path.join(GOLD, "closest_genes_golden.npz"), geneanno_csv=np.array(open(anno).read()),
                        vcf=np.array(open(vcf).read()), **out)
    print(f"[closest genes] {len(rows)} variants x {len(sub)} genes: reference outputs stored "
          f"({str(out['closest_tsv']).count(chr(10))} / {str(out['rf_tsv']).count(chr(10))} rows)")


def stage_consensus(work, sd):
    """geuvadis_predict_consensus.py as a subprocess on seeded consensus FASTAs (2 genes, 6 records, truncated at
    either end), then again with --exp_only; the restatements in beluga_oracle.py must reproduce both."""
    import torch
    pth = os.path.join(work, "resources", "deepsea.beluga.pth")
    if not os.path.exists(pth):
        os.makedirs(os.path.dirname(pth), exist_ok=True)
        torch.save({k: torch.from_numpy(v) for k, v in sd.items()}, pth)
    root = os.path.join(work, "consensus_in")
    shutil.rmtree(root, ignore_errors=True)
    shutil.rmtree(os.path.join(work, "out_consensus"), ignore_errors=True)
    cdir, genes_file, names = GI.write_consensus_dir(root)
    w, bias, base = GI.consensus_model()
    np.savez(os.path.join(work, "consensus_model.npz"), weights=w, bias=bias, base_score=base)
    run_ref_script("geuvadis_predict_consensus.py", ["consensus_model.npz ", cdir, genes_file, "-o", "out_consensus"], work)
    shifts = list(range(-20000, 20000, 200))
    keep = {}
    for gi, gene in enumerate(names):
        z = np.load(os.path.join(work, "out_consensus", gene, f"{gene}_chromatin.h5.npz"))
        e = np.load(os.path.join(work, "out_consensus", gene, f"{gene}.h5.npz"))
        preds, ids = z["chromatin_preds"], [x.decode() for x in z["record_ids"]]
        assert preds.dtype == np.float64 and preds.shape == (len(ids), 200, 2002)
        assert e["expecto_preds"].dtype == np.float32 and list(e["record_ids"]) == list(z["record_ids"])
        # restatement: records in the reference's own (glob) order, identified by "{record.id}|{stem}"
        recs = {f"{rid}|{stem}": (rid, seq) for stem, rs in GI.consensus_records(gi).items() for rid, seq in rs}
        assert sorted(ids) == sorted(recs), (ids, list(recs))
        strand = GI.CONSENSUS_GENES[gi][4]
        for k, rec in enumerate(ids):
            ours = O.consensus_chromatin(O.consensus_pad(*recs[rec]), strand, sd, shifts)
            err = np.abs(ours - preds[k]).max()
            assert err < 2e-6, f"consensus restatement differs for {rec}: {err}"
        feats = O.consensus_features(preds)
        score = O.gblinear_predict(feats, w, float(bias), float(base))
        assert np.abs(score - e["expecto_preds"]).max() < 1e-5
        order = np.argsort(ids)
        keep[f"ids_{gi}"] = np.array(ids)[order]
        keep[f"preds_{gi}"] = preds[order][:, GI.CONSENSUS_KEEP_SHIFTS][:, :, ::GI.CONSENSUS_FEATURE_STEP]
        keep[f"expecto_{gi}"] = e["expecto_preds"][order]
    # --exp_only re-scores the stored chromatin predictions: same numbers
    for gene in names:
        os.rename(os.path.join(work, "out_consensus", gene, f"{gene}.h5.npz"),
                  os.path.join(work, "out_consensus", gene, f"{gene}.first.npz"))
    run_ref_script("geuvadis_predict_consensus.py", ["consensus_model.npz", cdir, genes_file, "-o", "out_consensus",
                                                     "--exp_only"], work)
    for gene in names:
        a = np.load(os.path.join(work, "out_consensus", gene, f"{gene}.first.npz"))["expecto_preds"]
        b = np.load(os.path.join(work, "out_consensus", gene, f"{gene}.h5.npz"))["expecto_preds"]
        assert np.array_equal(a, b)
    np.savez_compressed(os.path.join(GOLD, "consensus_golden.npz"), genes=np.array(names), **keep)
    print(f"[consensus] {len(names)} genes, {sum(len(keep[f'ids_{i}']) for i in range(len(names)))} records: "
          f"oracle == reference (chromatin within 2e-6, scores within 1e-5); fixture written")


def stage_ref_all_genes(work, sd):
    """geuvadis_predict_ref_all_genes.py as a subprocess on one ref.fa per gene (stage_consensus's directory)."""
    import io
    import pandas as pd
    root = os.path.join(work, "consensus_in")
    cdir, genes_file = os.path.join(root, "consensus"), os.path.join(root, "genes.csv")
    if not os.path.exists(genes_file):
        cdir, genes_file, _ = GI.write_consensus_dir(root)
    GI.write_ref_fastas(cdir)
    w, bias, base = GI.consensus_model()
    np.savez(os.path.join(work, "consensus_model.npz"), weights=w, bias=bias, base_score=base)
    shutil.rmtree(os.path.join(work, "out_ref_all"), ignore_errors=True)
    run_ref_script("geuvadis_predict_ref_all_genes.py", ["consensus_model.npz", cdir, genes_file, "-o", "out_ref_all"], work)
    text = open(os.path.join(work, "out_ref_all", "ref_preds.csv")).read()
    df = pd.read_csv(io.StringIO(text))
    assert list(df.columns) == ["genes", "ref_preds"]
    assert list(df["genes"]) == [row[3] or row[0] for row in GI.CONSENSUS_GENES]      # original case, ens_id filled in
    ours = O.ref_all_genes_scores([GI.consensus_ref_record(i) for i in range(len(GI.CONSENSUS_GENES))],
                                  [row[4] for row in GI.CONSENSUS_GENES], sd, w, float(bias), float(base))
    err = np.abs(ours - df["ref_preds"].to_numpy()).max()
    assert err < 1e-5, f"ref_all_genes restatement differs: {err}"
    np.savez_compressed(os.path.join(GOLD, "ref_all_genes_golden.npz"), csv=np.array(text))
    print(f"[ref all genes] {len(df)} genes: oracle == reference (scores within 1e-5); fixture written")


def stage_num_repeats():
    """predict.py's get_num_repeats (:202-213), lifted out of the script with ``ast`` and exec'd (predict.py is
    module-level code), on seeded gene tables: runs of rows sharing columns 0-4, incl. a group that reappears later
    (counted again, the reference compares neighbours only) and numeric columns that pandas parses as floats."""
    import io
    import pandas as pd
    src = open(os.path.join(REF, "predict.py")).read()
    fn = [n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name == "get_num_repeats"][0]
    ns = {}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "predict.py", "exec"), ns)
    rng = np.random.default_rng(17)
    tables, expected = [], []
    for t in range(6):
        rows = []
        n_groups = int(rng.integers(1, 12))
        keys = [(f"chr{int(rng.integers(1, 4))}", int(rng.integers(1, 50)), "ACGT"[int(rng.integers(4))],
                 "ACGT"[int(rng.integers(4))]) for _ in range(n_groups)]
        if t == 3 and n_groups > 2:
            keys.append(keys[0])                       # the same variant again, not adjacent
        for c, p, r, a in keys:
            for k in range(int(rng.integers(1, 5))):
                pos = float(p) if t == 4 else p        # a float position column ("12.0" in the joined key)
                rows.append([c, pos, pos + 1, r, a, "+-"[int(rng.integers(2))], f"ENSG{int(rng.integers(1000))}",
                             int(rng.integers(-5000, 5000))])
        text = pd.DataFrame(rows).to_csv(sep="\t", header=False, index=False)
        df = pd.read_csv(io.StringIO(text), sep="\t", header=None, comment="#")
        tables.append(text)
        expected.append(",".join(str(x) for x in ns["get_num_repeats"](df)))
    np.savez_compressed(os.path.join(GOLD, "num_repeats_golden.npz"), tables=np.array(tables, dtype=str),
                        expected=np.array(expected, dtype=str))
    print(f"[num repeats] {len(tables)} gene tables: the reference's get_num_repeats outputs stored")


def stage_keep_mask():
    """cluster_utils.get_keep_mask (imported from /root/reference) on the reference's own 2002-feature table, for the
    16 combinations of --no_tf_features / --no_dnase_features / --no_histone_features / --no_pol2 (predict.py:139-147
    uses the mask to subset the model's columns).  The fixture keeps the two columns the masks depend on."""
    import contextlib
    import io
    import itertools
    import pandas as pd
    sys.path.insert(0, REF)
    import cluster_utils
    df = pd.read_csv(os.path.join(REF, "resources", "deepsea_beluga_2002_features.tsv"), sep="\t", index_col=0)
    combos, masks = [], []
    for tf, dn, hi, po in itertools.product([False, True], repeat=4):
        with contextlib.redirect_stdout(io.StringIO()):
            m = cluster_utils.get_keep_mask(df, tf, dn, hi, False, po, return_hgnc_df=True)[0]
        combos.append([tf, dn, hi, po])
        masks.append(np.packbits(np.asarray(m, dtype=bool)))
    np.savez_compressed(os.path.join(GOLD, "keep_mask_golden.npz"), assay=np.array([str(x) for x in df["Assay"]], dtype=str),
                        assay_type=np.array([str(x) for x in df["Assay type"]], dtype=str), combos=np.array(combos),
                        masks=np.stack(masks))
    print(f"[keep mask] {len(combos)} flag combinations on {len(df)} features: reference masks stored")


def lambert_tables(assays, seed=9):
    """Seeded stand-ins for resources/Lambert-hgnc-symbol-check.csv and resources/beluga_hgnc_mapping.csv (neither is
    in the reference repository): a mapping table whose index holds a third of the Beluga assay names — some twice with
    different ``Match type`` values, some without an approved symbol, some mapped to a differently-cased or different
    symbol — and a Lambert table listing about half of the resulting symbols plus unrelated ones."""
    import pandas as pd
    rng = np.random.default_rng(seed)
    uniq = sorted(set(assays))
    picked = [a for a in uniq if rng.random() < 0.35]
    rows = []
    for a in picked:
        u = rng.random()
        sym = a.upper() if u < 0.5 else (a.lower() + "x" if u < 0.7 else "SYM" + str(int(rng.integers(1000))))
        if u > 0.9:
            rows.append((a, "Unmatched", np.nan))
        elif rng.random() < 0.25:   # two rows: the "Approved symbol" one wins, wherever it stands
            other = (a, "Previous symbol", "OLD" + str(int(rng.integers(1000))))
            good = (a, "Approved symbol", sym)
            rows += [other, good] if rng.random() < 0.5 else [good, other]
        else:
            rows.append((a, "Approved symbol" if rng.random() < 0.8 else "Alias symbol", sym))
    mapping = pd.DataFrame(rows, columns=["Input", "Match type", "Approved symbol"]).set_index("Input")
    symbols = sorted({str(s).upper() for s in mapping["Approved symbol"].dropna()} | {a for a in uniq if a not in picked})
    keep = [s for s in symbols if rng.random() < 0.5] + ["ZNF9999", "NOTABELUGAASSAY"]
    lambert = pd.DataFrame({"Input": keep, "Approved symbol": keep})
    return lambert, mapping


def stage_keep_mask_lambert(work):
    """cluster_utils.get_keep_mask with intersect_with_lambert=True (cluster_utils.py:26-42), run from a scratch cwd
    holding the two seeded resource tables, alone and combined with --no_dnase_features / --no_pol2."""
    import contextlib
    import io
    import pandas as pd
    sys.path.insert(0, REF)
    import cluster_utils
    df = pd.read_csv(os.path.join(REF, "resources", "deepsea_beluga_2002_features.tsv"), sep="\t", index_col=0)
    lambert, mapping = lambert_tables([str(x) for x in df["Assay"]])
    os.makedirs(os.path.join(work, "resources"), exist_ok=True)
    lp, mp = (os.path.join(work, "resources", n) for n in ("Lambert-hgnc-symbol-check.csv", "beluga_hgnc_mapping.csv"))
    lambert.to_csv(lp)
    mapping.to_csv(mp)
    combos, masks = [], []
    cwd = os.getcwd()
    os.chdir(work)
    try:
        for dn, po in ((False, False), (True, False), (False, True), (True, True)):
            with contextlib.redirect_stdout(io.StringIO()):
                m = cluster_utils.get_keep_mask(df, False, dn, False, True, po, return_hgnc_df=True)[0]
            combos.append([dn, po])
            masks.append(np.packbits(np.asarray(m, dtype=bool)))
    finally:
        os.chdir(cwd)
    np.savez_compressed(os.path.join(GOLD, "keep_mask_lambert_golden.npz"),
                        assay=np.array([str(x) for x in df["Assay"]], dtype=str),
                        assay_type=np.array([str(x) for x in df["Assay type"]], dtype=str),
                        lambert_csv=np.array(open(lp).read()), mapping_csv=np.array(open(mp).read()),
                        combos=np.array(combos), masks=np.stack(masks))
    kept = int(np.unpackbits(masks[0])[:len(df)].sum())
    print(f"[keep mask / lambert] {len(mapping)} mapping rows, {len(lambert)} Lambert symbols: {kept} of {len(df)} "
          "features kept; reference masks stored")


def stage_replicate(work, sd, genome):
    """replicate_expecto_features.py as a subprocess on the two compute_expecto_features genes."""
    genes = GI.cfg4_genes()
    anno = os.path.join(work, "geneanno_small.csv")
    if not os.path.exists(anno):
        with open(anno, "w") as f:
            f.write("id,symbol,seqnames,strand,TSS,CAGE_representative_TSS,type\n")
            for gid, chrom, strand, tss in genes:
                f.write(f"{gid},SYM,{chrom},{strand},{tss},{tss},protein_coding\n")
    run_ref_script("replicate_expecto_features.py", ["geneanno_small.csv", "-o", "out_replicate"], work)
    g = O.DictGenome(genome)
    keep = {}
    for i, (gid, chrom, strand, tss) in enumerate(genes):
        ref = np.load(os.path.join(work, "out_replicate", f"{gid}.npy"))
        assert ref.dtype == np.float32 and ref.shape == (200, 2002)
        ours = O.replicate_prediction(g, chrom, tss, 1 if strand == "+" else -1, sd)
        err = np.abs(ours - ref).max()
        assert err < 2e-6, f"replicate restatement differs for {gid}: {err}"
        keep[f"pred_{i}"] = ref[GI.CONSENSUS_KEEP_SHIFTS]
    # the features file of compute_expecto_features is the pos_weights reduction of these dumps
    feats = np.load(os.path.join(GOLD, "features_golden.npz"))["features"]
    full = np.load(os.path.join(work, "out_replicate", f"{genes[0][0]}.npy"))
    red = np.sum(O.fixed_pos_weights()[:, :, None] * full[None, :, :], axis=1).flatten()
    assert np.abs(red - feats[0]).max() / np.abs(feats[0]).max() < 1e-6
    np.savez_compressed(os.path.join(GOLD, "replicate_golden.npz"),
                        genes=np.array([[str(x) for x in gg] for gg in genes]), **keep)
    print(f"[replicate] {len(genes)} genes: oracle == reference, dumps reduce to the features fixture")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workdir", default="/tmp/bx_golden")
    ap.add_argument("--skip-scripts", action="store_true", help="only the import-level stages (fast)")
    ap.add_argument("--only-closest", action="store_true", help="only the make_closest_genes_file.py stage")
    ap.add_argument("--only-keep-mask", action="store_true", help="only the cluster_utils.get_keep_mask stages")
    ap.add_argument("--only-features-override", action="store_true",
                    help="only the compute_expecto_features.py TSS-override stage")
    ap.add_argument("--only-ref-all-genes", action="store_true", help="only the geuvadis_predict_ref_all_genes.py stage")
    ap.add_argument("--only-consensus", action="store_true",
                    help="only the geuvadis_predict_consensus.py / replicate_expecto_features.py stages")
    args = ap.parse_args()
    os.makedirs(GOLD, exist_ok=True)
    if args.only_closest:
        os.makedirs(args.workdir, exist_ok=True)
        stage_closest_genes(args.workdir)
        return
    if args.only_keep_mask:
        os.makedirs(args.workdir, exist_ok=True)
        stage_keep_mask()
        stage_keep_mask_lambert(args.workdir)
        return
    if args.only_features_override:
        sd = synth.make_beluga_weights(GI.WEIGHT_SEED, round_dtype="f16")
        genome = prepare_workdir(args.workdir, sd)
        if not os.path.exists(os.path.join(args.workdir, "geneanno_small.csv")):
            stage_features(args.workdir, sd, genome)
        stage_features_override(args.workdir, sd, genome)
        return
    if args.only_ref_all_genes:
        sd = synth.make_beluga_weights(GI.WEIGHT_SEED, round_dtype="f16")
        prepare_workdir(args.workdir, sd)
        stage_ref_all_genes(args.workdir, sd)
        return
    if args.only_consensus:
        sd = synth.make_beluga_weights(GI.WEIGHT_SEED, round_dtype="f16")
        genome = prepare_workdir(args.workdir, sd)
        stage_consensus(args.workdir, sd)
        stage_ref_all_genes(args.workdir, sd)
        stage_replicate(args.workdir, sd, genome)
        return
    stage_encode()
    stage_keep_mask()
    os.makedirs(args.workdir, exist_ok=True)
    stage_keep_mask_lambert(args.workdir)
    stage_num_repeats()
    sd = synth.make_beluga_weights(GI.WEIGHT_SEED, round_dtype="f16")
    stage_beluga(sd)
    if not args.skip_scripts:
        genome = prepare_workdir(args.workdir, sd)
        stage_chromatin(args.workdir, sd, genome)
        stage_features(args.workdir, sd, genome)
        stage_features_override(args.workdir, sd, genome)
        stage_predict(args.workdir)
        stage_closest_genes(args.workdir)
        stage_consensus(args.workdir, sd)
        stage_ref_all_genes(args.workdir, sd)
        stage_replicate(args.workdir, sd, genome)
    print("golden fixtures written to", GOLD)


if __name__ == "__main__":
    main()
