#!/usr/bin/env python
"""BASELINE configs[4] at full size on ONE B200: a 100,000-variant synthetic VCF on the hg19-shaped packed genome ->
chromatin effects (9 shifts x ref/alt x fwd/rc, segment + delta mode) -> exp-decay spatial transform -> scoring with a
218-model synthetic gblinear list -> sed*.tsv, (a) in one pass through ``predict --vcf`` (no .diff.h5) and (b) as the
reference's two steps: chromatin CLI writing the nine .diff.h5 files, predict CLI reading them back."""
import argparse, json, os, shutil, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", type=int, default=100_000)
    ap.add_argument("--models", type=int, default=218)
    ap.add_argument("--two_step", action="store_true", help="also run chromatin CLI -> files -> predict CLI")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    from expecto_b200 import chromatin, gblinear, predict, synth
    from expecto_b200.engine import BelugaEngine

    tmp = tempfile.mkdtemp(prefix="bx_cfg5_")
    pg = synth.make_packed_genome(seed=3)
    chroms, pos, refs, alts = synth.make_vcf_rows_packed(pg, args.variants, seed=5)
    rng = np.random.default_rng(5)
    vcf = os.path.join(tmp, "in.vcf")
    with open(vcf, "w") as f:
        f.write("##fileformat=VCFv4.1\n#CHROM\tPOS\tID\tREF\tALT\n")
        f.write("".join(f"{c}\t{p}\trs{i}\t{r}\t{a}\n" for i, (c, p, r, a) in enumerate(zip(chroms, pos, refs, alts))))
    genefile = os.path.join(tmp, "closestgene")
    dist = rng.integers(-20000, 20000, size=args.variants)
    with open(genefile, "w") as f:
        f.write("".join(f"{c}\t{p - 1}\t{p}\t{r}\t{a}\t{'+' if i % 2 else '-'}\tENSG{i % 19997:011d}\t{d}\n"
                        for i, (c, p, r, a, d) in enumerate(zip(chroms, pos, refs, alts, dist))))
    mdir = os.path.join(tmp, "models")
    os.makedirs(mdir)
    mlist = os.path.join(tmp, "modellist")
    with open(mlist, "w") as f:
        for i in range(args.models):
            p = os.path.join(mdir, f"model_{i}.save")
            gblinear.save_binary(p, gblinear.GBLinearModel((rng.standard_normal(20020) * 0.01).astype(np.float32),
                                                            float(rng.standard_normal() * 0.1), 0.5))
            f.write(p + "\n")
    feats = os.path.join(tmp, "features.tsv")
    with open(feats, "w") as f:
        f.write("\tCell type\tAssay\tTreatment\tAssay type\tSource\n")
        f.write("".join(f"{i}\tcell\tassay{i}\tNone\tTF\tsrc\n" for i in range(2002)))
    eng = BelugaEngine(0, dtype="f16", capacity=4096)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
    eng.set_genome(pg)
    res = {"gpu": torch.cuda.get_device_name(0), "variants": args.variants, "models": args.models,
           "windows": 36 * args.variants}
    common = ["--modelList", mlist, "--belugaFeatures", feats, "--geneFile", genefile]
    # warm-up of the lazily allocated buffers
    k = 2000
    wv = os.path.join(tmp, "warm.vcf")
    open(wv, "w").write("".join(open(vcf).readlines()[:k + 2]))
    wg = os.path.join(tmp, "warm.gene")
    open(wg, "w").write("".join(open(genefile).readlines()[:k]))
    predict.run(predict.build_parser().parse_args(["--modelList", mlist, "--belugaFeatures", feats, "--geneFile", wg,
                                                   "--vcf", wv, "-o", os.path.join(tmp, "warm_out")]), engine=eng, genome=pg)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    info = predict.run(predict.build_parser().parse_args(common + ["--vcf", vcf, "-o", os.path.join(tmp, "fused")]),
                       engine=eng, genome=pg)
    dt = time.perf_counter() - t0
    res["one_pass"] = {"seconds": round(dt, 2), "variants_per_s": round(args.variants / dt),
                       "timing": {k: round(v, 2) for k, v in info["timing"].items()}, "rows": info["n_rows"]}
    if args.two_step:
        os.sync()
        out = os.path.join(tmp, "chromatin_out")
        t0 = time.perf_counter()
        chromatin.run(chromatin.build_parser().parse_args([vcf, "--output_dir", out]), engine=eng, genome=pg)
        t1 = time.perf_counter()
        info2 = predict.run(predict.build_parser().parse_args(common + [
            "--coorFile_chromatin", os.path.join(out, "snps_hg19.vcf"), "--snpEffectFilePattern",
            os.path.join(out, "snps.shift_SHIFT.diff.h5"), "-o", os.path.join(tmp, "two_step")]))
        t2 = time.perf_counter()
        same = all(open(os.path.join(tmp, "fused", n), "rb").read() == open(os.path.join(tmp, "two_step", n), "rb").read()
                   for n in ("sed.tsv", "sed_sorted_by_magnitude.tsv", "sed_sorted_by_proportion.tsv"))
        res["two_step"] = {"chromatin_cli_s": round(t1 - t0, 2), "predict_cli_s": round(t2 - t1, 2),
                           "seconds": round(t2 - t0, 2), "variants_per_s": round(args.variants / (t2 - t0)),
                           "h5_bytes": 9 * 3 * 2 * args.variants * 2002 * 4,
                           "predict_timing": {k: round(v, 2) for k, v in info2["timing"].items()},
                           "tables_byte_identical_to_one_pass": bool(same)}
    res["sed_tsv_bytes"] = os.path.getsize(os.path.join(tmp, "fused", "sed.tsv"))
    shutil.rmtree(tmp, ignore_errors=True)
    print(json.dumps(res, indent=1))
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
