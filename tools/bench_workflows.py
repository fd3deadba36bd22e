#!/usr/bin/env python
"""End-to-end timings of the drop-in workflows (BASELINE configs 3-5 at reduced, stated sizes) on one B200.

    python tools/bench_workflows.py [--variants 8192] [--genes 128] [--out gpurun_out/workflows.json]

* chromatin (cfg3 shape): synthetic VCF (90 % SNV / 5 % ref-mismatch / 5 % indel) on a synthetic genome,
  9 shifts x ref/alt x fwd/rc = 36 windows per variant; timed (a) host descriptors -> pinned host ref/alt/diff
  (no file) and (b) the whole CLI incl. HDF5 files.
* features (cfg4 shape): genes x 200 shifts x fwd/rc + fixed-weight transform -> .npy.
* predict (cfg5 shape): per-shift h5 -> fused transform + scoring over a synthetic 218-model list.
Prints one JSON object; the CPU reference (oracle port of the same path) is timed on a small sample."""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", type=int, default=8192)
    ap.add_argument("--genes", type=int, default=128)
    ap.add_argument("--models", type=int, default=218)
    ap.add_argument("--cpu_variants", type=int, default=2)
    ap.add_argument("--out", default=None)
    ap.add_argument("--tmp", default=None)
    args = ap.parse_args()

    import torch

    from expecto_b200 import chromatin, compute_expecto_features as cef, gblinear, predict, synth
    from expecto_b200.engine import BelugaEngine
    from expecto_b200.genome import PackedGenome
    from expecto_b200.windows import chromatin_window_descs, check_valid_bases

    res = {"gpu": torch.cuda.get_device_name(0), "host_cores": os.cpu_count()}
    tmp = args.tmp or tempfile.mkdtemp(prefix="bx_workflows_")
    t0 = time.perf_counter()
    contigs = [("chr1", 24_000_000), ("chr2", 20_000_000), ("chr7", 12_000_000), ("chrX", 8_000_000)]
    genome = synth.make_genome(contigs, seed=3)
    pg = PackedGenome.from_sequences(genome)
    res["setup_genome_s"] = time.perf_counter() - t0
    sd = synth.make_beluga_weights(0, round_dtype="f16")
    eng = BelugaEngine(0, dtype="f16", capacity=4096)
    eng.load_state_dict(sd)
    eng.set_genome(pg)

    # ---------------------------------------------------------------- chromatin (cfg3 shape)
    n = args.variants
    rows = synth.make_vcf_rows(genome, n, seed=3, odd_names=False)
    vcf = os.path.join(tmp, "synthetic.vcf")
    synth.write_vcf(vcf, rows)
    kept = [r for r in rows if r[0].replace("chr", "") in [str(i) for i in range(1, 23)] + ["X", "Y"]]
    chroms = ["chr" + str(r[0]).replace("chr", "") for r in kept]
    pos = np.array([r[1] for r in kept])
    refs, alts = [r[3] for r in kept], [r[4] for r in kept]
    nv = len(kept)
    pinned = tuple(torch.empty((2 * nv, 2002), dtype=torch.float32, pin_memory=True) for _ in range(3))
    shifts = chromatin.shift_list(800)
    for warm in (True, False):
        t_desc = t_gpu = 0.0
        t0 = time.perf_counter()
        for s in (shifts[:1] if warm else shifts):
            ta = time.perf_counter()
            descs, codes = chromatin_window_descs(pg, chroms, pos, refs, alts, s)
            check_valid_bases(pg, descs)
            tb = time.perf_counter()
            eng.chromatin_shift_host(descs, codes, nv, out=pinned)
            t_desc += tb - ta
            t_gpu += time.perf_counter() - tb
        total = time.perf_counter() - t0
    res["chromatin_no_file"] = {"variants": nv, "windows": nv * 36, "seconds": total, "host_descriptor_s": t_desc,
                                "gpu_pipeline_s": t_gpu, "windows_per_s": nv * 36 / total,
                                "variants_per_s": nv / total, "d2h_bytes": 9 * 3 * 2 * nv * 2002 * 4}
    # segment mode (shared activations across the 9 shifts), no file: device results copied to pinned host memory
    from expecto_b200.chromatin import predict_all_shifts_segments
    sink = lambda name, row0, arr: None
    predict_all_shifts_segments(eng, pg, chroms[:600], pos[:600], refs[:600], alts[:600], 800, {s: sink for s in shifts})
    t0 = time.perf_counter()
    sinfo = predict_all_shifts_segments(eng, pg, chroms, pos, refs, alts, 800, {s: sink for s in shifts})
    total = time.perf_counter() - t0
    res["chromatin_segments_no_file"] = {"variants": nv, "windows": nv * 36, "seconds": total,
                                         "windows_per_s": nv * 36 / total, "variants_per_s": nv / total, **sinfo}
    for mode in ("windows", "segments"):
        # fresh directory per mode and a sync before the clock starts: truncating the previous run's 1.6 GB files
        # and the write-back of its dirty pages would otherwise be charged to this run
        out_dir = os.path.join(tmp, f"chromatin_out_{mode}")
        os.sync()
        a = chromatin.build_parser().parse_args([vcf, "--output_dir", out_dir, "--mode", mode])
        t0 = time.perf_counter()
        info = chromatin.run(a, engine=eng, genome=pg)
        total = time.perf_counter() - t0
        if mode == "windows":
            import shutil
            shutil.rmtree(out_dir, ignore_errors=True)
        res[f"chromatin_cli_with_files_{mode}"] = {"variants": info["n_variants"], "seconds": total,
                                                   "windows_per_s": info["n_variants"] * 36 / total,
                                                   "bytes_written": 9 * 3 * 2 * info["n_variants"] * 2002 * 4,
                                                   **{k: info[k] for k in ("writer_busy_s", "writer_bytes",
                                                                           "producer_blocked_s", "chunk_variants",
                                                                           "host_descriptor_s", "host_enqueue_s",
                                                                           "host_gpu_wait_s")
                                                      if k in info}}

    # CPU reference (oracle port of chromatin.py's loop) on a tiny sample, all host threads
    from oracle import beluga_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    sdt = {k: torch.from_numpy(v) for k, v in sd.items()}
    dg = O.DictGenome(genome)
    sample = O.normalise_vcf_rows(rows[:args.cpu_variants])
    t0 = time.perf_counter()
    for s in shifts:
        O.chromatin_shift(dg, sample, sdt, s)
    dt = time.perf_counter() - t0
    res["chromatin_cpu_port"] = {"variants": len(sample), "windows": len(sample) * 36, "seconds": dt,
                                 "windows_per_s": len(sample) * 36 / dt, "cores": os.cpu_count()}

    # ---------------------------------------------------------------- predict (cfg5 shape)
    rng = np.random.default_rng(5)
    mdir = os.path.join(tmp, "models")
    os.makedirs(mdir, exist_ok=True)
    mlist = os.path.join(tmp, "modellist")
    with open(mlist, "w") as f:
        for i in range(args.models):
            p = os.path.join(mdir, f"model_{i}.save")
            gblinear.save_binary(p, gblinear.GBLinearModel((rng.standard_normal(20020) * 0.01).astype(np.float32),
                                                            float(rng.standard_normal() * 0.1), 0.5))
            f.write(p + "\n")
    genefile = os.path.join(tmp, "closestgene")
    with open(genefile, "w") as f:
        for i, r in enumerate(rows):
            f.write("\t".join([str(r[0]), str(r[1] - 1), str(r[1]), r[3], r[4], "+" if i % 2 else "-", f"ENSG{i % 997:05d}",
                               str(int(rng.integers(-20000, 20000)))]) + "\n")
    feats = os.path.join(tmp, "features.tsv")
    with open(feats, "w") as f:
        f.write("\tCell type\tAssay\tTreatment\tAssay type\tSource\n")
        for i in range(2002):
            f.write(f"{i}\tcell\tassay{i}\tNone\tTF\tsrc\n")
    pa = predict.build_parser().parse_args(["--modelList", mlist, "--belugaFeatures", feats, "--coorFile_chromatin",
                                            os.path.join(out_dir, "snps_hg19.vcf"), "--geneFile", genefile,
                                            "--snpEffectFilePattern", os.path.join(out_dir, "snps.shift_SHIFT.diff.h5"),
                                            "-o", os.path.join(tmp, "predict_out")])
    t0 = time.perf_counter()
    try:
        pinfo = predict.run(pa)
        res["predict_cli"] = {"rows": pinfo["n_rows"], "models": args.models, "seconds": time.perf_counter() - t0,
                              "variants_per_s": pinfo["n_rows"] / (time.perf_counter() - t0),
                              "timing": pinfo["timing"]}
    except Exception as e:  # the synthetic VCF has one filtered row: predict.py shares the reference's caveat
        res["predict_cli"] = {"error": repr(e)[:300]}

    # ---------------------------------------------------------------- features (cfg4 shape)
    g = args.genes
    genes = []
    for i in range(g):
        c = contigs[i % len(contigs)]
        genes.append((f"ENSG{i:05d}", c[0], int(rng.integers(30000, c[1] - 30000)), 1 if i % 2 else -1))
    from expecto_b200.transform import TSS_SHIFTS, fixed_pos_weights
    for mode in ("windows", "segments"):
        cef.gene_features(eng, pg, genes[:8], TSS_SHIFTS, fixed_pos_weights(), 8, mode=mode)  # warm-up
        t0 = time.perf_counter()
        f = cef.gene_features(eng, pg, genes, TSS_SHIFTS, fixed_pos_weights(), 32, mode=mode)
        dt = time.perf_counter() - t0
        res[f"features_{mode}"] = {"genes": g, "windows": g * 400, "seconds": dt, "windows_per_s": g * 400 / dt,
                                   "genes_per_s": g / dt, "checksum": float(f.sum())}
    print(json.dumps(res, indent=1))
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
