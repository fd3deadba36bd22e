#!/usr/bin/env python
"""BASELINE configs[2] at full size on ONE B200: an hg19-shaped synthetic genome (24 contigs, 3.1 Gb, packed) and a
synthetic VCF of --variants rows (default 1,000,000: 90 % SNV / 5 % ref-mismatch / 5 % indel), ref/alt x fwd/rc x
9 shifts = 36 windows per variant, through the drop-in's segment + delta mode with host descriptors in and pinned
host ref/alt/diff out (no files: 1 M variants would be 432 GB of HDF5).  Also checks, at this genome size (global
base indices beyond 2^31), that segment + delta results equal the per-window path bit for bit on a sample."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", type=int, default=1_000_000)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    from expecto_b200 import chromatin, synth
    from expecto_b200.engine import BelugaEngine, chromatin_outputs
    from expecto_b200.windows import allele_match_counts, chromatin_window_descs, check_valid_bases

    res = {"gpu": torch.cuda.get_device_name(0), "variants": args.variants}
    t0 = time.perf_counter()
    pg = synth.make_packed_genome(seed=3)
    res["genome_bases"] = int(pg.n_bases)
    res["genome_build_s"] = round(time.perf_counter() - t0, 2)
    t0 = time.perf_counter()
    chroms, pos, refs, alts = synth.make_vcf_rows_packed(pg, args.variants, seed=3)
    res["vcf_build_s"] = round(time.perf_counter() - t0, 2)
    eng = BelugaEngine(0, dtype="f16", capacity=4096)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
    t0 = time.perf_counter()
    eng.set_genome(pg)
    torch.cuda.synchronize()
    res["genome_upload_s"] = round(time.perf_counter() - t0, 2)
    res["genome_hbm_bytes"] = int(pg.bases2.nbytes + pg.nmask.nbytes)
    t0 = time.perf_counter()
    res["match_counters"] = [int(x) for x in allele_match_counts(pg, chroms, pos, refs, alts)]
    res["match_counters_s"] = round(time.perf_counter() - t0, 2)

    # parity sample: the last variants sit on chrX / chrY, global indices > 2^31
    n = len(pos)
    sample = np.concatenate([np.arange(0, 24), np.arange(n // 2, n // 2 + 24), np.arange(n - 48, n)])
    sc, sp = [chroms[i] for i in sample], pos[sample]
    sr, sa = [refs[i] for i in sample], [alts[i] for i in sample]
    shifts = chromatin.shift_list(800)
    got = {s: {} for s in shifts}

    def make(s):
        def sink(name, row0, arr):
            got[s].setdefault(name, np.zeros((2 * len(sample), 2002), np.float32))[row0:row0 + len(arr)] = arr
        return sink
    chromatin.predict_all_shifts_segments(eng, pg, sc, sp, sr, sa, 800, {s: make(s) for s in shifts})
    same = True
    for s in shifts:
        d, codes = chromatin_window_descs(pg, sc, sp, sr, sa, s)
        check_valid_bases(pg, d)
        ref, alt, diff = chromatin_outputs(eng.forward_windows(d, codes), len(sample))
        for name, t in (("ref", ref), ("alt", alt), ("diff", diff)):
            same &= bool(np.array_equal(got[s][name], t.cpu().numpy()))
    res["sample_bit_identical_to_window_path"] = bool(same)
    res["sample_max_global_index"] = int(pg.offsets[sc[-1]] + sp[-1])

    sinks = {s: (lambda *a: None) for s in shifts}
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    info = chromatin.predict_all_shifts_segments(eng, pg, chroms, pos, refs, alts, 800, sinks)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    res["chromatin"] = {"seconds": round(dt, 2), "windows": 36 * n, "windows_per_s": round(36 * n / dt),
                        "variants_per_s": round(n / dt), "d2h_bytes": 9 * 3 * 2 * n * 2002 * 4,
                        **{k: (round(v, 3) if isinstance(v, float) else v) for k, v in info.items()}}
    print(json.dumps(res, indent=1))
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)
    return 0 if same else 1


if __name__ == "__main__":
    sys.exit(main())
