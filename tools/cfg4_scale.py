#!/usr/bin/env python
"""BASELINE configs[3] at full size on ONE B200: 24,339 synthetic TSSs (the row count of resources/geneanno.csv) on the
hg19-shaped packed genome, 200 shifts x fwd/rc = 400 windows per gene + fixed-weight exp-decay transform ->
float64 [24339, 20020] (3.9 GB).  Checks a 16-gene sample against the per-window path (bit-identical)."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genes", type=int, default=24339)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    from expecto_b200 import compute_expecto_features as cef, synth
    from expecto_b200.engine import BelugaEngine
    from expecto_b200.transform import TSS_SHIFTS, fixed_pos_weights

    pg = synth.make_packed_genome(seed=3)
    rng = np.random.default_rng(4)
    names = pg.names
    w = np.array([pg.lengths[k] for k in names], dtype=np.float64)
    ci = rng.choice(len(names), size=args.genes, p=w / w.sum())
    genes = [(f"ENSG{i:011d}", names[c], int(rng.integers(21000, pg.lengths[names[c]] - 21000)), 1 if rng.random() < 0.5 else -1)
             for i, c in enumerate(ci)]
    eng = BelugaEngine(0, dtype="f16", capacity=4096)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
    eng.set_genome(pg)
    W = fixed_pos_weights(TSS_SHIFTS)
    cef.gene_features(eng, pg, genes[:32], TSS_SHIFTS, W, 32, mode="segments")  # warm-up
    a = cef.gene_features(eng, pg, genes[-16:], TSS_SHIFTS, W, 8, mode="segments")
    b = cef.gene_features(eng, pg, genes[-16:], TSS_SHIFTS, W, 8, mode="windows")
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    f = cef.gene_features(eng, pg, genes, TSS_SHIFTS, W, 32, mode="segments")
    dt = time.perf_counter() - t0
    res = {"gpu": torch.cuda.get_device_name(0), "genes": args.genes, "windows": 400 * args.genes, "seconds": round(dt, 2),
           "windows_per_s": round(400 * args.genes / dt), "genes_per_s": round(args.genes / dt, 1),
           "output_bytes_f64": int(f.nbytes), "sample_bit_identical_to_window_path": bool(np.array_equal(a, b)),
           "all_finite": bool(np.isfinite(f).all()), "checksum": float(f.sum())}
    print(json.dumps(res, indent=1))
    if args.out:
        json.dump(res, open(args.out, "w"), indent=1)
    return 0 if res["sample_bit_identical_to_window_path"] else 1


if __name__ == "__main__":
    sys.exit(main())
