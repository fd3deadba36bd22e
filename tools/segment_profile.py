#!/usr/bin/env python
"""Per-stage GPU time of segment-mode chromatin (9 shifts) and TSS features (200 shifts) on one B200."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def main():
    import torch
    from expecto_b200 import chromatin, compute_expecto_features as cef, synth
    from expecto_b200.engine import BelugaEngine
    from expecto_b200.genome import PackedGenome
    from expecto_b200.transform import TSS_SHIFTS, fixed_pos_weights
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    contigs = [("chr1", 24_000_000), ("chr2", 20_000_000)]
    genome = synth.make_genome(contigs, seed=3)
    pg = PackedGenome.from_sequences(genome)
    eng = BelugaEngine(0, dtype="f16", capacity=4096)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
    eng.set_genome(pg)
    rows = synth.make_vcf_rows(genome, n, seed=3, odd_names=False)
    chroms = [r[0] for r in rows]; pos = np.array([r[1] for r in rows]); refs = [r[3] for r in rows]; alts = [r[4] for r in rows]
    sinks = {s: (lambda *a: None) for s in chromatin.shift_list(800)}
    chromatin.predict_all_shifts_segments(eng, pg, chroms[:600], pos[:600], refs[:600], alts[:600], 800, sinks)
    res = {}
    for prof in (False, True):
        eng.profile(prof)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        info = chromatin.predict_all_shifts_segments(eng, pg, chroms, pos, refs, alts, 800, sinks)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        if prof:
            st = eng.profile_read()
            res["chromatin_stage_ms"] = {k: round(v[0], 2) for k, v in st.items()}
            res["chromatin_stage_sum_ms"] = round(sum(v[0] for v in st.values()), 2)
        res["chromatin_profiled_s" if prof else "chromatin_s"] = round(dt, 4)
        res["chromatin_windows_per_s"] = round(len(pos) * 36 / dt)
    eng.profile(False)
    rng = np.random.default_rng(5)
    genes = [(f"ENSG{i:05d}", contigs[i % 2][0], int(rng.integers(30000, contigs[i % 2][1] - 30000)), 1 if i % 2 else -1) for i in range(64)]
    cef.gene_features(eng, pg, genes[:8], TSS_SHIFTS, fixed_pos_weights(), 8, mode="segments")
    for prof in (False, True):
        eng.profile(prof)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        cef.gene_features(eng, pg, genes, TSS_SHIFTS, fixed_pos_weights(), 32, mode="segments")
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        if prof:
            st = eng.profile_read()
            res["features_stage_ms"] = {k: round(v[0], 2) for k, v in st.items()}
            res["features_stage_sum_ms"] = round(sum(v[0] for v in st.values()), 2)
        res["features_profiled_s" if prof else "features_s"] = round(dt, 4)
    print(json.dumps(res, indent=1))

if __name__ == "__main__":
    main()
