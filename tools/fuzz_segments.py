#!/usr/bin/env python
"""Randomised bit-identity sweep: segment + delta mode (chromatin loop) against the per-window path over random
variant mixes (SNVs, MNVs up to 20 bp, insertions up to 300 bp, deletions up to 120 bp, long ref alleles, lower case,
N alleles, ref mismatches, variants next to N runs and contig ends) and random shift ranges.

    python tools/fuzz_segments.py [--seconds 90] [--seed 0]"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def random_variants(rng, genome, n, margin):
    rows = []
    names = sorted(genome)
    for i in range(n):
        c = names[int(rng.integers(len(names)))]
        seq = genome[c]
        p = int(rng.integers(margin, len(seq) - margin))
        kind = rng.choice(["snv", "snv", "snv", "mnv", "ins", "del", "longref", "mismatch", "nallele"])
        g = lambda a, b: "".join(ch if ch.upper() in "ACGT" else "A" for ch in seq[a:b]).upper()
        rb = lambda k: "".join("ACGT"[int(x)] for x in rng.integers(0, 4, size=k))
        if kind == "snv":
            ref = g(p - 1, p); alt = rb(1)
        elif kind == "mnv":
            k = int(rng.integers(2, 21)); ref = g(p - 1, p - 1 + k); alt = rb(k)
        elif kind == "ins":
            ref = g(p - 1, p); alt = ref + rb(int(rng.integers(1, 301)))
        elif kind == "del":
            k = int(rng.integers(1, 121)); ref = g(p - 1, p + k); alt = ref[0]
        elif kind == "longref":
            k = int(rng.integers(100, 300)); ref = g(p - 1, p - 1 + k); alt = rb(int(rng.integers(1, 40)))
        elif kind == "mismatch":
            k = int(rng.integers(1, 5)); ref = rb(k); alt = rb(k)
        else:
            k = int(rng.integers(1, 6)); ref = g(p - 1, p - 1 + k); alt = "".join(rng.choice(list("ACGTN"), size=k))
        if rng.random() < 0.15:
            alt = alt.lower()
        if rng.random() < 0.1:
            ref = ref.lower()
        rows.append((c, p, f"v{i}", ref, alt))
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=90)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--variants", type=int, default=96)
    args = ap.parse_args()
    import torch
    from expecto_b200 import chromatin, synth
    from expecto_b200.engine import BelugaEngine, chromatin_outputs
    from expecto_b200.genome import PackedGenome
    from expecto_b200.windows import check_valid_bases, chromatin_window_descs

    eng = BelugaEngine(0, dtype="f16", capacity=1024)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
    rng = np.random.default_rng(args.seed)
    t_end = time.time() + args.seconds
    stats = {"iterations": 0, "variants": 0, "windows_compared": 0, "mismatches": 0, "delta_variants": 0,
             "window_path_variants": 0, "maxshifts": {}}
    while time.time() < t_end:
        genome = synth.make_genome([("chr1", int(rng.integers(60000, 120000))), ("chrX", int(rng.integers(50000, 90000)))],
                                   seed=int(rng.integers(1 << 30)))
        pg = PackedGenome.from_sequences(genome)
        eng.set_genome(pg)
        maxshift = int(rng.choice([0, 200, 400, 800, 800, 1000, 1200, 2000]))
        rows = random_variants(rng, genome, args.variants, margin=maxshift + 1300)
        chroms, pos = [r[0] for r in rows], np.array([r[1] for r in rows])
        refs, alts = [r[3] for r in rows], [r[4] for r in rows]
        n = len(rows)
        shifts = chromatin.shift_list(maxshift)
        store = {s: {k: np.full((2 * n, 2002), np.nan, np.float32) for k in ("diff", "ref", "alt")} for s in shifts}

        def make(s):
            def sink(name, row0, arr):
                store[s][name][row0:row0 + len(arr)] = arr
            return sink
        info = chromatin.predict_all_shifts_segments(eng, pg, chroms, pos, refs, alts, maxshift, {s: make(s) for s in shifts})
        for s in shifts:
            d, codes = chromatin_window_descs(pg, chroms, pos, refs, alts, s)
            check_valid_bases(pg, d)
            ref, alt, diff = chromatin_outputs(eng.forward_windows(d, codes), n)
            for name, t in (("ref", ref), ("alt", alt), ("diff", diff)):
                same = np.array_equal(store[s][name], t.cpu().numpy())
                stats["mismatches"] += 0 if same else 1
                if not same:
                    bad = np.flatnonzero((store[s][name] != t.cpu().numpy()).any(axis=1))
                    print("MISMATCH", maxshift, s, name, [rows[i % n] for i in bad[:3]], flush=True)
            stats["windows_compared"] += 4 * n
        stats["iterations"] += 1
        stats["variants"] += n
        stats["delta_variants"] += info["delta_variants"]
        stats["window_path_variants"] += info["window_variants"]
        stats["maxshifts"][str(maxshift)] = stats["maxshifts"].get(str(maxshift), 0) + 1
    print(json.dumps(stats))
    return 0 if stats["mismatches"] == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
