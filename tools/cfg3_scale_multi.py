#!/usr/bin/env python
"""BASELINE configs[2] at full size, sharded over the GPUs of one box (run under torchrun, one process per GPU):
the hg19-shaped synthetic genome (3.1 Gb packed, replicated), the synthetic VCF of --variants rows (default 1,000,000)
split into contiguous per-rank shards (expecto_b200/shard.py), segment + delta mode, host descriptors in, pinned host
ref/alt/diff out on every rank (no files, no collective on the path).  Timed from a barrier to the slowest rank's
finish; each rank also checks a sample of its shard bit for bit against the per-window path.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 \
        tools/cfg3_scale_multi.py [--variants 1000000] [--out file.json]"""
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
    import torch.distributed as dist
    from expecto_b200 import chromatin, shard, synth
    from expecto_b200.engine import BelugaEngine, chromatin_outputs
    from expecto_b200.windows import chromatin_window_descs, check_valid_bases

    rank, world, local = shard.init_distributed()
    dev = local if world > 1 else 0
    torch.cuda.set_device(dev)
    pg = synth.make_packed_genome(seed=3)
    chroms, pos, refs, alts = synth.make_vcf_rows_packed(pg, args.variants, seed=3)
    lo, hi = shard.shard_range(len(pos), rank, world)
    chroms, pos, refs, alts = chroms[lo:hi], pos[lo:hi], refs[lo:hi], alts[lo:hi]
    eng = BelugaEngine(dev, dtype="f16", capacity=4096)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
    eng.set_genome(pg)
    shifts = chromatin.shift_list(800)

    # parity sample of this rank's shard (its last variants: the highest global indices it sees)
    n = len(pos)
    sample = np.concatenate([np.arange(0, 12), np.arange(n - 12, n)])
    sc, sp = [chroms[i] for i in sample], pos[sample]
    sr, sa = [refs[i] for i in sample], [alts[i] for i in sample]
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

    sinks = {s: (lambda *a: None) for s in shifts}
    torch.cuda.synchronize()
    shard.barrier()
    t0 = time.perf_counter()
    info = chromatin.predict_all_shifts_segments(eng, pg, chroms, pos, refs, alts, 800, sinks)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    stats = torch.tensor([dt, float(n), float(same), float(info.get("delta_variants", 0))], dtype=torch.float64, device="cuda")
    if world > 1:
        all_stats = [torch.zeros_like(stats) for _ in range(world)]
        dist.all_gather(all_stats, stats)
    else:
        all_stats = [stats]
    if rank == 0:
        a = torch.stack(all_stats).cpu().numpy()
        slowest, total = float(a[:, 0].max()), int(a[:, 1].sum())
        res = {"gpu": torch.cuda.get_device_name(dev), "n_gpus": world, "variants": total, "windows": 36 * total,
               "seconds_slowest_rank": round(slowest, 3), "seconds_per_rank": [round(float(x), 3) for x in a[:, 0]],
               "windows_per_s": round(36 * total / slowest), "variants_per_s": round(total / slowest),
               "d2h_bytes": 9 * 3 * 2 * total * 2002 * 4, "delta_variants": int(a[:, 3].sum()),
               "every_rank_sample_bit_identical_to_window_path": bool(a[:, 2].min() == 1.0),
               "genome_bases": int(pg.n_bases)}
        print(json.dumps(res, indent=1))
        if args.out:
            json.dump(res, open(args.out, "w"), indent=1)
    shard.barrier()
    shard.finalize()
    return 0 if same else 1


if __name__ == "__main__":
    sys.exit(main())
