#!/usr/bin/env python
"""Sharded compute_expecto_features under torchrun (N GPUs of one box): the .npy gathered from N gene shards must be
bit-identical to a single-GPU run; then BASELINE configs[3] at full size (24,339 genes) is timed on the N GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 \
        tools/multi_gpu_features_check.py"""
import json, os, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def write_inputs(base, pg, n, seed):
    rng = np.random.default_rng(seed)
    names = pg.names
    w = np.array([pg.lengths[k] for k in names], dtype=np.float64)
    ci = rng.choice(len(names), size=n, p=w / w.sum())
    anno = os.path.join(base, f"geneanno_{n}.csv")
    with open(anno, "w") as f:
        f.write("id,symbol,seqnames,strand,TSS,CAGE_representative_TSS,type\n")
        for i, c in enumerate(ci):
            tss = int(rng.integers(21000, pg.lengths[names[c]] - 21000))
            f.write(f"ENSG{i:011d},G{i},{names[c]},{'+' if rng.random() < 0.5 else '-'},{tss},{tss},protein_coding\n")
    tss_file = os.path.join(base, "tss.tsv")
    with open(tss_file, "w") as f:
        f.write("\tens_id\tchrom\ttss\tstrand\tx\tis_default_tss\n0\tNOPE\tchr1\t1\t+\t0\tTrue\n")
    return anno, tss_file


def main():
    import torch
    import torch.distributed as dist
    from expecto_b200 import compute_expecto_features as cef, shard, synth
    from expecto_b200.engine import BelugaEngine

    rank, world, local = shard.init_distributed()
    base = os.environ.get("BX_CHECK_DIR") or os.path.join(tempfile.gettempdir(), "bx_multi_gpu_features")
    os.makedirs(base, exist_ok=True)
    pg = synth.make_packed_genome(seed=3)
    n_small, n_full = 203, int(os.environ.get("BX_CHECK_GENES", "24339"))
    if rank == 0:
        write_inputs(base, pg, n_small, 7)
        write_inputs(base, pg, n_full, 8)
    shard.barrier()
    eng = BelugaEngine(local, dtype="f16", capacity=4096)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
    eng.set_genome(pg)
    tss_file = os.path.join(base, "tss.tsv")
    res = {"world": world}
    args = cef.build_parser().parse_args([os.path.join(base, f"geneanno_{n_small}.csv"), tss_file, "-o",
                                          os.path.join(base, "sharded")])
    cef.run(args, engine=eng, genome=pg)
    ok = True
    if rank == 0:
        real = (shard.init_distributed, shard.barrier)
        shard.init_distributed = lambda backend=None: (0, 1, local)
        shard.barrier = lambda: None
        args1 = cef.build_parser().parse_args([os.path.join(base, f"geneanno_{n_small}.csv"), tss_file, "-o",
                                               os.path.join(base, "single"), "--mode", "windows"])
        cef.run(args1, engine=eng, genome=pg)
        shard.init_distributed, shard.barrier = real
        a = np.load(os.path.join(base, "sharded", "Xreducedall.2002.representative_tss_top.npy"))
        b = np.load(os.path.join(base, "single", "Xreducedall.2002.representative_tss_top.npy"))
        ok = bool(a.shape == b.shape == (n_small, 20020) and np.array_equal(a, b))
        res["sharded_bit_identical_to_single_gpu_window_path"] = ok
    shard.barrier()
    args = cef.build_parser().parse_args([os.path.join(base, f"geneanno_{n_full}.csv"), tss_file, "-o",
                                          os.path.join(base, "full")])
    torch.cuda.synchronize()
    shard.barrier()
    t0 = time.perf_counter()
    path = cef.run(args, engine=eng, genome=pg)
    dt = time.perf_counter() - t0
    if rank == 0:
        res["cfg4_full"] = {"genes": n_full, "windows": 400 * n_full, "seconds": round(dt, 2),
                            "windows_per_s": round(400 * n_full / dt), "genes_per_s": round(n_full / dt, 1),
                            "npy_bytes": os.path.getsize(path),
                            "note": "wall clock of the whole CLI run(): read genes, forward + transform on N GPUs, "
                                    "NCCL gather, np.save of the float64 matrix"}
    # the same job with --gather file: every rank writes its own rows of the .npy (no collective)
    shard.barrier()
    args = cef.build_parser().parse_args([os.path.join(base, f"geneanno_{n_full}.csv"), tss_file, "-o",
                                          os.path.join(base, "full_file"), "--gather", "file"])
    torch.cuda.synchronize()
    shard.barrier()
    t0 = time.perf_counter()
    path_f = cef.run(args, engine=eng, genome=pg)
    dt_f = time.perf_counter() - t0
    if rank == 0:
        same = open(path, "rb").read() == open(path_f, "rb").read() if world > 1 else True
        ok = ok and same
        res["cfg4_full_gather_file"] = {"seconds": round(dt_f, 2), "windows_per_s": round(400 * n_full / dt_f),
                                        "genes_per_s": round(n_full / dt_f, 1),
                                        "npy_byte_identical_to_nccl_gather": bool(same)}
        print(json.dumps(res), flush=True)
    if world > 1:
        flag = torch.tensor([1 if ok else 0], device=eng.device)
        dist.broadcast(flag, src=0)
        ok = bool(flag.item())
        dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
