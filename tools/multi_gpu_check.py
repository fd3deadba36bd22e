#!/usr/bin/env python
"""Multi-GPU check of the chromatin drop-in (run under torchrun on N GPUs of one box):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/multi_gpu_check.py

Every rank runs its shard; rank 0 compares the files written in ``--gather nccl`` and ``--gather file`` mode
with a single-GPU run of the same VCF on rank 0's GPU (bit-identical: results do not depend on batching)."""
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    from expecto_b200 import chromatin, hdf5, shard, synth
    from expecto_b200.engine import BelugaEngine
    from expecto_b200.genome import PackedGenome

    rank, world, local = shard.init_distributed()
    n_var = int(os.environ.get("BX_CHECK_VARIANTS", "301"))
    genome = synth.make_genome([("chr1", 400_000), ("chr2", 300_000), ("chrX", 200_000)], seed=21)
    pg = PackedGenome.from_sequences(genome)
    rows = synth.make_vcf_rows(genome, n_var, seed=3)
    base = os.environ.get("BX_CHECK_DIR") or os.path.join(tempfile.gettempdir(), "bx_multi_gpu_check")
    os.makedirs(base, exist_ok=True)
    vcf = os.path.join(base, "in.vcf")
    if rank == 0:
        synth.write_vcf(vcf, rows)
    shard.barrier()
    eng = BelugaEngine(local, dtype="f16", capacity=1024)
    eng.load_state_dict(synth.make_beluga_weights(0, round_dtype="f16"))
    res = {}
    for mode in ("nccl", "file"):
        args = chromatin.build_parser().parse_args([vcf, "--output_dir", os.path.join(base, mode), "--maxshift", "200",
                                                    "--gather", mode, "--gpu_batch", "64"])
        info = chromatin.run(args, engine=eng, genome=pg)
        res[mode] = info
    shard.barrier()
    ok = True
    if rank == 0:
        # single-process run on this GPU: temporarily hide the process group from chromatin.run
        env = {k: os.environ.pop(k) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK") if k in os.environ}
        import expecto_b200.shard as sh
        real = (sh.init_distributed, sh.barrier, sh.allreduce_sum_ints, sh.shard_range)
        sh.init_distributed = lambda backend=None: (0, 1, local)
        sh.barrier = lambda: None
        sh.allreduce_sum_ints = lambda v, device=None: [int(x) for x in v]
        args = chromatin.build_parser().parse_args([vcf, "--output_dir", os.path.join(base, "single"), "--maxshift",
                                                    "200", "--gpu_batch", "128", "--mode", "windows"])
        single = chromatin.run(args, engine=eng, genome=pg)
        sh.init_distributed, sh.barrier, sh.allreduce_sum_ints, sh.shard_range = real
        os.environ.update(env)
        report = {"world": world, "n_variants": single["n_variants"], "counters": list(single["counters"])}
        for mode in ("nccl", "file"):
            same = list(res[mode]["counters"]) == list(single["counters"])
            for s in (0, -200, 200):
                with hdf5.File(os.path.join(base, "single", f"snps.shift_{s}.diff.h5")) as a, \
                        hdf5.File(os.path.join(base, mode, f"snps.shift_{s}.diff.h5")) as b:
                    for k in ("diff", "ref", "alt"):
                        same &= bool(np.array_equal(np.asarray(a[k]), np.asarray(b[k])))
            report[mode + "_bit_identical_to_single_gpu"] = bool(same)
            ok &= same
        print(json.dumps(report), flush=True)
    if world > 1:
        flag = torch.tensor([1 if ok else 0], device=eng.device)
        dist.broadcast(flag, src=0)
        ok = bool(flag.item())
        dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
