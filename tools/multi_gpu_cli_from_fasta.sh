#!/bin/bash
# The chromatin CLI started cold from a FASTA (no .bx2.npz cache) on 1 and on N GPUs (torchrun): rank 0 packs the genome,
# the other ranks read the cache (shard.load_genome); the per-shift datasets must be bit-identical.
N=${1:-2}
mkdir -p gpurun_out
W=$(mktemp -d)
python - "$W" <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, ".")
from expecto_b200 import synth
w = sys.argv[1]
g = synth.make_genome([("chr1", 600_000), ("chr2", 400_000), ("chrX", 300_000)], seed=31)
synth.write_fasta(f"{w}/g.fa", g, width=50)
synth.write_vcf(f"{w}/in.vcf", synth.make_vcf_rows(g, 777, seed=5))
torch.save({k: torch.from_numpy(v) for k, v in synth.make_beluga_weights(0, "f16").items()}, f"{w}/beluga.pth")
PY
set -e
python -m expecto_b200.chromatin $W/in.vcf --genome $W/g.fa --weights $W/beluga.pth --output_dir $W/one > gpurun_out/cli_n1.log 2>&1
rm -f $W/g.fa.bx2.npz
for mode in nccl file; do
  rm -f $W/g.fa.bx2.npz
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    -m expecto_b200.chromatin $W/in.vcf --genome $W/g.fa --weights $W/beluga.pth --output_dir $W/many_$mode --gather $mode > gpurun_out/cli_n${N}_$mode.log 2>&1
  python - $W/one $W/many_$mode <<'PY'
import os, sys, numpy as np
sys.path.insert(0, ".")
from expecto_b200 import hdf5
a, b = sys.argv[1:3]
for f in sorted(os.listdir(a)):
    if f.endswith(".h5"):   # the object headers carry a modification time: compare the datasets bit for bit
        with hdf5.File(f"{a}/{f}") as x, hdf5.File(f"{b}/{f}") as y:
            assert sorted(x) == sorted(y)
            for k in x:
                assert np.array_equal(np.asarray(x[k][...]).view(np.uint32), np.asarray(y[k][...]).view(np.uint32)), (f, k)
    else:
        assert open(f"{a}/{f}", "rb").read() == open(f"{b}/{f}", "rb").read(), f
PY
  echo "N=$N --gather $mode: $(ls $W/one | wc -l) files bit-identical to the single-GPU run (cold FASTA start)"
done | tee gpurun_out/multi_gpu_cli_from_fasta_n$N.log
rm -rf $W
