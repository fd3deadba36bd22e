#!/bin/bash
# Weak-scaling run of bench.py at N=1,2,4,8 on one box, back to back, plus the sharded-CLI check at N=8.
mkdir -p gpurun_out
for N in 1 2 4 8; do
  if [ $N -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 20 --warmup 3 --no_cpu_baseline --no_transform --no_workflow > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 20 --warmup 3 --no_cpu_baseline --no_transform --no_workflow > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  fi
  echo "N=$N rc=$?"; tail -1 gpurun_out/scale_n$N.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'])"
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29711 tools/multi_gpu_check.py 2>&1 | tail -2 | tee gpurun_out/multi_gpu_check_n8.log
