#!/bin/bash
# A/B of an environment switch: tests with the default build, then bench default vs "$AB_ENV" (e.g. BX_CONV1_STG=1).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/pytest_gpu.log
for i in 1 2; do
timeout 600 python bench.py --steps 10 --warmup 3 --no_cpu_baseline --no_transform > gpurun_out/bench_new_$i.json 2> gpurun_out/bench_new.err; echo "bench new rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_new_$i.json'));print(d['value'],d['e2e']['value'],d['stage_ms_per_launch'],d['clocks'])"
env $AB_ENV timeout 600 python bench.py --steps 10 --warmup 3 --no_cpu_baseline --no_transform > gpurun_out/bench_old_$i.json 2> gpurun_out/bench_old.err; echo "bench old ($AB_ENV) rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_old_$i.json'));print(d['value'],d['e2e']['value'],d['stage_ms_per_launch'],d['clocks'])"
done
