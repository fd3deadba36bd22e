#!/bin/bash
# A/B of conv1 with in-kernel operand generation vs the im2col matrix: tests first, then bench both ways.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/summary.txt
tail -5 gpurun_out/pytest_gpu.log
for i in 1; do
timeout 600 python bench.py --steps 10 --warmup 3 --no_cpu_baseline --no_transform > gpurun_out/bench_gen_$i.json 2> gpurun_out/bench_gen.err; echo "bench gen rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_gen_$i.json'));print(d['value'],d['e2e']['value'],d['stage_ms_per_launch'],d['clocks'])"
BX_CONV1_IM2COL=1 timeout 600 python bench.py --steps 10 --warmup 3 --no_cpu_baseline --no_transform > gpurun_out/bench_im2col_$i.json 2> gpurun_out/bench_im2col.err; echo "bench im2col rc=$?"
python -c "import json;d=json.load(open('gpurun_out/bench_im2col_$i.json'));print(d['value'],d['e2e']['value'],d['stage_ms_per_launch'],d['clocks'])"
done
timeout 900 python tools/bench_workflows.py --variants 4096 --genes 64 --cpu_variants 1 --out gpurun_out/workflows_gen.json > gpurun_out/workflows_gen.log 2>&1; echo "workflows gen rc=$?"
BX_CONV1_IM2COL=1 timeout 900 python tools/bench_workflows.py --variants 4096 --genes 64 --cpu_variants 1 --out gpurun_out/workflows_im2col.json > gpurun_out/workflows_im2col.log 2>&1; echo "workflows im2col rc=$?"
python - <<'PY'
import json
for m in ("gen","im2col"):
    d=json.load(open(f"gpurun_out/workflows_{m}.json"))
    print(m,{k:round(v.get("windows_per_s",0)) for k,v in d.items() if isinstance(v,dict) and "windows_per_s" in v})
PY
