#!/bin/bash
# CLI-path check after host-side changes: GPU tests, then the workflow timings.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python tools/bench_workflows.py --variants ${VARIANTS:-8192} --genes 64 --cpu_variants 1 --out gpurun_out/workflows.json > gpurun_out/workflows.log 2>&1; echo "workflows rc=$?"
tail -3 gpurun_out/workflows.log
python - <<'PY'
import json
d=json.load(open("gpurun_out/workflows.json"))
for k,v in d.items():
    if isinstance(v,dict): print(k,{a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a!="checksum"})
PY
