#!/bin/bash
# A/B on one box: the r01 library (scratch copy under _r01/, not tracked) against the current tree, forward only.
set -x
for i in 1 2; do
  (cd _r01 && python bench.py --steps 10 --no_workflow --no_transform --no_cpu_baseline) > gpurun_out/ab_old_$i.json 2>/dev/null
  python bench.py --steps 10 --no_workflow --no_transform --no_cpu_baseline --no_sharded > gpurun_out/ab_new_$i.json 2>/dev/null
done
python - <<'PY'
import json
for n in ("old_1","new_1","old_2","new_2"):
    d=json.load(open(f"gpurun_out/ab_{n}.json"))
    print(n, round(d["value"]), d["stage_ms_per_launch"], d["clocks"]["sm_mhz"])
PY
