#!/bin/bash
# BASELINE configs[1]: Beluga forward batch sweep 64..8192 on one B200 (bench.py per batch size; device-resident and
# end-to-end windows/s, TFLOP/s, fraction of the measured bf16 peak, per-stage times).
mkdir -p gpurun_out
: > gpurun_out/batch_sweep.jsonl
for B in 64 128 256 512 1024 2048 4096 8192; do
  CAP=$(( B < 4096 ? B : 4096 ))
  timeout 300 python bench.py --batch $B --capacity $CAP --steps ${STEPS:-20} --warmup 5 --no_cpu_baseline --no_transform --no_workflow >> gpurun_out/batch_sweep.jsonl 2>> gpurun_out/batch_sweep.err || echo "B=$B failed"
done
python - <<'PY'
import json
rows = [json.loads(l) for l in open("gpurun_out/batch_sweep.jsonl") if l.strip()]
out = []
for d in rows:
    out.append({"batch": d["config"]["batch_per_gpu"], "windows_per_s": round(d["value"]), "e2e_windows_per_s": round(d["e2e"]["value"]),
                "ms_per_step": round(d["ms_per_step"], 3), "tflops": round(d["tflops_per_gpu"], 1),
                "frac_of_burst_peak": round(d["forward_roofline"]["frac"], 3), "sm_mhz": d["clocks"]["sm_mhz"],
                "stage_ms": d["stage_ms_per_launch"]})
    print(out[-1])
json.dump(out, open("gpurun_out/batch_sweep.json", "w"), indent=1)
PY
