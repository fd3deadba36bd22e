#!/bin/bash
# Does skipping the epilogue (BX_DEBUG_FLAGS=4) shorten conv layers in CYCLES or only in time (clock)?  ncu, clocks free.
mkdir -p gpurun_out
for F in 0 4; do
BX_DEBUG_FLAGS=$F timeout 600 ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.max,sm__cycles_elapsed.avg.per_second,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active \
  --clock-control none -k regex:gemm_tc_kernel -s 16 -c 8 --csv --log-file gpurun_out/epi_cycles_f$F.csv python tools/diag_flags.py stage 2048 > gpurun_out/epi_cycles_f$F.log 2>&1
echo "flags $F rc=$?"
done
python - <<'PY'
import csv
for F in (0,4):
    rows=[r for r in csv.reader(open(f"gpurun_out/epi_cycles_f{F}.csv")) if len(r)>10 and r[0].isdigit()]
    per={}
    for r in rows:
        per.setdefault(r[0],{"k":r[4][:60]})[r[-3]]=r[-1]
    print("flags",F)
    for k,v in per.items(): print(k,v)
PY
