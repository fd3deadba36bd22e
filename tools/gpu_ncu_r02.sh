#!/bin/bash
# Round-2 ncu passes (B200_PROFILING.md).  Each profiled command is run plainly first and must exit 0.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no_cpu_baseline --no_workflow --no_sharded --no_transform"
$CMD > gpurun_out/r02_ncu_plain.json 2> gpurun_out/ncu_plain.err || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_ncu_launches_bench.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
# conv2 = the 2nd GEMM launch of a forward chunk (8 GEMMs per chunk): skip one whole chunk, take the next conv2
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 9 -c 1 -f -o gpurun_out/r02_prof_conv2 $CMD > gpurun_out/ncu_conv2.log 2>&1
echo "ncu conv2 rc=$?"
ncu -i gpurun_out/r02_prof_conv2.ncu-rep --page raw --csv 2>/dev/null > gpurun_out/r02_ncu_conv2_raw.csv
# segment + delta mode through the in-library pipeline: which kernels remain
SEG="python tools/segment_profile.py 1052"
$SEG > gpurun_out/r02_segment_profile.json 2> gpurun_out/ncu_seg_plain.err || { echo "segment run failed"; tail -5 gpurun_out/ncu_seg_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_segments.csv $SEG > gpurun_out/ncu_seg.log 2>&1
echo "ncu segments rc=$?"
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/launches_segments.csv")) if len(r)>10 and r[0].isdigit()]
t=collections.defaultdict(float); n=collections.Counter()
for r in rows:
    name=r[4].split("(")[0].replace("void ","").replace("bx::","")[:90]
    v=float(r[-1].replace(",",""))
    if r[-3]=="gpu__time_duration.sum": t[name]+=v*{"ns":1e-6,"us":1e-3,"ms":1}.get(r[-2],1e-6); n[name]+=1
tot=sum(t.values())
with open("gpurun_out/r02_ncu_kernel_totals_segment_mode.csv","w") as f:
    f.write("kernel,launches,total_ms,share\n")
    for k,v in sorted(t.items(), key=lambda x:-x[1]):
        f.write(f"\"{k}\",{n[k]},{v:.3f},{v/tot:.4f}\n")
print(open("gpurun_out/r02_ncu_kernel_totals_segment_mode.csv").read())
PY
