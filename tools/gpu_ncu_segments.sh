#!/bin/bash
# ncu launch list (durations only) of a segment + delta mode run: which kernels the variant workflow spends its time in.
mkdir -p gpurun_out
CMD="python tools/segment_profile.py 1052"
$CMD > gpurun_out/ncu_seg_plain.json 2> gpurun_out/ncu_seg_plain.err || { echo "plain run failed"; tail -5 gpurun_out/ncu_seg_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_segments.csv $CMD > gpurun_out/ncu_seg.log 2>&1
echo "ncu rc=$?"
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/launches_segments.csv")) if len(r)>10 and r[0].isdigit()]
t=collections.defaultdict(float); n=collections.Counter(); rd=collections.defaultdict(float); wr=collections.defaultdict(float)
for r in rows:
    name=r[4].split("(")[0].replace("void ","").replace("bx::","")[:70]
    v=float(r[-1].replace(",",""))
    if r[-3]=="gpu__time_duration.sum": t[name]+=v*{"ns":1e-6,"us":1e-3,"ms":1}.get(r[-2],1e-6); n[name]+=1
    elif r[-3]=="dram__bytes_read.sum": rd[name]+=v*{"byte":1,"Kbyte":1e3,"Mbyte":1e6,"Gbyte":1e9}.get(r[-2],1)
    elif r[-3]=="dram__bytes_write.sum": wr[name]+=v*{"byte":1,"Kbyte":1e3,"Mbyte":1e6,"Gbyte":1e9}.get(r[-2],1)
tot=sum(t.values())
print("kernel,launches,total_ms,share,dram_read_GB,dram_write_GB")
for k,v in sorted(t.items(), key=lambda x:-x[1]):
    print(f"{k},{n[k]},{v:.3f},{v/tot:.3f},{rd[k]/1e9:.3f},{wr[k]/1e9:.3f}")
PY
