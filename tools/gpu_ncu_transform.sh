#!/bin/bash
# Kernel 4 under ncu: plain run first, then the launch list and one --set full capture per decay kernel.
mkdir -p gpurun_out
python tools/transform_bench.py 10 > gpurun_out/transform_plain.json 2> gpurun_out/transform_plain.err || { tail -5 gpurun_out/transform_plain.err; exit 1; }
cat gpurun_out/transform_plain.json
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:decay_ --csv --log-file gpurun_out/transform_launches.csv python tools/transform_bench.py 2 > gpurun_out/ncu_transform.log 2>&1
echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decay_fixed -s 3 -c 1 -f -o gpurun_out/prof_decay_fixed python tools/transform_bench.py 2 > gpurun_out/ncu_decay_fixed.log 2>&1
echo "ncu fixed rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decay_variant -s 3 -c 1 -f -o gpurun_out/prof_decay_variant python tools/transform_bench.py 2 > gpurun_out/ncu_decay_variant.log 2>&1
echo "ncu variant rc=$?"
