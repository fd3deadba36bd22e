#!/bin/bash
# ncu passes (B200_PROFILING.md): launch list of a short bench run, then --set full on the GEMM kernels of one forward.
# Run only after the same bench command exited 0 without ncu (done first below).
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --batch 2048 --capacity 2048 --no_cpu_baseline"
$CMD > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain.err; exit 1; }
cat gpurun_out/ncu_plain.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
# conv2 (2nd GEMM of the 2nd forward) with source; then all 8 GEMMs of that forward without source
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 9 -c 1 -f -o gpurun_out/prof_conv2 $CMD > gpurun_out/ncu_conv2.log 2>&1
echo "ncu conv2 rc=$?"
timeout 1200 ncu --set full --clock-control none -k regex:gemm_tc_kernel -s 8 -c 8 -f -o gpurun_out/prof_gemms $CMD > gpurun_out/ncu_gemms.log 2>&1
echo "ncu gemms rc=$?"
ls -la gpurun_out/
