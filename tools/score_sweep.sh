#!/bin/bash
# Times prebuilt variants of the scoring kernel (expecto_b200/lib/var/lib_*.so, built here with BX_NVCC_FLAGS).
set -u
mkdir -p gpurun_out
L=expecto_b200/lib/libbeluga_b200.so
cp $L /tmp/lib_keep.so
for v in expecto_b200/lib/var/lib_*.so; do
  cp $v $L; touch $L
  echo "== $v" >> gpurun_out/score_sweep.log
  for pitch in ${PITCHES:-2016}; do
    timeout 300 python tools/score_probe.py 4 $pitch 2>&1 | tail -1 >> gpurun_out/score_sweep.log
  done
done
cp ${TESTLIB:-expecto_b200/lib/var/lib_A.so} $L; touch $L
timeout 900 python -m pytest tests/test_gpu_transform.py -x -q -k "scoring or transform" 2>&1 | tail -5 >> gpurun_out/score_sweep.log
cp /tmp/lib_keep.so $L
