#!/bin/bash
# usage: gpu_sweep.sh [modes]   -- devbench check (atoms) + timing for every build/v_* variant
mkdir -p gpurun_out
rm -f gpurun_out/variants.log
MODES=${1:-atoms}
for v in build/v_*; do
  echo "== $v" >> gpurun_out/variants.log
  for mode in $MODES; do
    echo "-- rank=$mode" >> gpurun_out/variants.log
    LIBSORT_RANK=$mode timeout 300 $v/devbench check >> gpurun_out/variants.log 2>&1
    LIBSORT_RANK=$mode timeout 120 $v/devbench time 268435456 10 >> gpurun_out/variants.log 2>&1
  done
done
grep -E "==|--|check|full32" gpurun_out/variants.log
