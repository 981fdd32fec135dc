#!/bin/bash
# GPU session: tile-shape / look-back sweep of the single-buffer onesweep kernel, both rank modes.
mkdir -p gpurun_out
rm -f gpurun_out/variants.log
for v in build/v_*; do
  echo "== $v" >> gpurun_out/variants.log
  for mode in atoms ballot; do
    echo "-- rank=$mode" >> gpurun_out/variants.log
    if [ $mode = atoms ] || [ $v = build/v_256_32_4_lb2 ] || [ $v = build/v_512_16_2_lb2 ]; then
      LIBSORT_RANK=$mode timeout 300 $v/devbench check >> gpurun_out/variants.log 2>&1
    fi
    LIBSORT_RANK=$mode LIBSORT_DEBUG=1 timeout 120 $v/devbench time 268435456 10 >> gpurun_out/variants.log 2>&1
  done
done
echo "-- auto" >> gpurun_out/variants.log
LIBSORT_DEBUG=1 timeout 120 build/v_256_32_4_lb2/devbench time 268435456 10 >> gpurun_out/variants.log 2>&1
grep -E "==|--|check|full32|rank=" gpurun_out/variants.log
