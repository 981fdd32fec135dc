#!/bin/bash
# small-input path A/B: devbench check (all cases) + timings at small sizes, small tiles on/off
mkdir -p gpurun_out
D=gpu-radix-sort_b200/libsort/devbench
timeout 600 $D check > gpurun_out/small_check.log 2>&1; echo "check rc=$?"; tail -2 gpurun_out/small_check.log
rm -f gpurun_out/small_sizes.log
for sp in 0 1; do
  echo "== LIBSORT_SMALL_PATH=$sp" >> gpurun_out/small_sizes.log
  for n in 4096 65536 262144 1048576 2097152 4194304; do LIBSORT_SMALL_PATH=$sp $D time $n 20 2>&1 | grep full32 >> gpurun_out/small_sizes.log; done
done
cat gpurun_out/small_sizes.log
