#!/bin/bash
# usage: gpu_k3sweep.sh "<variants to time>" "<variants to check>" ["<variants to ncu>"]
# Times (and verifies at full size) the scatter2 variants selected with LIBSORT_K3; "base" = the shipped default.
mkdir -p gpurun_out
D=gpu-radix-sort_b200/libsort/devbench
L=gpurun_out/k3sweep.log
rm -f $L
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv >> $L 2>&1
for v in $1; do
  echo "== $v" >> $L
  if [ "$v" = "base" ]; then LIBSORT_DEBUG=1 timeout 120 $D time 268435456 10 >> $L 2>&1
  else LIBSORT_K3=$v LIBSORT_DEBUG=1 timeout 120 $D time 268435456 10 >> $L 2>&1; fi
done
for v in $2; do
  echo "== check $v" >> $L
  LIBSORT_K3=$v timeout 400 $D check >> $L 2>&1
done
if [ -x gpu-radix-sort_b200/libsort/microbench_tma_store ]; then
  echo "== microbench_tma_store" >> $L
  for t in "14336 3" "11264 4" "12288 2"; do timeout 60 gpu-radix-sort_b200/libsort/microbench_tma_store $t >> $L 2>&1; done
fi
for v in $3; do
  LIBSORT_K3=$v $D time 268435456 1 > gpurun_out/plain_$v.log 2>&1 && \
  LIBSORT_K3=$v timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_scatter -s 8 -c 1 -f -o gpurun_out/prof_$v $D time 268435456 1 > gpurun_out/ncu_$v.log 2>&1
  tail -2 gpurun_out/ncu_$v.log >> $L
done
grep -E "^==|full32|verify|check:|FAIL|K3 variant|bulk|lds" $L
