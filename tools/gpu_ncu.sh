#!/bin/bash
# usage: gpu_ncu.sh <out-name> [env...]  -- ncu --set full on one k_onesweep launch of the main build
mkdir -p gpurun_out
D=gpu-radix-sort_b200/libsort/devbench
NAME=${1:-prof}
$D time 268435456 1 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_onesweep -s 8 -c 1 -o gpurun_out/$NAME $D time 268435456 1 > gpurun_out/ncu.log 2>&1
cat gpurun_out/plain.log
