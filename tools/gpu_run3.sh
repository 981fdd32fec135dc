#!/bin/bash
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log )
D=gpu-radix-sort_b200/libsort/devbench
$D time 268435456 1 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_onesweep -s 8 -c 1 -o gpurun_out/prof_onesweep_atoms $D time 268435456 1 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/plain.log
