#!/bin/bash
# Round-1 GPU session A: parity tests, primitive rates, tile-shape sweep, bench, ncu capture.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
( timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log )
timeout 300 gpu-radix-sort_b200/libsort/microbench > gpurun_out/microbench.log 2>&1
for v in build/v_*; do
  echo "== $v" >> gpurun_out/variants.log
  timeout 300 $v/devbench check >> gpurun_out/variants.log 2>&1
  timeout 120 $v/devbench time 268435456 10 >> gpurun_out/variants.log 2>&1
done
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err
D=gpu-radix-sort_b200/libsort/devbench
$D time 268435456 1 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_onesweep -s 8 -c 1 -o gpurun_out/prof_onesweep $D time 268435456 1 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/microbench.log | grep "atoms lane"; cat gpurun_out/variants.log | grep -E "==|check|full32"; cat gpurun_out/bench.json
