#!/bin/bash
# Round evidence: GPU parity tests, bench line, ncu launch list of the bench command, ncu full capture of the scatter pass.
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log )
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err
python bench.py --steps 2 --warmup 1 > gpurun_out/plain_bench.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_launches.log 2>&1
D=gpu-radix-sort_b200/libsort/devbench
$D time 268435456 1 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_onesweep -s 8 -c 1 -o gpurun_out/prof_final $D time 268435456 1 > gpurun_out/ncu.log 2>&1
for n in 1048576 268435456 536870912; do $D time $n 10 >> gpurun_out/devbench_sizes.log 2>&1; done
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json gpurun_out/bench_ref.json; cat gpurun_out/devbench_sizes.log
