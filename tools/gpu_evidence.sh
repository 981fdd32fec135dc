#!/bin/bash
# Round evidence, one ncu per call:
#   gpu_evidence.sh tests  : GPU parity tests, both bench arms, ncu launch list of the bench command, devbench sizes
#   gpu_evidence.sh full   : ncu --set full capture of one scatter pass (after the same command ran plain)
mkdir -p gpurun_out
D=gpu-radix-sort_b200/libsort/devbench
if [ "${1:-tests}" = "tests" ]; then
( timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log )
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py --quick --steps 2 --warmup 1 > gpurun_out/plain_bench.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --quick --steps 2 --warmup 1 > gpurun_out/ncu_launches.log 2>&1
rm -f gpurun_out/devbench_sizes.log
for n in 1048576 268435456 536870912; do $D time $n 10 >> gpurun_out/devbench_sizes.log 2>&1; done
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json gpurun_out/bench_ref.json; cat gpurun_out/devbench_sizes.log
else
$D time 268435456 1 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_onesweep -s 8 -c 1 -f -o gpurun_out/prof_final $D time 268435456 1 > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log; ls -la gpurun_out/prof_final.ncu-rep
fi
