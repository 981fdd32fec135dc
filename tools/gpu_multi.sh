#!/bin/bash
# usage: gpu_multi.sh N  -- distributed tests and bench lines on N GPUs of one box
N=${1:-2}
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q -k "distrib or nccl or multi" > gpurun_out/pytest_multi_$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi_$N.log )
for ex in p2p nccl; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --exchange $ex > gpurun_out/bench_${N}gpu_$ex.json 2> gpurun_out/bench_${N}gpu_$ex.err
done
for d in zipf consthigh; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --dist $d > gpurun_out/bench_${N}gpu_$d.json 2> gpurun_out/bench_${N}gpu_$d.err
done
tail -3 gpurun_out/pytest_multi_$N.log; cat gpurun_out/bench_${N}gpu_*.json; tail -5 gpurun_out/bench_${N}gpu_p2p.err
