#!/bin/bash
# usage: gpu_multi_lite.sh N -- one bench line per distribution at N GPUs (p2p exchange)
N=${1:-8}
mkdir -p gpurun_out
for d in uniform consthigh zipf halfdup allequal; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 --dist $d > gpurun_out/bench_${N}gpu_$d.json 2> gpurun_out/bench_${N}gpu_$d.err
tail -2 gpurun_out/bench_${N}gpu_$d.err
done
cat gpurun_out/bench_${N}gpu_uniform.json
