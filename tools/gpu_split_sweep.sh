#!/bin/bash
# usage: gpu_split_sweep.sh N -- distributed bench for every build/s_* splitter-scatter shape
N=${1:-2}
mkdir -p gpurun_out
for v in build/s_*; do
LIBSORT_SO=$PWD/$v/libsort.so timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu_$(basename $v).json 2> gpurun_out/err.log || tail -3 gpurun_out/err.log
done
