#!/bin/bash
# usage: gpu_multi_ab.sh N -- distributed bench with the bulk-copy and the LDS+STG splitter scatter
N=${1:-2}
mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q -k "distrib or nccl or multi" > gpurun_out/pytest_multi_$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi_$N.log )
for b in 1 0; do
LIBSORT_BULK_EXCHANGE=$b timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu_bulk$b.json 2> gpurun_out/bench_${N}gpu_bulk$b.err
tail -2 gpurun_out/bench_${N}gpu_bulk$b.err
done
tail -2 gpurun_out/pytest_multi_$N.log
