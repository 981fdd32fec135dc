#!/bin/bash
# usage: gpu_multi_quick.sh N [dists] -- distributed tests + bench lines at N GPUs
N=${1:-2}
DISTS=${2:-uniform}
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_distrib.py -m gpu -x -q > gpurun_out/pytest_dist_${N}gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_dist_${N}gpu.log )
tail -3 gpurun_out/pytest_dist_${N}gpu.log
for d in $DISTS; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 --dist $d > gpurun_out/bench_${N}gpu_$d.json 2> gpurun_out/bench_${N}gpu_$d.err
tail -3 gpurun_out/bench_${N}gpu_$d.err
python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/bench_${N}gpu_$d.json").read().strip().splitlines()[-1])
    print("$d", l["value"], l["ms_per_step"], l["phases_ms"], l["verify"]["imbalance"], l["e2e"]["value"])
except Exception as e: print("no line", e)
PY
done
