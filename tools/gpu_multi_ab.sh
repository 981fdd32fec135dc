#!/bin/bash
# usage: gpu_multi_ab.sh N [dists] -- exchange plan A/B at N GPUs: remote cursors (default) vs counted layout
N=${1:-2}
DISTS=${2:-uniform}
mkdir -p gpurun_out
for plan in ${PLANS:-auto counts}; do
for d in $DISTS; do
LIBSORT_EXCHANGE_PLAN=$plan timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 --dist $d --quick > gpurun_out/ab_${N}gpu_${d}_$plan.json 2> gpurun_out/ab_${N}gpu_${d}_$plan.err
python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/ab_${N}gpu_${d}_$plan.json").read().strip().splitlines()[-1])
    print("$plan $d", round(l["value"],1), round(l["ms_per_step"],3), {k:round(v,3) for k,v in l["phases_ms"].items() if k.endswith("_ms")}, l["verify"]["ok"], round(l["verify"]["imbalance"],4))
except Exception as e: print("no line", e)
PY
done
done
