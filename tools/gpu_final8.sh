#!/bin/bash
# One lean N-GPU confirmation run: single-exchange sort (uniform, zipf, all-equal), step-wise mode, libsortDistribSort.
N=${1:-8}
mkdir -p gpurun_out
run() { # name, extra args
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 10 --warmup 3 $2 > gpurun_out/bench_${N}gpu_$1.json 2> gpurun_out/bench_${N}gpu_$1.err
  grep -h '^{' gpurun_out/bench_${N}gpu_$1.json | python -c "
import json,sys
for l in sys.stdin:
    j=json.loads(l); print('$1', round(j['value'],1),'Gkeys/s', round(j['ms_per_step'],2),'ms', {k: round(v,3) for k,v in j['phases_ms'].items() if v is not None}, j['verify'].get('imbalance'), j['nvlink'].get('gbs_per_direction'), (j.get('e2e') or {}).get('value'))" || tail -5 gpurun_out/bench_${N}gpu_$1.err
}
run uniform "--dist uniform"
run zipf "--dist zipf"
run allequal "--dist allequal"
run stepwise_uniform "--dist uniform --algo stepwise"
timeout 600 python tests/distrib_host_worker.py 27 > gpurun_out/dh_worker_${N}gpu.log 2>&1; echo "host worker rc=$?"; tail -2 gpurun_out/dh_worker_${N}gpu.log
