#!/bin/bash
# usage: gpu_multi_ab.sh N v_a v_b ... -- distributed bench (uniform) once per build variant under build/
N=${1:-2}; shift
mkdir -p gpurun_out
for v in "$@"; do
LIBSORT_SO=$PWD/build/$v/libsort.so timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${N}gpu_$v.json 2> gpurun_out/bench_${N}gpu_$v.err
grep -h '^{' gpurun_out/bench_${N}gpu_$v.json | python -c "
import json,sys
for l in sys.stdin:
    j=json.loads(l); print('$v', round(j['value'],1),'Gkeys/s', round(j['ms_per_step'],2),'ms', {k: round(x,3) for k,x in j['phases_ms'].items() if x is not None}, j['verify']['ok'])" || tail -3 gpurun_out/bench_${N}gpu_$v.err
done
