#!/bin/bash
# Step-wise distributed LSD sort + file arrays / FaaS worker on an N-GPU box.   gpurun --gpus N -- 'bash tools/gpu_stepwise.sh N'
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_distrib.py tests/test_distrib_array.py -q -m gpu \
  -k "stepwise or distributed_sort_on_gpus or faas or direct_invoke or splitter_scatter" > gpurun_out/step_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/step_pytest.log
tail -25 gpurun_out/step_pytest.log
for d in uniform zipf allequal; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 --dist $d --algo stepwise > gpurun_out/bench_${N}gpu_stepwise_$d.json 2> gpurun_out/bench_${N}gpu_stepwise_$d.err
grep -h '^{' gpurun_out/bench_${N}gpu_stepwise_$d.json | python -c "
import json,sys
for l in sys.stdin:
    j=json.loads(l); print('$d', round(j['value'],1),'Gkeys/s', round(j['ms_per_step'],2),'ms', j['phases_ms'], j['verify']['ok'], j['nvlink']['gbs_per_direction'])" || tail -5 gpurun_out/bench_${N}gpu_stepwise_$d.err
done
