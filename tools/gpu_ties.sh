#!/bin/bash
# Tie splitting on an N-GPU box: single-GPU splitter sweep (both exchange kernels), the NCCL/NVLink worker,
# libsortDistribSort with heavy duplicates, then bench lines.   gpurun --gpus N -- 'bash tools/gpu_ties.sh N'
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_distrib.py -q -m gpu -k "splitter_scatter or distributed_sort_on_gpus or distrib_sort_host_call" > gpurun_out/ties_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/ties_pytest.log
tail -15 gpurun_out/ties_pytest.log
timeout 600 python tests/distrib_host_worker.py 26 > gpurun_out/ties_host_worker.log 2>&1
echo "worker rc=$?" >> gpurun_out/ties_host_worker.log
tail -5 gpurun_out/ties_host_worker.log
bash tools/gpu_multi_lite.sh $N > gpurun_out/ties_bench.log 2>&1
for d in uniform consthigh zipf halfdup allequal; do python - <<PY
import json
try:
    j=[l for l in open("gpurun_out/bench_${N}gpu_$d.json") if l.startswith("{")][-1]; j=json.loads(j)
    print("$d", round(j["value"],1), "Gkeys/s", round(j["ms_per_step"],2), "ms", j["phases_ms"], j["verify"])
except Exception as e:
    print("$d", "no line", e); print(open("gpurun_out/bench_${N}gpu_$d.err").read()[-1500:])
PY
done
