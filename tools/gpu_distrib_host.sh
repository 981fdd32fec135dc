#!/bin/bash
# Single-process multi-GPU entry point (libsortDistribSort) on an N-GPU box: pytest cases, the C++ consumer,
# the worker's exact + full-size checks and timings.   gpurun --gpus N -- 'bash tools/gpu_distrib_host.sh'
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/dh_gpus.txt
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "distrib_sort_host_call or cxx_consumers" > gpurun_out/dh_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/dh_pytest.log
tail -3 gpurun_out/dh_pytest.log
timeout 600 python tests/distrib_host_worker.py ${1:-28} > gpurun_out/dh_worker.log 2>&1
echo "worker rc=$?" >> gpurun_out/dh_worker.log
tail -5 gpurun_out/dh_worker.log
timeout 300 gpu-radix-sort_b200/localTest/radix_sort bench 268435456 > gpurun_out/dh_cxx_bench.log 2>&1
tail -8 gpurun_out/dh_cxx_bench.log
