#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "splitter_scatter or stepwise or kernel_variants or distrib_sort_host" 2>&1 | tail -3
for v in v_walk v_new; do
  for P in 2 4 8 16; do
  echo "== $v P=$P"; LIBSORT_SO=$PWD/build/$v/libsort.so timeout 300 python tools/split_timing.py 28 $P 2>&1 | tail -4
  done
done | tee gpurun_out/split_variants.log
