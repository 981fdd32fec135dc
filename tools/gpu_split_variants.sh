#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "splitter_scatter or distrib_sort_host" 2>&1 | tail -3
for v in ${VARIANTS:-v_hyb}; do
  for P in 4 8 16; do
  echo "== $v P=$P"; LIBSORT_SO=$PWD/build/$v/libsort.so timeout 300 python tools/split_timing.py 28 $P 2>&1 | tail -4
  done
done | tee gpurun_out/split_variants_$$.log
