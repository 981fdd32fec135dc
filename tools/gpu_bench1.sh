#!/bin/bash
# usage: gpu_bench1.sh [sweep] -- the single-GPU bench line (with extras); "sweep": pageable-copy thread sweep too
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?"
tail -3 gpurun_out/bench1.err
python - <<'PY'
import json
l=json.loads(open("gpurun_out/bench1.json").read().strip().splitlines()[-1])
print("value", l["value"], "ms", l["ms_per_step"], "frac", l["roofline"]["frac"], "e2e", l["e2e"]["value"])
print(json.dumps(l["extras"], indent=1)[:7000])
PY
if [ "$1" = "sweep" ]; then
for t in 4 8 12 16; do for c in 2 4 8; do
LIBSORT_COPY_THREADS=$t LIBSORT_COPY_CHUNK_MB=$c python - <<'PY'
import importlib, numpy as np, time, os
pkg = importlib.import_module("gpu-radix-sort_b200"); pkg.binding.init_once(); lib = pkg.lib()
n = 1 << 28
a = np.empty(n, np.uint32); lib.populateInput(a.ctypes.data, n); w = a.copy()
lib.providedGpu(w.ctypes.data, n)
ts = []
for _ in range(3):
    w[:] = a; t0 = time.perf_counter(); lib.providedGpu(w.ctypes.data, n); ts.append(time.perf_counter() - t0)
print("threads", os.environ["LIBSORT_COPY_THREADS"], "chunkMB", os.environ["LIBSORT_COPY_CHUNK_MB"], "pageable providedGpu ms", [round(x * 1e3, 1) for x in ts])
PY
done; done
fi
