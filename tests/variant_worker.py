"""Run by test_gpu_parity.test_kernel_variants in a fresh process with LIBSORT_* set: a compact
parity sweep (C ABI vs the oracle) so that every compiled scatter kernel is exercised --
atomic / ballot ranking, LDS+STG / bulk-copy write-out, the heavy-tile path."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("gpu-radix-sort_b200")
from oracle import oracle as o  # noqa: E402

pkg.binding.init_once()
lib = pkg.lib()


def cases(n):
    x = o.pcg_stream(n)
    yield "uniform", x
    yield "consthigh", (np.uint32(0xAB000000) | (x & np.uint32(0x00FFFFFF))).astype(np.uint32)
    yield "zipf", (np.floor(2.0 ** (32.0 * x.astype(np.float64) / 2.0 ** 32)) - 1).astype(np.uint32)
    yield "allequal", np.full(n, 0x12345678, dtype=np.uint32)
    yield "two_values", np.where(x & 1, np.uint32(0x80000001), np.uint32(7)).astype(np.uint32)
    yield "sorted", np.sort(x)


bad = 0
for n in (1, 4099, 14336, 14337, 200003, (1 << 20) + 5, (1 << 22) + 4099):   # the last one is past the small-input path (2^22 keys)
    for name, k in cases(n):
        got = k.copy()
        assert lib.providedGpu(got.ctypes.data, n)
        if not np.array_equal(got, np.sort(k)):
            print("FAIL full", name, n)
            bad += 1
        for off, w in ((0, 8), (24, 8), (8, 16), (4, 10)):
            got = k.copy()
            b = np.zeros(1 << w, np.uint32)
            assert lib.gpuPartial(got.ctypes.data, b.ctypes.data, n, off, w)
            want, wb = o.gpu_partial(k, off, w)
            if not (np.array_equal(got, want) and np.array_equal(b, wb)):
                print("FAIL partial", name, n, off, w)
                bad += 1
print("variant sweep done, failures:", bad)
sys.exit(1 if bad else 0)
