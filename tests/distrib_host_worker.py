"""libsortDistribSort on every visible GPU from ONE process: correctness (ascending + multiset
fingerprint against the input, exact equality with numpy at the small size) and wall-clock time of the
host-pointer call with pinned memory.  Run on a multi-GPU box: python tests/distrib_host_worker.py [log2 keys per GPU]."""
import ctypes
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
binding = importlib.import_module("gpu-radix-sort_b200.binding")


def main():
    lg = int(sys.argv[1]) if len(sys.argv) > 1 else 28
    L = binding.lib()
    assert L.initLibSort() == 1
    P = L.libsortDeviceCount()
    res = {"gpus": P}
    # exact check, ragged size
    n = (1 << 22) * P + 12345
    k = np.empty(n, dtype=np.uint32)
    L.libsortPopulateInputAt(k.ctypes.data, n, 7)
    k[::5] &= np.uint32(0x0000FFFF)
    got = k.copy()
    assert L.libsortDistribSort(got.ctypes.data, n, 0) == 1
    assert np.array_equal(got, np.sort(k)), "small distributed sort differs from numpy"
    res["exact_small"] = True
    # heavy duplicates: one value in half of the keys, then every key equal (ties are cut between GPUs)
    for name, kk in (("halfdup", np.where(k & 1, np.uint32(0x40000000), k).astype(np.uint32)), ("allequal", np.full(n, 99, np.uint32))):
        got = kk.copy()
        assert L.libsortDistribSort(got.ctypes.data, n, 0) == 1
        assert np.array_equal(got, np.sort(kk)), name
    res["exact_duplicates"] = True
    # full size, pinned host memory
    n = (1 << lg) * P
    h = torch.empty(n, dtype=torch.int32).pin_memory()
    a = h.numpy().view(np.uint32)
    L.libsortPopulateInputAt(a.ctypes.data, n, 0)
    fp_in = (int(a.sum(dtype=np.uint64)), int(np.bitwise_xor.reduce(a)))
    orig = a.copy()
    times = []
    for it in range(4):
        a[:] = orig
        t0 = time.perf_counter()
        rc = L.libsortDistribSort(a.ctypes.data, n, 0)
        times.append(time.perf_counter() - t0)
        assert rc == 1
    assert bool(np.all(a[1:] >= a[:-1])), "not ascending"
    assert (int(a.sum(dtype=np.uint64)), int(np.bitwise_xor.reduce(a))) == fp_in, "multiset changed"
    best = min(times[1:])
    res.update({"keys": n, "ms": [round(t * 1e3, 2) for t in times], "e2e_gkeys_s": round(n / best / 1e9, 2),
                "host_bytes_moved_gbs": round(8.0 * n / best / 1e9, 1)})
    print(json.dumps(res))


if __name__ == "__main__":
    main()
