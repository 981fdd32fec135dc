"""Time the distributed sort's count and partition kernels on ONE GPU (all destinations local): no NVLink in the
way, so this is the kernels' own throughput.  usage: python tools/split_timing.py [log2 n] [nranks]"""
import ctypes
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("gpu-radix-sort_b200")
distrib = importlib.import_module("gpu-radix-sort_b200.distrib")
pkg.binding.init_once()
L = pkg.lib()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 28
P = int(sys.argv[2]) if len(sys.argv) > 2 else 8
n = 1 << lg
st = torch.cuda.current_stream().cuda_stream
S = 1 << 15
for kind in ("uniform", "zipf", "halfdup", "allequal"):
    keys = distrib.shard_of_pcg_stream(0, n, kind)
    i32 = torch.int32
    samples, ssorted, stmp = (torch.empty(S * P, dtype=i32, device="cuda") for _ in range(3))
    for r in range(P):
        pkg.binding.check(L.libsortDeviceSample(keys.data_ptr(), n, samples.data_ptr() + 4 * S * r, S, 7 + r, st), "sample")
    ws_s = torch.empty(L.libsortWorkspaceBytes(S * P), dtype=torch.uint8, device="cuda")
    pkg.binding.check(L.libsortDeviceSort(samples.data_ptr(), ssorted.data_ptr(), stmp.data_ptr(), S * P, 0, 32, None, 0,
                                          ws_s.data_ptr(), ws_s.numel(), st, None), "sort")
    spl = torch.zeros(32, dtype=i32, device="cuda")
    pkg.binding.check(L.libsortDevicePickSplitters(ssorted.data_ptr(), S * P, P, spl.data_ptr(), st), "pick")
    cnt = torch.zeros(16, dtype=i32, device="cuda")
    out = torch.empty(n + 64, dtype=i32, device="cuda")
    ws = torch.empty(L.libsortWorkspaceBytes(n), dtype=torch.uint8, device="cuda")

    def counts():
        pkg.binding.check(L.libsortDeviceSplitCounts(keys.data_ptr(), n, spl.data_ptr(), P - 1, cnt.data_ptr(), st), "counts")
    counts()
    c = cnt.cpu().numpy()[:P].astype(np.int64)
    offs = np.cumsum(c) - c
    ptrs = (ctypes.c_uint64 * P)(*[out.data_ptr() + 4 * int(o) for o in offs])

    def scatter():
        pkg.binding.check(L.libsortDeviceScatterToPeers(keys.data_ptr(), n, spl.data_ptr(), P - 1, ptrs, ws.data_ptr(), ws.numel(), st),
                          "scatter")
    res = {}
    for name, fn in (("counts", counts), ("scatter", scatter)):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / 5
    thr = spl.cpu().numpy().view(np.uint32)[16:16 + P - 1]
    print("%-9s P=%d counts %.3f ms  scatter %.3f ms  ties=%s  max/mean=%.4f" % (kind, P, res["counts"], res["scatter"], bool(thr.any()),
                                                                                 c.max() / c.mean()), flush=True)
