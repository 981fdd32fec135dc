#!/usr/bin/env python3
"""The exchange kernel alone on ONE GPU: W virtual ranks, every destination region local (no NVLink).
Used under ncu to get the LSU / stall breakdown of k_onesweep_bulk<SplitBinner, PTR_OUT, local cursors>
(a multi-rank command must not run under ncu).  usage: exchange_local.py [log2 keys] [W] [reps]"""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("gpu-radix-sort_b200")


def main():
    lg = int(sys.argv[1]) if len(sys.argv) > 1 else 28
    W = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    pkg.binding.init_once()
    L = pkg.lib()
    n = 1 << lg
    g = torch.Generator(device="cuda")
    g.manual_seed(5)
    keys = torch.randint(-2**31, 2**31 - 1, (n,), dtype=torch.int32, device="cuda", generator=g)
    spl = np.zeros(32, np.uint32)
    for r in range(W - 1):
        spl[r] = ((r + 1) << 32) // W
    spl_signed = spl.view(np.int32)
    d_spl = torch.from_numpy(spl_signed.copy()).cuda()
    cap = (n // W + n // (4 * W) + 4096) & ~3
    recv = torch.empty((W, cap), dtype=torch.int32, device="cuda")
    bases = torch.tensor([recv[r].data_ptr() for r in range(W)], dtype=torch.int64, device="cuda")
    cm = torch.zeros((W, 16), dtype=torch.int32, device="cuda")
    assert L.libsortDeviceSplitCounts(keys.data_ptr(), n, d_spl.data_ptr(), W - 1, cm[0].data_ptr(), None) == 0
    ws = torch.empty(L.libsortWorkspaceBytes(n), dtype=torch.uint8, device="cuda")
    nrecv = torch.zeros(1, dtype=torch.int32, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(reps):
        e0.record()
        assert L.libsortDeviceExchangePlanned(keys.data_ptr(), n, d_spl.data_ptr(), W, 0, cm.data_ptr(), bases.data_ptr(), cap,
                                              nrecv.data_ptr(), status.data_ptr(), ws.data_ptr(), ws.numel(), None) == 0
        e1.record()
        torch.cuda.synchronize()
        print("exchange (memset + plan + kernel), %d destinations, 2^%d keys: %.3f ms" % (W, lg, e0.elapsed_time(e1)))
    total = int(cm[0, :W].sum().item())
    got = sum(int((recv[r, :int(cm[0, r].item())] != 0).numel()) for r in range(W))
    print("status", int(status.item()), "keys placed", got, "of", total)


if __name__ == "__main__":
    main()
