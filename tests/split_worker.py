"""Run by test_gpu_parity.test_splitter_scatter_single_gpu in a fresh process (LIBSORT_BULK_EXCHANGE set):
the distributed sort's device building blocks on ONE GPU -- counts per destination and the stable
splitter scatter into per-destination regions of a local buffer -- for 2..16 destinations, against numpy."""
import ctypes
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("gpu-radix-sort_b200")
from oracle import oracle as o  # noqa: E402

plan = importlib.import_module("gpu-radix-sort_b200.distrib_plan")
pkg.binding.init_once()
L = pkg.lib()
torch.cuda.set_device(0)
bad = 0
for n in (1, 8191, 100003, 3000001):
    x = o.pcg_stream(n)
    inputs = {"uniform": x, "lowbits": (x & np.uint32(0xFFFF)).astype(np.uint32), "const": np.full(n, 77, np.uint32),
              # ties: splitters and thresholds picked on the device from a sorted sample of heavily repeated values
              "ties_const": np.full(n, 0xABCD1234, np.uint32), "ties_few": (x % np.uint32(3) * np.uint32(1000)).astype(np.uint32),
              "ties_ones": np.where(x & 1, np.uint32(0xFFFFFFFF), np.uint32(5)).astype(np.uint32)}
    for name, k in inputs.items():
        for nranks in (2, 3, 4, 8, 13, 16):
            ns = nranks - 1
            spl = np.sort(o.pcg_stream(64)[17:17 + ns]).astype(np.uint32)
            if name == "lowbits":
                spl = (spl >> 16).astype(np.uint32)
            thr = np.zeros(ns, np.uint32)
            d_keys = torch.from_numpy(k.view(np.int32)).cuda()
            d_spl = torch.zeros(32, dtype=torch.int32, device="cuda")
            d_spl[:ns] = torch.from_numpy(spl.view(np.int32)).cuda()
            if name.startswith("ties"):
                sample = np.sort(k[:: max(1, n // 4096)])
                d_sample = torch.from_numpy(sample.view(np.int32)).cuda()
                pkg.binding.check(L.libsortDevicePickSplitters(d_sample.data_ptr(), sample.size, nranks, d_spl.data_ptr(),
                                                               torch.cuda.current_stream().cuda_stream), "pick")
                table = d_spl.cpu().numpy().view(np.uint32)
                spl, thr = table[:ns].copy(), table[16:16 + ns].copy()
                wspl, wthr = plan.choose_splitters(sample, nranks, with_ties=True)
                if not (np.array_equal(spl, wspl) and np.array_equal(thr, wthr)):
                    print("FAIL pick_splitters", name, n, nranks, spl, thr, wspl, wthr)
                    bad += 1
                    continue
            d_cnt = torch.zeros(16, dtype=torch.int32, device="cuda")
            st = torch.cuda.current_stream().cuda_stream
            pkg.binding.check(L.libsortDeviceSplitCounts(d_keys.data_ptr(), n, d_spl.data_ptr(), ns, d_cnt.data_ptr(), st), "counts")
            dest = plan.destination(k, spl, thr)                    # number of splitters <= key, ties cut by position
            want_cnt = np.bincount(dest, minlength=16)[:16]
            got_cnt = d_cnt.cpu().numpy()
            if not np.array_equal(got_cnt, want_cnt):
                print("FAIL counts", name, n, nranks, got_cnt, want_cnt)
                bad += 1
                continue
            if name == "ties_const" and n > 100000 and want_cnt[:nranks].max() > 1.02 * n / nranks:
                print("FAIL all-equal input does not balance", n, nranks, want_cnt)
                bad += 1
            # destination regions: consecutive, each shifted by a few keys so that all 16-byte phases occur
            offs = np.cumsum(np.concatenate([[0], want_cnt[:-1] + 5])) + np.arange(16) % 4
            out = torch.full((int(offs[-1] + want_cnt[-1] + 16),), -1, dtype=torch.int32, device="cuda")
            ptrs = (ctypes.c_uint64 * nranks)(*[out.data_ptr() + 4 * int(offs[r]) for r in range(nranks)])
            ws = torch.empty(L.libsortWorkspaceBytes(n), dtype=torch.uint8, device="cuda")
            pkg.binding.check(L.libsortDeviceScatterToPeers(d_keys.data_ptr(), n, d_spl.data_ptr(), ns, ptrs, ws.data_ptr(),
                                                            ws.numel(), st), "scatter")
            torch.cuda.synchronize()
            got = out.cpu().numpy().view(np.uint32)
            for r in range(nranks):
                seg = got[offs[r]:offs[r] + want_cnt[r]]
                if not np.array_equal(seg, k[dest == r]):      # stable: input order kept inside a destination
                    print("FAIL scatter", name, n, nranks, "dest", r)
                    bad += 1
                    break
            used = np.zeros(got.size, bool)
            for r in range(nranks):
                used[offs[r]:offs[r] + want_cnt[r]] = True
            if not np.all(got[~used] == 0xFFFFFFFF):
                print("FAIL scatter wrote outside its regions", name, n, nranks)
                bad += 1
print("split sweep done, failures:", bad)
sys.exit(1 if bad else 0)
