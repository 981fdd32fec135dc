#!/usr/bin/env python3
"""bench.py -- headline benchmark of the B200-native libsort.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N = 1: BASELINE.json configs[1] -- full LSD sort of 2^28 uniform uint32 keys
       (first 2^28 outputs of the library's PCG stream) on one B200.
N > 1: the distributed sort (configs[3] shape, weak scaling): 2^28 keys per GPU,
       one rank per GPU under torchrun, MSD bucketing -> count all-gather ->
       key exchange over NVLink -> local LSD sort.

One JSON line on stdout (rank 0).  `value` is device-resident whole-job
Gkeys/s; `e2e` is the same sort through the reference-facing C ABI
(providedGpu, host buffers, H2D + D2H inside the timed region).
`--impl reference` times the reference's own CPU implementation of the path
(providedCpu = single-threaded std::sort, libsort/invokers.cu:68-71) from
oracle/_ref (or the oracle port) on a bounded sample of the same workload.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_PER_GPU = 1 << 28
METRIC = "Gkeys/s uint32 sort"
UNIT = "Gkeys/s"


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed regions (device-timed loop, per-kernel
    timing loop and the end-to-end loop, all of them the same sort back to back), every 25 ms."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "25"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def mark(self):
        """Number of samples taken so far (to report a sub-region separately)."""
        return len(self.lines)

    def stop(self, first=0, loop_end=None):
        """Summary of the samples from index `first` on; with loop_end, the samples [first, loop_end) -- taken
        while the device-timed loop ran -- are summarised separately under "device_loop"."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        self.t.join(2)
        out = self._summary(self.lines[first:])
        if loop_end is not None:
            out["device_loop"] = {k: v for k, v in self._summary(self.lines[first:loop_end]).items()
                                  if k in ("sm_mhz", "samples", "reasons")}
        return out

    @staticmethod
    def _summary(lines):
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                mx = float(p[1])
            except ValueError:
                continue
            for nm, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "samples": len(sm),
                "reasons": sorted(reasons)}


def cpu_reference_sorter():
    """(callable(np.uint32 array) -> None in place, kind)"""
    from oracle import oracle as o
    import ctypes
    if o.ref_available():
        try:
            r = o.ref()
            u32p = ctypes.POINTER(ctypes.c_uint32)
            return (lambda a: r.providedCpu(a.ctypes.data_as(u32p), a.size)), "reference"
        except OSError:
            pass
    L = o.lib()
    u32p = ctypes.POINTER(ctypes.c_uint32)
    return (lambda a: L.oracle_cpu_sort(a.ctypes.data_as(u32p), a.size)), "port"


def pcg_keys(n):
    """First n outputs of populateInput in a fresh process."""
    import numpy as np
    pkg = importlib.import_module("gpu-radix-sort_b200")
    a = np.empty(n, dtype=np.uint32)
    pkg.lib().populateInput(a.ctypes.data, n)
    return a


def time_cpu(sample_keys, reps):
    import numpy as np
    sorter, kind = cpu_reference_sorter()
    best = []
    for _ in range(reps):
        a = sample_keys.copy()
        t0 = time.perf_counter()
        sorter(a)
        best.append(time.perf_counter() - t0)
    assert bool(np.all(a[1:] >= a[:-1]))
    return sum(best) / len(best), kind


def run_reference(args, rank, world):
    """The reference's CPU path on the host cores; rank 0 only."""
    if rank != 0:
        return
    import numpy as np
    sample_n = 1 << 24   # bounded sample of the 2^28-key workload: ~1.5 s of std::sort per step
    keys = np.empty(sample_n, dtype=np.uint32)
    from oracle import oracle as o
    keys[:] = o.pcg_stream(sample_n)
    sorter, kind = cpu_reference_sorter()
    for _ in range(min(args.warmup, 1)):
        sorter(keys.copy())
    t = 0.0
    for _ in range(args.steps):
        a = keys.copy()
        t0 = time.perf_counter()
        sorter(a)
        t += time.perf_counter() - t0
    val = sample_n * args.steps / t / 1e9
    sample = "first 2^24 keys of the PCG stream per step (of the 2^28-key workload), std::sort, 1 thread"
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "256M uniform uint32 full sort (BASELINE configs[1])", "sample_keys": sample_n},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def _u(t):
    """uint32 view of an int32 CUDA tensor as int64 values (torch has no unsigned compare)."""
    return t.to(__import__("torch").int64) & 0xFFFFFFFF


def check_partial_on_device(torch, d_in, d_out, bounds, offset, width):
    """Size-independent checks of a partial sort at full size, on the device: groups ascend,
    boundaries are the exclusive prefix sums of the group histogram (localTest/tests.cpp:44-68),
    the multiset is preserved.  (Stability has its own leg: stability_leg.)"""
    n = d_out.numel()
    mask = (1 << width) - 1
    g_out = (_u(d_out) >> offset) & mask
    ok_groups = bool((g_out[1:] >= g_out[:-1]).all())
    del g_out
    g_in = ((_u(d_in) >> offset) & mask)
    hist = torch.bincount(g_in, minlength=1 << width)
    del g_in
    excl = torch.cumsum(hist, 0) - hist
    ok_bounds = bool((excl == _u(bounds)).all())
    a, b = _u(d_in), _u(d_out)
    ok_multiset = int(a.sum()) == int(b.sum()) and int((a * a % 1000003).sum()) == int((b * b % 1000003).sum())
    return {"groups_ascending": ok_groups, "boundaries_are_prefix_sums": ok_bounds, "multiset_preserved": ok_multiset,
            "keys": n}


def stability_leg(torch, dev, binding, width, n):
    """Full-size stability check of the default ranking path: keys (i << width) | h(i), partial sort
    on (0, width) -- inside every bucket the keys must still ascend, i.e. keep their input order
    (reference contract: stable LSD passes, libsort/sort.cu:322-346)."""
    i = torch.arange(n, dtype=torch.int64, device="cuda")
    h = (i * 2654435761 >> 7) & ((1 << width) - 1)
    keys64 = ((i << width) | h) & 0xFFFFFFFF
    keys = torch.where(keys64 >= (1 << 31), keys64 - (1 << 32), keys64).to(torch.int32)
    del i, h, keys64
    out = dev.sort_bits(keys, 0, width, boundary_mode=binding.BOUNDS_EXACT)
    o = _u(out)
    g = o & ((1 << width) - 1)
    same = g[1:] == g[:-1]
    ok = bool(((o[1:] > o[:-1]) | ~same).all()) and bool((g[1:] >= g[:-1]).all())
    return {"width": width, "keys": n, "stable": ok}


def single_gpu_extras(args, torch, np, pkg, dev, binding, h_keys, peak):
    """Everything BASELINE.json names besides the headline config, measured in the same run
    (extras never touch the headline keys; a failing extra is reported, not fatal)."""
    import ctypes
    lib = pkg.lib()
    ex = {}

    def timed(fn, reps=5, warm=2):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    # c1: localTest ./radix_sort -- 1M keys, GPU full sort == CPU sort (BASELINE configs[0])
    try:
        n1 = 1 << 20
        k1 = torch.from_numpy(h_keys[:n1].view(np.int32)).cuda()
        o1, t1 = torch.empty_like(k1), torch.empty_like(k1)
        ms1 = timed(lambda: dev.sort(k1, out=o1, tmp=t1), reps=20)
        want = np.sort(h_keys[:n1])
        ex["c1_localtest_1M"] = {"keys": n1, "ms": ms1, "gkeys_s": n1 / ms1 / 1e6,
                                 "equals_cpu_sort": bool(np.array_equal(o1.cpu().numpy().view(np.uint32), want))}
    except Exception as e:  # noqa: BLE001
        ex["c1_localtest_1M"] = {"error": repr(e)}

    # c3: partial sorts of 2^29 keys with boundaries (BASELINE configs[2]; analysis/libsort8b,16b.csv)
    try:
        n3 = 1 << 29
        h3 = np.empty(n3, dtype=np.uint32)
        lib.libsortPopulateInputAt(h3.ctypes.data, n3, 0)
        k3 = torch.from_numpy(h3.view(np.int32)).cuda()
        del h3
        o3, t3 = torch.empty_like(k3), torch.empty_like(k3)
        rows = []
        for off, w in ((0, 8), (0, 16), (24, 8), (16, 16)):
            kt = binding.KernelTimes()
            _, bnd = dev.sort_bits(k3, off, w, out=o3, tmp=t3, boundaries=True, boundary_mode=binding.BOUNDS_EXACT, times=kt)
            ms3 = timed(lambda: dev.sort_bits(k3, off, w, out=o3, tmp=t3, boundaries=True, boundary_mode=binding.BOUNDS_EXACT),
                        reps=5, warm=1)
            passes = [kt.pass_ms[i] for i in range(kt.npass)]
            nb = 4 + 8 * len(passes)
            row = {"offset": off, "width": w, "keys": n3, "ms": ms3, "gkeys_s": n3 / ms3 / 1e6,
                   "algorithmic_bytes_per_key": nb, "gbs": nb * n3 / ms3 / 1e6, "frac_of_peak": nb * n3 / ms3 / 1e6 / peak,
                   "pass_ms": passes, "pass_frac_of_peak": [8.0 * n3 / (p * 1e-3) / 1e9 / peak for p in passes],
                   "histogram_ms": kt.histogram_ms}
            row.update(check_partial_on_device(torch, k3, o3, bnd, off, w))
            rows.append(row)
        ex["c3_partial_512M"] = rows
        del k3, o3, t3
    except Exception as e:  # noqa: BLE001
        ex["c3_partial_512M"] = {"error": repr(e)}
    torch.cuda.empty_cache()

    # stability of the default (atomic-ranking) path at full size, and the per-tile audit flag
    try:
        ex["stability"] = [stability_leg(torch, dev, binding, 4, 1 << 28), stability_leg(torch, dev, binding, 8, 1 << 24)]
        ex["rank_audit_violations"] = int(lib.libsortRankAudit(torch.cuda.current_stream().cuda_stream))
    except Exception as e:  # noqa: BLE001
        ex["stability"] = {"error": repr(e)}
    torch.cuda.empty_cache()
    return ex


def host_side_extras(args, np, pkg, h_keys):
    """Host-pointer legs: pageable e2e (what every reference caller passes), the CPU baseline at the
    full 2^28 keys (one run), the reference's own GPU path on this box (the checker, timed alone)."""
    lib = pkg.lib()
    n = h_keys.size
    ex = {}
    try:
        work = h_keys.copy()            # ordinary pageable memory, like Go/Python/C++ callers hand in
        assert lib.providedGpu(work.ctypes.data, n)
        ts = []
        for _ in range(3):
            work[:] = h_keys
            t0 = time.perf_counter()
            ok = lib.providedGpu(work.ctypes.data, n)
            ts.append(time.perf_counter() - t0)
            assert ok
        assert bool(np.all(work[1:] >= work[:-1]))
        ex["e2e_pageable"] = {"value": n / (sum(ts) / len(ts)) / 1e9, "unit": UNIT, "ms_per_step": sum(ts) / len(ts) * 1e3,
                              "h2d_bytes_per_step": 4 * n, "d2h_bytes_per_step": 4 * n,
                              "api": "providedGpu (host pointers, pageable numpy memory)"}
        b = np.zeros(1 << 16, np.uint32)
        for (off, w) in ((0, 8), (0, 16)):
            work[:] = h_keys
            t0 = time.perf_counter()
            ok = lib.gpuPartial(work.ctypes.data, b.ctypes.data, n, off, w)
            ex["gpuPartial_%d_%d_pageable_ms" % (off, w)] = (time.perf_counter() - t0) * 1e3
            assert ok
        del work
    except Exception as e:  # noqa: BLE001
        ex["e2e_pageable"] = {"error": repr(e)}
    if not args.quick:
        try:
            s, kind = time_cpu(h_keys, 1)
            ex["cpu_baseline_full"] = {"value": n / s / 1e9, "unit": UNIT, "cores": 1, "kind": kind,
                                       "sample": "the whole 2^28-key workload, one std::sort run (%.1f s)" % s}
        except Exception as e:  # noqa: BLE001
            ex["cpu_baseline_full"] = {"error": repr(e)}
        try:
            ex["reference_gpu_same_box"] = reference_gpu_same_box(np, h_keys)
        except Exception as e:  # noqa: BLE001
            ex["reference_gpu_same_box"] = {"error": repr(e)}
    return ex


def reference_gpu_same_box(np, h_keys):
    """The reference's unmodified GPU path (oracle/_ref/libsort_ref.so, compiled for sm_100a) on the
    same B200 and the same keys: wall time of providedGpu and gpuPartial(0,8 | 0,16) through its
    host-pointer ABI (reference libsort/invokers.cu:15-64).  The checker's timing, reported beside
    the product's pageable e2e -- never inside a timed region of the product."""
    import ctypes
    from oracle import oracle as o
    if not o.ref_available():
        return {"unavailable": "oracle/_ref/libsort_ref.so not built"}
    r = o.ref()
    u32p = ctypes.POINTER(ctypes.c_uint32)
    if not getattr(r, "_bench_inited", False):
        if not r.initLibSort():
            return {"unavailable": "reference initLibSort failed"}
        r._bench_inited = True
    n = h_keys.size
    out = {"keys": n, "what": "unmodified reference libsort on this GPU, host-pointer ABI, pageable memory, wall clock"}
    work = h_keys.copy()
    r.providedGpu(work[: 1 << 20].ctypes.data_as(u32p), 1 << 20)   # context + first-call costs out of the way
    work[:] = h_keys
    t0 = time.perf_counter()
    ok = r.providedGpu(work.ctypes.data_as(u32p), n)
    out["providedGpu_ms"] = (time.perf_counter() - t0) * 1e3
    out["providedGpu_ok"] = bool(ok) and bool(np.all(work[1:] >= work[:-1]))
    out["providedGpu_gkeys_s"] = n / out["providedGpu_ms"] / 1e6
    b = np.zeros(1 << 16, np.uint32)
    for (off, w) in ((0, 8), (0, 16)):
        work[:] = h_keys
        t0 = time.perf_counter()
        ok = r.gpuPartial(work.ctypes.data_as(u32p), b.ctypes.data_as(u32p), n, off, w)
        out["gpuPartial_%d_%d_ms" % (off, w)] = (time.perf_counter() - t0) * 1e3
        out["gpuPartial_%d_%d_ok" % (off, w)] = bool(ok)
    return out


def run_single(args):
    import numpy as np
    import torch
    pkg = importlib.import_module("gpu-radix-sort_b200")
    dev = importlib.import_module("gpu-radix-sort_b200.device")
    binding = pkg.binding
    binding.init_once()
    torch.cuda.set_device(0)
    clocks = ClockSampler(0)
    clocks.start()                      # nvidia-smi needs ~0.1 s before its first sample: started ahead of the set-up
    n = args.keys
    h_keys = pcg_keys(n)
    d_in = torch.from_numpy(h_keys.view(np.int32)).cuda()
    d_out, d_tmp = torch.empty_like(d_in), torch.empty_like(d_in)

    for _ in range(max(args.warmup, 3)):
        dev.sort(d_in, out=d_out, tmp=d_tmp)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    clk_first = clocks.mark()
    e0.record()
    for _ in range(args.steps):
        dev.sort(d_in, out=d_out, tmp=d_tmp)   # input (1 GiB) is pristine every step and larger than L2
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    clk_mark = clocks.mark()

    # sortedness + multiset check of the last timed result (size-independent properties)
    out_u = d_out.view(torch.uint8).view(-1, 4)  # avoid signed compare: check via int64
    o64 = d_out.to(torch.int64) & 0xFFFFFFFF
    assert bool((o64[1:] >= o64[:-1]).all()), "device sort output not ascending"
    i64 = d_in.to(torch.int64) & 0xFFFFFFFF
    assert int(o64.sum()) == int(i64.sum()) and int((o64 * o64 % 1000003).sum()) == int((i64 * i64 % 1000003).sum())
    del o64, i64, out_u

    # per-kernel durations, CUDA events on the launching stream (live, same process)
    kt_acc, reps = None, 5
    for _ in range(reps):
        kt = binding.KernelTimes()
        dev.sort(d_in, out=d_out, tmp=d_tmp, times=kt)
        vals = [kt.memset_ms, kt.histogram_ms, kt.scan_ms] + [kt.pass_ms[i] for i in range(kt.npass)]
        kt_acc = vals if kt_acc is None else [a + b for a, b in zip(kt_acc, vals)]
    kt_avg = [v / reps for v in kt_acc]
    pass_ms = sum(kt_avg[3:]) / len(kt_avg[3:])
    peak, peak_src = measured_peak()
    achieved = 8.0 * n / (pass_ms * 1e-3) / 1e9
    traffic = None
    tfile = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tfile):
        try:
            traffic = json.load(open(tfile)).get("k_onesweep_dram_bytes_per_launch")
        except Exception:  # noqa: BLE001
            traffic = None

    # e2e: reference-facing C ABI, pinned host buffers, copies inside the timed region
    lib = pkg.lib()
    h_src = torch.from_numpy(h_keys.view(np.int32)).pin_memory()
    h_work = torch.empty_like(h_src).pin_memory()
    e2e_steps = max(1, min(args.steps, 5))
    for _ in range(2):
        h_work.copy_(h_src)
        assert lib.providedGpu(h_work.data_ptr(), n)
    t_e2e = 0.0
    for _ in range(e2e_steps):
        h_work.copy_(h_src)
        t0 = time.perf_counter()
        ok = lib.providedGpu(h_work.data_ptr(), n)
        t_e2e += time.perf_counter() - t0
        assert ok
    hw = h_work.numpy().view(np.uint32)
    assert bool(np.all(hw[1:] >= hw[:-1]))
    e2e_val = n * e2e_steps / t_e2e / 1e9
    clk = clocks.stop(first=clk_first, loop_end=clk_mark)   # all timed regions; "device_loop": the device-timed loop alone

    cpu_n = 1 << 26
    cpu_s, cpu_kind = time_cpu(h_keys[:cpu_n], 3)
    extras = {}
    if n == N_PER_GPU:
        del d_out, d_tmp, h_work
        torch.cuda.empty_cache()
        extras = single_gpu_extras(args, torch, np, pkg, dev, binding, h_keys, peak)
        extras.update(host_side_extras(args, np, pkg, h_keys))
    line = {
        "metric": METRIC, "value": n / (ms * 1e-3) / 1e9, "unit": UNIT, "n_gpus": 1, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": "256M uniform uint32 full LSD sort, 8-bit digits, 4 passes (BASELINE configs[1])",
                   "keys": n, "input": "first n outputs of populateInput (PCG32)", "tile_keys": lib.libsortTileKeys(),
                   "l2": "input 1 GiB per step, larger than the 126 MB L2; not flushed between steps"},
        "clocks": clk,
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": 4 * n, "d2h_bytes_per_step": 4 * n,
                "ms_per_step": t_e2e / e2e_steps * 1e3, "steps": e2e_steps, "api": "providedGpu (host pointers, pinned)"},
        "gpu_launches": 6 * args.steps,
        "roofline": {"bound": "hbm", "kernel": "k_onesweep (one 8-bit scatter pass)", "achieved": achieved,
                     "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "algorithmic_bytes_per_launch": 8 * n, "avg_launch_ms": pass_ms},
        "kernels_ms": {"memset": kt_avg[0], "k_histogram": kt_avg[1], "k_scan_hist": kt_avg[2],
                       "k_onesweep": kt_avg[3:]},
        "whole_sort_roofline": {"algorithmic_bytes": 36 * n, "achieved": 36.0 * n / (ms * 1e-3) / 1e9,
                                "frac": 36.0 * n / (ms * 1e-3) / 1e9 / peak},
        "cpu_baseline": {"value": cpu_n / cpu_s / 1e9, "unit": UNIT, "cores": 1, "kind": cpu_kind,
                         "sample": "first 2^26 keys of the same PCG stream, mean of 3 std::sort runs (%.2f s each)" % cpu_s},
        "traffic_source": "ncu --set full capture of this kernel (profiles/traffic.json); not re-measured in this run",
        "extras": extras,
    }
    print(json.dumps(line), flush=True)


def multi_extras(args, rank, world, sorter, distrib, torch, dist, np):
    """BASELINE configs[4] in the same run: 2^30 skewed keys over the N GPUs (Gkeys/s + imbalance per
    distribution), and an element-exact check at reduced size against np.sort on rank 0."""
    ex = {}
    n5 = min((1 << 30) // world, sorter.n_max)    # 2^30 keys in all once N >= 4; N = 2 is capped by the shard size
    rows = {}
    for d in ("zipf", "consthigh", "consthigh16", "allequal"):
        try:
            k = distrib.shard_of_pcg_stream(rank, n5, d)
            sorter.reset_plan()
            for _ in range(2):
                sorter.sort(k)
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                out, _ = sorter.sort_async(k) if args.exchange == "p2p" else sorter.sort(k)
            e1.record()
            torch.cuda.synchronize()
            cnt = sorter._finish() if args.exchange == "p2p" else _
            t = torch.tensor([e0.elapsed_time(e1) / 5], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            st = sorter.verify(k, out[:cnt], cnt)
            rows[d] = {"keys_total": n5 * world, "ms_per_step": float(t), "gkeys_s": n5 * world / float(t) / 1e6,
                       "imbalance": st["imbalance"], "verified": st["ok"], "plan": sorter.plan}
            del k
        except Exception as e:  # noqa: BLE001
            rows[d] = {"error": repr(e)}
    ex["c5_skew_1B"] = rows
    ex["c5_skew_1B"]["keys_total"] = n5 * world
    ns = 1 << 22
    L = ns * world          # a rank may end up with everything (first sort of a skewed input under plan="auto")
    exact = {}
    for d in ("uniform", "allequal", "consthigh"):
        try:
            k = distrib.shard_of_pcg_stream(rank, ns, d)
            sorter.reset_plan()
            sorter.sort(k)                      # plan="auto" settles on histogram or sampled splitters here
            out, cnt = sorter.sort(k)
            pad = torch.zeros(L, dtype=torch.int32, device="cuda")
            pad[:cnt] = out[:cnt]
            outs = torch.empty(L * world, dtype=torch.int32, device="cuda")
            ins = torch.empty(ns * world, dtype=torch.int32, device="cuda")
            cnts = torch.empty(world, dtype=torch.int64, device="cuda")
            dist.all_gather_into_tensor(outs, pad)
            dist.all_gather_into_tensor(ins, k)
            dist.all_gather_into_tensor(cnts, torch.tensor([cnt], dtype=torch.int64, device="cuda"))
            if rank == 0:
                c = cnts.cpu().tolist()
                o = outs.cpu().numpy().view(np.uint32).reshape(world, L)
                got = np.concatenate([o[r, :c[r]] for r in range(world)])
                want = np.sort(ins.cpu().numpy().view(np.uint32))
                exact[d] = bool(got.size == want.size and np.array_equal(got, want))
        except Exception as e:  # noqa: BLE001
            exact[d] = repr(e)
    ex["exact_small"] = {"keys_per_rank": ns, "equals_np_sort": exact}
    return ex


def run_multi(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    pkg = importlib.import_module("gpu-radix-sort_b200")
    distrib = importlib.import_module("gpu-radix-sort_b200.distrib")
    torch.cuda.set_device(local_rank)
    import datetime
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank), timeout=datetime.timedelta(seconds=180))
    n = args.keys
    # one continuous PCG stream; rank g owns elements [g*n, (g+1)*n)  
    clocks = ClockSampler(local_rank)
    clocks.start()                      # live well before the timed loop (set-up below takes seconds)
    d_in = distrib.shard_of_pcg_stream(rank, n, args.dist)
    sorter = distrib.DistributedSorter(n, dist.group.WORLD, mode=args.exchange)
    if args.algo == "stepwise":
        run_multi_stepwise(args, rank, world, local_rank, sorter, d_in, clocks)
        return
    for _ in range(max(args.warmup, 3)):
        sorter.sort(d_in)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clk_first = clocks.mark()
    e0.record()
    for _ in range(args.steps):
        if args.exchange == "p2p":
            out, _ = sorter.sort_async(d_in)        # nothing is read back inside the timed region
        else:
            out, cnt = sorter.sort(d_in)
    e1.record()
    torch.cuda.synchronize()
    if args.exchange == "p2p":
        cnt = sorter._finish()
        out = out[:cnt]
    clk_end = clocks.mark()
    dist.barrier()
    ms_local = torch.tensor([e0.elapsed_time(e1) / args.steps], device="cuda")
    dist.all_reduce(ms_local, op=dist.ReduceOp.MAX)
    ms = float(ms_local.item())
    plan_used = sorter.plan if args.exchange == "p2p" else "host"
    stats = sorter.verify(d_in, out, cnt)           # sortedness, rank ordering, multiset fingerprint
    stats["rank_audit_violation"] = bool(getattr(sorter, "rank_violation", False))   # max over ranks (DistributedSorter._finish)
    phases = sorter.phase_times(d_in)               # per-phase CUDA-event times, max over ranks
    e2e = sorter.e2e(args, d_in)                    # collective: every rank takes part
    clk = clocks.stop(first=clk_first, loop_end=clk_end)
    extras = {}
    if args.dist == "uniform" and n == N_PER_GPU and not args.quick:
        del d_in
        torch.cuda.empty_cache()
        extras = multi_extras(args, rank, world, sorter, distrib, torch, dist, np)
        stats["exact_small"] = extras["exact_small"]["equals_np_sort"]
    if rank == 0:
        peak, peak_src = measured_peak()
        total = n * world
        exch_bytes = 4.0 * n * (world - 1) / world
        line = {
            "metric": METRIC, "value": total / (ms * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": "distributed uint32 sort, 2^28 keys per GPU (BASELINE configs[3] shape)",
                       "keys_per_gpu": n, "keys_total": total, "exchange": args.exchange, "distribution": args.dist,
                       "exchange_plan": plan_used,
                       "parallelism": "key-range sharding, 1 rank/GPU, MSD bucketing + exchange + local LSD",
                       "l2": "1 GiB shard per step per GPU, larger than L2"},
            "clocks": clk, "verify": stats, "phases_ms": phases,
            "gpu_launches": sorter.launches_per_sort * args.steps,
            "roofline": {"bound": "hbm", "kernel": "k_onesweep (local sort pass)", "achieved": phases.get("pass_gbs"),
                         "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                         "frac": (phases.get("pass_gbs") or 0) / peak, "traffic": None},
            "nvlink": {"bytes_out_per_gpu": exch_bytes, "exchange_ms": phases.get("exchange_ms"),
                       "gbs_per_direction": exch_bytes / (phases["exchange_ms"] * 1e-3) / 1e9 if phases.get("exchange_ms") else None,
                       "peak_gbs": 770.0, "peak_source": "measured peer copy per direction (B200_PROFILING.md)"},
            "e2e": e2e,
            "extras": extras,
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()


def run_multi_stepwise(args, rank, world, local_rank, sorter, d_in, clocks):
    """--algo stepwise: the reference's own distributed algorithm (one exchange per 8-bit step) on the
    NVLink backend, beside the default single-exchange sort -- the like-for-like number."""
    import torch
    import torch.distributed as dist
    n = d_in.numel()
    for _ in range(max(args.warmup, 3)):
        sorter.sort_stepwise(d_in)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clk_first = clocks.mark()
    e0.record()
    for _ in range(args.steps):
        out, cnt = sorter.sort_stepwise(d_in)
    e1.record()
    torch.cuda.synchronize()
    clk_end = clocks.mark()
    dist.barrier()
    ms_local = torch.tensor([e0.elapsed_time(e1) / args.steps], device="cuda")
    dist.all_reduce(ms_local, op=dist.ReduceOp.MAX)
    ms = float(ms_local.item())
    clk = clocks.stop(first=clk_first, loop_end=clk_end)
    stats = sorter.verify(d_in, out, cnt)
    marks = []

    def ev(name):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        marks.append((name, e))
    dist.barrier()
    torch.cuda.synchronize()
    ev("start")
    sorter.sort_stepwise(d_in, ev=ev)
    torch.cuda.synchronize()
    vals = torch.tensor([marks[i][1].elapsed_time(marks[i + 1][1]) for i in range(len(marks) - 1)], dtype=torch.float64,
                        device="cuda")
    dist.all_reduce(vals, op=dist.ReduceOp.MAX)
    phases = {m[0] + "_ms": float(v) for m, v in zip(marks[1:], vals.tolist())}
    if rank == 0:
        total = n * world
        step_bytes = 4.0 * n * (world - 1) / world          # uniform digits: (P-1)/P of a slice leaves the GPU per step
        step_ms = sum(phases.values()) / max(len(phases), 1)
        line = {
            "metric": METRIC, "value": total / (ms * 1e-3) / 1e9, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": "step-wise distributed LSD sort (reference distribSort protocol, 4 x 8-bit steps), 2^28 keys per GPU",
                       "keys_per_gpu": n, "keys_total": total, "exchange": "p2p", "distribution": args.dist, "algo": "stepwise",
                       "parallelism": "equal slices, 1 rank/GPU, one global counting sort per digit, stores into peer slices",
                       "l2": "1 GiB slice per step per GPU, larger than L2"},
            "clocks": clk, "verify": stats, "phases_ms": phases,
            "gpu_launches": sorter.launches_per_stepwise_sort * args.steps,
            "nvlink": {"bytes_out_per_gpu_per_step": step_bytes, "step_ms": step_ms,
                       "gbs_per_direction": step_bytes / (step_ms * 1e-3) / 1e9 if step_ms else None, "peak_gbs": 770.0,
                       "peak_source": "measured peer copy per direction (B200_PROFILING.md)"},
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--quick", action="store_true", help="skip the slow extras (full-size CPU baseline, reference GPU run)")
    ap.add_argument("--keys", type=int, default=N_PER_GPU, help="keys per GPU")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="multi-GPU key exchange: fused scatter into peer memory, or NCCL all-to-all")
    ap.add_argument("--dist", default="uniform", choices=["uniform", "zipf", "consthigh", "consthigh16", "allequal", "halfdup"],
                    help="multi-GPU input distribution (BASELINE configs[4] skew cases)")
    ap.add_argument("--algo", default="samplesort", choices=["samplesort", "stepwise"],
                    help="multi-GPU algorithm: one exchange by sampled splitters (default), or the reference's step-wise LSD protocol")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world == 1 and args.gpus == 1:
        run_single(args)
    elif world == 1:
        raise SystemExit("for --gpus N > 1 launch with: python -m torch.distributed.run --nnodes=1 "
                         "--nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...")
    else:
        run_multi(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
