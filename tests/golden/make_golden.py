#!/usr/bin/env python3
"""Generate tests/golden/ref_*.npz by running the UNMODIFIED reference libsort
(oracle/_ref/libsort_ref.so, built from /root/reference by oracle/Makefile) on a
GPU box:   python tests/golden/make_golden.py gpurun_out/golden
then copy the .npz files into tests/golden/.  Inputs are deterministic (PCG
stream of the oracle + fixed transforms), outputs are whatever the reference's
providedGpu / gpuPartial return.
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as o  # noqa: E402


def inputs():
    base = o.pcg_stream(1 << 15)
    out = {}
    for n in (1, 2, 127, 128, 129, 255, 1021, 1111, 4099):
        out["uniform%d" % n] = base[:n].copy()
    x = base[:3000]
    out["consthigh"] = (np.uint32(0xAB000000) | (x & np.uint32(0x00FFFFFF))).astype(np.uint32)
    out["consthigh16"] = (np.uint32(0xABCD0000) | (x & np.uint32(0xFFFF))).astype(np.uint32)
    out["zipf"] = (np.floor(2.0 ** (32.0 * x.astype(np.float64) / 2.0 ** 32)) - 1).astype(np.uint32)
    out["allequal"] = np.full(777, 0x12345678, dtype=np.uint32)
    out["nogroup0"] = (x | np.uint32(0x01010101)).astype(np.uint32)
    out["nogroup1"] = (x & ~np.uint32(0x01010101)).astype(np.uint32)   # low bit of every byte clear
    out["descending"] = np.arange(5000, 0, -1, dtype=np.uint32)
    return out


def main():
    outdir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(outdir, exist_ok=True)
    r = o.ref()
    assert r.initLibSort(), "reference initLibSort failed (no GPU?)"
    u32p = ctypes.POINTER(ctypes.c_uint32)
    full, part = {}, {}
    ins = inputs()
    for name, keys in ins.items():
        k = keys.copy()
        assert r.providedGpu(k.ctypes.data_as(u32p), k.size)
        full["in__" + name] = keys
        full["full_%s__out" % name] = k
    combos = [(0, 8), (4, 8), (0, 4), (8, 8), (16, 8), (24, 8), (0, 16), (16, 16), (0, 2), (30, 2), (12, 10)]
    for name in ("uniform1021", "uniform1111", "uniform4099", "uniform129", "consthigh", "consthigh16", "zipf",
                 "allequal", "nogroup0", "nogroup1"):
        keys = ins[name]
        for off, w in combos:
            k = keys.copy()
            b = np.zeros(1 << w, dtype=np.uint32)
            assert r.gpuPartial(k.ctypes.data_as(u32p), b.ctypes.data_as(u32p), k.size, off, w)
            tag = "part_%s_o%d_w%d" % (name, off, w)
            part["in__" + name] = keys          # inputs stored once per name
            part[tag + "__out"] = k
            part[tag + "__bounds"] = b
            part[tag + "__ow"] = np.array([off, w], dtype=np.uint32)
    np.savez_compressed(os.path.join(outdir, "ref_full.npz"), **full)
    np.savez_compressed(os.path.join(outdir, "ref_partial.npz"), **part)
    print("wrote", sum(k.endswith("__out") for k in full), "full and", sum(k.endswith("__out") for k in part), "partial cases to", outdir)


if __name__ == "__main__":
    main()
