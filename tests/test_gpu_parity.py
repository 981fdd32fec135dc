"""GPU parity: the CUDA path, called through the C ABI, against the oracle
(bit-exact: keys and boundaries), the committed golden vectors of the reference
GPU path and -- when oracle/_ref loads -- the reference itself run side by side."""
import ctypes
import glob
import os
import re
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

SIZES = [1, 2, 127, 128, 129, 255, 1021, 1111, 4099, 8191, 8192, 8193, 100003, (1 << 20) + 5]
COMBOS = [(0, 8), (4, 8), (0, 4), (8, 8), (16, 8), (24, 8), (0, 16), (16, 16), (0, 2), (30, 2), (6, 10)]


def provided_gpu(lib, keys):
    x = np.ascontiguousarray(keys, dtype=np.uint32).copy()
    assert lib.providedGpu(x.ctypes.data, x.size) is True
    return x


def gpu_partial(lib, keys, off, w):
    x = np.ascontiguousarray(keys, dtype=np.uint32).copy()
    b = np.full(1 << w, 0xDEADBEEF, dtype=np.uint32)
    assert lib.gpuPartial(x.ctypes.data, b.ctypes.data, x.size, off, w) is True
    return x, b


def skewed(oracle, n):
    x = oracle.pcg_stream(n)
    return {
        "consthigh": (np.uint32(0xAB000000) | (x & np.uint32(0x00FFFFFF))).astype(np.uint32),
        "consthigh16": (np.uint32(0xABCD0000) | (x & np.uint32(0xFFFF))).astype(np.uint32),
        "zipf": (np.floor(2.0 ** (32.0 * x.astype(np.float64) / 2.0 ** 32)) - 1).astype(np.uint32),
        "allequal": np.full(n, 0x12345678, dtype=np.uint32),
        "allones": np.full(n, 0xFFFFFFFF, dtype=np.uint32),
        "nogroup0": (x | np.uint32(0x01010101)).astype(np.uint32),
        "nogroup1": (x & ~np.uint32(0x01010101)).astype(np.uint32),
        "descending": np.arange(n, 0, -1, dtype=np.uint32),
    }


@pytest.mark.parametrize("n", SIZES)
def test_provided_gpu_equals_oracle(gpu_lib, oracle, n):
    k = oracle.pcg_stream(n)
    assert np.array_equal(provided_gpu(gpu_lib, k), oracle.full_sort(k))


def test_empty_input(gpu_lib):
    x = np.empty(0, np.uint32)
    assert gpu_lib.providedGpu(x.ctypes.data, 0) is True
    b = np.full(256, 7, np.uint32)
    assert gpu_lib.gpuPartial(x.ctypes.data, b.ctypes.data, 0, 0, 8) is True
    assert not b.any()


@pytest.mark.parametrize("n", [1, 129, 1021, 1111, 4099, 8193, 100003])
@pytest.mark.parametrize("off,w", COMBOS)
def test_gpu_partial_equals_oracle(gpu_lib, oracle, n, off, w):
    k = oracle.pcg_stream(n)
    got, gb = gpu_partial(gpu_lib, k, off, w)
    want, wb = oracle.gpu_partial(k, off, w)
    assert np.array_equal(got, want)
    assert np.array_equal(gb, wb)


@pytest.mark.parametrize("off,w", [(0, 8), (16, 8), (24, 8), (16, 16), (0, 16), (0, 4), (28, 4), (0, 2)])
def test_skewed_inputs_reference_boundaries(gpu_lib, oracle, off, w):
    """Empty leading groups: boundaries must reproduce the reference's
    gpu_groups/back-fill output (sort.cu:14-27,381-387), not the prefix sums."""
    for name, k in skewed(oracle, 5000).items():
        got, gb = gpu_partial(gpu_lib, k, off, w)
        want, wb = oracle.gpu_partial(k, off, w)
        assert np.array_equal(got, want), name
        assert np.array_equal(gb, wb), name
        assert np.array_equal(provided_gpu(gpu_lib, k), np.sort(k)), name


def test_exact_boundary_mode_device_api(gpu_lib, oracle, pkg):
    import torch
    dev = importlib_device(pkg)
    for name, k in list(skewed(oracle, 30000).items()) + [("uniform", oracle.pcg_stream(30000))]:
        t = torch.from_numpy(k.view(np.int32)).cuda()
        for off, w in [(0, 8), (24, 8), (16, 16), (8, 12)]:
            out, b = dev.sort_bits(t, off, w, boundaries=True, boundary_mode=pkg.binding.BOUNDS_EXACT)
            assert np.array_equal(out.cpu().numpy().view(np.uint32), oracle.step(k, off, w)), name
            assert np.array_equal(b.cpu().numpy().view(np.uint32), oracle.boundaries_exact(k, off, w)), name
        assert np.array_equal(t.cpu().numpy().view(np.uint32), k)   # input untouched


def importlib_device(pkg):
    import importlib
    return importlib.import_module(pkg.__name__ + ".device")


def test_device_api_in_place_and_misaligned(gpu_lib, oracle, pkg):
    import torch
    dev = importlib_device(pkg)
    k = oracle.pcg_stream(50001)
    big = torch.from_numpy(k.view(np.int32)).cuda()
    for off in (0, 1, 3):   # element offsets 1 and 3: pointer not 16 B aligned -> non-TMA path
        t = big[off:].clone() if off == 0 else big[off:]
        src = k[off:]
        out = dev.sort(t.contiguous() if off == 0 else t)
        assert np.array_equal(out.cpu().numpy().view(np.uint32), np.sort(src))
    t = big.clone()
    out = dev.sort_bits(t, 0, 8, out=t)      # in place with an odd pass count
    assert np.array_equal(out.cpu().numpy().view(np.uint32), oracle.step(k, 0, 8))
    t = big.clone()
    out = dev.sort_bits(t, 0, 32, out=t)     # in place, even pass count
    assert np.array_equal(out.cpu().numpy().view(np.uint32), np.sort(k))


def test_lookback_portions(gpu_lib, oracle):
    """Force several look-back portions (carry of bucket bases between launches)."""
    tile = gpu_lib.libsortTileKeys()
    k = oracle.pcg_stream(tile * 17 + 123)
    gpu_lib.libsortDebugSetPortionTiles(3)
    try:
        assert np.array_equal(provided_gpu(gpu_lib, k), np.sort(k))
        got, gb = gpu_partial(gpu_lib, k, 8, 16)
        want, wb = oracle.gpu_partial(k, 8, 16)
        assert np.array_equal(got, want) and np.array_equal(gb, wb)
    finally:
        gpu_lib.libsortDebugSetPortionTiles(0)


def test_golden_reference_gpu_outputs(gpu_lib):
    files = sorted(glob.glob(os.path.join(GOLDEN, "ref_*.npz")))
    assert files, "golden fixtures missing"
    for f in files:
        z = np.load(f)
        for name in sorted(k[:-5] for k in z.files if k.endswith("__out")):
            keys = z["in__" + re.sub(r"^(full_|part_)|_o\d+_w\d+$", "", name)]
            if name.startswith("full"):
                assert np.array_equal(provided_gpu(gpu_lib, keys), z[name + "__out"]), name
            else:
                off, w = int(z[name + "__ow"][0]), int(z[name + "__ow"][1])
                got, gb = gpu_partial(gpu_lib, keys, off, w)
                assert np.array_equal(got, z[name + "__out"]), name
                assert np.array_equal(gb, z[name + "__bounds"]), name


def test_side_by_side_with_compiled_reference(gpu_lib, oracle):
    """The unmodified reference (oracle/_ref, sm_100a build) on the same GPU."""
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not present")
    r = oracle.ref()
    if not getattr(test_side_by_side_with_compiled_reference, "_inited", False):
        assert r.initLibSort()
        test_side_by_side_with_compiled_reference._inited = True
    u32p = ctypes.POINTER(ctypes.c_uint32)
    cases = [("uniform", oracle.pcg_stream(200000))] + list(skewed(oracle, 20000).items())
    for name, k in cases:
        a = k.copy()
        assert r.providedGpu(a.ctypes.data_as(u32p), a.size)
        assert np.array_equal(provided_gpu(gpu_lib, k), a), name
        for off, w in [(0, 8), (8, 8), (24, 8), (0, 16), (16, 16), (4, 8)]:
            a = k.copy()
            rb = np.zeros(1 << w, np.uint32)
            assert r.gpuPartial(a.ctypes.data_as(u32p), rb.ctypes.data_as(u32p), a.size, off, w)
            got, gb = gpu_partial(gpu_lib, k, off, w)
            assert np.array_equal(got, a), (name, off, w)
            assert np.array_equal(gb, rb), (name, off, w)


def test_concurrent_callers(gpu_lib, oracle):
    """16 host threads at once, full and partial, inside the reference's 2 s bound
    (TestParallel, benchmark/pkg/sort/libsort_test.go:35-87, timeout at :60); no two callers ever
    hold the same GPU (libsort/utils.cu:33-61) and each call is served by a GPU of the pool."""
    import time
    k = oracle.pcg_stream(4099)
    want_full, (want_part, want_b) = np.sort(k), oracle.gpu_partial(k, 0, 8)
    provided_gpu(gpu_lib, k)                                   # buffers of this size exist on one GPU at least
    errs, devs = [], []

    def worker(i):
        try:
            if i % 2:
                assert np.array_equal(provided_gpu(gpu_lib, k), want_full)
            else:
                got, gb = gpu_partial(gpu_lib, k, 0, 8)
                assert np.array_equal(got, want_part) and np.array_equal(gb, want_b)
            devs.append(gpu_lib.libsortDebugLastDevice())
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))

    th = [threading.Thread(target=worker, args=(i,)) for i in range(16)]
    t0 = time.monotonic()
    [t.start() for t in th]
    [t.join(max(0.0, 2.0 - (time.monotonic() - t0))) for t in th]
    assert not any(t.is_alive() for t in th), "16 concurrent callers did not finish within 2 s"
    assert not errs, errs
    ndev = gpu_lib.libsortDeviceCount()
    assert len(devs) == 16 and all(0 <= d < ndev for d in devs)
    assert gpu_lib.libsortDebugPoolOverlaps() == 0


def test_concurrent_callers_spread_over_the_pool(gpu_lib, oracle):
    """Long enough calls (2^24 keys) from 2 x ndev threads: every GPU of the pool serves some, never
    two callers on one GPU at a time.  On the driver's one-GPU box this is the serialised case."""
    ndev = gpu_lib.libsortDeviceCount()
    k = oracle.pcg_stream(1 << 24)
    fp = (int(k.astype(np.uint64).sum()), int(np.bitwise_xor.reduce(k)))
    devs, errs = [], []

    def worker():
        try:
            x = provided_gpu(gpu_lib, k)
            assert bool(np.all(x[1:] >= x[:-1])) and (int(x.astype(np.uint64).sum()), int(np.bitwise_xor.reduce(x))) == fp
            devs.append(gpu_lib.libsortDebugLastDevice())
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))

    for _ in range(2):                                        # first round allocates per-GPU buffers
        devs.clear()
        th = [threading.Thread(target=worker) for _ in range(2 * ndev)]
        [t.start() for t in th]
        [t.join(120) for t in th]
        assert not errs and not any(t.is_alive() for t in th), errs
    assert set(devs) == set(range(ndev)), devs
    assert gpu_lib.libsortDebugPoolOverlaps() == 0


def test_second_init_fails_and_oversize_rejected(gpu_lib):
    assert gpu_lib.initLibSort() is False                      # utils.cu:14-17
    x = np.zeros(4, np.uint32)
    assert gpu_lib.providedGpu(x.ctypes.data, (1 << 32) + 5) is False   # invokers.cu:54-57


def test_pylibsort_wrapper(gpu_lib, oracle, pkg):
    """Mirror of faasTest/testPyLibSort.py:195-245 (testSortFull / testSortPartial)."""
    import importlib
    pls = importlib.import_module(pkg.__name__ + ".pylibsort")
    k = oracle.pcg_stream(1021)
    buf = bytearray(k.tobytes())
    pls.sortFull(buf)
    pls.checkSortFull(pls.bytesToInts(buf), k)
    pls.checkOrder(pls.bytesToInts(buf))
    buf = bytearray(k.tobytes())
    bounds = pls.sortPartial(buf, 4, 8)
    want, wb = oracle.gpu_partial(k, 4, 8)
    assert np.array_equal(pls.bytesToInts(buf), want) and bounds == wb.tolist()
    caps = np.diff(np.array(bounds, dtype=np.int64) * 4, append=len(buf))    # data.py:301-304
    pls.checkPartial(k.tobytes(), buf, caps, 4, 8)
    assert len(pls.generateInputs(10)) == 40


def test_cxx_consumers(gpu_lib, pkg):
    """BASELINE configs[0]: localTest ./radix_sort, 1M keys, GPU sort vs the CPU local sort;
    and the reference's own unmodified harness linked against our libsort.so."""
    import subprocess
    exe = os.path.join(os.path.dirname(pkg.__file__), "localTest", "radix_sort")
    for n in ("1048576", "1111"):
        r = subprocess.run([exe, n], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0 and "Success!" in r.stdout, r.stdout + r.stderr
    ref_exe = os.path.join(os.path.dirname(os.path.dirname(__file__)), "oracle", "_ref", "radix_sort_reference_harness")
    if os.path.exists(ref_exe):
        r = subprocess.run([ref_exe], capture_output=True, text=True, timeout=300)   # runBenches(): 16-bit partial, 2^28 keys
        assert r.returncode == 0 and "Partial sort (16b)" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("env", [{}, {"LIBSORT_RANK": "ballot"}, {"LIBSORT_BULK_OUT": "1"}, {"LIBSORT_RANK": "atoms"},
                                 {"LIBSORT_PDL": "0"}, {"LIBSORT_SMALL_PATH": "0"}, {"LIBSORT_K3": "inp256x56p"},
                                 {"LIBSORT_K3": "pipe42w"}],
                         ids=["default", "ballot", "bulk_out", "atoms_forced", "no_dependent_launch", "no_small_tiles",
                              "k3_bulk_inplace", "k3_pipeline"])
def test_kernel_variants(env):
    """Every compiled scatter kernel (ranking mode x write-out) against the oracle, multi-tile
    skewed inputs included; the variant is fixed when the library is first used, hence a fresh process."""
    import subprocess
    import sys
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "variant_worker.py")], env=e,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("bulk", ["1", "0"], ids=["bulk_copy", "lds_stg"])
def test_splitter_scatter_single_gpu(bulk):
    """Counts per destination and the stable splitter scatter (both write-out kernels, 2..16
    destinations, every compiled splitter capacity) on one GPU against numpy."""
    import subprocess
    import sys
    e = dict(os.environ)
    e["LIBSORT_BULK_EXCHANGE"] = bulk
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "split_worker.py")], env=e,
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("n", [0, 1, 3, 1000, (1 << 20) + 7, (1 << 24) + 5])
def test_distrib_sort_host_call(gpu_lib, oracle, n):
    """libsortDistribSort (the whole distributed sort behind one host-pointer call, every GPU the
    pool has -- one on the driver's box, where it degenerates to partition-to-self + local sort)
    against the oracle's full sort; the multi-GPU run is tools/gpu_multi.sh / tests/distrib_host_worker.py."""
    k = oracle.pcg_stream(n).copy()
    if n >= 1000:
        k[::3] &= np.uint32(0x00FF00FF)            # duplicates + skew
    want = oracle.full_sort(k)
    got = k.copy()
    assert gpu_lib.libsortDistribSort(got.ctypes.data, n, 0) == 1
    assert np.array_equal(got, want)
    assert gpu_lib.libsortDistribSort(got.ctypes.data, n, 99) == 0      # more GPUs than any pool has


@pytest.mark.parametrize("lens", [[40000], [30000, 30000], [20000, 31111, 15000, 7], [14336, 14336, 14336, 14336, 1]])
def test_stepwise_blocks_virtual_ranks(gpu_lib, oracle, pkg, lens):
    """The step-wise distributed LSD building blocks (libsortStep*) with several VIRTUAL ranks on one GPU:
    each rank's scatter stores every key at its global position inside whichever slice owns it.  After step
    k the whole array must be the input stably sorted by its low 8k bits -- the reference distribSort's
    intermediate state (localTest/benchmarks.cpp:70-160) -- for uniform and skewed keys."""
    import ctypes
    import torch
    L, P, total = gpu_lib, len(lens), sum(lens)
    starts = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    st = torch.cuda.current_stream().cuda_stream
    x = oracle.pcg_stream(total)
    cases = {"uniform": x, "lowbyte": (x & np.uint32(0xFF0000FF)).astype(np.uint32), "equal": np.full(total, 0xA5A5A5A5, np.uint32)}
    tb = L.libsortStepTableBytes()
    ws = torch.empty(L.libsortWorkspaceBytes(max(lens)), dtype=torch.uint8, device="cuda")
    for name, keys in cases.items():
        bufs = [torch.from_numpy(keys.view(np.int32)).cuda(), torch.empty(total, dtype=torch.int32, device="cuda")]
        tables = [torch.zeros(tb, dtype=torch.uint8, device="cuda") for _ in range(2)]
        for k in range(2):
            peers = (ctypes.c_uint64 * P)(*[bufs[k].data_ptr() + 4 * int(starts[r]) for r in range(P)])
            pkg.binding.check(L.libsortStepTableInit(tables[k].data_ptr(), peers, (ctypes.c_uint64 * (P + 1))(*starts.tolist()), P, st),
                              "table")
        gathered = torch.zeros(P * 256, dtype=torch.int32, device="cuda")
        state = keys
        for step in range(4):
            src, dst = bufs[step & 1], (step + 1) & 1
            for r in range(P):
                pkg.binding.check(L.libsortStepCounts(src.data_ptr() + 4 * int(starts[r]), lens[r], 8 * step, 8,
                                                      gathered.data_ptr() + 1024 * r, st), "counts")
            for r in range(P):      # every virtual rank: bases, then its scatter into the other buffer's slices
                pkg.binding.check(L.libsortStepBases(gathered.data_ptr(), P, r, tables[dst].data_ptr(), st), "bases")
                pkg.binding.check(L.libsortStepScatter(src.data_ptr() + 4 * int(starts[r]), lens[r], 8 * step, 8,
                                                       tables[dst].data_ptr(), ws.data_ptr(), ws.numel(), st), "scatter")
            state = state[np.argsort((state >> np.uint32(8 * step)) & np.uint32(0xFF), kind="stable")]
            got = bufs[dst].cpu().numpy().view(np.uint32)
            assert np.array_equal(got, state), (name, lens, "step", step + 1)
        assert np.array_equal(got, np.sort(keys))
    # a middle slice shorter than one tile cannot be addressed: rejected, not mis-sorted
    if P >= 3:
        bad = np.array([0, 20000, 20010, 60000], dtype=np.int64)
        peers = (ctypes.c_uint64 * 3)(*[bufs[0].data_ptr()] * 3)
        assert L.libsortStepTableInit(tables[0].data_ptr(), peers, (ctypes.c_uint64 * 4)(*bad.tolist()), 3, st) != 0
