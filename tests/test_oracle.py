"""The oracle against the reference's known answers, its own compiled host
code (oracle/_ref) and the committed golden vectors of the reference GPU path."""
import glob
import os
import re

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

# first eight outputs of populateInput in a fresh process, produced by the
# reference libsort.so itself (libsort/utils.cu:63-79)
PCG_KAT = [0x285594EA, 0x190CA349, 0xCBC42FF2, 0xD6508153, 0xC2A8052F, 0x0F55AC5F, 0xD6FF3E32, 0x4F46689C]


def test_pcg_known_answer(oracle):
    assert oracle.pcg_stream(8).tolist() == PCG_KAT
    assert oracle.np_pcg_stream(8).tolist() == PCG_KAT


def test_pcg_c_vs_python(oracle):
    assert np.array_equal(oracle.pcg_stream(3000), oracle.np_pcg_stream(3000))
    p = oracle.Pcg()
    a, b = p.fill(100), p.fill(50)   # the state keeps advancing across calls (utils.cu:66)
    assert np.array_equal(np.concatenate([a, b]), oracle.pcg_stream(150))


@pytest.mark.parametrize("n", [0, 1, 2, 127, 128, 129, 255, 1021, 1111, 4099])
def test_full_sort_is_ascending_sort(oracle, n):
    k = oracle.pcg_stream(n)
    assert np.array_equal(oracle.full_sort(k), np.sort(k))
    assert np.array_equal(oracle.cpu_sort(k), np.sort(k))


@pytest.mark.parametrize("off,w", [(0, 8), (4, 8), (0, 4), (8, 8), (16, 16), (0, 16), (24, 8), (0, 2), (30, 2), (6, 10)])
def test_step_c_vs_numpy(oracle, off, w):
    k = oracle.pcg_stream(3001)
    s, b = oracle.gpu_partial(k, off, w)
    assert np.array_equal(s, oracle.np_step(k, off, w))
    assert np.array_equal(b, oracle.np_boundaries_ref(s, off, w))
    assert np.array_equal(oracle.boundaries_exact(k, off, w), oracle.np_boundaries_exact(k, off, w))
    # the reference's own test contract (tests.cpp:44-68): uniform keys => ref == exact
    if (1 << w) <= 256:
        assert np.array_equal(b, oracle.boundaries_exact(k, off, w))


def test_boundary_quirks(oracle):
    """Empty leading groups: reference output deviates from the prefix sums
    (sort.cu:19-23 thread 0 never writes; sort.cu:381 back-fill stops at 2)."""
    k = np.full(7, 3, dtype=np.uint32)              # all keys in group 3 of 4
    s, b = oracle.gpu_partial(k, 0, 2)
    assert b.tolist() == [0, 0, 7, 7]               # exact would be [0,0,0,0]
    assert oracle.boundaries_exact(k, 0, 2).tolist() == [0, 0, 0, 0]
    k = np.array([0, 0, 2, 2, 2], dtype=np.uint32)  # group 1 empty, group 0 not
    s, b = oracle.gpu_partial(k, 0, 2)
    assert b.tolist() == [0, 0, 2, 5]               # exact: [0,2,2,5]
    assert oracle.boundaries_exact(k, 0, 2).tolist() == [0, 2, 2, 5]
    for arr in ([5, 5, 9], [1, 1, 1], [0], [255], [2, 3], []):
        a = np.array(arr, dtype=np.uint32)
        for w in (2, 4, 8):
            s, b = oracle.gpu_partial(a, 0, w)
            assert np.array_equal(b, oracle.np_boundaries_ref(s, 0, w)), (arr, w)


@pytest.mark.parametrize("n,p,w", [(1111, 2, 8), (1021, 2, 16), (4099, 3, 8), (1000, 8, 8), (5, 2, 8)])
def test_distrib_oracle(oracle, n, p, w):
    k = oracle.pcg_stream(n)
    assert np.array_equal(oracle.distrib_sort(k, p, w, cxx_split=(p == 2)), np.sort(k))
    assert np.array_equal(oracle.distrib_sort(k, p, w, cxx_split=False), np.sort(k))


def test_against_compiled_reference_host_code(oracle):
    """populateInput / providedCpu of the unmodified reference (oracle/_ref)."""
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    import ctypes
    r = oracle.ref()
    x = np.empty(5000, np.uint32)
    r.populateInput(x.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32)), x.size)
    # NB: only valid the first time this process calls the reference's generator
    assert np.array_equal(x, oracle.pcg_stream(5000))
    assert np.array_equal(oracle.ref_cpu_sort(x), oracle.cpu_sort(x))
    assert np.array_equal(oracle.ref_cpu_sort(x), oracle.full_sort(x))


def test_golden_reference_gpu_outputs(oracle):
    """Outputs of the reference's own GPU path, captured on a B200 by
    tests/golden/make_golden.py, must equal the oracle bit for bit."""
    files = sorted(glob.glob(os.path.join(GOLDEN, "ref_*.npz")))
    assert files, "golden fixtures missing"
    ncases = 0
    for f in files:
        z = np.load(f)
        for name in sorted(k[:-5] for k in z.files if k.endswith("__out")):
            keys = z["in__" + re.sub(r"^(full_|part_)|_o\d+_w\d+$", "", name)]
            if name.startswith("full"):
                assert np.array_equal(z[name + "__out"], oracle.full_sort(keys)), name
            else:
                off, w = int(z[name + "__ow"][0]), int(z[name + "__ow"][1])
                s, b = oracle.gpu_partial(keys, off, w)
                assert np.array_equal(z[name + "__out"], s), name
                assert np.array_equal(z[name + "__bounds"], b), name
            ncases += 1
    assert ncases >= 20
