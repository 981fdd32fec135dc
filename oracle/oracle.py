"""Python face of the test oracle (TEST INFRASTRUCTURE ONLY).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import this module; the product package never
does.  Two independent restatements of the reference hot path live here:

* the C restatement in ``oracle.c`` (loaded from ``oracle/_build/liboracle.so``)
* a numpy restatement (``np_*`` functions) used to cross-check the C one

plus a loader for ``oracle/_ref/libsort_ref.so`` -- the unmodified reference
``libsort`` compiled for sm_100a by ``oracle/Makefile`` -- which is the real
reference whenever it is present (its host-only entry points run anywhere,
its GPU entry points only on a GPU box).

Reference citations are relative to the reference repo root.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "liboracle.so")
_REF = os.path.join(_HERE, "_ref", "libsort_ref.so")

PCG_INITIAL_STATE = 0x4D595DF4D0F33173  # libsort/utils.cu:66


def build(quiet=True):
    """Compile oracle.c (and oracle/_ref when /root/reference is present)."""
    res = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + res.stdout + res.stderr)
    if not quiet:
        print(res.stdout)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = ctypes.CDLL(_LIB)
        u32p = ctypes.POINTER(ctypes.c_uint32)
        L.oracle_pcg_fill.argtypes = [ctypes.POINTER(ctypes.c_uint64), u32p, ctypes.c_size_t]
        L.oracle_pcg_fill.restype = None
        L.oracle_step.argtypes = [u32p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_uint32]
        L.oracle_full_sort.argtypes = [u32p, ctypes.c_size_t]
        L.oracle_cpu_sort.argtypes = [u32p, ctypes.c_size_t]
        L.oracle_boundaries_ref.argtypes = [u32p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_uint32, u32p]
        L.oracle_boundaries_ref.restype = None
        L.oracle_boundaries_exact.argtypes = [u32p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_uint32, u32p]
        L.oracle_boundaries_exact.restype = None
        L.oracle_gpu_partial.argtypes = [u32p, u32p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_uint32]
        L.oracle_distrib_sort.argtypes = [u32p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint32, ctypes.c_int]
        L.oracle_fingerprint.argtypes = [u32p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint64)]
        L.oracle_fingerprint.restype = None
        _lib = L
    return _lib


def _p(a):
    assert a.dtype == np.uint32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32))


# ---------------------------------------------------------------- C oracle --

class Pcg:
    """populateInput stream (libsort/utils.cu:63-79) with an explicit state."""

    def __init__(self, state=PCG_INITIAL_STATE):
        self.state = ctypes.c_uint64(state)

    def fill(self, n):
        out = np.empty(n, dtype=np.uint32)
        lib().oracle_pcg_fill(ctypes.byref(self.state), _p(out), n)
        return out


def pcg_stream(n, state=PCG_INITIAL_STATE):
    return Pcg(state).fill(n)


def step(keys, offset, width):
    """SortState::Step (libsort/sort.cu:322-346); returns a new array."""
    out = np.ascontiguousarray(keys, dtype=np.uint32).copy()
    assert lib().oracle_step(_p(out), out.size, offset, width) == 0
    return out


def full_sort(keys):
    """providedGpu (libsort/invokers.cu:45-64)."""
    return step(keys, 0, 32)


def cpu_sort(keys):
    """providedCpu (libsort/invokers.cu:68-71) -- std::sort port."""
    out = np.ascontiguousarray(keys, dtype=np.uint32).copy()
    lib().oracle_cpu_sort(_p(out), out.size)
    return out


def boundaries_ref(sorted_keys, offset, width):
    """gpu_groups + back-fill, literally (libsort/sort.cu:14-27,367-394)."""
    k = np.ascontiguousarray(sorted_keys, dtype=np.uint32)
    b = np.empty(1 << width, dtype=np.uint32)
    lib().oracle_boundaries_ref(_p(k), k.size, offset, width, _p(b))
    return b


def boundaries_exact(keys, offset, width):
    """Exclusive scan of the group histogram (localTest/tests.cpp:44-68)."""
    k = np.ascontiguousarray(keys, dtype=np.uint32)
    b = np.empty(1 << width, dtype=np.uint32)
    lib().oracle_boundaries_exact(_p(k), k.size, offset, width, _p(b))
    return b


def gpu_partial(keys, offset, width):
    """gpuPartial (libsort/invokers.cu:15-41): returns (sorted keys, boundaries)."""
    out = np.ascontiguousarray(keys, dtype=np.uint32).copy()
    b = np.empty(1 << width, dtype=np.uint32)
    assert lib().oracle_gpu_partial(_p(out), _p(b), out.size, offset, width) == 0
    return out, b


def distrib_sort(keys, nworker=2, width=8, cxx_split=True):
    """distribSort (localTest/benchmarks.cpp:70-160 / distrib.go:90-179)."""
    out = np.ascontiguousarray(keys, dtype=np.uint32).copy()
    assert lib().oracle_distrib_sort(_p(out), out.size, nworker, width, int(cxx_split)) == 0
    return out


def fingerprint(keys):
    k = np.ascontiguousarray(keys, dtype=np.uint32)
    out = (ctypes.c_uint64 * 3)()
    lib().oracle_fingerprint(_p(k), k.size, out)
    return tuple(int(v) for v in out)


# ------------------------------------------------------------ numpy oracle --

def np_group_bits(v, pos, width):
    """group_bits (libsort/sort.cu:9; pylibsort/sort.py:36-37)."""
    v = np.asarray(v, dtype=np.uint64)
    return ((v >> np.uint64(pos)) & np.uint64((1 << width) - 1)).astype(np.uint32)


def np_pcg_stream(n, state=PCG_INITIAL_STATE):
    """populateInput in pure Python ints (small n only)."""
    mult, inc, mask = 6364136223846793005, 1442695040888963407, (1 << 64) - 1
    out = np.empty(n, dtype=np.uint32)
    s = state
    for i in range(n):
        x = s
        rot = x >> 59
        s = (x * mult + inc) & mask
        x ^= x >> 18
        xs = (x >> 27) & 0xFFFFFFFF
        out[i] = ((xs >> rot) | (xs << ((32 - rot) & 31))) & 0xFFFFFFFF
    return out


def np_step(keys, offset, width):
    """Step as 2-bit stable passes with numpy's stable argsort."""
    k = np.asarray(keys, dtype=np.uint32).copy()
    shift = offset
    while shift < offset + width:
        d = np_group_bits(k, shift, 2) if shift < 32 else np.zeros(k.size, np.uint32)
        k = k[np.argsort(d, kind="stable")]
        shift += 2
    return k


def np_boundaries_exact(keys, offset, width):
    g = np_group_bits(keys, offset, width)
    cnt = np.bincount(g, minlength=1 << width).astype(np.uint64)
    return (np.cumsum(cnt) - cnt).astype(np.uint32)


def np_boundaries_ref(sorted_keys, offset, width):
    """Closed form of gpu_groups + back-fill (boundary quirks)."""
    k = np.asarray(sorted_keys, dtype=np.uint32)
    nb = 1 << width
    n = k.size
    g = np_group_bits(k, offset, width)
    cnt = np.bincount(g, minlength=nb).astype(np.uint64)
    exact = np.concatenate([np.cumsum(cnt) - cnt, [n]]).astype(np.uint64)
    nonempty = np.nonzero(cnt)[0]
    g0 = int(nonempty[0]) if nonempty.size else nb
    out = np.zeros(nb, dtype=np.uint64)
    for grp in range(2, nb):
        out[grp] = exact[min(g0, nb - 1) + 1] if grp <= g0 else exact[grp]
    if nb > 1 and cnt[1] > 0 and cnt[0] > 0:
        out[1] = exact[1]
    return (out & 0xFFFFFFFF).astype(np.uint32)


# --------------------------------------------------- the compiled reference --

_ref = None


def ref_available():
    return os.path.exists(_REF)


def ref():
    """The unmodified reference libsort built for sm_100a (oracle/_ref)."""
    global _ref
    if _ref is None:
        if not os.path.exists(_REF):
            raise RuntimeError("oracle/_ref/libsort_ref.so not built (run make -C oracle)")
        L = ctypes.CDLL(_REF)
        u32p = ctypes.POINTER(ctypes.c_uint32)
        L.initLibSort.restype = ctypes.c_bool
        L.providedCpu.argtypes = [u32p, ctypes.c_size_t]
        L.providedCpu.restype = ctypes.c_bool
        L.providedGpu.argtypes = [u32p, ctypes.c_size_t]
        L.providedGpu.restype = ctypes.c_bool
        L.gpuPartial.argtypes = [u32p, u32p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_uint32]
        L.gpuPartial.restype = ctypes.c_bool
        L.populateInput.argtypes = [u32p, ctypes.c_size_t]
        L.populateInput.restype = None
        _ref = L
    return _ref


def ref_cpu_sort(keys):
    out = np.ascontiguousarray(keys, dtype=np.uint32).copy()
    assert ref().providedCpu(_p(out), out.size)
    return out
