"""ctypes binding of libsort.so (include/libsort.h, include/libsort_b200.h)."""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_HERE, "csrc")
_LIBDIR = os.path.join(_HERE, "libsort")

LEGACY_SYMBOLS = ["initLibSort", "gpuPartial", "providedGpu", "providedCpu", "populateInput",
                  "gpuPartialProfile", "providedGpuProfile"]
ADDITIVE_SYMBOLS = ["libsortTileKeys", "libsortWorkspaceBytes", "libsortDeviceSort", "libsortDeviceCount",
                    "libsortReleaseCaches", "libsortErrorString", "libsortDebugSetPortionTiles",
                    "libsortDeviceSample", "libsortDevicePickSplitters", "libsortDeviceSplitCounts", "libsortPopulateInputAt", "libsortDeviceScatterToPeers",
                    "libsortIpcAlloc", "libsortIpcFree", "libsortIpcExport", "libsortIpcOpen", "libsortIpcClose", "libsortDistribSort",
                    "libsortStepTableBytes", "libsortStepTableInit", "libsortStepCounts", "libsortStepBases", "libsortStepScatter",
                    "libsortRankAudit", "libsortDeviceExchangePlanned", "libsortDeviceSortDyn",
                    "libsortDebugLastDevice", "libsortDebugPoolOverlaps", "libsortDeviceExchangeCursor",
                    "libsortDeviceTopHistogram", "libsortDeviceSplittersFromHistograms", "libsortDeviceSortNoAlloc"]

BOUNDS_REFERENCE = 0
BOUNDS_EXACT = 1


class KernelTimes(ctypes.Structure):
    _fields_ = [("memset_ms", ctypes.c_float), ("histogram_ms", ctypes.c_float), ("scan_ms", ctypes.c_float),
                ("pass_ms", ctypes.c_float * 4), ("boundaries_ms", ctypes.c_float), ("npass", ctypes.c_int),
                ("launches", ctypes.c_int)]


def lib_path():
    # LIBSORT_SO: developer override used by the kernel-shape sweeps (tools/gpu_*.sh)
    return os.environ.get("LIBSORT_SO") or os.path.join(_LIBDIR, "libsort.so")


def build(verbose=False):
    """Compile csrc/ for sm_100a into libsort/libsort.so (nvcc; no GPU needed)."""
    res = subprocess.run(["make", "-C", _CSRC, "all", "tools"], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("libsort build failed:\n" + res.stdout[-4000:] + res.stderr[-4000:])
    res2 = subprocess.run(["make", "-C", os.path.join(_HERE, "localTest")], capture_output=True, text=True)
    if res2.returncode != 0:
        raise RuntimeError("localTest build failed:\n" + res2.stdout[-2000:] + res2.stderr[-2000:])
    if verbose:
        print(res.stdout, res2.stdout)
    return lib_path()


_lib = None


def lib():
    """Load libsort.so.  Raises if it is missing: the CUDA library IS the product."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError(
            "%s not found: build it with `make -C %s` (or __graft_entry__.build()); "
            "there is no CPU fallback" % (path, _CSRC))
    L = ctypes.CDLL(path)
    u32p = ctypes.POINTER(ctypes.c_uint32)
    vp, sz, u32, ci = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int
    L.initLibSort.argtypes = []
    L.initLibSort.restype = ctypes.c_bool
    for name in ("gpuPartial", "gpuPartialProfile"):
        f = getattr(L, name)
        f.argtypes = [vp, vp, sz, u32, u32]
        f.restype = ctypes.c_bool
    for name in ("providedGpu", "providedCpu", "providedGpuProfile"):
        f = getattr(L, name)
        f.argtypes = [vp, sz]
        f.restype = ctypes.c_bool
    L.populateInput.argtypes = [vp, sz]
    L.populateInput.restype = None
    L.libsortTileKeys.restype = ci
    L.libsortWorkspaceBytes.argtypes = [sz]
    L.libsortWorkspaceBytes.restype = sz
    L.libsortDeviceSort.argtypes = [vp, vp, vp, sz, u32, u32, vp, ci, vp, sz, vp, ctypes.POINTER(KernelTimes)]
    L.libsortDeviceSort.restype = ci
    L.libsortDeviceCount.restype = ci
    L.libsortReleaseCaches.restype = None
    L.libsortErrorString.argtypes = [ci]
    L.libsortErrorString.restype = ctypes.c_char_p
    L.libsortDebugSetPortionTiles.argtypes = [u32]
    L.libsortDebugSetPortionTiles.restype = None
    L.libsortDeviceSample.argtypes = [vp, sz, vp, u32, u32, vp]
    L.libsortDevicePickSplitters.argtypes = [vp, sz, ci, vp, vp]
    L.libsortDevicePickSplitters.restype = ci
    L.libsortPopulateInputAt.argtypes = [vp, sz, ctypes.c_uint64]
    L.libsortPopulateInputAt.restype = None
    L.libsortDeviceSplitCounts.argtypes = [vp, sz, vp, ci, vp, vp]
    L.libsortDeviceScatterToPeers.argtypes = [vp, sz, vp, ci, vp, vp, sz, vp]
    L.libsortIpcAlloc.argtypes = [sz, ctypes.POINTER(vp)]
    L.libsortIpcFree.argtypes = [vp]
    L.libsortIpcExport.argtypes = [vp, vp]
    L.libsortIpcOpen.argtypes = [vp, ctypes.POINTER(vp)]
    L.libsortIpcClose.argtypes = [vp]
    for name in ("libsortDeviceSample", "libsortDevicePickSplitters", "libsortDeviceSplitCounts", "libsortPopulateInputAt", "libsortDeviceScatterToPeers", "libsortIpcAlloc",
                 "libsortIpcFree", "libsortIpcExport", "libsortIpcOpen", "libsortIpcClose"):
        getattr(L, name).restype = ci
    L.libsortDebugLastDevice.argtypes = []
    L.libsortDebugLastDevice.restype = ci
    L.libsortDebugPoolOverlaps.argtypes = []
    L.libsortDebugPoolOverlaps.restype = ctypes.c_uint
    L.libsortRankAudit.argtypes = [vp]
    L.libsortRankAudit.restype = ci
    L.libsortDeviceExchangePlanned.argtypes = [vp, sz, vp, ci, ci, vp, vp, sz, vp, vp, vp, sz, vp]
    L.libsortDeviceExchangePlanned.restype = ci
    L.libsortDeviceSortNoAlloc.argtypes = [vp, vp, vp, sz, vp]
    L.libsortDeviceSortNoAlloc.restype = ci
    L.libsortDeviceTopHistogram.argtypes = [vp, sz, vp, vp]
    L.libsortDeviceTopHistogram.restype = ci
    L.libsortDeviceSplittersFromHistograms.argtypes = [vp, ci, vp, vp, vp, vp]
    L.libsortDeviceSplittersFromHistograms.restype = ci
    L.libsortDeviceExchangeCursor.argtypes = [vp, sz, vp, ci, vp, sz, vp, vp, sz, vp]
    L.libsortDeviceExchangeCursor.restype = ci
    L.libsortDeviceSortDyn.argtypes = [vp, vp, vp, sz, vp, vp, sz, vp, ctypes.POINTER(KernelTimes)]
    L.libsortDeviceSortDyn.restype = ci
    L.libsortDistribSort.argtypes = [vp, sz, ci]
    L.libsortDistribSort.restype = ci
    L.libsortStepTableBytes.argtypes = []
    L.libsortStepTableBytes.restype = sz
    L.libsortStepTableInit.argtypes = [vp, vp, vp, ci, vp]
    L.libsortStepCounts.argtypes = [vp, sz, ctypes.c_uint32, ctypes.c_uint32, vp, vp]
    L.libsortStepBases.argtypes = [vp, ci, ci, vp, vp]
    L.libsortStepScatter.argtypes = [vp, sz, ctypes.c_uint32, ctypes.c_uint32, vp, vp, sz, vp]
    for name in ("libsortStepTableInit", "libsortStepCounts", "libsortStepBases", "libsortStepScatter"):
        getattr(L, name).restype = ci
    _lib = L
    return L


_inited = False


def init_once():
    """initLibSort exactly once per process (reference libsort.go:22-33 guards the same way)."""
    global _inited
    if not _inited:
        if not lib().initLibSort():
            raise RuntimeError("initLibSort failed (no usable CUDA device?)")
        _inited = True


def check(code, what="libsort"):
    if code != 0:
        raise RuntimeError("%s failed: %s (%d)" % (what, lib().libsortErrorString(code).decode(), code))
