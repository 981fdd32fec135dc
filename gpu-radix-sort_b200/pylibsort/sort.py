"""Mirror of reference faasTest/pylibsort/sort.py: same names, arguments and errors."""
import ctypes

import numpy as np

from .. import binding
from ..distrib_array import exactBoundaries  # noqa: F401  (pure Python; re-exported under pylibsort)


class sortException(Exception):
    def __init__(self, msg):
        self.msg = msg

    def __str__(self):
        return 'Sort error: {}'.format(self.msg)


class sortResultException(Exception):
    def __init__(self, idx, expect, got, msg=None):
        self.idx, self.expect, self.got, self.msg = idx, expect, got, msg

    def __str__(self):
        desc = "Sorted Incorrectly at index {}, Expected {:#08x}, Got {:#08x}".format(self.idx, self.expect, self.got)
        if self.msg is not None:
            desc += " ({})".format(self.msg)
        return desc


def groupBits(v, pos, width):
    """Group ID of v for `width` bits starting at `pos` (reference sort.py:36-37)."""
    return (v >> pos) & ((1 << width) - 1)


def bytesToInts(barr):
    return np.frombuffer(barr, dtype=np.uint32)


def checkOrder(arr):
    """Verify that arr is in order (reference sort.py:50-56)."""
    a = np.asarray(arr)
    bad = np.nonzero(a[1:] < a[:-1])[0]
    if bad.size:
        i = int(bad[0]) + 1
        raise sortResultException(i, int(a[i - 1]), int(a[i]), "value should be >= to expected")


def checkSortFull(new, orig):
    """Verify that new is a correctly sorted copy of orig (reference sort.py:59-64)."""
    want = np.sort(np.asarray(orig, dtype=np.uint32))
    got = np.asarray(new, dtype=np.uint32)
    bad = np.nonzero(want != got)[0]
    if bad.size:
        i = int(bad[0])
        raise sortResultException(i, int(want[i]), int(got[i]))


def checkPartial(refBytes, testBytes, caps, pos, width):
    """Reference sort.py:67-91, with the group taken at `pos` (the reference
    masks the low bits regardless of pos, which is only right for pos == 0)."""
    refInts, testInts = bytesToInts(refBytes), bytesToInts(testBytes)
    caps = np.array(caps)
    if len(refInts) != len(testInts):
        raise sortException("test length doesnt match reference: expected {}, got {}".format(len(refInts), len(testInts)))
    if len(caps) != (1 << width):
        raise sortException("Not enough output buckets: expected {}, got {}".format(len(caps), (1 << width)))
    caps = caps // 4
    expectGroups = np.repeat(np.arange(len(caps)), caps)
    testGroups = groupBits(testInts, pos, width)
    if not np.array_equal(testGroups, expectGroups):
        raise sortException("Output does not have expected groups")


def sortFull(buf: bytearray):
    """Interpret buf as C uint32s and sort it in place (reference sort.py:94-105)."""
    nElem = int(len(buf) / 4)
    if nElem == 0:
        return
    cRaw = (ctypes.c_uint8 * len(buf)).from_buffer(buf)
    res = binding.lib().providedGpu(ctypes.cast(cRaw, ctypes.c_void_p), ctypes.c_size_t(nElem))
    if not res:
        raise RuntimeError("Libsort had an internal error")


def sortPartial(buf: bytearray, offset, width):
    """Partial sort of buf in place on bits [offset, offset+width); returns the
    list of group boundaries in elements (reference sort.py:108-126)."""
    nElem = int(len(buf) / 4)
    boundaries = (ctypes.c_uint32 * (1 << width))()
    if nElem:
        cRaw = (ctypes.c_uint8 * len(buf)).from_buffer(buf)
        ptr = ctypes.cast(cRaw, ctypes.c_void_p)
    else:
        ptr = ctypes.c_void_p(0)
    res = binding.lib().gpuPartial(ptr, ctypes.cast(boundaries, ctypes.c_void_p), ctypes.c_size_t(nElem),
                                   ctypes.c_uint32(offset), ctypes.c_uint32(width))
    if not res:
        raise RuntimeError("Libsort had an internal error")
    return list(boundaries)


def sortDistrib(buf: bytearray, ngpu=0):
    """Sort buf (C uint32s) in place across `ngpu` GPUs of this box (0 = all of the pool):
    shards exchanged GPU-to-GPU over NVLink instead of through the host shuffle of
    reference sort.py:129-175 / benchmark/pkg/sort/distrib.go:183-248 (additive API)."""
    nElem = int(len(buf) / 4)
    if nElem == 0:
        return
    cRaw = (ctypes.c_uint8 * len(buf)).from_buffer(buf)
    res = binding.lib().libsortDistribSort(ctypes.cast(cRaw, ctypes.c_void_p), ctypes.c_size_t(nElem), ctypes.c_int(ngpu))
    if not res:
        raise RuntimeError("Libsort had an internal error")
