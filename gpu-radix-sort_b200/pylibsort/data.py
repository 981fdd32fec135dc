"""Data side of the reference's ``pylibsort`` (faasTest/pylibsort/data.py): the file-backed
distributed arrays and request helpers live in ``distrib_array`` (pure Python, importable without a
GPU) and are re-exported here under the reference's names; ``generateInputs`` is the one helper that
calls into libsort (data.py:313-317).
"""
import ctypes

from .. import binding
from ..distrib_array import (ArrayShape, DistribArray, DistribArrayError, SetDistribMount, closeOpenArrays,  # noqa: F401
                             fileDistribArray, getOutputArray, getPartRefs, openArrs, partRef, readPartRefs, writeOutput)


def generateInputs(n):
    """n pseudo-random uint32 from the library's PCG stream, as a bytearray."""
    b = bytearray(n * 4)
    if n:
        cInts = (ctypes.c_uint32 * n).from_buffer(b)
        binding.lib().populateInput(cInts, n)
    return b
