"""The one data helper on the hot path (reference faasTest/pylibsort/data.py:313-317).

The reference's file-backed distributed arrays (data.py:1-309) are FaaS
exchange plumbing and out of scope.
"""
import ctypes

from .. import binding


def generateInputs(n):
    """n pseudo-random uint32 from the library's PCG stream, as a bytearray."""
    b = bytearray(n * 4)
    if n:
        cInts = (ctypes.c_uint32 * n).from_buffer(b)
        binding.lib().populateInput(cInts, n)
    return b
