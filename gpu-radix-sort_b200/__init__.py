"""gpu-radix-sort_b200 -- B200-native radix sort behind the libsort.h C ABI.

The product is the shared library ``libsort/libsort.so`` built from ``csrc/``
(hand-written CUDA for sm_100a).  This package is the Python host side:

* ``binding``   ctypes loader for the C ABI (legacy seven symbols + additive
                device-pointer entry points); fails loudly when the library
                has not been built -- there is no CPU or PyTorch fallback.
* ``pylibsort`` mirror of the reference's ctypes wrapper
                (reference faasTest/pylibsort/{__init__,sort,data}.py).
* ``device``    torch-tensor front end over the device-pointer ABI.
* ``distrib``   one-process-per-GPU distributed sort (torch.distributed/NCCL).

The directory name carries a hyphen, so import it with
``importlib.import_module("gpu-radix-sort_b200")``.
"""
from . import binding  # noqa: F401
from .binding import build, lib, lib_path  # noqa: F401
