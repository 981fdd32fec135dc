"""torch-tensor front end over the device-pointer ABI.

PyTorch supplies device memory and the stream; every byte of sorting is done
by libsort.so's sm_100a kernels.  Keys are uint32, passed as torch.int32 /
torch.uint32 CUDA tensors (same bits).
"""
import ctypes

import torch

from . import binding
from .binding import BOUNDS_EXACT, BOUNDS_REFERENCE, KernelTimes  # noqa: F401

_ws_cache = {}


def _ws(n, device):
    need = binding.lib().libsortWorkspaceBytes(n)
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.empty(max(need, 1 << 20), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


def _check_keys(t):
    if not t.is_cuda or t.dtype not in (torch.int32, torch.uint32) or not t.is_contiguous() or t.dim() != 1:
        raise ValueError("keys must be a contiguous 1-D CUDA int32/uint32 tensor")


def sort_bits(keys, offset=0, width=32, out=None, tmp=None, boundaries=False, boundary_mode=BOUNDS_REFERENCE,
              times=None):
    """Stable sort of `keys` by bits [offset, offset+width) on the current stream.

    Returns `out` (and the boundaries tensor when boundaries=True).  `keys` is
    left untouched unless out is keys.  `times`: a KernelTimes to fill (syncs).
    """
    _check_keys(keys)
    n = keys.numel()
    dev = keys.device
    if out is None:
        out = torch.empty_like(keys)
    if tmp is None:
        tmp = torch.empty_like(keys)
    bnd = torch.empty(1 << width, dtype=torch.int32, device=dev) if boundaries else None
    ws = _ws(n, dev)
    with torch.cuda.device(dev):
        rc = binding.lib().libsortDeviceSort(
            keys.data_ptr(), out.data_ptr(), tmp.data_ptr(), n, offset, width,
            bnd.data_ptr() if bnd is not None else None, boundary_mode, ws.data_ptr(), ws.numel(),
            torch.cuda.current_stream(dev).cuda_stream, ctypes.byref(times) if times is not None else None)
    binding.check(rc, "libsortDeviceSort")
    return (out, bnd) if boundaries else out


def sort(keys, **kw):
    """Full ascending uint32 sort (device-resident providedGpu)."""
    return sort_bits(keys, 0, 32, **kw)


def sort_dyn(keys, d_n, out=None, tmp=None, times=None):
    """Full sort of keys[0 .. d_n[0]) where the count lives on the DEVICE (libsortDeviceSortDyn):
    nothing is read back, launches are sized for keys.numel().  out[d_n[0]:] is left alone."""
    _check_keys(keys)
    if not d_n.is_cuda or d_n.dtype not in (torch.int32, torch.uint32) or d_n.numel() < 1:
        raise ValueError("d_n must be a CUDA int32/uint32 tensor holding the key count")
    cap = keys.numel()
    dev = keys.device
    out = torch.empty_like(keys) if out is None else out
    tmp = torch.empty_like(keys) if tmp is None else tmp
    ws = _ws(cap, dev)
    with torch.cuda.device(dev):
        rc = binding.lib().libsortDeviceSortDyn(
            keys.data_ptr(), out.data_ptr(), tmp.data_ptr(), cap, d_n.data_ptr(), ws.data_ptr(), ws.numel(),
            torch.cuda.current_stream(dev).cuda_stream, ctypes.byref(times) if times is not None else None)
    binding.check(rc, "libsortDeviceSortDyn")
    return out


def rank_audit(device=None):
    """libsortRankAudit on the current stream of `device`: True = a ranking violation was seen
    (results since the last audit must be sorted again; the device now uses the ballot kernels)."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    with torch.cuda.device(dev):
        rc = binding.lib().libsortRankAudit(torch.cuda.current_stream(dev).cuda_stream)
    if rc < 0:
        binding.check(rc, "libsortRankAudit")
    return rc == 1
