"""libsortDeviceSortNoAlloc: the sort entry point that allocates nothing (its look-back workspace is a
file-scope device array inside libsort.so; include/libsort_b200.h).  Checked through the C ABI against
the existing device entry point libsortDeviceSort, bit for bit, on calls in a row that share the
scratch array -- a big input first, then a smaller and a skewed one, so stale contents would show.
Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
MAX_KEYS = 1 << 22      # LIBSORT_NOALLOC_MAX_KEYS


def _no_alloc_sort(L, torch, keys):
    n = keys.numel()
    out = torch.full_like(keys, 0x5A5A5A5A)
    tmp = torch.empty_like(keys)
    st = torch.cuda.current_stream().cuda_stream
    assert L.libsortDeviceSortNoAlloc(keys.data_ptr(), out.data_ptr(), tmp.data_ptr(), n, st) == 0
    return out


def _reference_entry_point(L, torch, keys):
    n = keys.numel()
    out, tmp = torch.empty_like(keys), torch.empty_like(keys)
    ws = torch.empty(max(L.libsortWorkspaceBytes(n), 1 << 20), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert L.libsortDeviceSort(keys.data_ptr(), out.data_ptr(), tmp.data_ptr(), n, 0, 32, None, 0, ws.data_ptr(), ws.numel(),
                               st, None) == 0
    return out


def test_no_alloc_sort_matches_device_sort_on_calls_in_a_row(gpu_lib, oracle):
    import torch
    L = gpu_lib
    x = oracle.pcg_stream(MAX_KEYS)
    inputs = [
        x,                                                          # the maximum size
        x[: (1 << 20) + 12345].copy(),                              # smaller: most of the scratch holds the last call's data
        (x[:300001] & np.uint32(0x00FF00FF)).astype(np.uint32),     # two constant digits (copy passes) + duplicates
        np.full(70000, 0xFFFFFFFF, dtype=np.uint32),                # all keys equal to the padding value
        x[:1].copy(),
    ]
    for k in inputs:
        d = torch.from_numpy(k.view(np.int32)).cuda()
        got = _no_alloc_sort(L, torch, d)                           # the new entry point goes first: it needs no set-up call
        want = _reference_entry_point(L, torch, d)
        torch.cuda.synchronize()
        assert torch.equal(got, want), k.size                       # integer keys: bit-exact
        assert np.array_equal(d.cpu().numpy().view(np.uint32), k)   # input untouched
    # in place (d_out == d_in) and two identical calls in a row
    d = torch.from_numpy(inputs[1].view(np.int32)).cuda()
    tmp = torch.empty_like(d)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        d.copy_(torch.from_numpy(inputs[1].view(np.int32)))
        assert L.libsortDeviceSortNoAlloc(d.data_ptr(), d.data_ptr(), tmp.data_ptr(), d.numel(), st) == 0
        torch.cuda.synchronize()
        assert np.array_equal(d.cpu().numpy().view(np.uint32), np.sort(inputs[1]))


def test_no_alloc_sort_limits(gpu_lib):
    import torch
    L = gpu_lib
    d = torch.zeros(16, dtype=torch.int32, device="cuda")
    t = torch.zeros(16, dtype=torch.int32, device="cuda")
    assert L.libsortDeviceSortNoAlloc(d.data_ptr(), d.data_ptr(), t.data_ptr(), MAX_KEYS + 1, None) != 0    # beyond the scratch array
    assert L.libsortDeviceSortNoAlloc(d.data_ptr(), d.data_ptr(), d.data_ptr(), 16, None) != 0              # d_tmp must be distinct
    assert L.libsortDeviceSortNoAlloc(d.data_ptr(), d.data_ptr(), t.data_ptr(), 0, None) == 0               # nothing to do
    torch.cuda.synchronize()
