"""GPU tests of the device-planned entry points behind the multi-GPU sort
(libsortDeviceSortDyn, libsortDeviceExchangePlanned, libsortRankAudit) and of
stability at sizes the CPU oracle cannot reach: all checked bit-exactly, the big
ones through size-independent properties computed on the device."""
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _dev(pkg):
    return importlib.import_module(pkg.__name__ + ".device")


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


def test_sort_dyn_reads_count_on_device(gpu_lib, oracle, pkg):
    """Key count read by the kernels (0, 1, tile edges, capacity); the tail of `out` stays untouched."""
    import torch
    dev = _dev(pkg)
    tile = gpu_lib.libsortTileKeys()
    cap = tile * 5 + 77
    k = oracle.pcg_stream(cap)
    k[::5] &= np.uint32(0xFF00FFFF)
    buf = torch.from_numpy(k.view(np.int32)).cuda()
    for n in [0, 1, 2, tile - 1, tile, tile + 1, 3 * tile + 9, cap - 1, cap]:
        d_n = torch.tensor([n], dtype=torch.int32, device="cuda")
        out = torch.full((cap,), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
        got = _u32(dev.sort_dyn(buf, d_n, out=out))
        assert np.array_equal(got[:n], oracle.full_sort(k[:n])), n
        assert (got[n:] == 0x5A5A5A5A).all(), n
        assert np.array_equal(_u32(buf), k)                 # input untouched
    assert dev.rank_audit() is False


def test_sort_dyn_rejects_oversize_capacity(gpu_lib):
    import torch
    x = torch.zeros(16, dtype=torch.int32, device="cuda")
    rc = gpu_lib.libsortDeviceSortDyn(x.data_ptr(), x.data_ptr(), x.data_ptr(), 1 << 30, x.data_ptr(), x.data_ptr(), 64,
                                      None, None)
    assert rc != 0


def _stable_by_construction(n, w, off, torch):
    """keys whose bits outside [off, off+w) hold the original index: a stable partial sort leaves
    every digit bucket strictly ascending."""
    idx = torch.arange(n, dtype=torch.int64, device="cuda")
    digit = (idx * 2654435761 >> 7) & ((1 << w) - 1)
    lo = idx & ((1 << off) - 1)
    hi = idx >> off
    return ((hi << (off + w)) | (digit << off) | lo).to(torch.int32)


@pytest.mark.parametrize("n,off,w", [(1 << 26, 0, 4), (1 << 24, 0, 8), (1 << 24, 4, 8), (1 << 16, 8, 16), (1 << 28, 0, 4)])
def test_stability_at_scale(gpu_lib, pkg, n, off, w):
    """Stability of every pass at sizes far above the oracle's reach."""
    import torch
    dev = _dev(pkg)
    keys = _stable_by_construction(n, w, off, torch)
    out = dev.sort_bits(keys, off, w)
    assert dev.rank_audit() is False
    u = out.to(torch.int64) & 0xFFFFFFFF
    del out, keys
    digit = (u >> off) & ((1 << w) - 1)
    ok = (digit[1:] > digit[:-1]) | ((digit[1:] == digit[:-1]) & (u[1:] > u[:-1]))
    assert bool(ok.all())
    del ok, digit
    # a permutation of the input: the original indices sum to n(n-1)/2
    orig = ((u >> (off + w)) << off) | (u & ((1 << off) - 1))
    assert int(orig.sum()) == n * (n - 1) // 2


@pytest.mark.parametrize("W", [1, 2, 4, 8, 16])
def test_planned_exchange_virtual_ranks(gpu_lib, oracle, W):
    """W virtual ranks on one GPU: each runs the planned exchange into W receive regions; every
    region ends up with exactly the keys of its splitter range (any order), nrecv matches."""
    import torch
    L = gpu_lib
    n = 50000 + 17 * W
    stream = oracle.pcg_stream(n * W)
    shards = [stream[r * n:(r + 1) * n].copy() for r in range(W)]
    for s in shards:
        s[::7] &= np.uint32(0x0FFFFFFF)
    allk = np.sort(np.concatenate(shards))
    spl = np.zeros(32, np.uint32)                           # 16 values + 16 tie thresholds (0: plain <=)
    for r in range(W - 1):
        spl[r] = allk[(r + 1) * allk.size // W]
    d_spl = torch.from_numpy(spl.view(np.int32)).cuda()
    cap = int(n * 2 + 1024)
    recv = torch.full((W, cap), -1, dtype=torch.int32, device="cuda")
    bases = torch.tensor([recv[r].data_ptr() for r in range(W)], dtype=torch.int64, device="cuda")
    d_keys = [torch.from_numpy(s.view(np.int32)).cuda() for s in shards]
    cm = torch.zeros((W, 16), dtype=torch.int32, device="cuda")
    for r in range(W):
        assert L.libsortDeviceSplitCounts(d_keys[r].data_ptr(), n, d_spl.data_ptr(), W - 1, cm[r].data_ptr(), None) == 0
    ws = torch.empty(max(L.libsortWorkspaceBytes(n), 1 << 20), dtype=torch.uint8, device="cuda")
    nrecv = torch.zeros(W, dtype=torch.int32, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    for r in range(W):
        rc = L.libsortDeviceExchangePlanned(d_keys[r].data_ptr(), n, d_spl.data_ptr(), W, r, cm.data_ptr(), bases.data_ptr(),
                                            cap, nrecv[r:].data_ptr(), status.data_ptr(), ws.data_ptr(), ws.numel(), None)
        assert rc == 0
    torch.cuda.synchronize()
    assert int(status.item()) == 0
    h = recv.cpu().numpy().view(np.uint32)
    got_n = nrecv.cpu().numpy()
    for r in range(W):
        assert np.all(h[r, got_n[r]:] == 0xFFFFFFFF)        # nothing beyond the announced count
    assert int(got_n.sum()) == n * W
    parts = [np.sort(h[r, :got_n[r]]) for r in range(W)]
    for a, b in zip(parts[:-1], parts[1:]):
        if a.size and b.size:
            assert a[-1] <= b[0]
    assert np.array_equal(np.concatenate(parts), allk)


def test_planned_exchange_overflow_sets_status(gpu_lib, oracle):
    import torch
    L = gpu_lib
    n = 40000
    k = oracle.pcg_stream(n)
    d_k = torch.from_numpy(k.view(np.int32)).cuda()
    d_spl = torch.zeros(32, dtype=torch.int32, device="cuda")
    cm = torch.zeros((1, 16), dtype=torch.int32, device="cuda")
    assert L.libsortDeviceSplitCounts(d_k.data_ptr(), n, d_spl.data_ptr(), 0, cm.data_ptr(), None) == 0
    cap = n - 1
    recv = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    bases = torch.tensor([recv.data_ptr()], dtype=torch.int64, device="cuda")
    ws = torch.empty(max(L.libsortWorkspaceBytes(n), 1 << 20), dtype=torch.uint8, device="cuda")
    nrecv = torch.full((1,), 123, dtype=torch.int32, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    assert L.libsortDeviceExchangePlanned(d_k.data_ptr(), n, d_spl.data_ptr(), 1, 0, cm.data_ptr(), bases.data_ptr(), cap,
                                          nrecv.data_ptr(), status.data_ptr(), ws.data_ptr(), ws.numel(), None) == 0
    torch.cuda.synchronize()
    assert int(status.item()) == 1 and int(nrecv.item()) == 0
    assert bool((recv == -1).all())                        # nothing stored


@pytest.mark.parametrize("W", [1, 2, 4, 8, 16])
def test_cursor_exchange_virtual_ranks(gpu_lib, oracle, W):
    """libsortDeviceExchangeCursor, W virtual ranks on one GPU: no counts anywhere -- every tile reserves
    its room with an atomicAdd on the destination's cursor word (behind the `cap` keys of each receive
    region).  Afterwards cursor r = keys of range r, the regions hold exactly those keys (any order)."""
    import torch
    L = gpu_lib
    n = 70001 + 13 * W
    stream = oracle.pcg_stream(n * W)
    shards = [stream[r * n:(r + 1) * n].copy() for r in range(W)]
    for s in shards:
        s[::5] &= np.uint32(0x3FFFFFFF)
    allk = np.sort(np.concatenate(shards))
    spl = np.zeros(32, np.uint32)
    for r in range(W - 1):
        spl[r] = allk[(r + 1) * allk.size // W]
    d_spl = torch.from_numpy(spl.view(np.int32)).cuda()
    cap = int(n * 2 + 1024)
    recv = torch.full((W, cap + 16), -1, dtype=torch.int32, device="cuda")
    recv[:, cap:] = 0                                              # the cursor words
    bases = torch.tensor([recv[r].data_ptr() for r in range(W)], dtype=torch.int64, device="cuda")
    ws = torch.empty(max(L.libsortWorkspaceBytes(n), 1 << 20), dtype=torch.uint8, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    for r in range(W):
        k = torch.from_numpy(shards[r].view(np.int32)).cuda()
        assert L.libsortDeviceExchangeCursor(k.data_ptr(), n, d_spl.data_ptr(), W, bases.data_ptr(), cap, status.data_ptr(),
                                             ws.data_ptr(), ws.numel(), None) == 0
    torch.cuda.synchronize()
    assert int(status.item()) == 0
    h = recv.cpu().numpy().view(np.uint32)
    got_n = h[:, cap].astype(np.int64)
    assert int(got_n.sum()) == n * W
    for r in range(W):
        assert np.all(h[r, got_n[r]:cap] == 0xFFFFFFFF)           # nothing beyond the cursor
    parts = [np.sort(h[r, :got_n[r]]) for r in range(W)]
    for a, b in zip(parts[:-1], parts[1:]):
        if a.size and b.size:
            assert a[-1] <= b[0]
    assert np.array_equal(np.concatenate(parts), allk)
    # the cursor word is the device-side key count of the local sort
    dev_out = torch.empty(cap, dtype=torch.int32, device="cuda")
    dev_tmp = torch.empty(cap, dtype=torch.int32, device="cuda")
    ws2 = torch.empty(L.libsortWorkspaceBytes(cap), dtype=torch.uint8, device="cuda")
    for r in range(W):
        assert L.libsortDeviceSortDyn(recv[r].data_ptr(), dev_out.data_ptr(), dev_tmp.data_ptr(), cap,
                                      recv[r][cap:].data_ptr(), ws2.data_ptr(), ws2.numel(), None, None) == 0
        assert np.array_equal(dev_out[:got_n[r]].cpu().numpy().view(np.uint32), parts[r])


def test_cursor_exchange_overflow(gpu_lib, oracle):
    """A destination that would overflow: status raised, nothing written past cap, cursor > cap, and the
    dynamic sort clamps to its capacity instead of running off the buffer."""
    import torch
    L = gpu_lib
    n = 50000
    k = oracle.pcg_stream(n)
    d_k = torch.from_numpy(k.view(np.int32)).cuda()
    d_spl = torch.zeros(32, dtype=torch.int32, device="cuda")
    cap = 30000                                                    # a multiple of 4
    recv = torch.full((cap + 16 + 4096,), -1, dtype=torch.int32, device="cuda")
    recv[cap:cap + 16] = 0
    bases = torch.tensor([recv.data_ptr()], dtype=torch.int64, device="cuda")
    ws = torch.empty(max(L.libsortWorkspaceBytes(n), 1 << 20), dtype=torch.uint8, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    assert L.libsortDeviceExchangeCursor(d_k.data_ptr(), n, d_spl.data_ptr(), 1, bases.data_ptr(), cap, status.data_ptr(),
                                         ws.data_ptr(), ws.numel(), None) == 0
    torch.cuda.synchronize()
    assert int(status.item()) == 1
    h = recv.cpu().numpy().view(np.uint32)
    assert int(h[cap]) == n                                        # every reservation was counted
    assert np.all(h[cap + 16:] == 0xFFFFFFFF)                      # nothing stored behind the buffer
    out = torch.empty(cap, dtype=torch.int32, device="cuda")
    tmp = torch.empty(cap, dtype=torch.int32, device="cuda")
    ws2 = torch.empty(L.libsortWorkspaceBytes(cap), dtype=torch.uint8, device="cuda")
    assert L.libsortDeviceSortDyn(recv.data_ptr(), out.data_ptr(), tmp.data_ptr(), cap, recv[cap:].data_ptr(),
                                  ws2.data_ptr(), ws2.numel(), None, None) == 0
    torch.cuda.synchronize()                                       # memory-safe: clamped to cap keys


@pytest.mark.parametrize("W,kind", [(1, "uniform"), (2, "uniform"), (8, "uniform"), (16, "uniform"), (4, "lowbits"), (8, "clustered")])
def test_splitters_from_histograms(gpu_lib, oracle, W, kind):
    """plan="hist": 4096-bin top-bit histograms -> bin-aligned splitters + the exact count matrix, then
    the planned exchange with them (virtual ranks).  Checked against numpy on the same keys."""
    import torch
    L = gpu_lib
    n = 60000 + 11 * W
    stream = oracle.pcg_stream(n * W)
    if kind == "lowbits":
        stream = stream & np.uint32(0x000FFFFF)                    # one top-bit bin holds everything
    elif kind == "clustered":
        stream = (stream >> np.uint32(3)) + np.uint32(0x60000000)  # an eighth of the key space
    shards = [stream[r * n:(r + 1) * n].copy() for r in range(W)]
    d_keys = [torch.from_numpy(s.view(np.int32)).cuda() for s in shards]
    hists = torch.zeros((W, 4096), dtype=torch.int32, device="cuda")
    for r in range(W):
        hists[r] = 77                                              # the call zeroes its output first
        assert L.libsortDeviceTopHistogram(d_keys[r].data_ptr(), n, hists[r].data_ptr(), None) == 0
    want_h = np.stack([np.bincount(s >> np.uint32(20), minlength=4096) for s in shards])
    assert np.array_equal(hists.cpu().numpy(), want_h)
    spl = torch.full((32,), -1, dtype=torch.int32, device="cuda")
    cm = torch.zeros((W, 16), dtype=torch.int32, device="cuda")
    imb = torch.zeros(1, dtype=torch.int32, device="cuda")
    assert L.libsortDeviceSplittersFromHistograms(hists.data_ptr(), W, spl.data_ptr(), cm.data_ptr(), imb.data_ptr(), None) == 0
    h_spl = spl.cpu().numpy().view(np.uint32)
    s_val = h_spl[:W - 1].astype(np.uint64)
    assert np.all(h_spl[16:] == 0) and np.all(s_val % (1 << 20) == 0) and np.all(np.diff(s_val.astype(np.int64)) >= 0)
    want_cm = np.zeros((W, 16), np.int64)
    for r in range(W):
        dest = np.searchsorted(s_val, shards[r].astype(np.uint64), side="right")   # number of splitters <= key
        want_cm[r, :W] = np.bincount(dest, minlength=W)
    assert np.array_equal(cm.cpu().numpy().astype(np.int64), want_cm)
    totals = want_cm.sum(axis=0)[:W]
    assert int(imb.item()) == int(totals.max()) * 1000 * W // (n * W)
    if kind == "uniform":
        assert int(imb.item()) <= 1030                             # 4096 bins balance uniform keys
    elif kind == "lowbits" and W > 1:
        assert int(imb.item()) == 1000 * W                         # cannot be balanced: the caller falls back to sampling
    # the planned exchange accepts table and matrix as they are
    cap = n * W + 64
    recv = torch.full((W, cap), -1, dtype=torch.int32, device="cuda")
    bases = torch.tensor([recv[r].data_ptr() for r in range(W)], dtype=torch.int64, device="cuda")
    ws = torch.empty(max(L.libsortWorkspaceBytes(n), 1 << 20), dtype=torch.uint8, device="cuda")
    nrecv = torch.zeros(W, dtype=torch.int32, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    for r in range(W):
        assert L.libsortDeviceExchangePlanned(d_keys[r].data_ptr(), n, spl.data_ptr(), W, r, cm.data_ptr(), bases.data_ptr(),
                                              cap, nrecv[r:].data_ptr(), status.data_ptr(), ws.data_ptr(), ws.numel(), None) == 0
    torch.cuda.synchronize()
    assert int(status.item()) == 0 and np.array_equal(nrecv.cpu().numpy().astype(np.int64), totals)
    h = recv.cpu().numpy().view(np.uint32)
    got = np.concatenate([np.sort(h[r, :totals[r]]) for r in range(W)])
    assert np.array_equal(got, np.sort(np.concatenate(shards)))
