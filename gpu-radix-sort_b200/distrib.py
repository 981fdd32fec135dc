"""One-rank-per-GPU distributed uint32 sort over NVLink.

Replaces the reference's host-mediated, 32/w-step LSD shuffle
(localTest/benchmarks.cpp:70-160; benchmark/pkg/sort/distrib.go:90-179) with a
single exchange:

  1. sample the shard (k_sample), all-gather the samples (NCCL), sort them with
     the device sort and take the g/P quantiles as splitters (on device)
  2. count keys per destination (k_dest_hist), all-gather the P x P count
     matrix (NCCL) -> receive layout
  3. exchange, one of
       p2p : the receive layout is derived ON THE DEVICE from the gathered matrix
             (k_plan_exchange), then ONE kernel partitions the shard and stores every
             key straight into its destination GPU's receive buffer through
             cudaIpc/NVLink-mapped pointers (k_onesweep_bulk<SplitBinner, PTR_OUT>) --
             partition + all-to-all fused, no host round trip anywhere in the sort
             plan="cursor" (LIBSORT_EXCHANGE_PLAN=cursor) drops step 2: every receive buffer carries a
             cursor word, a tile reserves room for its run at the destination with one system-scope
             atomicAdd on that word over NVLink (libsortDeviceExchangeCursor), and after the barrier
             the word is the number of keys received.  Measured: 2 GPUs 154.8 vs 152.4 Gkeys/s, but
             8 GPUs 445.7 vs 504.1 -- 350 K atomics per sort on ONE word per destination serialise at
             its home L2 (exchange 2.28 ms vs 1.48 ms) -- so the counted layout stays the default.
       nccl: host-planned: partition locally with the same kernel, then NCCL all_to_all_single
  4. local LSD sort of what arrived (4 passes); p2p: libsortDeviceSortDyn, the number of
     keys that arrived is read by the kernels from device memory

sort_stepwise() is the reference's OWN distributed algorithm on the same plumbing (SURVEY.md 8(f)
N2): one global counting sort per 8-bit digit, 32/8 exchanges, every key stored straight into the
slice of the GPU that owns its global position.  Same intermediate states as the reference after
every step, perfectly balanced output (slices keep their sizes) for any input, no sampling -- but
four trips over NVLink instead of one.

Rank r ends up with the r-th value range, ascending.  torch.distributed is the
plumbing (collectives, rendezvous); all key movement and sorting on the GPU is
libsort.so.  A key value heavy enough to span a splitter has its copies cut
between the neighbouring ranks by position (tie thresholds, see SplitBinner in
csrc/kernels.cuh), so even an all-equal input balances; verify() reports
`imbalance` = largest range / mean.
"""
import ctypes
import os

import numpy as np
import torch
import torch.distributed as dist

from . import binding, distrib_plan
from .binding import KernelTimes

SAMPLES_PER_RANK = 1 << 13   # 8 K per rank: the quantiles of P x 8 K samples balance to about 1 %
HIST_IMBALANCE_LIMIT = 1.03  # plan="auto": above this the histogram plan gives way to sampled splitters


def _u32_to_i32(v64):
    return torch.where(v64 >= (1 << 31), v64 - (1 << 32), v64).to(torch.int32)


def shard_of_pcg_stream(rank, n, transform=None):
    """Rank's shard [rank*n, (rank+1)*n) of the one continuous PCG input stream, on the GPU."""
    h = np.empty(n, dtype=np.uint32)
    binding.lib().libsortPopulateInputAt(h.ctypes.data, n, rank * n)
    t = torch.from_numpy(h.view(np.int32)).cuda()
    x = t.to(torch.int64) & 0xFFFFFFFF
    if transform == "zipf":        # log-uniform: k = floor(2^(32 x / 2^32)) - 1   (SURVEY.md 8(d) c5 i)
        k = torch.floor(torch.exp2(32.0 * x.to(torch.float64) / 4294967296.0)) - 1
        t = _u32_to_i32(k.clamp_(0, 4294967295).to(torch.int64))
    elif transform == "consthigh":  # 0xAB000000 | (x & 0x00FFFFFF)                 (c5 ii)
        t = _u32_to_i32((x & 0x00FFFFFF) | 0xAB000000)
    elif transform == "consthigh16":
        t = _u32_to_i32((x & 0xFFFF) | 0xABCD0000)
    elif transform == "allequal":   # every key the same: ties must be cut between ranks
        t = torch.full_like(t, 0x12345678)
    elif transform == "halfdup":    # every other key is one value, the rest uniform
        t = torch.where((x & 1) == 1, torch.full_like(t, 0x40000000), t)
    elif transform not in (None, "uniform"):
        raise ValueError(transform)
    return t.contiguous()


class _RawCuda:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i4", "data": (ptr, False), "version": 2}


def _as_tensor(ptr, n, device):
    return torch.as_tensor(_RawCuda(ptr, n), device=device)


class DistributedSorter:
    def __init__(self, n_local_max, group=None, mode="p2p", slack=1.25, plan=None):
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        if self.world > distrib_plan.MAX_RANKS:
            raise ValueError("at most %d ranks" % distrib_plan.MAX_RANKS)
        if mode not in ("p2p", "nccl"):
            raise ValueError(mode)
        self.mode = mode
        # "auto" (default): splitters and counts from one 4096-bin histogram per shard (plan "hist") as
        # long as that balances the ranks to within HIST_IMBALANCE_LIMIT; otherwise, from the next sort
        # on, sampled splitters with tie thresholds + a counting pass ("counts").  "cursor": see above.
        self.plan_arg = plan or os.environ.get("LIBSORT_EXCHANGE_PLAN", "auto")
        if self.plan_arg not in ("auto", "hist", "cursor", "counts"):
            raise ValueError(self.plan_arg)
        self.plan = "hist" if self.plan_arg == "auto" else self.plan_arg
        self.lib = binding.lib()
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.n_max = n_local_max
        self.cap = int(n_local_max * slack) + (1 << 16)
        L, dev, i32 = self.lib, self.dev, torch.int32
        # receive buffer: own cudaMalloc so that its cudaIpc handle maps at offset 0
        p = ctypes.c_void_p()
        binding.check(L.libsortIpcAlloc((self.cap + 16) * 4, ctypes.byref(p)), "libsortIpcAlloc")
        self.recv_ptr = p.value
        self.recv = _as_tensor(self.recv_ptr, self.cap, dev)
        self.cursor = _as_tensor(self.recv_ptr + 4 * self.cap, 16, dev)   # word 0: keys received (plan="cursor")
        self.cursor.zero_()
        self.peer_ptrs = [self.recv_ptr] * self.world
        self._opened = []
        if mode == "p2p" and self.world > 1:
            handle = (ctypes.c_uint8 * 64)()
            binding.check(L.libsortIpcExport(self.recv_ptr, handle), "libsortIpcExport")
            mine = torch.tensor(list(handle), dtype=torch.uint8, device=dev)
            allh = torch.empty(self.world * 64, dtype=torch.uint8, device=dev)
            dist.all_gather_into_tensor(allh, mine, group=self.group)
            allh = allh.cpu().numpy().reshape(self.world, 64)
            for r in range(self.world):
                if r == self.rank:
                    continue
                q = ctypes.c_void_p()
                buf = (ctypes.c_uint8 * 64)(*allh[r].tolist())
                binding.check(L.libsortIpcOpen(buf, ctypes.byref(q)), "libsortIpcOpen(rank %d)" % r)
                self.peer_ptrs[r] = q.value
                self._opened.append(q.value)
        if self.world > 1:
            torch.cuda.current_stream(dev).synchronize()
            dist.barrier(group=self.group)          # every cursor is zero before anyone's first exchange
        self.part = torch.empty(self.n_max, dtype=i32, device=dev) if mode == "nccl" else None
        self.out = torch.empty(self.cap, dtype=i32, device=dev)
        self.tmp = torch.empty(self.cap, dtype=i32, device=dev)
        self.ws = torch.empty(L.libsortWorkspaceBytes(self.cap), dtype=torch.uint8, device=dev)
        S = SAMPLES_PER_RANK
        self.samples = torch.empty(S, dtype=i32, device=dev)
        self.all_samples = torch.empty(S * self.world, dtype=i32, device=dev)
        self.samples_sorted = torch.empty_like(self.all_samples)
        self.samples_tmp = torch.empty_like(self.all_samples)
        self.ws_small = torch.empty(L.libsortWorkspaceBytes(S * self.world), dtype=torch.uint8, device=dev)
        self.splitters = torch.zeros(32, dtype=i32, device=dev)   # 16 values + 16 tie thresholds
        self.counts = torch.zeros(16, dtype=i32, device=dev)
        self.count_matrix = torch.zeros(16 * self.world, dtype=i32, device=dev)
        self.token = torch.zeros(1, dtype=i32, device=dev)
        self.h_counts = torch.empty(16 * self.world, dtype=i32).pin_memory()
        self.d_peer_bases = torch.tensor(self.peer_ptrs, dtype=torch.int64, device=dev)
        self.top_hist = torch.zeros(4096, dtype=i32, device=dev)
        self.all_top_hists = torch.zeros(4096 * self.world, dtype=i32, device=dev)
        self.d_imbalance = torch.full((1,), 1000, dtype=i32, device=dev)
        self.d_nrecv = torch.zeros(1, dtype=i32, device=dev)
        self.d_status = torch.zeros(1, dtype=i32, device=dev)
        self.seed = 1
        self._step = None      # lazily built state of sort_stepwise()
        # kernels per sort: sample, sample sort (hist, scan, 4 passes), pick, dest hist, store ptrs,
        # scatter, local sort (hist, scan, 4 passes)
        # p2p: sample, sample sort (hist, scan, 4 passes), pick, dest hist, plan, scatter, local sort (hist, scan, 4 passes)
        self.launches_per_stepwise_sort = 4 * 3    # per step: histogram, bases, scatter
        self.last_totals = None
        self.rank_violation = False
        self.hist_imbalance = None

    @property
    def launches_per_sort(self):
        """Kernels of ours per sort in the current plan (bench.py's gpu_launches)."""
        local = 6                                   # histogram, scan, 4 scatter passes
        if self.mode != "p2p" or self.plan == "counts":
            return 1 + 6 + 1 + 1 + 1 + 1 + local    # sample, sample sort, pick, dest hist, plan / store ptrs, scatter
        if self.plan == "cursor":
            return 1 + 6 + 1 + 1 + 1 + local        # sample, sample sort, pick, store plan, scatter
        return 1 + 1 + 1 + 1 + local                # top histogram, splitters from histograms, plan, scatter

    def reset_plan(self):
        """plan="auto": forget that an earlier input needed sampled splitters (new data, new try)."""
        self.plan = "hist" if self.plan_arg == "auto" else self.plan_arg

    def close(self):
        if self._step is not None:
            for p in self._step["opened"]:
                self.lib.libsortIpcClose(p)
            self.lib.libsortIpcFree(self._step["recv2_ptr"])
            self._step = None
        for p in self._opened:
            self.lib.libsortIpcClose(p)
        self._opened = []
        if self.recv_ptr:
            self.recv = None
            self.lib.libsortIpcFree(self.recv_ptr)
            self.recv_ptr = None

    # ------------------------------------------------------------------ sort --
    def _stream(self):
        return torch.cuda.current_stream(self.dev).cuda_stream

    def _splitters(self, keys, ev=None):
        """Step 1: sample, all-gather, sort the sample, pick the quantiles (all on the device)."""
        L, st, n, W = self.lib, self._stream(), keys.numel(), self.world
        S = SAMPLES_PER_RANK
        self.seed += 1
        binding.check(L.libsortDeviceSample(keys.data_ptr(), n, self.samples.data_ptr(), S, self.seed, st), "sample")
        if W > 1:
            dist.all_gather_into_tensor(self.all_samples, self.samples, group=self.group)
        else:
            self.all_samples.copy_(self.samples)
        binding.check(L.libsortDeviceSort(self.all_samples.data_ptr(), self.samples_sorted.data_ptr(),
                                          self.samples_tmp.data_ptr(), S * W, 0, 32, None, 0, self.ws_small.data_ptr(),
                                          self.ws_small.numel(), st, None), "sample sort")
        binding.check(L.libsortDevicePickSplitters(self.samples_sorted.data_ptr(), S * W, W, self.splitters.data_ptr(), st),
                      "pick splitters")
        if ev:
            ev("splitters")

    def _plan(self, keys, ev=None, on_host=True):
        """Steps 1-2: splitters and the gathered count matrix on the device; on_host: also read the
        matrix back (one host sync -- only the NCCL all-to-all mode needs it)."""
        L, st, n, W = self.lib, self._stream(), keys.numel(), self.world
        self._splitters(keys, ev)
        binding.check(L.libsortDeviceSplitCounts(keys.data_ptr(), n, self.splitters.data_ptr(), W - 1,
                                                 self.counts.data_ptr(), st), "split counts")
        if W > 1:
            dist.all_gather_into_tensor(self.count_matrix, self.counts, group=self.group)
        else:
            self.count_matrix.copy_(self.counts)
        if not on_host:
            if ev:
                ev("counts")
            return None
        self.h_counts.copy_(self.count_matrix, non_blocking=True)
        torch.cuda.current_stream(self.dev).synchronize()
        cm = self.h_counts.numpy().reshape(W, 16)[:, :W].astype(np.int64)
        if ev:
            ev("counts")
        return cm

    def sort_async(self, keys, ev=None, times=None):
        """The p2p sort with nothing read back: returns (buffer holding this rank's range, DEVICE tensor
        with its length).  Everything -- receive layout, exchange, local sort on a device-side key
        count -- is enqueued on the current stream; the host never waits."""
        if self.mode != "p2p":
            raise ValueError("sort_async needs the p2p exchange")
        if keys.dtype not in (torch.int32, torch.uint32) or not keys.is_cuda or not keys.is_contiguous():
            raise ValueError("keys must be a contiguous CUDA int32/uint32 tensor")
        L, st, n, W, me = self.lib, self._stream(), keys.numel(), self.world, self.rank
        if n > self.n_max:
            raise ValueError("shard larger than the sorter was built for")
        if self.plan == "cursor":
            return self._sort_async_cursor(keys, ev, times)
        if self.plan == "hist":
            self._plan_from_histograms(keys, ev)
        else:
            self._plan(keys, ev, on_host=False)
        binding.check(L.libsortDeviceExchangePlanned(keys.data_ptr(), n, self.splitters.data_ptr(), W, me,
                                                     self.count_matrix.data_ptr(), self.d_peer_bases.data_ptr(), self.cap,
                                                     self.d_nrecv.data_ptr(), self.d_status.data_ptr(),
                                                     self.ws.data_ptr(), self.ws.numel(), st), "planned exchange")
        if W > 1:
            # stream-ordered barrier: nobody starts sorting before every rank's stores have landed
            dist.all_reduce(self.token, group=self.group)
        if ev:
            ev("exchange")
        binding.check(L.libsortDeviceSortDyn(self.recv_ptr, self.out.data_ptr(), self.tmp.data_ptr(), self.cap,
                                             self.d_nrecv.data_ptr(), self.ws.data_ptr(), self.ws.numel(), st,
                                             ctypes.byref(times) if times is not None else None), "local sort")
        # No fence here: the receive buffer is next written by the NEXT sort's exchange, and every exchange
        # comes after that sort's all-gather (histograms or samples), which no rank can leave before every
        # rank -- stream-ordered behind its local sort above -- has entered it.
        if ev:
            ev("local_sort")
        return self.out, self.d_nrecv

    def _plan_from_histograms(self, keys, ev=None):
        """plan="hist": one read of the shard (4096-bin histogram of the top 12 bits), one all-gather,
        one small kernel -> splitters on bin boundaries + the exact count matrix.  No sample, no sample
        sort, no counting pass."""
        L, st, n, W = self.lib, self._stream(), keys.numel(), self.world
        binding.check(L.libsortDeviceTopHistogram(keys.data_ptr(), n, self.top_hist.data_ptr(), st), "top histogram")
        if W > 1:
            dist.all_gather_into_tensor(self.all_top_hists, self.top_hist, group=self.group)
        else:
            self.all_top_hists.copy_(self.top_hist)
        if ev:
            ev("splitters")
        binding.check(L.libsortDeviceSplittersFromHistograms(self.all_top_hists.data_ptr(), W, self.splitters.data_ptr(),
                                                             self.count_matrix.data_ptr(), self.d_imbalance.data_ptr(), st),
                      "splitters from histograms")
        if ev:
            ev("counts")

    def _sort_async_cursor(self, keys, ev, times):
        """splitters -> exchange with remote cursors -> local sort on the cursor's value.  No counting
        pass, no count all-gather, nothing read back."""
        L, st, n, W = self.lib, self._stream(), keys.numel(), self.world
        self._splitters(keys, ev)
        if ev:
            ev("counts")        # (phase kept for the reports: empty in this mode)
        binding.check(L.libsortDeviceExchangeCursor(keys.data_ptr(), n, self.splitters.data_ptr(), W,
                                                    self.d_peer_bases.data_ptr(), self.cap, self.d_status.data_ptr(),
                                                    self.ws.data_ptr(), self.ws.numel(), st), "cursor exchange")
        if W > 1:
            dist.all_reduce(self.token, group=self.group)   # every rank's stores (and cursor updates) have landed
        if ev:
            ev("exchange")
        binding.check(L.libsortDeviceSortDyn(self.recv_ptr, self.out.data_ptr(), self.tmp.data_ptr(), self.cap,
                                             self.cursor.data_ptr(), self.ws.data_ptr(), self.ws.numel(), st,
                                             ctypes.byref(times) if times is not None else None), "local sort")
        self.d_nrecv.copy_(self.cursor[:1])
        self.cursor[:1].zero_()                             # ready for the next exchange, which nobody starts before
        #                                                     the next sort's sample all-gather (see sort_async)
        if ev:
            ev("local_sort")
        return self.out, self.d_nrecv

    def _finish(self):
        """Read back what sort_async left on the device: this rank's key count, every rank's count,
        the overflow flag and the ranking audit (one synchronisation, after everything has been
        enqueued; the ranks agree on the two verdicts so that they raise or repeat together)."""
        W = self.world
        # atomic-ranking audit (include/libsort_b200.h libsortRankAudit): a violation on ANY rank makes
        # every rank's result suspect
        rc = self.lib.libsortRankAudit(self._stream())
        if rc < 0:
            binding.check(rc, "libsortRankAudit")
        switched = False
        if self.plan == "hist" and self.plan_arg == "auto":
            # every rank computed the same figure from the same gathered histograms: no collective needed
            self.hist_imbalance = int(self.d_imbalance.item()) / 1000.0
            if self.hist_imbalance > HIST_IMBALANCE_LIMIT:
                self.plan = "counts"        # bin-aligned splitters cannot balance this input: sample from now on
                switched = True
        mine = torch.stack([self.d_nrecv[0], self.d_status[0], torch.tensor(rc, dtype=torch.int32, device=self.dev)])
        if W > 1:
            allv = torch.empty(3 * W, dtype=torch.int32, device=self.dev)
            dist.all_gather_into_tensor(allv, mine, group=self.group)
        else:
            allv = mine
        v = allv.cpu().numpy().reshape(W, 3).astype(np.int64) & 0xFFFFFFFF
        self.last_totals = v[:, 0].copy()
        nrecv = int(v[self.rank, 0])
        if v[:, 1].any() or int(v[:, 0].max()) > self.cap:
            self.d_status.zero_()
            if switched:
                return -1               # overflow under bin-aligned splitters: sort() repeats with sampled ones
            raise RuntimeError("receive buffer too small: %d keys arrive at one rank, capacity %d (raise slack)"
                               % (int(v[:, 0].max()), self.cap))
        self.rank_violation = bool(v[:, 2].any())
        return nrecv

    def sort(self, keys, ev=None, times=None):
        """Sort the union of all ranks' `keys`; returns (tensor view of this rank's range, its length)."""
        if keys.dtype not in (torch.int32, torch.uint32) or not keys.is_cuda or not keys.is_contiguous():
            raise ValueError("keys must be a contiguous CUDA int32/uint32 tensor")
        L, st, n, W, me = self.lib, self._stream(), keys.numel(), self.world, self.rank
        if n > self.n_max:
            raise ValueError("shard larger than the sorter was built for")
        if self.mode == "p2p":
            # `keys` is untouched, so the sort can simply be repeated: after a ranking violation (ballot
            # kernels from then on) or when the histogram plan overflowed a receive buffer (sampled
            # splitters from then on)
            for _attempt in range(3):
                out, _ = self.sort_async(keys, ev=ev, times=times)
                nrecv = self._finish()
                if nrecv >= 0 and not self.rank_violation:
                    break
            if nrecv < 0:
                raise RuntimeError("receive buffer too small (raise slack)")
            return out[:nrecv], nrecv
        cm = self._plan(keys, ev)
        offsets, totals = distrib_plan.recv_layout(cm)
        self.last_totals = totals
        if int(totals.max()) > self.cap:
            raise RuntimeError("receive buffer too small: %d keys arrive, capacity %d (raise slack)" % (int(totals.max()), self.cap))
        nrecv = int(totals[me])
        ptrs = (ctypes.c_uint64 * W)()
        if self.mode == "p2p":
            for r in range(W):
                ptrs[r] = self.peer_ptrs[r] + 4 * int(offsets[me][r])
        else:
            send_off = np.cumsum(cm[me]) - cm[me]
            for r in range(W):
                ptrs[r] = self.part.data_ptr() + 4 * int(send_off[r])
        binding.check(L.libsortDeviceScatterToPeers(keys.data_ptr(), n, self.splitters.data_ptr(), W - 1, ptrs,
                                                    self.ws.data_ptr(), self.ws.numel(), st), "scatter")
        if W > 1:
            if self.mode == "p2p":
                # stream-ordered barrier: nobody starts sorting before every rank's stores have landed
                dist.all_reduce(self.token, group=self.group)
            else:
                dist.all_to_all_single(self.recv[:nrecv], self.part[:n], output_split_sizes=cm[:, me].tolist(),
                                       input_split_sizes=cm[me].tolist(), group=self.group)
        if ev:
            ev("exchange")
        binding.check(L.libsortDeviceSort(self.recv_ptr, self.out.data_ptr(), self.tmp.data_ptr(), nrecv, 0, 32, None, 0,
                                          self.ws.data_ptr(), self.ws.numel(), st,
                                          ctypes.byref(times) if times is not None else None), "local sort")
        if W > 1 and self.mode == "p2p":
            # the receive buffer is overwritten by the next sort's remote stores: fence it
            dist.all_reduce(self.token, group=self.group)
        if ev:
            ev("local_sort")
        return self.out[:nrecv], nrecv

    # ------------------------------------------------------------- step-wise --
    def _map_on_peers(self, ptr):
        """cudaIpc-export a local allocation and open everybody else's: [address of rank r's buffer as seen here]."""
        L, W = self.lib, self.world
        ptrs, opened = [ptr] * W, []
        if W > 1:
            handle = (ctypes.c_uint8 * 64)()
            binding.check(L.libsortIpcExport(ptr, handle), "libsortIpcExport")
            mine = torch.tensor(list(handle), dtype=torch.uint8, device=self.dev)
            allh = torch.empty(W * 64, dtype=torch.uint8, device=self.dev)
            dist.all_gather_into_tensor(allh, mine, group=self.group)
            allh = allh.cpu().numpy().reshape(W, 64)
            for r in range(W):
                if r != self.rank:
                    q = ctypes.c_void_p()
                    binding.check(L.libsortIpcOpen((ctypes.c_uint8 * 64)(*allh[r].tolist()), ctypes.byref(q)),
                                  "libsortIpcOpen(rank %d)" % r)
                    ptrs[r] = q.value
                    opened.append(q.value)
        return ptrs, opened

    def _enable_stepwise(self):
        if self._step is not None:
            return self._step
        if self.mode != "p2p":
            raise ValueError("sort_stepwise needs the p2p exchange (peer-mapped slices)")
        L, dev, i32, W = self.lib, self.dev, torch.int32, self.world
        p = ctypes.c_void_p()
        binding.check(L.libsortIpcAlloc(self.cap * 4, ctypes.byref(p)), "libsortIpcAlloc")
        peers2, opened = self._map_on_peers(p.value)
        tb = L.libsortStepTableBytes()
        self._step = {
            "recv2_ptr": p.value, "recv2": _as_tensor(p.value, self.cap, dev), "opened": opened,
            "peers": [self.peer_ptrs, peers2],              # ping-pong: slices live in recv, then in recv2, ...
            "tables": [torch.zeros(tb, dtype=torch.uint8, device=dev) for _ in range(2)],
            "counts": torch.zeros(256, dtype=i32, device=dev), "gathered": torch.zeros(256 * W, dtype=i32, device=dev),
            "lens": torch.zeros(W, dtype=torch.int64, device=dev), "layout": None,
        }
        return self._step

    def sort_stepwise(self, keys, ev=None, on_step=None):
        """The reference's step-wise distributed LSD sort (8-bit steps) with device-resident slices.
        Rank r keeps exactly keys.numel() keys: afterwards they are the r-th slice of the sorted whole.
        Returns (tensor view of the slice, its length); valid until the next sort on this object.
        on_step(k, view) is called after step k with this rank's slice of the intermediate array."""
        if keys.dtype not in (torch.int32, torch.uint32) or not keys.is_cuda or not keys.is_contiguous():
            raise ValueError("keys must be a contiguous CUDA int32/uint32 tensor")
        L, st, n, W, me = self.lib, self._stream(), keys.numel(), self.world, self.rank
        if n > self.cap:
            raise ValueError("slice larger than the sorter was built for")
        S = self._enable_stepwise()
        if W > 1:
            dist.all_gather_into_tensor(S["lens"], torch.tensor([n], dtype=torch.int64, device=self.dev), group=self.group)
            lens = S["lens"].cpu().tolist()
        else:
            lens = [n]
        if S["layout"] != lens:
            starts = [0]
            for x in lens:
                starts.append(starts[-1] + x)
            arr_s = (ctypes.c_uint64 * (W + 1))(*starts)
            for k in range(2):
                arr_p = (ctypes.c_uint64 * W)(*S["peers"][k])
                binding.check(L.libsortStepTableInit(S["tables"][k].data_ptr(), arr_p, arr_s, W, st),
                              "slice table (every slice but the last needs >= %d keys)" % L.libsortTileKeys())
            S["layout"] = lens
        src, bufs = keys.data_ptr(), [self.recv, S["recv2"]]
        shift = k = 0
        while shift < 32:
            nb = min(8, 32 - shift)
            table = S["tables"][k & 1]
            binding.check(L.libsortStepCounts(src, n, shift, nb, S["counts"].data_ptr(), st), "step counts")
            if W > 1:
                dist.all_gather_into_tensor(S["gathered"], S["counts"], group=self.group)
            else:
                S["gathered"].copy_(S["counts"])
            binding.check(L.libsortStepBases(S["gathered"].data_ptr(), W, me, table.data_ptr(), st), "step bases")
            binding.check(L.libsortStepScatter(src, n, shift, nb, table.data_ptr(), self.ws.data_ptr(), self.ws.numel(), st),
                          "step scatter")
            if W > 1:
                # every rank's stores of this step have landed, and nobody still reads the buffer the
                # next step overwrites
                dist.all_reduce(self.token, group=self.group)
            out = bufs[k & 1]
            src, shift, k = out.data_ptr(), shift + nb, k + 1
            if ev:
                ev("step%d" % k)
            if on_step:
                on_step(k, out[:n])
        self.last_totals = np.asarray(lens)
        return out[:n], n

    # ---------------------------------------------------------- verification --
    def verify(self, keys, out, nrecv):
        """Size-independent properties of a distributed sort: every rank's range is
        ascending, ranges are ordered across ranks, and the multiset is preserved
        (sum / sum of squares mod p / count, all-reduced)."""
        W = self.world
        o64 = out.to(torch.int64) & 0xFFFFFFFF
        i64 = keys.to(torch.int64) & 0xFFFFFFFF
        asc = bool((o64[1:] >= o64[:-1]).all()) if nrecv > 1 else True
        lo = int(o64[0]) if nrecv else (1 << 32)
        hi = int(o64[-1]) if nrecv else -1
        edge = torch.tensor([lo, hi, nrecv, int(asc)], dtype=torch.int64, device=self.dev)
        edges = torch.empty(4 * W, dtype=torch.int64, device=self.dev)
        fp = torch.stack([i64.sum(), (i64 * i64 % 1000003).sum(), torch.tensor(keys.numel(), device=self.dev),
                          o64.sum(), (o64 * o64 % 1000003).sum(), torch.tensor(nrecv, device=self.dev)]).to(torch.int64)
        if W > 1:
            dist.all_gather_into_tensor(edges, edge, group=self.group)
            dist.all_reduce(fp, group=self.group)
        else:
            edges.copy_(edge)
        e = edges.cpu().numpy().reshape(W, 4)
        fp = fp.cpu().numpy()
        ordered, prev_hi = True, -1
        for r in range(W):
            if e[r][2] == 0:
                continue
            ordered &= bool(e[r][0] >= prev_hi)
            prev_hi = int(e[r][1])
        ok = bool(e[:, 3].all()) and ordered and bool((fp[:3] == fp[3:]).all())
        if not ok:
            raise AssertionError("distributed sort verification failed: edges=%s fp=%s" % (e.tolist(), fp.tolist()))
        counts = e[:, 2].tolist()
        return {"ok": True, "keys_per_rank": counts, "imbalance": distrib_plan.imbalance(counts)}

    def phase_times(self, keys):
        """CUDA-event time of each phase of one sort (ms, max over ranks)."""
        marks = []

        def ev(name):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            marks.append((name, e))
        if self.world > 1:
            dist.barrier(group=self.group)
        torch.cuda.synchronize()
        ev("start")
        kt = KernelTimes()
        self.sort(keys, ev=ev, times=kt)
        torch.cuda.synchronize()
        names = [m[0] for m in marks[1:]]
        vals = torch.tensor([marks[i][1].elapsed_time(marks[i + 1][1]) for i in range(len(marks) - 1)],
                            dtype=torch.float64, device=self.dev)
        passes = [kt.pass_ms[i] for i in range(kt.npass)]
        nrecv = int(self.last_totals[self.rank])
        extra = torch.tensor([sum(passes) / max(len(passes), 1), kt.histogram_ms], dtype=torch.float64, device=self.dev)
        if self.world > 1:
            dist.all_reduce(vals, op=dist.ReduceOp.MAX, group=self.group)
            dist.all_reduce(extra, op=dist.ReduceOp.MAX, group=self.group)
        res = {k + "_ms": float(v) for k, v in zip(names, vals.tolist())}
        pass_ms = float(extra[0])
        res["local_pass_ms"] = pass_ms
        res["pass_gbs"] = 8.0 * nrecv / (pass_ms * 1e-3) / 1e9 if pass_ms > 0 else None
        return res

    def e2e(self, args, keys):
        """Same sort with HOST shards: pinned H2D of the shard and D2H of the sorted range
        inside the timed region (wall clock, max over ranks)."""
        import time
        n = keys.numel()
        h_in = keys.cpu().pin_memory()
        h_out = torch.empty(self.cap, dtype=torch.int32).pin_memory()
        d_in = torch.empty_like(keys)
        steps = max(1, min(args.steps, 5))
        tot = 0.0
        for it in range(steps + 1):
            if self.world > 1:
                dist.barrier(group=self.group)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            d_in.copy_(h_in, non_blocking=True)
            out, nrecv = self.sort(d_in)
            h_out[:nrecv].copy_(out, non_blocking=True)
            torch.cuda.synchronize()
            dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=self.dev)
            if self.world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX, group=self.group)
            if it > 0:
                tot += float(dt)
        total_keys = n * self.world
        return {"value": total_keys * steps / tot / 1e9, "unit": "Gkeys/s", "h2d_bytes_per_step": 4 * total_keys,
                "d2h_bytes_per_step": 4 * total_keys, "ms_per_step": tot / steps * 1e3, "steps": steps,
                "api": "DistributedSorter.sort with pinned host shards"}
