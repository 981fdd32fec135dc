"""Host-side planning of the distributed sort (pure numpy; no GPU needed).

Replaces the bookkeeping of the reference's distributed driver
(benchmark/pkg/sort/distrib.go:107-158: slices per worker, bucket sizes from
boundaries, bucket-major regather order in helpers.go:67-78) for the
single-exchange design: keys are cut into `world` contiguous value ranges by
world-1 splitters; rank r receives every key whose destination is r.
"""
import numpy as np

MAX_RANKS = 16   # SplitBinner holds at most 15 splitters


TIE_MIN_RUN = 8          # k_pick_splitters: runs of equal samples shorter than this are not split
TIE_HASH = 0x9E3779B1    # tie_hash(position) = position * TIE_HASH mod 2^32


def choose_splitters(sorted_sample, world, with_ties=False):
    """world-1 ascending uint32 splitters: the g/world quantiles of a sorted sample
    (mirrors k_pick_splitters on the device).  with_ties=True also returns the tie
    thresholds: where quantile index q falls inside a run [r0, r1) of >= TIE_MIN_RUN
    equal samples, threshold = floor((q - r0) * 2^32 / (r1 - r0)), else 0."""
    s = np.asarray(sorted_sample, dtype=np.uint32)
    if world < 1 or world > MAX_RANKS:
        raise ValueError("world size must be in 1..%d" % MAX_RANKS)
    if s.size == 0:
        z = np.zeros(world - 1, dtype=np.uint32)
        return (z, z.copy()) if with_ties else z
    idx = ((np.arange(1, world, dtype=np.uint64) * np.uint64(s.size)) // np.uint64(world)).astype(np.int64)
    spl = s[idx]
    if not with_ties:
        return spl
    r0 = np.searchsorted(s, spl, side="left")
    r1 = np.searchsorted(s, spl, side="right")
    thr = np.zeros(world - 1, dtype=np.uint32)
    for j in range(world - 1):
        run, below = int(r1[j] - r0[j]), int(idx[j] - r0[j])
        if run >= TIE_MIN_RUN and below > 0 and int(spl[j]) != 0xFFFFFFFF:
            thr[j] = (below << 32) // run
    return spl, thr


def destination(keys, splitters, thresholds=None, positions=None):
    """Destination rank of every key (SplitBinner::bin): the number of splitters s_j with
    key > s_j, or key == s_j and tie_hash(position) >= threshold_j.  Without thresholds
    that is the number of splitters <= key."""
    k = np.asarray(keys, dtype=np.uint32)
    spl = np.asarray(splitters, dtype=np.uint32)
    if thresholds is None or not np.any(thresholds):
        return np.searchsorted(spl, k, side="right")
    thr = np.asarray(thresholds, dtype=np.uint32)
    pos = np.arange(k.size, dtype=np.uint64) if positions is None else np.asarray(positions, dtype=np.uint64)
    h = ((pos * np.uint64(TIE_HASH)) & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    d = np.zeros(k.size, dtype=np.int64)
    for j in range(spl.size):
        d += (k > spl[j]) | ((k == spl[j]) & (h >= thr[j]))
    return d


def recv_layout(count_matrix):
    """count_matrix[s][r] = keys rank s sends to rank r.

    Returns (offsets, totals): at destination r, the block from source s starts
    at element offsets[s][r] (sources in rank order, no gaps); totals[r] keys arrive."""
    c = np.asarray(count_matrix, dtype=np.int64)
    offsets = np.cumsum(c, axis=0) - c
    return offsets, c.sum(axis=0)


def imbalance(totals):
    t = np.asarray(totals, dtype=np.float64)
    mean = t.mean() if t.size else 0.0
    return float(t.max() / mean) if mean > 0 else 1.0


# ---- step-wise distributed LSD sort (DistributedSorter.sort_stepwise; k_step_bases / OUT_SLICED on the device) ----
def step_bases(count_matrix, me):
    """count_matrix[k][d] = keys of digit d on rank k.  After the step the whole array is bucket-major,
    rank-minor (reference localTest/benchmarks.cpp:120-136, benchmark/pkg/sort/helpers.go:67-78), so rank
    `me`'s first key of digit d goes to global position
        sum over digits < d of every rank's count  +  sum over ranks < me of their digit-d count."""
    c = np.asarray(count_matrix, dtype=np.int64)
    per_digit = c.sum(axis=0)
    return (np.cumsum(per_digit) - per_digit) + c[:me].sum(axis=0)


def step_positions(keys, shift, nbits, bases):
    """Global position of every key of this rank's slice after a step: bases[digit] + its rank among
    the slice's keys of the same digit (stable).  Mirrors the OUT_SLICED scatter pass."""
    k = np.asarray(keys, dtype=np.uint32)
    d = ((k >> np.uint32(shift)) & np.uint32((1 << nbits) - 1)).astype(np.int64)
    order = np.argsort(d, kind="stable")
    counts = np.bincount(d, minlength=1 << nbits)
    first = np.cumsum(counts) - counts
    pos = np.empty(k.size, dtype=np.int64)
    pos[order] = np.asarray(bases, dtype=np.int64)[d[order]] + (np.arange(k.size) - first[d[order]])
    return pos


def slice_of(positions, slice_starts):
    """(owner, offset inside the owner's slice) of global positions; slice j holds [slice_starts[j], slice_starts[j+1])."""
    starts = np.asarray(slice_starts, dtype=np.int64)
    owner = np.searchsorted(starts[1:], positions, side="right")
    return owner, np.asarray(positions, dtype=np.int64) - starts[owner]
