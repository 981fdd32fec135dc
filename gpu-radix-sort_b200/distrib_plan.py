"""Host-side planning of the distributed sort (pure numpy; no GPU needed).

Replaces the bookkeeping of the reference's distributed driver
(benchmark/pkg/sort/distrib.go:107-158: slices per worker, bucket sizes from
boundaries, bucket-major regather order in helpers.go:67-78) for the
single-exchange design: keys are cut into `world` contiguous value ranges by
world-1 splitters; rank r receives every key whose destination is r.
"""
import numpy as np

MAX_RANKS = 16   # SplitBinner holds at most 15 splitters


def choose_splitters(sorted_sample, world):
    """world-1 ascending uint32 splitters: the g/world quantiles of a sorted sample.
    (Mirrors k_pick_splitters on the device.)"""
    s = np.asarray(sorted_sample, dtype=np.uint32)
    if world < 1 or world > MAX_RANKS:
        raise ValueError("world size must be in 1..%d" % MAX_RANKS)
    if s.size == 0:
        return np.zeros(world - 1, dtype=np.uint32)
    idx = (np.arange(1, world, dtype=np.uint64) * np.uint64(s.size)) // np.uint64(world)
    return s[idx.astype(np.int64)]


def destination(keys, splitters):
    """Destination rank of every key = number of splitters <= key (SplitBinner::bin)."""
    return np.searchsorted(np.asarray(splitters, dtype=np.uint32), np.asarray(keys, dtype=np.uint32), side="right")


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
