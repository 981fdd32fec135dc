// engine.cuh -- pass driver over the sm_100a kernels (device-resident data).
// Replaces SortState::{Step,GetBoundaries} (reference libsort/sort.cu:322-394).
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace b200sort {

enum BoundaryMode : int {
    BOUNDS_REFERENCE = 0,  // bit-exact with the reference's gpu_groups + back-fill
    BOUNDS_EXACT = 1,      // exclusive prefix sums of the group histogram
};

struct KernelTimes {        // milliseconds, filled when requested
    float memset_ms;
    float histogram_ms;
    float scan_ms;
    float pass_ms[4];
    float boundaries_ms;
    int npass;
    int launches;           // kernels launched by this call
};

// Audit of the atomic-ranking assumption: true if a kernel on the current device
// flagged an out-of-order rank since the last audit; the device then falls back to ballot ranking.
cudaError_t rank_audit(cudaStream_t stream, bool *violated);

// Keys per scatter tile of the compiled-in kernel variant.
int tile_keys();

// Scratch bytes needed to sort n keys (histograms, bucket bases, tile
// tickets, look-back descriptors).
size_t workspace_bytes(size_t n);

// Stable sort of `in` by bits [offset, offset + 2*ceil(width/2)) -- the net
// effect of SortState::Step(offset,width) -- using `a` and `b` (n keys each)
// as ping-pong buffers; `in` may be `a` or `b`, or a third read-only buffer.
// *result receives the buffer (a or b, or `in` itself when no pass is needed)
// that holds the sorted keys when the stream reaches this point; if `want` is
// non-null the result is forced into `want` (one extra device copy when the
// pass parity does not land there).  d_bounds (nullable, 1<<width entries)
// receives the group boundaries.  All work is enqueued on `stream`; nothing
// synchronises unless `times` is non-null.  dyn_n (nullable, DEVICE word): the key count is read on
// the device when the kernels run, `n` is then only the capacity the launches and the look-back
// descriptors are sized for (n < 2^30, no boundaries) -- the distributed sort's local sort, whose
// input size is known to the GPU (count matrix) before it is known to the host.
cudaError_t sort_bits(const uint32_t *in, uint32_t *a, uint32_t *b, uint32_t *want, size_t n, uint32_t offset,
                      uint32_t width, uint32_t *d_bounds, int bounds_mode, void *workspace, size_t workspace_len,
                      cudaStream_t stream, uint32_t **result, KernelTimes *times, const uint32_t *dyn_n = nullptr);

// Distributed-sort building blocks (one rank per GPU).
// sample_keys: nsamples stratified pseudo-random keys of the shard.
// pick_splitters: d_splitters[g-1] = g/nranks quantile of the sorted sample.
// split_counts: d_counts[0..nsplit] = keys per destination, destination of a
//   key = number of splitters <= key (DEVICE array, ascending, nsplit <= 15).
// scatter_by_splitters: stable partition of the shard by destination, written
//   straight to dest_ptrs[d] (device pointers, local or NVLink-mapped peer
//   memory) -- the fused partition + key exchange; n < 2^30.
cudaError_t sample_keys(const uint32_t *in, size_t n, uint32_t *out, uint32_t nsamples, uint32_t seed, cudaStream_t st);
cudaError_t pick_splitters(const uint32_t *sorted, size_t nsorted, int nranks, uint32_t *d_splitters, cudaStream_t st);
cudaError_t split_counts(const uint32_t *in, size_t n, const uint32_t *splitters, int nsplit, uint32_t *d_counts,
                         cudaStream_t st);
cudaError_t scatter_by_splitters(const uint32_t *in, size_t n, const uint32_t *splitters, int nsplit,
                                 const unsigned long long *dest_ptrs, void *workspace, size_t workspace_len,
                                 cudaStream_t st);

// Step-wise distributed LSD sort: the reference's P-worker protocol -- every
// step, each worker sorts its slice by one digit and the slices are regathered bucket-major,
// worker-minor (localTest/benchmarks.cpp:70-160, benchmark/pkg/sort/distrib.go:119-175) -- is one
// GLOBAL stable counting sort per digit.  Here the scatter pass of rank `me` writes each key
// straight to its global position, i.e. into the slice of the GPU that owns that position:
//   step_counts : digit histogram of the local slice (256 words)
//   (all-gather of the P x 256 counts: the caller's collective)
//   step_bases  : global position of this rank's first key of every digit -> table
//   step_scatter: the scatter pass with sliced output (peer stores over NVLink)
// The table (step_table_bytes() of device memory) maps positions to slices: slice j holds global
// positions [slice_starts[j], slice_starts[j+1]) at device address peers[j]; every slice but the
// last must be at least one scatter tile long.  n < 2^30 per rank, total < 2^32.
size_t step_table_bytes();
cudaError_t step_table_init(void *d_table, const unsigned long long *peers, const unsigned long long *slice_starts,
                            int nslice, cudaStream_t st);
cudaError_t step_counts(const uint32_t *in, size_t n, uint32_t shift, uint32_t nbits, uint32_t *d_counts, cudaStream_t st);
cudaError_t step_bases(const uint32_t *d_gathered, int nranks, int me, void *d_table, cudaStream_t st);
cudaError_t step_scatter(const uint32_t *in, size_t n, uint32_t shift, uint32_t nbits, const void *d_table,
                         void *workspace, size_t workspace_len, cudaStream_t st);

// The exchange with its receive layout planned on the device: d_count_matrix = the all-gathered
// nranks x 16 per-destination counts (split_counts of every rank), d_peer_bases[r] = rank r's
// receive buffer as mapped here (DEVICE array), cap = keys a receive buffer holds.  Writes
// *d_nrecv (keys arriving at this rank; 0 on overflow) and raises *d_status on overflow (then no
// key is stored anywhere).  No host synchronisation.
cudaError_t exchange_planned(const uint32_t *in, size_t n, const uint32_t *splitters, int nranks, int me,
                             const uint32_t *d_count_matrix, const unsigned long long *d_peer_bases, size_t cap,
                             uint32_t *d_nrecv, uint32_t *d_status, void *workspace, size_t workspace_len,
                             cudaStream_t st);

// Splitters and the count matrix from histograms instead of a sample + counting pass (kernels.cuh
// "splitters straight from a histogram"): top_histogram = 4096 bins of key >> 20 of one shard (zeroes
// d_hist first); splitters_from_histograms = the all-gathered nranks x 4096 histograms -> 32-word splitter
// table, nranks x 16 count matrix (the layout exchange_planned takes) and the imbalance those
// bin-aligned splitters give (largest range * 1000 / mean).
cudaError_t top_histogram(const uint32_t *in, size_t n, uint32_t *d_hist, cudaStream_t st);
cudaError_t splitters_from_histograms(const uint32_t *d_all_hists, int nranks, uint32_t *d_splitters, uint32_t *d_count_matrix,
                                      uint32_t *d_imbalance_permille, cudaStream_t st);

// The exchange without any plan: every receive buffer is followed by its cursor word (at key index
// `cap`), zero before the exchange; sources reserve room for each tile's run with a system-scope
// atomicAdd on the destination's cursor (NVLink), so no rank needs to know any counts beforehand.
// Afterwards (all ranks done) the cursor = keys received.  *d_status is raised when a reservation
// would pass `cap` (those keys are dropped; the destination's cursor then exceeds cap).
cudaError_t exchange_cursor(const uint32_t *in, size_t n, const uint32_t *splitters, int nranks,
                            const unsigned long long *d_peer_bases, size_t cap, uint32_t *d_status, void *workspace,
                            size_t workspace_len, cudaStream_t st);

// Full 32-bit sort of in[0..n) into out (tmp: n keys of scratch; in may equal out) that allocates nothing:
// its look-back workspace is a file-scope device array in engine.cu, n <= noalloc_max_keys().  Calls
// sharing a GPU must not overlap.  See libsortDeviceSortNoAlloc in include/libsort_b200.h.
size_t noalloc_max_keys();
cudaError_t sort_noalloc(const uint32_t *in, uint32_t *out, uint32_t *tmp, size_t n, cudaStream_t st);

// Test hook: cap the number of tiles per look-back portion (0 = default,
// i.e. the largest portion whose key count fits the 30-bit descriptors).
void debug_set_portion_tiles(uint32_t tiles);

}  // namespace b200sort
