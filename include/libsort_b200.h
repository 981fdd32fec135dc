/*
 * libsort_b200.h -- additive C ABI of the B200-native libsort.so.
 *
 * Nothing here exists in the reference; these entry points expose the same
 * sort on DEVICE-resident keys (for roofline measurement, the PyTorch-facing
 * wrapper and the multi-GPU sort, SURVEY.md 8(b) "Additive") and never
 * collide with the seven legacy names in libsort.h.  Plain pointers and
 * sizes only; `stream` is a cudaStream_t passed as void*.  Unless stated
 * otherwise a function returns 0 on success, a positive cudaError_t value on
 * a CUDA failure, or a negative LIBSORT_E* code; nothing synchronises the
 * stream unless stated.
 */
#ifndef LIBSORT_B200_H
#define LIBSORT_B200_H
#include "libsort.h"

#define LIBSORT_EINVAL (-1)   /* bad argument */
#define LIBSORT_ENOINIT (-2)  /* initLibSort has not succeeded */
#define LIBSORT_ENOMEM (-3)

/* boundary_mode */
#define LIBSORT_BOUNDS_REFERENCE 0 /* bit-exact with reference libsort/sort.cu:14-27,367-394 */
#define LIBSORT_BOUNDS_EXACT 1     /* exclusive prefix sums (localTest/tests.cpp:44-54) */

typedef struct LibsortKernelTimes {
    float memset_ms;     /* workspace clear */
    float histogram_ms;  /* K1: upsweep histogram, all digits */
    float scan_ms;       /* bucket-base scan */
    float pass_ms[4];    /* K3: one scatter pass each */
    float boundaries_ms;
    int npass;
    int launches;        /* kernels launched by the call */
} LibsortKernelTimes;

/* Keys per scatter tile of the built kernel (informational). */
LIBSORT_API int libsortTileKeys(void);

/* Device scratch bytes libsortDeviceSort needs for n keys. */
LIBSORT_API size_t libsortWorkspaceBytes(size_t n);

/*
 * Device-resident equivalent of SortState::Step (+ GetBoundaries)
 * (reference libsort/sort.cu:322-394): stable sort of d_in[0..n) by bits
 * [offset, offset+width) into d_out, d_tmp (n keys) as the ping-pong buffer.
 * d_in may equal d_out (in place) and is otherwise left untouched.
 * d_boundaries: NULL or 1<<width device words.  times: NULL, or receives the
 * per-kernel CUDA-event durations (this variant synchronises the stream).
 * Runs on the calling thread's current device; does not use the device pool.
 */
LIBSORT_API int libsortDeviceSort(const uint32_t *d_in, uint32_t *d_out, uint32_t *d_tmp, size_t n, uint32_t offset,
                                  uint32_t width, uint32_t *d_boundaries, int boundary_mode, void *d_workspace,
                                  size_t workspace_bytes, void *stream, LibsortKernelTimes *times);

/*
 * Building blocks of the one-rank-per-GPU distributed sort that replaces the
 * reference's host-mediated LSD shuffle (localTest/benchmarks.cpp:70-160,
 * benchmark/pkg/sort/distrib.go:90-179).  The collectives themselves (count
 * all-gather, optional NCCL all-to-all) are issued by the host side over
 * torch.distributed; see gpu-radix-sort_b200/distrib.py.
 *
 * libsortDeviceSample:        nsamples stratified pseudo-random keys of the shard.
 * The splitter table is a DEVICE array of LIBSORT_SPLIT_WORDS (32) words: [0, 16) the
 *     splitter values (ascending, nsplit <= 15 used), [16, 32) their tie thresholds.
 *     Destination of the key at position i of the shard = number of splitters s_j with
 *     key > s_j, or key == s_j and (i * 0x9E3779B1 mod 2^32) >= threshold_j.  With all
 *     thresholds 0 that is "number of splitters <= key"; non-zero thresholds cut the copies
 *     of a heavily repeated key value between neighbouring ranks (keys-only sort, so any
 *     copy may go anywhere in the value's range) -- even an all-equal input balances.
 * libsortDevicePickSplitters: d_splitters[g-1] = g/nranks quantile of a sorted sample;
 *     thresholds from where the quantile falls inside a run of >= 8 equal samples.
 * libsortDeviceSplitCounts:   d_counts[0..nsplit] (16 words) = keys per destination.
 * libsortDeviceScatterToPeers: stable partition by destination written straight
 *     to dest_ptrs[0..nsplit] (splitters: DEVICE array; dest_ptrs: HOST array of device addresses, local or
 *     cudaIpc/NVLink-mapped peer memory): partition + key exchange in one
 *     kernel.  n < 2^30; workspace as for libsortDeviceSort.
 * libsortIpc*: peer-mappable allocations and cudaIpc handle plumbing (64-byte
 *     handles) so ranks in different processes can map each other's receive
 *     buffers.
 */
#define LIBSORT_SPLIT_WORDS 32
#define LIBSORT_SPLIT_TIE_OFFSET 16
LIBSORT_API int libsortDeviceSample(const uint32_t *d_in, size_t n, uint32_t *d_out, uint32_t nsamples, uint32_t seed,
                                    void *stream);
LIBSORT_API int libsortDevicePickSplitters(const uint32_t *d_sorted, size_t nsorted, int nranks, uint32_t *d_splitters,
                                           void *stream);
LIBSORT_API int libsortDeviceSplitCounts(const uint32_t *d_in, size_t n, const uint32_t *splitters, int nsplit,
                                         uint32_t *d_counts, void *stream);
LIBSORT_API int libsortDeviceScatterToPeers(const uint32_t *d_in, size_t n, const uint32_t *splitters, int nsplit,
                                            const unsigned long long *dest_ptrs, void *d_workspace,
                                            size_t workspace_bytes, void *stream);
/*
 * The same exchange and local sort without any host round trip (what DistributedSorter.sort uses):
 * libsortDeviceExchangePlanned: d_count_matrix = the all-gathered nranks x 16 counts of
 *     libsortDeviceSplitCounts (DEVICE), d_peer_bases[r] = rank r's receive buffer as mapped on this
 *     GPU (DEVICE array of nranks addresses), cap = keys a receive buffer holds.  A one-warp kernel
 *     derives where this rank's keys for every destination go and *d_nrecv (keys arriving here),
 *     then the fused partition + exchange kernel runs.  If some rank would receive more than cap
 *     keys, *d_status becomes 1 (sticky; zero it beforehand), *d_nrecv 0 and nothing is stored.
 * libsortDeviceSortDyn: full 32-bit sort of d_in[0..*d_n) into d_out, the key count read ON THE
 *     DEVICE when the kernels run; launches, workspace and d_tmp are sized for `capacity` (< 2^30);
 *     times as for libsortDeviceSort (non-NULL synchronises the stream).
 */
LIBSORT_API int libsortDeviceExchangePlanned(const uint32_t *d_in, size_t n, const uint32_t *splitters, int nranks,
                                             int me, const uint32_t *d_count_matrix,
                                             const unsigned long long *d_peer_bases, size_t cap, uint32_t *d_nrecv,
                                             uint32_t *d_status, void *d_workspace, size_t workspace_bytes,
                                             void *stream);
/*
 * libsortDeviceSortNoAlloc: full ascending 32-bit sort of d_in[0..n) into d_out that ALLOCATES NOTHING.
 * It has no counterpart in the reference; it was added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  The caller passes the input, the output and n keys of scratch (d_tmp, distinct from
 * both; d_in may equal d_out); the look-back workspace that libsortDeviceSort takes from the caller is a
 * file-scope device array inside the library (about 2 MB, held by every GPU that loads the library until
 * exit), hence the limit: n <= LIBSORT_NOALLOC_MAX_KEYS, larger sizes are rejected with
 * cudaErrorInvalidValue and a message on stderr.  No cudaMalloc, cudaFree, stream creation or self-test
 * runs on this path (it uses the ballot-ranking kernels, which need neither self-test nor audit; set-up
 * once per device only queries attributes and the array's address).  initLibSort() is not required.
 * Every call on a GPU shares that one array: it keeps the previous call's contents (the part a call
 * uses is cleared at its start, in stream order), and calls that overlap on the GPU -- different streams
 * without ordering, different host threads -- race and give wrong results; serialise them.
 */
#define LIBSORT_NOALLOC_MAX_KEYS ((size_t)1 << 22)
LIBSORT_API int libsortDeviceSortNoAlloc(const uint32_t *d_in, uint32_t *d_out, uint32_t *d_tmp, size_t n, void *stream);

/*
 * Splitters without a sample and without a counting pass (DistributedSorter plan="hist"):
 * libsortDeviceTopHistogram: d_hist4096[b] = keys of the shard with key >> 20 == b (zeroed first).
 * libsortDeviceSplittersFromHistograms: from the all-gathered nranks x 4096 histograms (DEVICE) ->
 *     the 32-word splitter table (splitters on bin boundaries, no tie thresholds), the nranks x 16
 *     count matrix in the layout libsortDeviceExchangePlanned takes, and *d_imbalance_permille =
 *     largest range * 1000 / mean under those splitters.  Exact for any input; balanced only when
 *     the keys spread over enough of the 4096 top-bit bins -- callers fall back to sampled splitters
 *     (libsortDeviceSample / PickSplitters / SplitCounts) when the imbalance is too high.
 */
LIBSORT_API int libsortDeviceTopHistogram(const uint32_t *d_in, size_t n, uint32_t *d_hist4096, void *stream);
LIBSORT_API int libsortDeviceSplittersFromHistograms(const uint32_t *d_all_hists, int nranks, uint32_t *d_splitters,
                                                     uint32_t *d_count_matrix, uint32_t *d_imbalance_permille,
                                                     void *stream);
/*
 * libsortDeviceExchangeCursor: the exchange with NO plan at all (DistributedSorter plan="cursor"; faster
 *     than the counted layout at 2 GPUs, slower from 4 on: the atomics on one word per destination
 *     serialise, see profiles/r02b_exchange_plan_ab.log).  Every rank's receive buffer holds `cap` keys followed by one cursor word (key
 *     index cap; allocate at least (cap + 16) * 4 bytes), zero when the exchange starts on any rank.
 *     d_peer_bases[r] = rank r's buffer as mapped on this GPU (DEVICE array of nranks addresses, own
 *     buffer included).  Each tile reserves room for its run at destination r with one system-scope
 *     atomicAdd on r's cursor over NVLink and bulk-copies the run there: no counting pass, no count
 *     exchange, arrival order arbitrary.  Once every rank's exchange has completed (barrier), the
 *     cursor word IS the number of keys received -- pass its address to libsortDeviceSortDyn as d_n.
 *     A reservation that would pass cap is dropped and raises *d_status (sticky; zero it beforehand);
 *     the destination's cursor then exceeds cap (libsortDeviceSortDyn clamps to its capacity).
 */
LIBSORT_API int libsortDeviceExchangeCursor(const uint32_t *d_in, size_t n, const uint32_t *splitters, int nranks,
                                            const unsigned long long *d_peer_bases, size_t cap, uint32_t *d_status,
                                            void *d_workspace, size_t workspace_bytes, void *stream);
LIBSORT_API int libsortDeviceSortDyn(const uint32_t *d_in, uint32_t *d_out, uint32_t *d_tmp, size_t capacity,
                                     const uint32_t *d_n, void *d_workspace, size_t workspace_bytes, void *stream,
                                     LibsortKernelTimes *times);
LIBSORT_API int libsortIpcAlloc(size_t bytes, void **ptr);
LIBSORT_API int libsortIpcFree(void *ptr);
LIBSORT_API int libsortIpcExport(void *ptr, void *handle64);
LIBSORT_API int libsortIpcOpen(const void *handle64, void **ptr);
LIBSORT_API int libsortIpcClose(void *ptr);

/*
 * The whole distributed sort behind one call, for a caller that drives every
 * GPU of the box from ONE process -- what the reference's Go benchmark driver
 * does with its worker pool (benchmark/pkg/sort/distrib.go:183-248) and its C++
 * twin with host threads (localTest/benchmarks.cpp:70-160).  h_keys (HOST,
 * pinned or pageable) is cut into ngpu contiguous shards (ngpu <= 0: every GPU
 * of the pool; at most 16), sampled, partitioned by value range and exchanged
 * GPU-to-GPU over NVLink peer stores in one fused kernel per GPU, sorted locally
 * and written back ascending into h_keys.  n may exceed 2^32 as long as every
 * shard stays below 2^30 keys.  Blocks until every requested GPU is free, like
 * the legacy calls.  Returns bool like them (1 = ok, 0 = failed, message on
 * stderr).  Requires initLibSort() and peer access between the GPUs.
 */
LIBSORT_API int libsortDistribSort(uint32_t *h_keys, size_t n, int ngpu);

/*
 * Step-wise distributed LSD sort on device buffers: the reference's own distributed
 * algorithm (P workers; every step each worker sorts its slice by one digit, then the
 * slices are regathered bucket-major / worker-minor: localTest/benchmarks.cpp:70-160,
 * benchmark/pkg/sort/distrib.go:119-175, helpers.go:67-78) with the host shuffle replaced
 * by peer stores.  A step is one GLOBAL stable counting sort by the digit:
 *   libsortStepCounts   digit histogram of this rank's slice -> d_counts[256]
 *   (all-gather the P x 256 counts with the caller's collective)
 *   libsortStepBases    d_gathered[P][256] -> where this rank's first key of every digit goes
 *   libsortStepScatter  the scatter pass; each key is stored at its global position, i.e. into
 *                       the slice of whichever GPU owns that position (local or NVLink-mapped)
 * The slice table (libsortStepTableBytes() of device memory, filled by libsortStepTableInit
 * from HOST arrays) says where slice j lives (peers[j]) and which global positions it holds
 * ([slice_starts[j], slice_starts[j+1])); every slice but the last must be at least
 * libsortTileKeys() long, n < 2^30 per rank, total < 2^32.  Intermediate states equal the
 * reference's after every step; after the last step slice r holds the r-th part of the sorted
 * whole.  See gpu-radix-sort_b200/distrib.py (DistributedSorter.sort_stepwise).
 */
LIBSORT_API size_t libsortStepTableBytes(void);
LIBSORT_API int libsortStepTableInit(void *d_table, const unsigned long long *peers,
                                     const unsigned long long *slice_starts, int nslice, void *stream);
LIBSORT_API int libsortStepCounts(const uint32_t *d_in, size_t n, uint32_t shift, uint32_t nbits, uint32_t *d_counts,
                                  void *stream);
LIBSORT_API int libsortStepBases(const uint32_t *d_gathered, int nranks, int me, void *d_table, void *stream);
LIBSORT_API int libsortStepScatter(const uint32_t *d_in, size_t n, uint32_t shift, uint32_t nbits, const void *d_table,
                                   void *d_workspace, size_t workspace_bytes, void *stream);

/* populateInput at an explicit stream position: arr[i] = element first+i of the
 * fresh-process PCG sequence (reference libsort/utils.cu:63-79).  Thread safe. */
LIBSORT_API void libsortPopulateInputAt(uint32_t *arr, size_t nelem, uint64_t first);

/*
 * Audit of the default ranking mode.  The scatter pass ranks keys with returning shared-memory
 * atomics, which is stable only while the hardware serves colliding lanes of one instruction in
 * ascending lane order (true on sm_100, self-tested per device, not an architectural promise).
 * Every tile therefore re-checks one of its rows with ballots; a mismatch raises a device flag.
 * libsortRankAudit synchronises `stream` and returns 0 = no violation since the last audit,
 * 1 = violation seen: the device has been switched to the ballot kernels and everything sorted
 * on it since the last audit must be sorted again; < 0 = CUDA error.  The host-pointer entry
 * points (gpuPartial, providedGpu) audit and repeat by themselves.
 */
LIBSORT_API int libsortRankAudit(void *stream);

/* Number of GPUs in the pool (0 before initLibSort). */
LIBSORT_API int libsortDeviceCount(void);

/* Free the per-device buffers the host-pointer entry points cache. */
LIBSORT_API void libsortReleaseCaches(void);

/* Message for a code returned by this header's functions. */
LIBSORT_API const char *libsortErrorString(int code);

/* Test hook: cap tiles per look-back portion (0 restores the default). */
LIBSORT_API void libsortDebugSetPortionTiles(uint32_t tiles);
/* Pool test hooks: the GPU that served the calling thread's last host-pointer call (-1: none), and
 * how many times two callers were found holding the same GPU at once (always 0; the reference
 * reserves a device per caller, libsort/utils.cu:33-61). */
LIBSORT_API int libsortDebugLastDevice(void);
LIBSORT_API unsigned libsortDebugPoolOverlaps(void);

#endif /* LIBSORT_B200_H */
