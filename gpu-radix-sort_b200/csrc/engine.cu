// engine.cu -- see engine.cuh.
#include "engine.cuh"
#include "kernels.cuh"
#include "onesweep_bulk.cuh"
#include "scatter2.cuh"
#include "scatter3.cuh"

#include <algorithm>
#include <atomic>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace b200sort {

// ---- compiled-in kernel variant -------------------------------------------
// Two tile shapes.  A: the atomic-ranking scatter (few registers per key, big
// tiles).  B: the ballot-ranking kernels -- the fallback for devices that fail the
// lane-order self-test, and the <= 16-way splitter scatter of the distributed sort,
// whose few bins would serialise atomic ranking.
#ifndef SORT_THREADS
#define SORT_THREADS 256
#endif
#ifndef SORT_KPT
#define SORT_KPT 56
#endif
#ifndef SORT_MINB
#define SORT_MINB 3
#endif
#ifndef SORT_LB
#define SORT_LB 4   // look-back descriptors fetched per round trip
#endif
#ifndef BALLOT_KPT
#define BALLOT_KPT 32
#endif
#ifndef BALLOT_MINB
#define BALLOT_MINB 4
#endif
// The bulk-copy splitter scatter (the NVLink exchange): 24 keys per thread, 4 CTAs per SM.  More CTAs in
// flight hide the peer-store latency better than longer runs per bulk copy do (exchange at 2 GPUs:
// 16/24/32/48/56 keys per thread -> 1.03 / 0.88 / 0.94 / 1.06 / 1.08 ms).
#ifndef SPLIT_KPT
#define SPLIT_KPT 24
#endif
#ifndef SPLIT_BULK_MINB
#define SPLIT_BULK_MINB 4
#endif
#ifndef SPLIT_MINB
#define SPLIT_MINB 3   // the splitter scatter keeps up to 15 splitters in registers
#endif
#ifndef HIST_THREADS
#define HIST_THREADS 512
#endif
#ifndef HIST_REPL
#define HIST_REPL 2
#endif

// Small inputs (<= kSmallNMax keys): the same scatter kernel with 16 keys per thread.  A 1 M-key
// sort is 73 big tiles -- half the SMs idle, every pass one long tile -- but 256 small ones.
#ifndef SMALL_KPT
#define SMALL_KPT 16
#endif
#ifndef SMALL_MINB
#define SMALL_MINB 4
#endif
constexpr int kTileS = SORT_THREADS * SMALL_KPT;
constexpr size_t kSmallNMax = (size_t)1 << 23;  // workspace is sized for short tiles up to here; LIBSORT_SMALL_MAX picks the actual limit
constexpr int kTileA = SORT_THREADS * SORT_KPT;
constexpr int kTileB = SORT_THREADS * BALLOT_KPT;
constexpr int kTileMin = kTileA < kTileB ? kTileA : kTileB;
using SmemA = OnesweepSmem<SORT_THREADS, SORT_KPT, false, RadixBinner::Arg>;
using SmemB = OnesweepSmem<SORT_THREADS, BALLOT_KPT, false, RadixBinner::Arg>;
using SmemSplit = OnesweepSmem<SORT_THREADS, BALLOT_KPT, true, SplitArg>;

#define KERNEL_A(NBITS) k_onesweep<SORT_THREADS, SORT_KPT, SORT_MINB, RANK_ATOMS, NBITS, SORT_LB, RadixBinner, false>
#define KERNEL_S(NBITS) k_onesweep<SORT_THREADS, SMALL_KPT, SMALL_MINB, RANK_ATOMS, NBITS, SORT_LB, RadixBinner, false>
using SmemS = OnesweepSmem<SORT_THREADS, SMALL_KPT, false, RadixBinner::Arg>;
#define KERNEL_BULK k_onesweep_bulk<SORT_THREADS, SORT_KPT, SORT_MINB, SORT_LB, RANK_ATOMS, RadixBinner, false>
using SmemBulk = BulkSmem<SORT_THREADS, SORT_KPT, RadixBinner::Arg>;
#define KERNEL_SPLIT_BULK(NS) k_onesweep_bulk<SORT_THREADS, SPLIT_KPT, SPLIT_BULK_MINB, SORT_LB, RANK_BALLOT, SplitBinner<NS>, true>
#define KERNEL_SPLIT_BULK_UNORDERED(NS) \
    k_onesweep_bulk<SORT_THREADS, SPLIT_KPT, SPLIT_BULK_MINB, SORT_LB, RANK_BALLOT, SplitBinner<NS>, true, EXCH_LOCAL_CURSOR>
#define KERNEL_SPLIT_BULK_REMOTE(NS) \
    k_onesweep_bulk<SORT_THREADS, SPLIT_KPT, SPLIT_BULK_MINB, SORT_LB, RANK_BALLOT, SplitBinner<NS>, true, EXCH_REMOTE_CURSOR>
using SmemSplitBulk = BulkSmem<SORT_THREADS, SPLIT_KPT, SplitArg>;
constexpr int kTileSplitBulk = SORT_THREADS * SPLIT_KPT;
#define KERNEL_B(NBITS) k_onesweep<SORT_THREADS, BALLOT_KPT, BALLOT_MINB, RANK_BALLOT, NBITS, SORT_LB, RadixBinner, false>
#define KERNEL_SPLIT(NS) k_onesweep<SORT_THREADS, BALLOT_KPT, SPLIT_MINB, RANK_BALLOT, 0, SORT_LB, SplitBinner<NS>, true>
// the radix scatter writing into an array sliced over several GPUs (step-wise distributed LSD sort)
#define KERNEL_SLICED_A k_onesweep<SORT_THREADS, SORT_KPT, SORT_MINB, RANK_ATOMS, 0, SORT_LB, RadixBinner, OUT_SLICED>
#define KERNEL_SLICED_B k_onesweep<SORT_THREADS, BALLOT_KPT, SPLIT_MINB, RANK_BALLOT, 0, SORT_LB, RadixBinner, OUT_SLICED>
#define KERNEL_SLICED_BULK \
    k_onesweep_bulk<SORT_THREADS, SORT_KPT, SORT_MINB, SORT_LB, RANK_ATOMS, RadixBinner, false, EXCH_ORDERED, true>
using SmemSlicedA = OnesweepSmem<SORT_THREADS, SORT_KPT, OUT_SLICED, RadixBinner::Arg>;
using SmemSlicedB = OnesweepSmem<SORT_THREADS, BALLOT_KPT, OUT_SLICED, RadixBinner::Arg>;

// ---- second-generation scatter pass (scatter2.cuh): developer-selectable variants ----------
typedef void (*Scatter2Fn)(const uint32_t *, uint32_t *, uint32_t, RadixBinner::Arg, const uint32_t *, uint32_t *, uint32_t *,
                           uint32_t *, uint32_t, int);
struct K3Variant {
    const char *name;
    Scatter2Fn fn;
    int threads, tile;
    size_t smem;
};
#define K3V(NAME, T, K, M, LBW, LD, RG, E, PF) \
    {NAME, k_scatter2<T, K, M, LBW, LD, RG, E, PF>, T, (T) * (K), sizeof(Scatter2Smem<T, K, LD, RadixBinner::Arg>)}
#define K3P(NAME, K, LBW, BATCH, PF, WO) {NAME, k_scatter3<K, LBW, BATCH, PF, WO>, S3_THREADS, S3_COMPUTE * (K), sizeof(Scatter3Smem<K>)}
static const K3Variant g_k3[] = {
    K3V("ldg256x56p", 256, 56, 3, 4, LOAD_LDG, RAG_PAIR, false, true),
    K3V("ldg256x56m", 256, 56, 3, 4, LOAD_LDG, RAG_MASK, false, true),
    K3V("ldg256x56s", 256, 56, 3, 4, LOAD_LDG, RAG_SCALAR, false, true),
    K3V("ldg256x56p_nopf", 256, 56, 3, 4, LOAD_LDG, RAG_PAIR, false, false),
    K3V("ldg256x56p_lb8", 256, 56, 3, 8, LOAD_LDG, RAG_PAIR, false, true),
    K3V("inp256x56p", 256, 56, 3, 4, LOAD_TMA_INPLACE, RAG_PAIR, false, true),
    K3V("inp256x56p_nopf", 256, 56, 3, 4, LOAD_TMA_INPLACE, RAG_PAIR, false, false),
    K3V("inp256x56m", 256, 56, 3, 4, LOAD_TMA_INPLACE, RAG_MASK, false, true),
    K3V("sep512x22p", 512, 22, 2, 4, LOAD_TMA_SEP, RAG_PAIR, true, false),
    K3V("sep512x22p_late", 512, 22, 2, 4, LOAD_TMA_SEP, RAG_PAIR, false, true),
    K3V("sep512x22m", 512, 22, 2, 4, LOAD_TMA_SEP, RAG_MASK, true, false),
    K3V("sep512x22p_lb8", 512, 22, 2, 8, LOAD_TMA_SEP, RAG_PAIR, true, false),
    K3V("sep384x30p", 384, 30, 2, 4, LOAD_TMA_SEP, RAG_PAIR, true, false),
    K3V("ldg512x28p", 512, 28, 2, 4, LOAD_LDG, RAG_PAIR, false, true),
    K3V("ldg256x40p", 256, 40, 4, 4, LOAD_LDG, RAG_PAIR, false, true),
    K3V("ldg256x48p", 256, 48, 3, 4, LOAD_LDG, RAG_PAIR, false, true),
    K3P("pipe44", 44, 4, 4, true, 0),
    K3P("pipe44w", 44, 4, 4, true, 1),
    K3P("pipe44w_b2", 44, 4, 2, true, 1),
    K3P("pipe44w_lb8", 44, 8, 4, true, 1),
    K3P("pipe42w", 42, 4, 2, true, 1),
    K3P("pipe40w", 40, 4, 4, true, 1),
    K3P("pipe36w", 36, 4, 4, true, 1),
};
constexpr int kNumK3 = (int)(sizeof(g_k3) / sizeof(g_k3[0]));
static int k3_selected()
{
    static const int sel = [] {
        const char *e = getenv("LIBSORT_K3");
        if (!e || !*e) return -1;
        for (int i = 0; i < kNumK3; i++)
            if (!strcmp(e, g_k3[i].name)) return i;
        fprintf(stderr, "libsort: unknown LIBSORT_K3=%s\n", e);
        return -1;
    }();
    return sel;
}



// workspace layout, in uint32 words
constexpr size_t WS_HIST = 0;                                   // [4][256]
constexpr size_t WS_BASE = WS_HIST + MAX_PASSES * RADIX;        // [2][4][256]
constexpr size_t WS_TICKET = WS_BASE + 2 * MAX_PASSES * RADIX;  // [4 passes][8 portions]
constexpr size_t WS_DESC = WS_TICKET + 64;                      // [4][ntiles][256]
constexpr int MAX_PORTIONS = 8;

static std::atomic<uint32_t> g_portion_tiles{0};
void debug_set_portion_tiles(uint32_t tiles) { g_portion_tiles.store(tiles); }

static uint32_t portion_tiles(int tile)
{
    uint32_t t = g_portion_tiles.load();
    return t ? t : (uint32_t)(((1u << 30) - 1) / tile);
}

static size_t num_tiles(size_t n, int tile) { return (n + tile - 1) / tile; }

size_t workspace_bytes(size_t n)
{
    const size_t tiles = std::max(num_tiles(n, kTileMin), n <= kSmallNMax ? num_tiles(n, kTileS) : (size_t)0);
    return (WS_DESC + (size_t)MAX_PASSES * tiles * RADIX) * sizeof(uint32_t);
}

// ---- per-device launch parameters -----------------------------------------
struct DevInfo {
    bool ready = false;
    int sms = 0;
    int occ_s = 0;  // the small-input scatter kernel
    int occ_a = 0, occ_b = 0, occ_split = 0, occ_bulk = 0, occ_split_bulk = 0;  // resident CTAs per SM of the three scatter kernels
    int occ_sliced_a = 0, occ_sliced_b = 0, occ_sliced_bulk = 0;
    int occ_k3[32] = {};
    int hist_occ = 0;
    bool rank_atoms = false;  // rank keys with returning shared atomics (self-tested) instead of ballots
};
static DevInfo g_dev[64];
static std::mutex g_dev_mu;

static int env_int(const char *name, int dflt)
{
    const char *e = getenv(name);
    return e ? atoi(e) : dflt;
}
// 1 = bank-private K1 (one 1024-thread CTA per SM), 0 = replicated K1
static const int g_hist_mode = env_int("LIBSORT_HIST_MODE", 1);
// 1 = bulk-copy write-out of the scatter pass (k_onesweep_bulk), 0 = LDS + STG write-out (k_onesweep)
static const int g_bulk_out = env_int("LIBSORT_BULK_OUT", 0);
static const int g_pdl = env_int("LIBSORT_PDL", 1);  // programmatic dependent launch between the kernels of one sort
static const int g_bulk_step = env_int("LIBSORT_BULK_STEP", 0);  // 1: step-wise distributed sort writes its runs through the bulk-copy engine (slower: 8.09 vs 6.55 ms at 2 GPUs)
static const int g_small_max = env_int("LIBSORT_SMALL_MAX", 1 << 22);  // inputs up to this many keys use the 4096-key tiles (break-even ~4 M keys)
static const int g_small_path = env_int("LIBSORT_SMALL_PATH", 1);  // 0: big tiles at every size (A/B measurements)
// the same choice for the splitter scatter of the distributed sort (runs of kilobytes, peer destinations)
static const int g_bulk_exchange = env_int("LIBSORT_BULK_EXCHANGE", 1);
// 1 = the planned exchange keeps the tiles of a destination in shard order (look-back), 0 = cursor atomics
static const int g_exchange_ordered = env_int("LIBSORT_EXCHANGE_ORDERED", 0);

template <int NPASS>
static cudaError_t launch_hist(const uint32_t *in, size_t n, const PassPlan &plan, uint32_t *ghist, int sms, int occ,
                               cudaStream_t st, const uint32_t *dyn_n = nullptr)
{
    if (g_hist_mode == 1 && n >= (1u << 16)) {
        auto kern = k_histogram_lanes<NPASS, 1024>;  // shared-memory attribute: set per device in dev_info()
        const int smem = NPASS * RADIX * 32 * (int)sizeof(uint32_t);
        const size_t want = (n / 4 + 1024 * 4 - 1) / (1024 * 4);
        const int grid = (int)std::max<size_t>(1, std::min<size_t>((size_t)sms, want));
        kern<<<grid, 1024, smem, st>>>(in, n, plan, ghist, dyn_n);
    } else {
        const size_t want = (n / 4 + HIST_THREADS * 4 - 1) / (HIST_THREADS * 4);
        const int grid = (int)std::max<size_t>(1, std::min<size_t>((size_t)sms * occ, want));
        k_histogram<NPASS, HIST_THREADS, HIST_REPL><<<grid, HIST_THREADS, 0, st>>>(in, n, plan, ghist, dyn_n);
    }
    return cudaSuccess;
}

// LIBSORT_RANK = ballot | atoms | auto (default).  "atoms" ranks a key with the
// value its warp's shared-memory atomicAdd returns, which is stable only if the
// hardware serves the lanes of one instruction that hit the same address in
// ascending lane order.  That is how sm_100 behaves (profiles/r01_atoms_lane_order.log:
// 0 exceptions in 2x10^10 collisions) but it is not an architectural promise,
// so `auto` first proves it on the device with k_selftest_atoms_order and
// falls back to the ballot kernel (same results, more instructions) otherwise.
static cudaError_t choose_rank_mode(int dev, bool *atoms)
{
    const char *e = getenv("LIBSORT_RANK");
    if (e && !strcmp(e, "ballot")) { *atoms = false; return cudaSuccess; }
    if (e && !strcmp(e, "atoms")) { *atoms = true; return cudaSuccess; }
    unsigned long long *d_bad = nullptr, h_bad[2] = {0, 0};
    cudaError_t err = cudaMalloc(&d_bad, sizeof(h_bad));
    if (err != cudaSuccess) return err;
    err = cudaMemset(d_bad, 0, sizeof(h_bad));
    int sms = 1;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (err == cudaSuccess) {
        cudaStream_t st = nullptr;  // private stream: the legacy default stream would serialise with the caller's work
        err = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
        if (err == cudaSuccess) {
            err = cudaMemsetAsync(d_bad, 0, sizeof(h_bad), st);
            k_selftest_atoms_order<<<sms, 1024, 0, st>>>(d_bad, 256);
            if (err == cudaSuccess) err = cudaMemcpyAsync(h_bad, d_bad, sizeof(h_bad), cudaMemcpyDeviceToHost, st);
            if (err == cudaSuccess) err = cudaStreamSynchronize(st);
            cudaStreamDestroy(st);
        }
    }
    cudaFree(d_bad);
    if (err != cudaSuccess) return err;
    *atoms = (h_bad[0] == 0 && h_bad[1] > 0);
    if (!*atoms)
        fprintf(stderr, "libsort: device %d serves colliding shared atomics out of lane order (%llu of %llu); using ballot ranking\n",
                dev, h_bad[0], h_bad[1]);
    return cudaSuccess;
}

// Launch `kern` so that it may begin while the previous kernel of the stream is still draining
// (programmatic dependent launch): the kernel itself calls pdl_wait() before it reads anything the
// predecessor wrote.  Cuts the ~4 us launch gap between the 6-7 kernels of one sort, which is
// most of a small sort's time.
template <typename... KArgs, typename... Args>
static cudaError_t launch_after(void (*kern)(KArgs...), int grid, int block, size_t smem, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3((unsigned)block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = g_pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

static cudaError_t dev_info(DevInfo **out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_dev_mu);
    DevInfo &d = g_dev[dev];
    if (!d.ready) {
        e = cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        struct { const void *fn; int smem; } kernels[] = {
            {(const void *)KERNEL_A(8), (int)sizeof(SmemA)},     {(const void *)KERNEL_A(0), (int)sizeof(SmemA)},
            {(const void *)KERNEL_S(8), (int)sizeof(SmemS)},     {(const void *)KERNEL_S(0), (int)sizeof(SmemS)},
            {(const void *)KERNEL_BULK, (int)sizeof(SmemBulk)},
            {(const void *)KERNEL_SPLIT_BULK(1), (int)sizeof(SmemSplitBulk)}, {(const void *)KERNEL_SPLIT_BULK(3), (int)sizeof(SmemSplitBulk)},
            {(const void *)KERNEL_SPLIT_BULK(7), (int)sizeof(SmemSplitBulk)}, {(const void *)KERNEL_SPLIT_BULK(15), (int)sizeof(SmemSplitBulk)},
            {(const void *)KERNEL_SPLIT_BULK_UNORDERED(1), (int)sizeof(SmemSplitBulk)}, {(const void *)KERNEL_SPLIT_BULK_UNORDERED(3), (int)sizeof(SmemSplitBulk)},
            {(const void *)KERNEL_SPLIT_BULK_UNORDERED(7), (int)sizeof(SmemSplitBulk)}, {(const void *)KERNEL_SPLIT_BULK_UNORDERED(15), (int)sizeof(SmemSplitBulk)},
            {(const void *)KERNEL_SPLIT_BULK_REMOTE(1), (int)sizeof(SmemSplitBulk)}, {(const void *)KERNEL_SPLIT_BULK_REMOTE(3), (int)sizeof(SmemSplitBulk)},
            {(const void *)KERNEL_SPLIT_BULK_REMOTE(7), (int)sizeof(SmemSplitBulk)}, {(const void *)KERNEL_SPLIT_BULK_REMOTE(15), (int)sizeof(SmemSplitBulk)},
            {(const void *)KERNEL_B(8), (int)sizeof(SmemB)},     {(const void *)KERNEL_B(0), (int)sizeof(SmemB)},
            {(const void *)KERNEL_SPLIT(1), (int)sizeof(SmemSplit)},  {(const void *)KERNEL_SPLIT(3), (int)sizeof(SmemSplit)},
            {(const void *)KERNEL_SPLIT(7), (int)sizeof(SmemSplit)},  {(const void *)KERNEL_SPLIT(15), (int)sizeof(SmemSplit)},
            {(const void *)KERNEL_SLICED_BULK, (int)sizeof(SmemBulk)},
            {(const void *)KERNEL_SLICED_A, (int)sizeof(SmemSlicedA)}, {(const void *)KERNEL_SLICED_B, (int)sizeof(SmemSlicedB)}};
        for (auto &k : kernels) {
            e = cudaFuncSetAttribute(k.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, k.smem);
            if (e != cudaSuccess) return e;
        }
        {   // K1, bank-private layout: NPASS x 32 KB of dynamic shared memory (this device only)
            const void *hk[] = {(const void *)k_histogram_lanes<1, 1024>, (const void *)k_histogram_lanes<2, 1024>,
                                (const void *)k_histogram_lanes<3, 1024>, (const void *)k_histogram_lanes<4, 1024>};
            for (int np = 1; np <= 4; np++) {
                e = cudaFuncSetAttribute(hk[np - 1], cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         np * RADIX * 32 * (int)sizeof(uint32_t));
                if (e != cudaSuccess) return e;
            }
        }
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_a, KERNEL_A(8), SORT_THREADS, sizeof(SmemA));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_s, KERNEL_S(8), SORT_THREADS, sizeof(SmemS));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_bulk, KERNEL_BULK, SORT_THREADS, sizeof(SmemBulk));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_split_bulk, KERNEL_SPLIT_BULK(15), SORT_THREADS, sizeof(SmemSplitBulk));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_b, KERNEL_B(8), SORT_THREADS, sizeof(SmemB));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_split, KERNEL_SPLIT(15), SORT_THREADS, sizeof(SmemSplit));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_sliced_bulk, KERNEL_SLICED_BULK, SORT_THREADS, sizeof(SmemBulk));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_sliced_a, KERNEL_SLICED_A, SORT_THREADS, sizeof(SmemSlicedA));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_sliced_b, KERNEL_SLICED_B, SORT_THREADS, sizeof(SmemSlicedB));
        if (e != cudaSuccess) return e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.hist_occ, k_histogram<4, HIST_THREADS, HIST_REPL>,
                                                          HIST_THREADS, 0);
        if (e != cudaSuccess) return e;
        if (d.occ_a < 1 || d.occ_b < 1 || d.occ_split < 1 || d.hist_occ < 1) return cudaErrorLaunchOutOfResources;
        for (int i = 0; i < kNumK3; i++) {
            e = cudaFuncSetAttribute((const void *)g_k3[i].fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g_k3[i].smem);
            if (e != cudaSuccess) { d.occ_k3[i] = 0; (void)cudaGetLastError(); continue; }
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ_k3[i], (const void *)g_k3[i].fn, g_k3[i].threads, g_k3[i].smem);
            if (e != cudaSuccess) { d.occ_k3[i] = 0; (void)cudaGetLastError(); }
        }
        e = choose_rank_mode(dev, &d.rank_atoms);
        if (e != cudaSuccess) return e;
        if (const char *dbg = getenv("LIBSORT_DEBUG"))
            if (dbg[0] == '1')
                fprintf(stderr, "libsort: dev %d sms=%d rank=%s tiles A/B=%d/%d occ A/B/split=%d/%d/%d hist_occ=%d smem A/B=%zu/%zu\n", dev,
                        d.sms, d.rank_atoms ? "atoms" : "ballot", kTileA, kTileB, d.occ_a, d.occ_b, d.occ_split, d.hist_occ,
                        sizeof(SmemA), sizeof(SmemB));
        if (const char *dbg = getenv("LIBSORT_DEBUG"))
            if (dbg[0] == '1' && k3_selected() >= 0)
                fprintf(stderr, "libsort: dev %d K3 variant %s tile=%d smem=%zu occ=%d\n", dev, g_k3[k3_selected()].name,
                        g_k3[k3_selected()].tile, g_k3[k3_selected()].smem, d.occ_k3[k3_selected()]);
        d.ready = true;
    }
    *out = &d;
    return cudaSuccess;
}

int tile_keys()
{
    DevInfo *dev = nullptr;  // the tile of the kernels this device runs (atomic or ballot ranking)
    if (dev_info(&dev) != cudaSuccess) return kTileA;
    const int k3 = k3_selected();
    if (dev->rank_atoms && k3 >= 0 && dev->occ_k3[k3] > 0) return g_k3[k3].tile;
    return dev->rank_atoms ? kTileA : kTileB;
}

static PassPlan make_plan(uint32_t offset, uint32_t width)
{
    // SortState::Step walks 2-bit digits from `offset` while shift < offset+width
    // (reference sort.cu:323); shifts >= 32 see digit 0 and are identity passes.
    PassPlan p{};
    if (offset >= 32) return p;
    uint64_t eff = ((uint64_t)width + 1) & ~1ull;
    if (eff > 32 - offset) eff = 32 - offset;
    uint32_t done = 0;
    while (done < eff) {
        uint32_t bits = std::min<uint32_t>(8, (uint32_t)eff - done);
        p.shift[p.npass] = offset + done;
        p.mask[p.npass] = (1u << bits) - 1;
        p.npass++;
        done += bits;
    }
    return p;
}

struct Timer {
    cudaStream_t st;
    bool on;
    cudaEvent_t ev[16];
    int n = 0;
    Timer(cudaStream_t s, bool enable) : st(s), on(enable) {}
    void mark()
    {
        if (!on) return;
        cudaEventCreate(&ev[n]);
        cudaEventRecord(ev[n], st);
        n++;
    }
    float span(int i)
    {
        float ms = 0;
        cudaEventElapsedTime(&ms, ev[i], ev[i + 1]);
        return ms;
    }
    ~Timer()
    {
        for (int i = 0; i < n; i++) cudaEventDestroy(ev[i]);
    }
};

#define CK(x)                                \
    do {                                     \
        cudaError_t e_ = (x);                \
        if (e_ != cudaSuccess) return e_;    \
    } while (0)

// Boundaries from the sorted keys: exact by binary search, then (reference
// mode) the gpu_groups/back-fill transform.
static cudaError_t finish_bounds(const uint32_t *sorted, size_t n, uint32_t offset, size_t ngroups, uint32_t *d_bounds,
                                 int bounds_mode, cudaStream_t st, int *launches)
{
    const int grid = (int)std::min<size_t>((ngroups + 255) / 256, 4096);
    k_bounds_search<<<grid, 256, 0, st>>>(sorted, (uint32_t)n, offset >= 32 ? 31 : offset,
                                          offset >= 32 ? 0 : (uint32_t)(ngroups - 1), d_bounds, ngroups);
    if (launches) ++*launches;
    if (bounds_mode == BOUNDS_REFERENCE) {
        k_bounds_refquirk<<<1, 1024, 0, st>>>(d_bounds, ngroups, (uint32_t)n);
        if (launches) ++*launches;
    }
    return cudaGetLastError();
}

cudaError_t sort_bits(const uint32_t *in, uint32_t *a, uint32_t *b, uint32_t *want, size_t n, uint32_t offset,
                      uint32_t width, uint32_t *d_bounds, int bounds_mode, void *workspace, size_t workspace_len,
                      cudaStream_t st, uint32_t **result, KernelTimes *times, const uint32_t *dyn_n)
{
    if (n > 0xffffffffull) return cudaErrorInvalidValue;
    if (d_bounds && width > 31) return cudaErrorInvalidValue;
    const PassPlan plan = make_plan(offset, width);
    const int P = plan.npass;
    const size_t ngroups = (size_t)1 << (width > 31 ? 0 : width);
    if (times) *times = KernelTimes{};

    uint32_t *res = const_cast<uint32_t *>(in);
    if (n == 0 || P == 0) {
        // no key moves (empty input, width 0, or offset past bit 31)
        if (want && want != in && n) {
            CK(cudaMemcpyAsync(want, in, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
            res = want;
        }
        if (d_bounds) {
            if (n == 0) CK(cudaMemsetAsync(d_bounds, 0, ngroups * sizeof(uint32_t), st));
            else CK(finish_bounds(res, n, offset, ngroups, d_bounds, bounds_mode, st, nullptr));
        }
        if (result) *result = res;
        return cudaGetLastError();
    }
    if (workspace_len < workspace_bytes(n) || !workspace) return cudaErrorInvalidValue;
    if (!a || !b || a == b) return cudaErrorInvalidValue;

    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    uint32_t *ws = static_cast<uint32_t *>(workspace);
    uint32_t *ghist = ws + WS_HIST, *base = ws + WS_BASE, *tickets = ws + WS_TICKET, *desc = ws + WS_DESC;
    // dyn_n (key count read on the device, n = capacity): plain kernels, one look-back portion, no boundaries
    if (dyn_n && (d_bounds || n >= (1ull << 30))) return cudaErrorInvalidValue;
    const int k3 = (!dyn_n && dev->rank_atoms && k3_selected() >= 0 && dev->occ_k3[k3_selected()] > 0) ? k3_selected() : -1;
    // small input: more, shorter tiles (not with the debug portion override, whose tests count big tiles)
    const bool small = g_small_path && !dyn_n && k3 < 0 && dev->rank_atoms && dev->occ_s > 0 && n <= kSmallNMax &&
                       n <= (size_t)g_small_max && g_portion_tiles.load() == 0;
    const int tile = small ? kTileS : k3 >= 0 ? g_k3[k3].tile : dev->rank_atoms ? kTileA : kTileB;
    const size_t ntiles = num_tiles(n, tile);
    const uint32_t ptiles = portion_tiles(tile);
    const size_t nportions = (ntiles + ptiles - 1) / ptiles;
    if (nportions > MAX_PORTIONS) return cudaErrorInvalidValue;

    Timer tm(st, times != nullptr);
    int launches = 0;
    tm.mark();
    const bool bounds_from_hist = d_bounds && P == 1 && plan.mask[0] == ngroups - 1;
    CK(cudaMemsetAsync(ws, 0, (WS_DESC + (size_t)P * ntiles * RADIX) * sizeof(uint32_t), st));
    tm.mark();

    // K1: one upsweep, all digits
    switch (P) {
        case 1: CK(launch_hist<1>(in, n, plan, ghist, dev->sms, dev->hist_occ, st, dyn_n)); break;
        case 2: CK(launch_hist<2>(in, n, plan, ghist, dev->sms, dev->hist_occ, st, dyn_n)); break;
        case 3: CK(launch_hist<3>(in, n, plan, ghist, dev->sms, dev->hist_occ, st, dyn_n)); break;
        default: CK(launch_hist<4>(in, n, plan, ghist, dev->sms, dev->hist_occ, st, dyn_n)); break;
    }
    launches++;
    tm.mark();
    // constant-digit passes become copies (single-portion inputs only: the copy's destination is `dst` itself)
    CK(launch_after(k_scan_hist, P, RADIX, 0, st, ghist, base, bounds_from_hist ? d_bounds : nullptr, (uint32_t)ngroups, (uint32_t)n,
                    nportions == 1 ? tickets + SKIP_OFFSET : nullptr, dyn_n));
    launches++;
    tm.mark();

    // pick the ping-pong order
    uint32_t *dst;
    if (in == a) dst = b;
    else if (in == b) dst = a;
    else {
        uint32_t *last = want ? want : a;
        if (last != a && last != b) last = a;
        dst = (P & 1) ? last : (last == a ? b : a);
    }
    const uint32_t *src = in;
    const bool aligned = true;
    for (int p = 0; p < P; p++) {
        const int use_tma = aligned && ((reinterpret_cast<uintptr_t>(src) & 15u) == 0);
        for (size_t q = 0; q < nportions; q++) {
            const size_t t0 = q * ptiles, t1 = std::min(ntiles, t0 + ptiles);
            const size_t k0 = t0 * tile, k1 = std::min(n, t1 * (size_t)tile);
            const uint32_t nt = (uint32_t)(t1 - t0);
            const int grid = (int)std::min<size_t>(nt, (size_t)dev->sms * (dev->rank_atoms ? dev->occ_a : dev->occ_b));
            uint32_t *bcur = base + ((q & 1) * MAX_PASSES + p) * RADIX;
            uint32_t *bnext = (q + 1 < nportions) ? base + (((q + 1) & 1) * MAX_PASSES + p) * RADIX : nullptr;
            const bool full_digit = plan.mask[p] == 0xffu;
            const RadixBinner::Arg barg{plan.shift[p], plan.mask[p]};
            uint32_t *dsc = desc + ((size_t)p * ntiles + t0) * RADIX, *tk = tickets + p * MAX_PORTIONS + q;
            if (small) {
                auto kern = full_digit ? KERNEL_S(8) : KERNEL_S(0);
                CK(launch_after(kern, (int)std::min<size_t>(nt, (size_t)dev->sms * dev->occ_s), SORT_THREADS, sizeof(SmemS), st,
                                src + k0, dst, (uint32_t)(k1 - k0), barg, (const void *)bcur, bnext, dsc, tk, nt, use_tma,
                                (const uint32_t *)nullptr));
            } else if (k3 >= 0) {
                const int gridk = (int)std::min<size_t>(nt, (size_t)dev->sms * dev->occ_k3[k3]);
                g_k3[k3].fn<<<gridk, g_k3[k3].threads, g_k3[k3].smem, st>>>(src + k0, dst, (uint32_t)(k1 - k0), barg, bcur, bnext, dsc, tk,
                                                                           nt, use_tma);
            } else if (!dyn_n && dev->rank_atoms && g_bulk_out && dev->occ_bulk > 0) {
                const int gridb = (int)std::min<size_t>(nt, (size_t)dev->sms * dev->occ_bulk);
                KERNEL_BULK<<<gridb, SORT_THREADS, sizeof(SmemBulk), st>>>(src + k0, dst, (uint32_t)(k1 - k0), barg, bcur, bnext, dsc, tk,
                                                                          nt, use_tma);
            } else {
                auto kern = dev->rank_atoms ? (full_digit ? KERNEL_A(8) : KERNEL_A(0)) : (full_digit ? KERNEL_B(8) : KERNEL_B(0));
                const size_t smem = dev->rank_atoms ? sizeof(SmemA) : sizeof(SmemB);
                CK(launch_after(kern, grid, SORT_THREADS, smem, st, src + k0, dst, (uint32_t)(k1 - k0), barg, (const void *)bcur, bnext,
                                dsc, tk, nt, use_tma, dyn_n));
            }
            launches++;
        }
        tm.mark();
        src = dst;
        dst = (dst == a) ? b : a;
    }
    res = const_cast<uint32_t *>(src);
    if (want && res != want) {
        CK(cudaMemcpyAsync(want, res, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
        res = want;
    }

    if (d_bounds) {
        if (bounds_from_hist) {
            if (bounds_mode == BOUNDS_REFERENCE) {
                k_bounds_refquirk<<<1, 1024, 0, st>>>(d_bounds, ngroups, (uint32_t)n);
                launches++;
            }
        } else {
            CK(finish_bounds(res, n, offset, ngroups, d_bounds, bounds_mode, st, &launches));
        }
    }
    tm.mark();
    CK(cudaGetLastError());
    if (result) *result = res;
    if (times) {
        CK(cudaStreamSynchronize(st));
        times->memset_ms = tm.span(0);
        times->histogram_ms = tm.span(1);
        times->scan_ms = tm.span(2);
        for (int p = 0; p < P; p++) times->pass_ms[p] = tm.span(3 + p);
        times->boundaries_ms = tm.span(3 + P);
        times->npass = P;
        times->launches = launches;
    }
    return cudaSuccess;
}

// Did any atomic-ranking kernel on the current device see lanes served out of order since the
// last audit?  Synchronises `st`.  On a violation the device is switched to the ballot kernels for
// the rest of the process; the caller repeats whatever it sorted since the previous audit.
cudaError_t rank_audit(cudaStream_t st, bool *violated)
{
    *violated = false;
    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    if (!dev->rank_atoms) return cudaSuccess;
    unsigned int flag = 0;
    CK(cudaMemcpyFromSymbolAsync(&flag, g_rank_violation, sizeof(flag), 0, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (flag) {
        int d = 0;
        cudaGetDevice(&d);
        fprintf(stderr, "libsort: device %d served colliding shared atomics out of lane order; switching to ballot ranking\n", d);
        {
            std::lock_guard<std::mutex> lk(g_dev_mu);
            dev->rank_atoms = false;
        }
        flag = 0;
        CK(cudaMemcpyToSymbolAsync(g_rank_violation, &flag, sizeof(flag), 0, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        *violated = true;
    }
    return cudaSuccess;
}

// ---- distributed-sort building blocks ---------------------------------------
cudaError_t sample_keys(const uint32_t *in, size_t n, uint32_t *out, uint32_t nsamples, uint32_t seed, cudaStream_t st)
{
    if (n == 0 || n > 0xffffffffull || nsamples == 0) return cudaErrorInvalidValue;
    k_sample<<<(nsamples + 255) / 256, 256, 0, st>>>(in, (uint32_t)n, out, nsamples, seed);
    return cudaGetLastError();
}

cudaError_t pick_splitters(const uint32_t *sorted, size_t nsorted, int nranks, uint32_t *d_splitters, cudaStream_t st)
{
    if (nranks < 1 || nranks > MAX_SPLITTERS + 1 || nsorted == 0 || nsorted > 0xffffffffull) return cudaErrorInvalidValue;
    k_pick_splitters<<<1, 32, 0, st>>>(sorted, (uint32_t)nsorted, nranks, d_splitters);
    return cudaGetLastError();
}

cudaError_t split_counts(const uint32_t *in, size_t n, const uint32_t *splitters, int nsplit, uint32_t *d_counts,
                         cudaStream_t st)
{
    if (nsplit < 0 || nsplit > MAX_SPLITTERS || n > 0xffffffffull) return cudaErrorInvalidValue;
    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    CK(cudaMemsetAsync(d_counts, 0, (MAX_SPLITTERS + 1) * sizeof(uint32_t), st));
    if (n == 0) return cudaSuccess;
    const size_t want = (n / 4 + 4095) / 4096;
    const int grid = (int)std::max<size_t>(1, std::min<size_t>((size_t)dev->sms * 2, want));
    const SplitArg arg{splitters, nsplit};
    if (nsplit <= 1) k_dest_hist<1><<<grid, 1024, 0, st>>>(in, n, arg, d_counts);
    else if (nsplit <= 3) k_dest_hist<3><<<grid, 1024, 0, st>>>(in, n, arg, d_counts);
    else if (nsplit <= 7) k_dest_hist<7><<<grid, 1024, 0, st>>>(in, n, arg, d_counts);
    else k_dest_hist<15><<<grid, 1024, 0, st>>>(in, n, arg, d_counts);
    return cudaGetLastError();
}

cudaError_t scatter_by_splitters(const uint32_t *in, size_t n, const uint32_t *splitters, int nsplit,
                                 const unsigned long long *dest_ptrs, void *workspace, size_t workspace_len,
                                 cudaStream_t st)
{
    if (nsplit < 0 || nsplit > MAX_SPLITTERS || n >= (1ull << 30)) return cudaErrorInvalidValue;
    if (n == 0) return cudaSuccess;
    if (workspace_len < workspace_bytes(n) || !workspace) return cudaErrorInvalidValue;
    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    uint32_t *ws = static_cast<uint32_t *>(workspace);
    const bool bulk = g_bulk_exchange && dev->occ_split_bulk > 0;
    const size_t ntiles = num_tiles(n, bulk ? kTileSplitBulk : kTileB);
    CK(cudaMemsetAsync(ws, 0, (WS_DESC + ntiles * RADIX) * sizeof(uint32_t), st));
    PtrTable t{};
    for (int i = 0; i <= nsplit; i++) t.p[i] = dest_ptrs[i];
    unsigned long long *d_ptrs = reinterpret_cast<unsigned long long *>(ws + WS_HIST);
    k_store_ptrs<<<1, 32, 0, st>>>(t, d_ptrs);
    const int grid = (int)std::min<size_t>(ntiles, (size_t)dev->sms * dev->occ_split);
    const int use_tma = (reinterpret_cast<uintptr_t>(in) & 15u) == 0;
    const SplitArg arg{splitters, nsplit};
    if (bulk) {
        auto kern = nsplit <= 1 ? KERNEL_SPLIT_BULK(1) : nsplit <= 3 ? KERNEL_SPLIT_BULK(3) : nsplit <= 7 ? KERNEL_SPLIT_BULK(7)
                                                                                                        : KERNEL_SPLIT_BULK(15);
        const int gridb = (int)std::min<size_t>(ntiles, (size_t)dev->sms * dev->occ_split_bulk);
        kern<<<gridb, SORT_THREADS, sizeof(SmemSplitBulk), st>>>(in, nullptr, (uint32_t)n, arg, d_ptrs, nullptr, ws + WS_DESC,
                                                                 ws + WS_TICKET, (uint32_t)ntiles, use_tma);
    } else {
        auto kern = nsplit <= 1 ? KERNEL_SPLIT(1) : nsplit <= 3 ? KERNEL_SPLIT(3) : nsplit <= 7 ? KERNEL_SPLIT(7) : KERNEL_SPLIT(15);
        kern<<<grid, SORT_THREADS, sizeof(SmemSplit), st>>>(in, nullptr, (uint32_t)n, arg, d_ptrs, nullptr, ws + WS_DESC,
                                                            ws + WS_TICKET, (uint32_t)ntiles, use_tma, nullptr);
    }
    return cudaGetLastError();
}

cudaError_t exchange_planned(const uint32_t *in, size_t n, const uint32_t *splitters, int nranks, int me,
                             const uint32_t *d_count_matrix, const unsigned long long *d_peer_bases, size_t cap,
                             uint32_t *d_nrecv, uint32_t *d_status, void *workspace, size_t workspace_len, cudaStream_t st)
{
    if (nranks < 1 || nranks > MAX_SPLITTERS + 1 || me < 0 || me >= nranks || n >= (1ull << 30)) return cudaErrorInvalidValue;
    if (workspace_len < workspace_bytes(n) || !workspace) return cudaErrorInvalidValue;
    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    uint32_t *ws = static_cast<uint32_t *>(workspace);
    const bool bulk = g_bulk_exchange && dev->occ_split_bulk > 0;
    const size_t ntiles = num_tiles(n, bulk ? kTileSplitBulk : kTileB);
    CK(cudaMemsetAsync(ws, 0, (WS_DESC + ntiles * RADIX) * sizeof(uint32_t), st));
    unsigned long long *d_ptrs = reinterpret_cast<unsigned long long *>(ws + WS_HIST);
    k_plan_exchange<<<1, 32, 0, st>>>(d_count_matrix, nranks, me, d_peer_bases, (unsigned long long)cap, d_ptrs, d_nrecv,
                                      d_status, ws + WS_TICKET);
    if (n == 0) return cudaGetLastError();
    const int nsplit = nranks - 1;
    const int use_tma = (reinterpret_cast<uintptr_t>(in) & 15u) == 0;
    const SplitArg arg{splitters, nsplit};
    if (bulk) {
        // a full local sort follows the exchange: the tiles need not arrive in order (one cursor
        // atomic per destination instead of the look-back); LIBSORT_EXCHANGE_ORDERED=1 keeps the chain
        auto kern = g_exchange_ordered
                        ? (nsplit <= 1 ? KERNEL_SPLIT_BULK(1) : nsplit <= 3 ? KERNEL_SPLIT_BULK(3) : nsplit <= 7 ? KERNEL_SPLIT_BULK(7)
                                                                                                                 : KERNEL_SPLIT_BULK(15))
                        : (nsplit <= 1   ? KERNEL_SPLIT_BULK_UNORDERED(1)
                           : nsplit <= 3 ? KERNEL_SPLIT_BULK_UNORDERED(3)
                           : nsplit <= 7 ? KERNEL_SPLIT_BULK_UNORDERED(7)
                                         : KERNEL_SPLIT_BULK_UNORDERED(15));
        const int gridb = (int)std::min<size_t>(ntiles, (size_t)dev->sms * dev->occ_split_bulk);
        kern<<<gridb, SORT_THREADS, sizeof(SmemSplitBulk), st>>>(in, nullptr, (uint32_t)n, arg, d_ptrs, nullptr, ws + WS_DESC,
                                                                 ws + WS_TICKET, (uint32_t)ntiles, use_tma);
    } else {
        auto kern = nsplit <= 1 ? KERNEL_SPLIT(1) : nsplit <= 3 ? KERNEL_SPLIT(3) : nsplit <= 7 ? KERNEL_SPLIT(7) : KERNEL_SPLIT(15);
        const int grid = (int)std::min<size_t>(ntiles, (size_t)dev->sms * dev->occ_split);
        kern<<<grid, SORT_THREADS, sizeof(SmemSplit), st>>>(in, nullptr, (uint32_t)n, arg, d_ptrs, nullptr, ws + WS_DESC,
                                                            ws + WS_TICKET, (uint32_t)ntiles, use_tma, nullptr);
    }
    return cudaGetLastError();
}

cudaError_t top_histogram(const uint32_t *in, size_t n, uint32_t *d_hist, cudaStream_t st)
{
    if (!d_hist || n >= (1ull << 32)) return cudaErrorInvalidValue;
    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    CK(cudaMemsetAsync(d_hist, 0, TOP_BINS * sizeof(uint32_t), st));
    if (n == 0) return cudaSuccess;
    const size_t want = (n / 4 + 512 * 4 - 1) / (512 * 4);
    const int grid = (int)std::max<size_t>(1, std::min<size_t>((size_t)dev->sms * 4, want));
    k_hist_top12<512><<<grid, 512, 0, st>>>(in, n, d_hist);
    return cudaGetLastError();
}

cudaError_t splitters_from_histograms(const uint32_t *d_all_hists, int nranks, uint32_t *d_splitters, uint32_t *d_count_matrix,
                                      uint32_t *d_imbalance_permille, cudaStream_t st)
{
    if (nranks < 1 || nranks > MAX_SPLITTERS + 1 || !d_all_hists || !d_splitters || !d_count_matrix || !d_imbalance_permille)
        return cudaErrorInvalidValue;
    k_splitters_from_hist<<<1, 1024, 0, st>>>(d_all_hists, nranks, d_splitters, d_count_matrix, d_imbalance_permille);
    return cudaGetLastError();
}

cudaError_t exchange_cursor(const uint32_t *in, size_t n, const uint32_t *splitters, int nranks,
                            const unsigned long long *d_peer_bases, size_t cap, uint32_t *d_status, void *workspace,
                            size_t workspace_len, cudaStream_t st)
{
    if (nranks < 1 || nranks > MAX_SPLITTERS + 1 || n >= (1ull << 30) || cap >= (1ull << 32)) return cudaErrorInvalidValue;
    if (workspace_len < workspace_bytes(n) || !workspace) return cudaErrorInvalidValue;
    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    if (dev->occ_split_bulk < 1) return cudaErrorLaunchOutOfResources;
    uint32_t *ws = static_cast<uint32_t *>(workspace);
    const size_t ntiles = num_tiles(n, kTileSplitBulk);
    static_assert(sizeof(CursorPlan) <= (WS_TICKET - WS_HIST) * sizeof(uint32_t), "the plan lives in the histogram words");
    CK(cudaMemsetAsync(ws + WS_TICKET, 0, 64 * sizeof(uint32_t), st));  // the tile ticket; no descriptors in this mode
    CursorPlan *plan = reinterpret_cast<CursorPlan *>(ws + WS_HIST);
    k_store_cursor_plan<<<1, 32, 0, st>>>(d_peer_bases, nranks, (unsigned long long)cap, d_status, plan);
    if (n == 0) return cudaGetLastError();
    const int nsplit = nranks - 1;
    const int use_tma = (reinterpret_cast<uintptr_t>(in) & 15u) == 0;
    const SplitArg arg{splitters, nsplit};
    auto kern = nsplit <= 1   ? KERNEL_SPLIT_BULK_REMOTE(1)
                : nsplit <= 3 ? KERNEL_SPLIT_BULK_REMOTE(3)
                : nsplit <= 7 ? KERNEL_SPLIT_BULK_REMOTE(7)
                              : KERNEL_SPLIT_BULK_REMOTE(15);
    const int gridb = (int)std::min<size_t>(ntiles, (size_t)dev->sms * dev->occ_split_bulk);
    kern<<<gridb, SORT_THREADS, sizeof(SmemSplitBulk), st>>>(in, nullptr, (uint32_t)n, arg, plan, nullptr, ws + WS_DESC, ws + WS_TICKET,
                                                             (uint32_t)ntiles, use_tma);
    return cudaGetLastError();
}

// ---- step-wise distributed LSD sort (one global counting sort per digit) -----------
size_t step_table_bytes() { return sizeof(SliceTable); }

cudaError_t step_table_init(void *d_table, const unsigned long long *peers, const unsigned long long *slice_starts,
                            int nslice, cudaStream_t st)
{
    if (nslice < 1 || nslice > MAX_SLICES || !d_table) return cudaErrorInvalidValue;
    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    const size_t tile = dev->rank_atoms ? kTileA : kTileB;
    SliceTable h{};
    for (int j = 0; j < nslice; j++) {
        // a bucket's run inside one tile may cross at most one slice boundary
        if (slice_starts[j + 1] < slice_starts[j] || slice_starts[j + 1] > 0xffffffffull) return cudaErrorInvalidValue;
        if (j + 1 < nslice && slice_starts[j + 1] - slice_starts[j] < tile) return cudaErrorInvalidValue;
        h.peer[j] = peers[j];
        h.slice_start[j] = (uint32_t)slice_starts[j];
    }
    for (int j = nslice; j <= MAX_SLICES; j++) h.slice_start[j] = (uint32_t)slice_starts[nslice];
    h.nslice = (uint32_t)nslice;
    // everything up to gbase (that part is written by step_bases); pageable source: staged before return
    return cudaMemcpyAsync(d_table, &h, offsetof(SliceTable, gbase), cudaMemcpyHostToDevice, st);
}

cudaError_t step_counts(const uint32_t *in, size_t n, uint32_t shift, uint32_t nbits, uint32_t *d_counts, cudaStream_t st)
{
    if (nbits < 1 || nbits > 8 || shift + nbits > 32 || n > 0xffffffffull) return cudaErrorInvalidValue;
    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    CK(cudaMemsetAsync(d_counts, 0, RADIX * sizeof(uint32_t), st));
    if (n == 0) return cudaSuccess;
    PassPlan plan{};
    plan.npass = 1;
    plan.shift[0] = shift;
    plan.mask[0] = (1u << nbits) - 1;
    CK(launch_hist<1>(in, n, plan, d_counts, dev->sms, dev->hist_occ, st));
    return cudaGetLastError();
}

cudaError_t step_bases(const uint32_t *d_gathered, int nranks, int me, void *d_table, cudaStream_t st)
{
    if (nranks < 1 || nranks > MAX_SLICES || me < 0 || me >= nranks) return cudaErrorInvalidValue;
    k_step_bases<<<1, RADIX, 0, st>>>(d_gathered, nranks, me, static_cast<SliceTable *>(d_table));
    return cudaGetLastError();
}

cudaError_t step_scatter(const uint32_t *in, size_t n, uint32_t shift, uint32_t nbits, const void *d_table,
                         void *workspace, size_t workspace_len, cudaStream_t st)
{
    if (nbits < 1 || nbits > 8 || shift + nbits > 32 || n >= (1ull << 30)) return cudaErrorInvalidValue;
    if (n == 0) return cudaSuccess;
    if (workspace_len < workspace_bytes(n) || !workspace || !d_table) return cudaErrorInvalidValue;
    DevInfo *dev = nullptr;
    CK(dev_info(&dev));
    uint32_t *ws = static_cast<uint32_t *>(workspace);
    const int tile = dev->rank_atoms ? kTileA : kTileB;
    const size_t ntiles = num_tiles(n, tile);
    CK(cudaMemsetAsync(ws + WS_TICKET, 0, (WS_DESC - WS_TICKET + ntiles * RADIX) * sizeof(uint32_t), st));
    const RadixBinner::Arg arg{shift, (1u << nbits) - 1};
    const int use_tma = (reinterpret_cast<uintptr_t>(in) & 15u) == 0;
    if (dev->rank_atoms && g_bulk_step && dev->occ_sliced_bulk > 0) {
        const int grid = (int)std::min<size_t>(ntiles, (size_t)dev->sms * dev->occ_sliced_bulk);
        KERNEL_SLICED_BULK<<<grid, SORT_THREADS, sizeof(SmemBulk), st>>>(in, nullptr, (uint32_t)n, arg, d_table, nullptr, ws + WS_DESC,
                                                                        ws + WS_TICKET, (uint32_t)ntiles, use_tma);
    } else if (dev->rank_atoms) {
        const int grid = (int)std::min<size_t>(ntiles, (size_t)dev->sms * dev->occ_sliced_a);
        KERNEL_SLICED_A<<<grid, SORT_THREADS, sizeof(SmemSlicedA), st>>>(in, nullptr, (uint32_t)n, arg, d_table, nullptr,
                                                                         ws + WS_DESC, ws + WS_TICKET, (uint32_t)ntiles, use_tma, nullptr);
    } else {
        const int grid = (int)std::min<size_t>(ntiles, (size_t)dev->sms * dev->occ_sliced_b);
        KERNEL_SLICED_B<<<grid, SORT_THREADS, sizeof(SmemSlicedB), st>>>(in, nullptr, (uint32_t)n, arg, d_table, nullptr,
                                                                         ws + WS_DESC, ws + WS_TICKET, (uint32_t)ntiles, use_tma, nullptr);
    }
    return cudaGetLastError();
}

// ---- allocation-free full sort (libsortDeviceSortNoAlloc, include/libsort_b200.h) --------------
// For callers that must not allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc): the look-back workspace of the scatter passes lives in a file-scope device array
// instead of caller- or pool-allocated memory.  Every GPU that loads the library holds the array
// (about 2 MB) until exit.  The path is self-contained: the ballot-ranking kernels (architecturally
// ordered, so no self-test and no audit), its own per-device set-up -- which only queries attributes
// and the array's device address -- and no cudaMalloc / cudaFree / stream creation anywhere.
// All calls share the array: it keeps the last call's contents (the words in use are cleared at the
// start of every call, in stream order) and two calls that overlap on the GPU race.
constexpr size_t kNoAllocMaxKeys = (size_t)1 << 22;
constexpr size_t kNoAllocTiles = (kNoAllocMaxKeys + kTileB - 1) / kTileB;
constexpr size_t kNoAllocWords = WS_DESC + (size_t)MAX_PASSES * kNoAllocTiles * RADIX;
static __device__ __align__(256) uint32_t g_noalloc_scratch[kNoAllocWords];

struct NoAllocDev {
    bool ready = false;
    int sms = 0, occ = 0, hist_occ = 0;
    uint32_t *scratch = nullptr;  // DEVICE address of g_noalloc_scratch on this GPU (the symbol itself is a host-side stand-in)
};
static NoAllocDev g_noalloc_dev[64];
static std::mutex g_noalloc_mu;

static cudaError_t noalloc_setup(NoAllocDev **out)
{
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    NoAllocDev &d = g_noalloc_dev[dev];
    if (!d.ready) {  // once per device, with that device current
        CK(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev));
        CK(cudaFuncSetAttribute((const void *)KERNEL_B(8), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SmemB)));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.occ, KERNEL_B(8), SORT_THREADS, sizeof(SmemB)));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&d.hist_occ, k_histogram<4, HIST_THREADS, HIST_REPL>, HIST_THREADS, 0));
        if (d.occ < 1 || d.hist_occ < 1) return cudaErrorLaunchOutOfResources;
        void *ptr = nullptr;
        CK(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        if (!ptr) return cudaErrorInvalidSymbol;
        d.scratch = static_cast<uint32_t *>(ptr);
        d.ready = true;
    }
    *out = &d;
    return cudaSuccess;
}

size_t noalloc_max_keys() { return kNoAllocMaxKeys; }

cudaError_t sort_noalloc(const uint32_t *in, uint32_t *out, uint32_t *tmp, size_t n, cudaStream_t st)
{
    if (n > kNoAllocMaxKeys) return cudaErrorInvalidValue;
    if (n == 0) return cudaSuccess;
    if (!in || !out || !tmp || tmp == out || tmp == in) return cudaErrorInvalidValue;
    NoAllocDev *dev = nullptr;
    CK(noalloc_setup(&dev));
    const PassPlan plan = make_plan(0, 32);
    const int P = plan.npass;  // 4
    const size_t ntiles = num_tiles(n, kTileB);
    uint32_t *ws = dev->scratch;
    uint32_t *ghist = ws + WS_HIST, *base = ws + WS_BASE, *tickets = ws + WS_TICKET, *desc = ws + WS_DESC;
    // histograms, tickets, skip flags and the descriptors this call uses start at zero
    CK(cudaMemsetAsync(ws, 0, (WS_DESC + (size_t)P * ntiles * RADIX) * sizeof(uint32_t), st));
    const size_t want = (n / 4 + HIST_THREADS * 4 - 1) / (HIST_THREADS * 4);
    const int hgrid = (int)std::max<size_t>(1, std::min<size_t>((size_t)dev->sms * dev->hist_occ, want));
    k_histogram<4, HIST_THREADS, HIST_REPL><<<hgrid, HIST_THREADS, 0, st>>>(in, n, plan, ghist, nullptr);
    k_scan_hist<<<P, RADIX, 0, st>>>(ghist, base, nullptr, 0u, (uint32_t)n, tickets + SKIP_OFFSET, nullptr);
    const int grid = (int)std::min<size_t>(ntiles, (size_t)dev->sms * dev->occ);
    const uint32_t *src = in;
    uint32_t *dst = tmp;  // in -> tmp -> out -> tmp -> out
    for (int p = 0; p < P; p++) {
        const int use_tma = (reinterpret_cast<uintptr_t>(src) & 15u) == 0;
        const RadixBinner::Arg barg{plan.shift[p], plan.mask[p]};
        KERNEL_B(8)<<<grid, SORT_THREADS, sizeof(SmemB), st>>>(src, dst, (uint32_t)n, barg, (const void *)(base + p * RADIX), nullptr,
                                                               desc + (size_t)p * ntiles * RADIX, tickets + p * MAX_PORTIONS,
                                                               (uint32_t)ntiles, use_tma, nullptr);
        src = dst;
        dst = dst == tmp ? out : tmp;
    }
    return cudaGetLastError();
}

}  // namespace b200sort
