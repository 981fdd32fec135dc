// abi.cu -- the C ABI: the seven legacy libsort symbols (include/libsort.h,
// reference libsort/libsort.h:12-47) and the additive device-pointer entry
// points (include/libsort_b200.h).
//
// This file deliberately does NOT include libsort.h: the legacy symbols are
// DEFINED here as returning `int` (0/1 in the whole of eax) while the header
// declares them `bool`.  Both are returned in al on x86-64/aarch64, so C, C++
// and cgo callers see a bool, and ctypes callers that never set restype
// (reference faasTest/pylibsort/sort.py:101,118 read a c_int) see clean upper
// bits.
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include <cuda_profiler_api.h>
#include <cuda_runtime.h>

#include "engine.cuh"
#include "host_copy.h"

using namespace b200sort;

#define API extern "C" __attribute__((visibility("default")))

namespace {

// ---- device pool (reference libsort/utils.cu:6-61, utils.h:19-68) ----------
// A counting semaphore over the GPUs plus one busy flag per GPU; a caller
// blocks while every GPU is taken and gets exclusive use of one otherwise.
struct DeviceCtx {
    int dev = -1;
    bool busy = false;
    cudaStream_t stream = nullptr;
    uint32_t *a = nullptr, *b = nullptr;
    size_t cap_keys = 0;
    void *ws = nullptr;
    size_t ws_cap = 0;
    uint32_t *d_bounds = nullptr;
    size_t bounds_cap = 0;
    HostCopier copier;
    // single-process distributed sort (libsortDistribSort)
    uint32_t *recv = nullptr;   // keys arriving from every GPU
    size_t recv_cap = 0;
    uint32_t *small = nullptr;  // samples | splitters + tie thresholds (32 words) | counts
    uint32_t *gather = nullptr; // device 0: all samples | sorted | scratch
    size_t gather_cap = 0;
    void *ws_small = nullptr;
    size_t ws_small_cap = 0;
};

struct Pool {
    std::mutex mu;
    std::condition_variable cv;
    int free_count = 0;
    std::vector<DeviceCtx> ctx;
    bool inited = false;
};
Pool g_pool;
// test hooks (libsortDebugLastDevice / libsortDebugPoolOverlaps): which GPU served this thread's last
// host-pointer call, and how often two callers held the same GPU at once (must stay 0) -- counted
// outside the pool mutex so that the check does not lean on the thing it checks
thread_local int t_last_device = -1;
std::atomic<int> g_holders[64];
std::atomic<unsigned> g_pool_overlaps{0};

bool cuda_ok(cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return true;
    fprintf(stderr, "libsort: %s failed: %s\n", what, cudaGetErrorString(e));
    return false;
}

// RAII reservation, released on every return path (reference invokers.cu:17).
struct Reservation {
    DeviceCtx *c = nullptr;
    bool acquire()
    {
        std::unique_lock<std::mutex> lk(g_pool.mu);
        if (!g_pool.inited) {
            fprintf(stderr, "libsort: initLibSort() has not been called\n");
            return false;
        }
        g_pool.cv.wait(lk, [] { return g_pool.free_count > 0; });
        for (auto &x : g_pool.ctx) {
            if (!x.busy) {
                x.busy = true;
                g_pool.free_count--;
                c = &x;
                break;
            }
        }
        lk.unlock();
        if (!c) {
            fprintf(stderr, "libsort: no free device after semaphore (this should not happen)\n");
            return false;
        }
        t_last_device = c->dev;
        if (c->dev < 64 && g_holders[c->dev].fetch_add(1) != 0) g_pool_overlaps.fetch_add(1);
        // the device is set on the calling thread, as in utils.cu:49
        if (!cuda_ok(cudaSetDevice(c->dev), "cudaSetDevice")) return false;
        if (!c->stream && !cuda_ok(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "cudaStreamCreate"))
            return false;
        return true;
    }
    ~Reservation()
    {
        if (!c) return;
        if (c->dev < 64) g_holders[c->dev].fetch_sub(1);
        {
            std::lock_guard<std::mutex> lk(g_pool.mu);
            c->busy = false;
            g_pool.free_count++;
        }
        g_pool.cv.notify_all();
    }
};

// All of the first `want` GPUs at once (want <= 0: the whole pool), taken in
// index order so that concurrent callers cannot hold each other's devices.
struct MultiReservation {
    std::vector<DeviceCtx *> cs;
    bool acquire(int want)
    {
        std::unique_lock<std::mutex> lk(g_pool.mu);
        if (!g_pool.inited) {
            fprintf(stderr, "libsort: initLibSort() has not been called\n");
            return false;
        }
        const int have = (int)g_pool.ctx.size();
        if (want <= 0) want = have;
        if (want > have) {
            fprintf(stderr, "libsort: %d GPUs requested, the pool has %d\n", want, have);
            return false;
        }
        for (int i = 0; i < want; i++) {
            DeviceCtx *x = &g_pool.ctx[i];
            g_pool.cv.wait(lk, [x] { return !x->busy; });
            x->busy = true;
            g_pool.free_count--;
            cs.push_back(x);
        }
        lk.unlock();
        for (DeviceCtx *c : cs) {
            if (!cuda_ok(cudaSetDevice(c->dev), "cudaSetDevice")) return false;
            if (!c->stream && !cuda_ok(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "cudaStreamCreate"))
                return false;
        }
        return true;
    }
    ~MultiReservation()
    {
        if (cs.empty()) return;
        {
            std::lock_guard<std::mutex> lk(g_pool.mu);
            for (DeviceCtx *c : cs) {
                c->busy = false;
                g_pool.free_count++;
            }
        }
        g_pool.cv.notify_all();
    }
};

void free_ctx_buffers(DeviceCtx &c)
{
    if (c.dev < 0) return;
    cudaSetDevice(c.dev);
    if (c.a) cudaFree(c.a);
    if (c.b) cudaFree(c.b);
    if (c.ws) cudaFree(c.ws);
    if (c.d_bounds) cudaFree(c.d_bounds);
    if (c.recv) cudaFree(c.recv);
    if (c.small) cudaFree(c.small);
    if (c.gather) cudaFree(c.gather);
    if (c.ws_small) cudaFree(c.ws_small);
    c.a = c.b = c.recv = c.small = c.gather = nullptr;
    c.ws = c.ws_small = nullptr;
    c.d_bounds = nullptr;
    c.cap_keys = c.ws_cap = c.bounds_cap = c.recv_cap = c.gather_cap = c.ws_small_cap = 0;
    c.copier.release();
}

// Per-device cached buffers: steady-state calls allocate nothing (the
// reference spends most of its API time in cudaMalloc/cudaFree,
// analysis/libsort8b.csv:15-17).
bool ensure_capacity(DeviceCtx &c, size_t n, size_t ngroups)
{
    if (n > c.cap_keys) {
        if (c.a) cudaFree(c.a);
        if (c.b) cudaFree(c.b);
        c.a = c.b = nullptr;
        c.cap_keys = 0;
        size_t cap = (n + 1023) & ~(size_t)1023;
        if (!cuda_ok(cudaMalloc(&c.a, cap * sizeof(uint32_t)), "cudaMalloc(keys)")) return false;
        if (!cuda_ok(cudaMalloc(&c.b, cap * sizeof(uint32_t)), "cudaMalloc(keys)")) return false;
        c.cap_keys = cap;
    }
    size_t need = workspace_bytes(n);
    if (need > c.ws_cap) {
        if (c.ws) cudaFree(c.ws);
        c.ws = nullptr;
        c.ws_cap = 0;
        if (!cuda_ok(cudaMalloc(&c.ws, need), "cudaMalloc(workspace)")) return false;
        c.ws_cap = need;
    }
    if (ngroups > c.bounds_cap) {
        if (c.d_bounds) cudaFree(c.d_bounds);
        c.d_bounds = nullptr;
        c.bounds_cap = 0;
        if (!cuda_ok(cudaMalloc(&c.d_bounds, ngroups * sizeof(uint32_t)), "cudaMalloc(boundaries)")) return false;
        c.bounds_cap = ngroups;
    }
    return true;
}

int boundary_mode_from_env()
{
    const char *e = getenv("LIBSORT_BOUNDARIES");
    if (e && (strcmp(e, "exact") == 0 || strcmp(e, "EXACT") == 0)) return BOUNDS_EXACT;
    return BOUNDS_REFERENCE;
}

// Host-pointer sort shared by gpuPartial and providedGpu:
// reserve -> H2D -> K1/K3 passes -> D2H keys (+ boundaries).
int host_sort(uint32_t *h_keys, size_t n, uint32_t offset, uint32_t width, uint32_t *h_bounds)
{
    Reservation r;
    if (!r.acquire()) return 0;
    // the sort internally only supports 32-bit sizes (reference invokers.cu:30-33,54-57)
    if (n > UINT32_MAX) {
        fprintf(stderr, "Input array length must be less than 2^32\n");
        return 0;
    }
    if (h_bounds && width > 31) {
        fprintf(stderr, "libsort: group width %u needs 2^%u boundaries; unsupported\n", width, width);
        return 0;
    }
    DeviceCtx &c = *r.c;
    const size_t ngroups = h_bounds ? ((size_t)1 << width) : 0;
    if (n == 0) {
        if (h_bounds) memset(h_bounds, 0, ngroups * sizeof(uint32_t));
        return 1;
    }
    if (!ensure_capacity(c, n, ngroups)) return 0;
    uint32_t *res = nullptr;
    for (int attempt = 0; attempt < 2; attempt++) {
        if (!c.copier.to_device(c.a, h_keys, n * sizeof(uint32_t), c.stream)) return 0;
        cudaError_t e = sort_bits(c.a, c.a, c.b, nullptr, n, offset, width, h_bounds ? c.d_bounds : nullptr,
                                  boundary_mode_from_env(), c.ws, c.ws_cap, c.stream, &res, nullptr);
        if (!cuda_ok(e, "sort")) return 0;
        // the caller's keys are still intact: if the atomic-ranking audit tripped, sort them again
        // with the ballot kernels before anything is copied back
        bool violated = false;
        if (!cuda_ok(rank_audit(c.stream, &violated), "rank audit")) return 0;
        if (!violated) break;
    }
    if (!c.copier.to_host(h_keys, res, n * sizeof(uint32_t), c.stream)) return 0;
    if (h_bounds &&
        !cuda_ok(cudaMemcpyAsync(h_bounds, c.d_bounds, ngroups * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream),
                 "cudaMemcpy(boundaries)"))
        return 0;
    if (!cuda_ok(cudaStreamSynchronize(c.stream), "cudaStreamSynchronize")) return 0;
    return 1;
}

}  // namespace

// ---- legacy ABI -------------------------------------------------------------

// reference libsort/utils.cu:10-32
API int initLibSort(void)
{
    std::lock_guard<std::mutex> lk(g_pool.mu);
    if (g_pool.inited) {
        fprintf(stderr, "Libsort: attempted to initialize multiple times!\n");
        return 0;
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) {
        fprintf(stderr, "Failed to get device count (%d): %s\n", ndev, cudaGetErrorString(e));
        return 0;
    }
    if (const char *lim = getenv("LIBSORT_DEVICES")) {
        int v = atoi(lim);
        if (v > 0 && v < ndev) ndev = v;
    }
    g_pool.ctx.resize(ndev);
    for (int i = 0; i < ndev; i++) g_pool.ctx[i].dev = i;
    g_pool.free_count = ndev;
    g_pool.inited = true;
    return 1;
}

// reference libsort/invokers.cu:15-41
API int gpuPartial(uint32_t *h_in, uint32_t *boundaries, size_t h_in_len, uint32_t offset, uint32_t width)
{
    return host_sort(h_in, h_in_len, offset, width, boundaries);
}

// reference libsort/invokers.cu:45-64
API int providedGpu(unsigned int *h_in, size_t len) { return host_sort(h_in, len, 0, 32, nullptr); }

// reference libsort/invokers.cu:68-71
API int providedCpu(unsigned int *in, size_t len)
{
    std::sort(in, in + len);
    return 1;
}

// reference libsort/utils.cu:63-79 -- PCG-XSH-RR 64/32, process-wide state
API void populateInput(uint32_t *arr, size_t nelem)
{
    static uint64_t state = 0x4d595df4d0f33173ull;
    uint64_t s = state;
    for (size_t i = 0; i < nelem; i++) {
        const uint64_t old = s;
        s = old * 6364136223846793005ull + 1442695040888963407ull;
        const uint32_t xorshifted = (uint32_t)((old ^ (old >> 18)) >> 27);
        const uint32_t rot = (uint32_t)(old >> 59);
        arr[i] = (xorshifted >> rot) | (xorshifted << ((0u - rot) & 31u));
    }
    state = s;
}

// reference libsort/invokers.cu:73-85
API int gpuPartialProfile(uint32_t *h_in, uint32_t *boundaries, size_t h_in_len, uint32_t offset, uint32_t width)
{
    cudaProfilerStart();
    int ret = gpuPartial(h_in, boundaries, h_in_len, offset, width);
    cudaProfilerStop();
    return ret;
}

API int providedGpuProfile(unsigned int *h_in, size_t h_in_len)
{
    cudaProfilerStart();
    int ret = providedGpu(h_in, h_in_len);
    cudaProfilerStop();
    return ret;
}

// ---- additive ABI (include/libsort_b200.h) -----------------------------------

struct LibsortKernelTimes {
    float memset_ms, histogram_ms, scan_ms, pass_ms[4], boundaries_ms;
    int npass, launches;
};
static_assert(sizeof(LibsortKernelTimes) == sizeof(KernelTimes), "ABI struct mirrors engine struct");

API int libsortTileKeys(void) { return tile_keys(); }
API size_t libsortWorkspaceBytes(size_t n) { return workspace_bytes(n); }

API int libsortDeviceSort(const uint32_t *d_in, uint32_t *d_out, uint32_t *d_tmp, size_t n, uint32_t offset,
                          uint32_t width, uint32_t *d_boundaries, int boundary_mode, void *d_workspace,
                          size_t workspace_len, void *stream, LibsortKernelTimes *times)
{
    if (n > UINT32_MAX) return -1;
    if (n && (!d_in || !d_out)) return -1;
    if (d_tmp == d_out || (d_tmp && d_tmp == d_in)) return -1;
    uint32_t *res = nullptr;
    cudaError_t e = sort_bits(d_in, d_out, d_tmp, d_out, n, offset, width, d_boundaries, boundary_mode, d_workspace,
                              workspace_len, (cudaStream_t)stream, &res, reinterpret_cast<KernelTimes *>(times));
    return (int)e;
}

API int libsortRankAudit(void *stream)
{
    bool violated = false;
    cudaError_t e = rank_audit((cudaStream_t)stream, &violated);
    if (e != cudaSuccess) return -(int)e - 100;
    return violated ? 1 : 0;
}

API int libsortDeviceCount(void)
{
    std::lock_guard<std::mutex> lk(g_pool.mu);
    return g_pool.inited ? (int)g_pool.ctx.size() : 0;
}

API void libsortReleaseCaches(void)
{
    std::unique_lock<std::mutex> lk(g_pool.mu);
    if (!g_pool.inited) return;
    int prev = 0;
    cudaGetDevice(&prev);
    for (auto &c : g_pool.ctx)
        if (!c.busy) free_ctx_buffers(c);
    cudaSetDevice(prev);
}

API const char *libsortErrorString(int code)
{
    if (code == 0) return "success";
    if (code == -1) return "invalid argument";
    if (code == -2) return "initLibSort has not succeeded";
    if (code == -3) return "out of memory";
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "unknown libsort error";
}

API void libsortDebugSetPortionTiles(uint32_t tiles) { debug_set_portion_tiles(tiles); }
API int libsortDebugLastDevice(void) { return t_last_device; }
API unsigned libsortDebugPoolOverlaps(void) { return g_pool_overlaps.load(); }

// ---- distributed-sort building blocks (include/libsort_b200.h) ----------------

API int libsortDeviceSample(const uint32_t *d_in, size_t n, uint32_t *d_out, uint32_t nsamples, uint32_t seed,
                            void *stream)
{
    return (int)sample_keys(d_in, n, d_out, nsamples, seed, (cudaStream_t)stream);
}

API int libsortDevicePickSplitters(const uint32_t *d_sorted, size_t nsorted, int nranks, uint32_t *d_splitters,
                                   void *stream)
{
    return (int)pick_splitters(d_sorted, nsorted, nranks, d_splitters, (cudaStream_t)stream);
}

// populateInput with an explicit position in the (fresh-process) stream:
// arr[i] = element (first + i) of the PCG sequence of libsort/utils.cu:63-79.
// Lets every rank generate its own shard of one continuous stream.
API void libsortPopulateInputAt(uint32_t *arr, size_t nelem, uint64_t first)
{
    const uint64_t a = 6364136223846793005ull, c = 1442695040888963407ull;
    uint64_t acc_mult = 1, acc_plus = 0, cur_mult = a, cur_plus = c, delta = first;
    while (delta) {  // LCG jump-ahead (Brown, "Random number generation with arbitrary strides")
        if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
        cur_plus = (cur_mult + 1) * cur_plus;
        cur_mult *= cur_mult;
        delta >>= 1;
    }
    uint64_t s = acc_mult * 0x4d595df4d0f33173ull + acc_plus;
    for (size_t i = 0; i < nelem; i++) {
        const uint64_t old = s;
        s = old * a + c;
        const uint32_t xorshifted = (uint32_t)((old ^ (old >> 18)) >> 27);
        const uint32_t rot = (uint32_t)(old >> 59);
        arr[i] = (xorshifted >> rot) | (xorshifted << ((0u - rot) & 31u));
    }
}

API int libsortDeviceSplitCounts(const uint32_t *d_in, size_t n, const uint32_t *splitters, int nsplit,
                                 uint32_t *d_counts, void *stream)
{
    return (int)split_counts(d_in, n, splitters, nsplit, d_counts, (cudaStream_t)stream);
}

API int libsortDeviceScatterToPeers(const uint32_t *d_in, size_t n, const uint32_t *splitters, int nsplit,
                                    const unsigned long long *dest_ptrs, void *d_workspace, size_t workspace_len,
                                    void *stream)
{
    return (int)scatter_by_splitters(d_in, n, splitters, nsplit, dest_ptrs, d_workspace, workspace_len,
                                     (cudaStream_t)stream);
}

API int libsortDeviceExchangePlanned(const uint32_t *d_in, size_t n, const uint32_t *splitters, int nranks, int me,
                                     const uint32_t *d_count_matrix, const unsigned long long *d_peer_bases, size_t cap,
                                     uint32_t *d_nrecv, uint32_t *d_status, void *d_workspace, size_t workspace_len,
                                     void *stream)
{
    return (int)exchange_planned(d_in, n, splitters, nranks, me, d_count_matrix, d_peer_bases, cap, d_nrecv, d_status,
                                 d_workspace, workspace_len, (cudaStream_t)stream);
}

API int libsortDeviceSortNoAlloc(const uint32_t *d_in, uint32_t *d_out, uint32_t *d_tmp, size_t n, void *stream)
{
    if (n > noalloc_max_keys()) {
        fprintf(stderr, "libsortDeviceSortNoAlloc: %zu keys exceed the compiled-in maximum of %zu (LIBSORT_NOALLOC_MAX_KEYS)\n", n,
                noalloc_max_keys());
        return (int)cudaErrorInvalidValue;
    }
    return (int)sort_noalloc(d_in, d_out, d_tmp, n, (cudaStream_t)stream);
}

API int libsortDeviceTopHistogram(const uint32_t *d_in, size_t n, uint32_t *d_hist4096, void *stream)
{
    return (int)top_histogram(d_in, n, d_hist4096, (cudaStream_t)stream);
}
API int libsortDeviceSplittersFromHistograms(const uint32_t *d_all_hists, int nranks, uint32_t *d_splitters,
                                             uint32_t *d_count_matrix, uint32_t *d_imbalance_permille, void *stream)
{
    return (int)splitters_from_histograms(d_all_hists, nranks, d_splitters, d_count_matrix, d_imbalance_permille,
                                          (cudaStream_t)stream);
}

API int libsortDeviceExchangeCursor(const uint32_t *d_in, size_t n, const uint32_t *splitters, int nranks,
                                    const unsigned long long *d_peer_bases, size_t cap, uint32_t *d_status,
                                    void *d_workspace, size_t workspace_len, void *stream)
{
    return (int)exchange_cursor(d_in, n, splitters, nranks, d_peer_bases, cap, d_status, d_workspace, workspace_len,
                                (cudaStream_t)stream);
}

API int libsortDeviceSortDyn(const uint32_t *d_in, uint32_t *d_out, uint32_t *d_tmp, size_t capacity, const uint32_t *d_n,
                             void *d_workspace, size_t workspace_len, void *stream, LibsortKernelTimes *times)
{
    if (!d_in || !d_out || !d_tmp || !d_n || d_tmp == d_out || d_tmp == d_in || capacity == 0) return -1;
    uint32_t *res = nullptr;
    return (int)sort_bits(d_in, d_out, d_tmp, d_out, capacity, 0, 32, nullptr, 0, d_workspace, workspace_len,
                          (cudaStream_t)stream, &res, reinterpret_cast<KernelTimes *>(times), d_n);
}

// Peer-mappable device memory for the exchange: plain cudaMalloc (one
// allocation per handle, so cudaIpc maps it at offset 0).
API int libsortIpcAlloc(size_t bytes, void **ptr) { return (int)cudaMalloc(ptr, bytes); }
API int libsortIpcFree(void *ptr) { return (int)cudaFree(ptr); }
API int libsortIpcExport(void *ptr, void *handle64)
{
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    return (int)cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t *>(handle64), ptr);
}
API int libsortIpcOpen(const void *handle64, void **ptr)
{
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    return (int)cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
}
API int libsortIpcClose(void *ptr) { return (int)cudaIpcCloseMemHandle(ptr); }

// ---- single-process multi-GPU sort ---------------------------------------------
// The whole distributed sort behind one host-pointer call, for callers that own
// every GPU of the box from one process (the cgo benchmark driver, reference
// benchmark/pkg/sort/distrib.go:183-248; the C++ twin localTest/benchmarks.cpp:
// 70-160).  Same algorithm as the one-rank-per-GPU Python driver (distrib.py):
// sample -> splitters -> destination counts -> ONE fused partition + exchange
// kernel per GPU storing straight into its peers' receive buffers over NVLink
// -> local 4-pass sort.  The tiny all-gathers (samples, count matrix) are peer
// copies / host reads because one thread drives every device.
namespace {

constexpr uint32_t kDistSamples = 1u << 15;  // per GPU
constexpr int kDistMaxGpus = 16;             // SplitBinner holds 15 splitters

bool dist_small_buffers(DeviceCtx &c, int P, bool root)
{
    if (!c.small && !cuda_ok(cudaMalloc(&c.small, (kDistSamples + 64) * sizeof(uint32_t)), "cudaMalloc(samples)"))
        return false;
    if (!root) return true;
    const size_t all = (size_t)P * kDistSamples;
    if (all > c.gather_cap) {
        if (c.gather) cudaFree(c.gather);
        c.gather = nullptr;
        c.gather_cap = 0;
        if (!cuda_ok(cudaMalloc(&c.gather, 3 * all * sizeof(uint32_t)), "cudaMalloc(sample gather)")) return false;
        c.gather_cap = all;
    }
    const size_t need = workspace_bytes(all);
    if (need > c.ws_small_cap) {
        if (c.ws_small) cudaFree(c.ws_small);
        c.ws_small = nullptr;
        c.ws_small_cap = 0;
        if (!cuda_ok(cudaMalloc(&c.ws_small, need), "cudaMalloc(sample workspace)")) return false;
        c.ws_small_cap = need;
    }
    return true;
}

bool dist_recv_buffer(DeviceCtx &c, size_t keys)
{
    if (keys <= c.recv_cap) return true;
    if (c.recv) cudaFree(c.recv);
    c.recv = nullptr;
    c.recv_cap = 0;
    const size_t cap = (keys + keys / 16 + 1023) & ~(size_t)1023;
    if (!cuda_ok(cudaMalloc(&c.recv, cap * sizeof(uint32_t)), "cudaMalloc(receive buffer)")) return false;
    c.recv_cap = cap;
    return true;
}

bool dist_enable_peers(const std::vector<DeviceCtx *> &cs)
{
    static std::mutex mu;
    static bool done[kDistMaxGpus][kDistMaxGpus] = {};
    std::lock_guard<std::mutex> lk(mu);
    for (DeviceCtx *a : cs)
        for (DeviceCtx *b : cs) {
            if (a == b || done[a->dev][b->dev]) continue;
            int can = 0;
            if (!cuda_ok(cudaDeviceCanAccessPeer(&can, a->dev, b->dev), "cudaDeviceCanAccessPeer")) return false;
            if (!can) {
                fprintf(stderr, "libsort: GPU %d cannot address GPU %d's memory (no NVLink/P2P)\n", a->dev, b->dev);
                return false;
            }
            if (!cuda_ok(cudaSetDevice(a->dev), "cudaSetDevice")) return false;
            cudaError_t e = cudaDeviceEnablePeerAccess(b->dev, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
                e = cudaSuccess;
            }
            if (!cuda_ok(e, "cudaDeviceEnablePeerAccess")) return false;
            done[a->dev][b->dev] = true;
        }
    return true;
}

bool dist_sync_all(const std::vector<DeviceCtx *> &cs, int P, const char *what)
{
    for (int g = 0; g < P; g++) {
        if (!cuda_ok(cudaSetDevice(cs[g]->dev), "cudaSetDevice")) return false;
        if (!cuda_ok(cudaStreamSynchronize(cs[g]->stream), what)) return false;
    }
    return true;
}

}  // namespace

API int libsortDistribSort(uint32_t *h_keys, size_t n, int ngpu)
{
    if (ngpu > kDistMaxGpus) {
        fprintf(stderr, "libsort: at most %d GPUs per distributed sort\n", kDistMaxGpus);
        return 0;
    }
    MultiReservation r;
    if (!r.acquire(ngpu)) return 0;
    if (n == 0) return 1;
    if (!h_keys) return 0;
    int P = std::min<int>((int)r.cs.size(), kDistMaxGpus);
    // contiguous shards, starts 16-byte aligned; fewer GPUs when there are fewer keys than that
    const size_t shard = (((n + P - 1) / P) + 3) & ~(size_t)3;
    P = (int)((n + shard - 1) / shard);
    if (shard >= ((size_t)1 << 30)) {
        fprintf(stderr, "libsort: %zu keys over %d GPUs: shards must stay below 2^30 keys\n", n, P);
        return 0;
    }
    auto &cs = r.cs;
    auto shard_len = [&](int g) { return std::min(shard, n - (size_t)g * shard); };
    if (P > 1 && !dist_enable_peers(std::vector<DeviceCtx *>(cs.begin(), cs.begin() + P))) return 0;

    // 1. shards to the GPUs, a stratified sample of each
    static std::atomic<uint32_t> seed{1};
    const uint32_t sd = seed.fetch_add(1);
    // (one host thread per GPU: a pageable caller buffer is staged chunk by chunk and to_device blocks
    // until its GPU has everything, so a single thread would feed the GPUs one after the other)
    for (int g = 0; g < P; g++) {  // buffers first, GPU 0's gather buffer is everybody's target
        if (!cuda_ok(cudaSetDevice(cs[g]->dev), "cudaSetDevice")) return 0;
        if (!ensure_capacity(*cs[g], shard + shard / 4, 0) || !dist_small_buffers(*cs[g], P, g == 0)) return 0;
    }
    auto per_gpu = [&](auto &&fn) {
        std::atomic<int> ok{1};
        std::vector<std::thread> th;
        for (int g = 0; g < P; g++)
            th.emplace_back([&, g] {
                if (!cuda_ok(cudaSetDevice(cs[g]->dev), "cudaSetDevice") || !fn(g)) ok.store(0);
            });
        for (auto &t : th) t.join();
        return ok.load() != 0;
    };
    if (!per_gpu([&](int g) {
            DeviceCtx &c = *cs[g];
            if (!c.copier.to_device(c.a, h_keys + (size_t)g * shard, shard_len(g) * sizeof(uint32_t), c.stream)) return false;
            if (!cuda_ok(sample_keys(c.a, shard_len(g), c.small, kDistSamples, sd, c.stream), "sample")) return false;
            return cuda_ok(cudaMemcpyPeerAsync(cs[0]->gather + (size_t)g * kDistSamples, cs[0]->dev, c.small, c.dev,
                                               kDistSamples * sizeof(uint32_t), c.stream),
                           "cudaMemcpyPeerAsync(samples)");
        }))
        return 0;
    if (!dist_sync_all(cs, P, "sync(samples)")) return 0;

    // 2. splitters = quantiles of the sorted sample (GPU 0), handed to every GPU
    uint32_t *splitters[kDistMaxGpus], *counts[kDistMaxGpus];
    for (int g = 0; g < P; g++) {
        splitters[g] = cs[g]->small + kDistSamples;
        counts[g] = cs[g]->small + kDistSamples + 32;
    }
    {
        DeviceCtx &c = *cs[0];
        const size_t all = (size_t)P * kDistSamples;
        if (!cuda_ok(cudaSetDevice(c.dev), "cudaSetDevice")) return 0;
        uint32_t *res = nullptr;
        if (!cuda_ok(sort_bits(c.gather, c.gather + all, c.gather + 2 * all, c.gather + all, all, 0, 32, nullptr,
                               BOUNDS_EXACT, c.ws_small, c.ws_small_cap, c.stream, &res, nullptr),
                     "sample sort"))
            return 0;
        if (!cuda_ok(pick_splitters(c.gather + all, all, P, splitters[0], c.stream), "pick splitters")) return 0;
        for (int g = 1; g < P; g++)
            if (!cuda_ok(cudaMemcpyPeerAsync(splitters[g], cs[g]->dev, splitters[0], c.dev, 32 * sizeof(uint32_t), c.stream),
                         "cudaMemcpyPeerAsync(splitters)"))
                return 0;
        if (!cuda_ok(cudaStreamSynchronize(c.stream), "sync(splitters)")) return 0;
    }

    // 3. keys per destination on every GPU -> receive layout (sources in GPU order, no gaps)
    uint32_t h_counts[kDistMaxGpus][16];
    for (int g = 0; g < P; g++) {
        DeviceCtx &c = *cs[g];
        if (!cuda_ok(cudaSetDevice(c.dev), "cudaSetDevice")) return 0;
        if (!cuda_ok(split_counts(c.a, shard_len(g), splitters[g], P - 1, counts[g], c.stream), "destination counts"))
            return 0;
        if (!cuda_ok(cudaMemcpyAsync(h_counts[g], counts[g], 16 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c.stream),
                     "cudaMemcpy(counts)"))
            return 0;
    }
    if (!dist_sync_all(cs, P, "sync(counts)")) return 0;
    size_t totals[kDistMaxGpus] = {}, offs[kDistMaxGpus][kDistMaxGpus];
    for (int d = 0; d < P; d++)
        for (int g = 0; g < P; g++) {
            offs[g][d] = totals[d];
            totals[d] += h_counts[g][d];
        }
    for (int d = 0; d < P; d++) {
        if (totals[d] > UINT32_MAX) {
            fprintf(stderr, "libsort: %zu keys fall into one GPU's range; the local sort is 32-bit\n", totals[d]);
            return 0;
        }
        if (!cuda_ok(cudaSetDevice(cs[d]->dev), "cudaSetDevice")) return 0;
        if (!dist_recv_buffer(*cs[d], totals[d])) return 0;
    }

    // 4. fused partition + exchange: each GPU stores its keys into its peers' receive buffers
    for (int g = 0; g < P; g++) {
        DeviceCtx &c = *cs[g];
        unsigned long long ptrs[kDistMaxGpus];
        for (int d = 0; d < P; d++) ptrs[d] = reinterpret_cast<unsigned long long>(cs[d]->recv + offs[g][d]);
        if (!cuda_ok(cudaSetDevice(c.dev), "cudaSetDevice")) return 0;
        if (!cuda_ok(scatter_by_splitters(c.a, shard_len(g), splitters[g], P - 1, ptrs, c.ws, c.ws_cap, c.stream),
                     "partition + exchange"))
            return 0;
    }
    if (!dist_sync_all(cs, P, "sync(exchange)")) return 0;  // every store has landed before anyone sorts

    // 5. local sort of what arrived (GPU d owns the d-th value range): all GPUs sort at once, then each result is
    //    audited (atomic-ranking check, include/libsort_b200.h libsortRankAudit; the received keys are still
    //    intact in c.recv, so a violation just repeats the sort with the ballot kernels) and goes straight
    //    back into the caller's array
    uint32_t *sorted[kDistMaxGpus] = {};
    auto local_sort = [&](int d) {
        DeviceCtx &c = *cs[d];
        return cuda_ok(sort_bits(c.recv, c.a, c.b, nullptr, totals[d], 0, 32, nullptr, BOUNDS_EXACT, c.ws, c.ws_cap, c.stream,
                                 &sorted[d], nullptr),
                       "local sort");
    };
    for (int d = 0; d < P; d++) {
        if (!totals[d]) continue;
        if (!cuda_ok(cudaSetDevice(cs[d]->dev), "cudaSetDevice")) return 0;
        if (!ensure_capacity(*cs[d], totals[d], 0)) return 0;  // the shard in c.a is no longer needed
        if (!local_sort(d)) return 0;
    }
    size_t first = 0, firsts[kDistMaxGpus];
    for (int d = 0; d < P; d++) {
        firsts[d] = first;
        first += totals[d];
    }
    if (!per_gpu([&](int d) {
            DeviceCtx &c = *cs[d];
            if (!totals[d]) return true;
            bool violated = false;
            if (!cuda_ok(rank_audit(c.stream, &violated), "rank audit")) return false;
            if (violated && !local_sort(d)) return false;
            return c.copier.to_host(h_keys + firsts[d], sorted[d], totals[d] * sizeof(uint32_t), c.stream);
        }))
        return 0;
    if (first != n) {
        fprintf(stderr, "libsort: distributed sort lost keys (%zu of %zu arrived)\n", first, n);
        return 0;
    }
    if (!dist_sync_all(cs, P, "sync(result)")) return 0;
    return 1;
}

// ---- step-wise distributed LSD sort: device building blocks (include/libsort_b200.h) ----
API size_t libsortStepTableBytes(void) { return step_table_bytes(); }
API int libsortStepTableInit(void *d_table, const unsigned long long *peers, const unsigned long long *slice_starts,
                             int nslice, void *stream)
{
    return (int)step_table_init(d_table, peers, slice_starts, nslice, (cudaStream_t)stream);
}
API int libsortStepCounts(const uint32_t *d_in, size_t n, uint32_t shift, uint32_t nbits, uint32_t *d_counts, void *stream)
{
    return (int)step_counts(d_in, n, shift, nbits, d_counts, (cudaStream_t)stream);
}
API int libsortStepBases(const uint32_t *d_gathered, int nranks, int me, void *d_table, void *stream)
{
    return (int)step_bases(d_gathered, nranks, me, d_table, (cudaStream_t)stream);
}
API int libsortStepScatter(const uint32_t *d_in, size_t n, uint32_t shift, uint32_t nbits, const void *d_table,
                           void *d_workspace, size_t workspace_len, void *stream)
{
    return (int)step_scatter(d_in, n, shift, nbits, d_table, d_workspace, workspace_len, (cudaStream_t)stream);
}
