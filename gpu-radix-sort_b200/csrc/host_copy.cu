// host_copy.cu -- see host_copy.h.
#include "host_copy.h"

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

namespace b200sort {

namespace {
constexpr size_t kDirectBelow = 4u << 20;  // small copies: let the driver stage them
constexpr size_t kChunk = 2u << 20;   // LIBSORT_COPY_CHUNK_MB overrides

bool ok(cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return true;
    fprintf(stderr, "libsort: %s failed: %s\n", what, cudaGetErrorString(e));
    return false;
}

bool is_pinned(const void *p)
{
    cudaPointerAttributes at{};
    cudaError_t e = cudaPointerGetAttributes(&at, p);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

int worker_count()
{
    int w = 16;  // one memcpy thread moves ~5 GB/s next to the DMA traffic: 16 x 2 MB chunks measured best (profiles/r02_pageable_copy_sweep.log)
    if (const char *e = getenv("LIBSORT_COPY_THREADS")) w = atoi(e);
    unsigned hc = std::thread::hardware_concurrency();
    if (hc && (int)hc < w) w = (int)hc;
    return std::max(1, std::min(w, HostCopier::kMaxWorkers));
}
}  // namespace

bool HostCopier::ensure(int workers)
{
    if (nworkers_ >= workers) return true;
    release();
    chunk_ = kChunk;
    if (const char *e = getenv("LIBSORT_COPY_CHUNK_MB"))
        if (atoi(e) > 0) chunk_ = (size_t)atoi(e) << 20;
    if (!ok(cudaHostAlloc((void **)&pinned_, (size_t)workers * kSlots * chunk_, cudaHostAllocDefault), "cudaHostAlloc"))
        return false;
    for (int w = 0; w < workers; w++) {
        if (!ok(cudaStreamCreateWithFlags(&wstream_[w], cudaStreamNonBlocking), "cudaStreamCreate")) return false;
        for (int s = 0; s < kSlots; s++)
            if (!ok(cudaEventCreateWithFlags(&wevent_[w][s], cudaEventDisableTiming), "cudaEventCreate")) return false;
        if (!ok(cudaEventCreateWithFlags(&done_[w], cudaEventDisableTiming), "cudaEventCreate")) return false;
    }
    if (!ok(cudaEventCreateWithFlags(&ready_, cudaEventDisableTiming), "cudaEventCreate")) return false;
    nworkers_ = workers;
    return true;
}

void HostCopier::release()
{
    for (int w = 0; w < nworkers_; w++) {
        if (wstream_[w]) cudaStreamDestroy(wstream_[w]);
        wstream_[w] = nullptr;
        for (int s = 0; s < kSlots; s++) {
            if (wevent_[w][s]) cudaEventDestroy(wevent_[w][s]);
            wevent_[w][s] = nullptr;
        }
        if (done_[w]) cudaEventDestroy(done_[w]);
        done_[w] = nullptr;
    }
    if (ready_) cudaEventDestroy(ready_);
    ready_ = nullptr;
    if (pinned_) cudaFreeHost(pinned_);
    pinned_ = nullptr;
    nworkers_ = 0;
}

// Worker w owns chunks w, w+W, w+2W, ... and two staging slots.
bool HostCopier::staged(void *dst, const void *src, size_t bytes, cudaStream_t stream, bool h2d)
{
    const int W = worker_count();
    if (!ensure(W)) return false;
    int dev = 0;
    if (!ok(cudaGetDevice(&dev), "cudaGetDevice")) return false;
    const size_t nchunks = (bytes + chunk_ - 1) / chunk_;
    if (!h2d) {
        // the DMA engines may start only after the sort has finished
        if (!ok(cudaEventRecord(ready_, stream), "cudaEventRecord")) return false;
    }
    std::atomic<bool> failed{false};
    auto body = [&](int w) {
        if (cudaSetDevice(dev) != cudaSuccess) { failed = true; return; }
        cudaStream_t ws = wstream_[w];
        unsigned char *slot[kSlots];
        for (int s = 0; s < kSlots; s++) slot[s] = pinned_ + ((size_t)w * kSlots + s) * chunk_;
        if (h2d) {
            int use = 0;
            bool used[kSlots] = {false, false};
            for (size_t c = w; c < nchunks; c += W, use ^= 1) {
                const size_t off = c * chunk_, len = std::min(chunk_, bytes - off);
                if (used[use] && cudaEventSynchronize(wevent_[w][use]) != cudaSuccess) { failed = true; return; }
                memcpy(slot[use], (const unsigned char *)src + off, len);
                if (cudaMemcpyAsync((unsigned char *)dst + off, slot[use], len, cudaMemcpyHostToDevice, ws) != cudaSuccess ||
                    cudaEventRecord(wevent_[w][use], ws) != cudaSuccess) { failed = true; return; }
                used[use] = true;
            }
            if (cudaEventRecord(done_[w], ws) != cudaSuccess) failed = true;
            // staging slots are reused by the next call: drain before returning
            if (cudaStreamSynchronize(ws) != cudaSuccess) failed = true;
        } else {
            if (cudaStreamWaitEvent(ws, ready_, 0) != cudaSuccess) { failed = true; return; }
            // software pipeline: DMA chunk k+1 while copying chunk k out of staging
            const size_t step = W;
            size_t c = w;
            int use = 0;
            auto start = [&](size_t cc, int s) {
                const size_t off = cc * chunk_, len = std::min(chunk_, bytes - off);
                return cudaMemcpyAsync(slot[s], (const unsigned char *)src + off, len, cudaMemcpyDeviceToHost, ws) == cudaSuccess &&
                       cudaEventRecord(wevent_[w][s], ws) == cudaSuccess;
            };
            if (c < nchunks && !start(c, use)) { failed = true; return; }
            for (; c < nchunks; c += step, use ^= 1) {
                const size_t nxt = c + step;
                if (nxt < nchunks && !start(nxt, use ^ 1)) { failed = true; return; }
                if (cudaEventSynchronize(wevent_[w][use]) != cudaSuccess) { failed = true; return; }
                const size_t off = c * chunk_, len = std::min(chunk_, bytes - off);
                memcpy((unsigned char *)dst + off, slot[use], len);
            }
        }
    };
    const int active = (int)std::min<size_t>(W, nchunks);
    std::vector<std::thread> th;
    for (int w = 1; w < active; w++) th.emplace_back(body, w);
    body(0);
    for (auto &t : th) t.join();
    if (failed.load()) {
        fprintf(stderr, "libsort: staged host copy failed: %s\n", cudaGetErrorString(cudaGetLastError()));
        return false;
    }
    if (h2d)
        for (int w = 0; w < active; w++)
            if (!ok(cudaStreamWaitEvent(stream, done_[w], 0), "cudaStreamWaitEvent")) return false;
    return true;
}

bool HostCopier::to_device(void *dst, const void *src, size_t bytes, cudaStream_t stream)
{
    if (bytes == 0) return true;
    if (bytes < kDirectBelow || is_pinned(src))
        return ok(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream), "cudaMemcpyAsync(H2D)");
    return staged(dst, src, bytes, stream, true);
}

bool HostCopier::to_host(void *dst, const void *src, size_t bytes, cudaStream_t stream)
{
    if (bytes == 0) return true;
    if (bytes < kDirectBelow || is_pinned(dst))
        return ok(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream), "cudaMemcpyAsync(D2H)");
    return staged(dst, src, bytes, stream, false);
}

}  // namespace b200sort
