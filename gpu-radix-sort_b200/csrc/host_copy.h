// host_copy.h -- host<->device key movement for the host-pointer ABI.
// Replaces the reference's blocking pageable cudaMemcpy pair
// (libsort/sort.cu:275 H2D, sort.cu:397 D2H; 327 ms + 321 ms per GiB in
// analysis/libsort8b.csv:7-8).
//
// Pinned caller memory is DMA'd directly.  Pageable caller memory is moved
// through a small ring of cached pinned staging chunks by a few worker
// threads, each overlapping its memcpy with the DMA of its previous chunk.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace b200sort {

class HostCopier {
  public:
    // Enqueue host -> device; on return `stream` is ordered after the copy and
    // the host buffer may be reused.
    bool to_device(void *dst, const void *src, size_t bytes, cudaStream_t stream);
    // device -> host, ordered after prior work on `stream`; on return the data
    // is in `dst` for pinned or staged copies alike.
    bool to_host(void *dst, const void *src, size_t bytes, cudaStream_t stream);
    void release();

    static constexpr int kMaxWorkers = 16;
    static constexpr int kSlots = 2;

  private:
    bool ensure(int workers);
    bool staged(void *dst, const void *src, size_t bytes, cudaStream_t stream, bool h2d);
    int nworkers_ = 0;
    size_t chunk_ = 0;
    unsigned char *pinned_ = nullptr;  // nworkers * kSlots * chunk
    cudaStream_t wstream_[kMaxWorkers] = {};
    cudaEvent_t wevent_[kMaxWorkers][kSlots] = {};
    cudaEvent_t done_[kMaxWorkers] = {};
    cudaEvent_t ready_ = nullptr;
};

}  // namespace b200sort
