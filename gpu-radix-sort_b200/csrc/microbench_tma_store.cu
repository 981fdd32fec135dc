// microbench_tma_store.cu -- developer tool: can the bulk-copy engine write the
// per-bucket runs of a digit-sorted tile (about 256 runs of ~200 bytes, each 16 B
// aligned) from shared memory to global memory fast enough to replace the
// LDS + STG write-out of the scatter kernel?  Prints GB/s and runs/us.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace b200sort;

__device__ __forceinline__ void bulk_store(void *dst_gmem, const void *src_smem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_store_masked16(void *dst_gmem, const void *src_smem, uint32_t mask)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.cp_mask [%0], [%1], 16, %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
                 "h"((unsigned short)mask)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// MODE 0: one bulk store per bucket thread.  MODE 2: three (masked head, body, masked tail).  MODE 1: classic write-out (LDS + STG, coalesced per run).
template <int MODE>
__global__ void __launch_bounds__(256) k_store(uint32_t *out, size_t bucket_stride_words, int tiles_per_cta, int tile_words)
{
    extern __shared__ __align__(128) uint32_t tile[];
    __shared__ uint32_t s_start[257];
    const int tid = threadIdx.x;
    for (int i = tid; i < tile_words; i += 256) tile[i] = i * 2654435761u;
    // run lengths: multiples of 4 words, mean tile_words/256, varying +-50%
    const int mean4 = tile_words / 256 / 4;
    uint32_t h = (tid * 2654435761u) >> 16;
    int len4 = mean4 / 2 + (int)(h % (uint32_t)(mean4 + 1));
    __shared__ uint32_t s_len[256];
    s_len[tid] = len4 * 4;
    __syncthreads();
    if (tid == 0) {
        uint32_t run = 0;
        for (int d = 0; d < 256; d++) { s_start[d] = run; run += s_len[d]; if (run > (uint32_t)tile_words) { s_len[d] -= run - tile_words; run = tile_words; } }
        s_start[256] = run;
    }
    __syncthreads();
    const uint32_t start = s_start[tid], len = s_len[tid];
    uint32_t *dst = out + (size_t)tid * bucket_stride_words + (size_t)blockIdx.x * tiles_per_cta * (mean4 * 8);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    for (int t = 0; t < tiles_per_cta; t++) {
        if (MODE == 0) {
            if (len) bulk_store(dst, tile + start, len * 4);
            bulk_commit();
            bulk_wait_read();
        } else if (MODE == 2) {
            // ragged run: masked 16 B head, plain body, masked 16 B tail
            if (len >= 12) {
                bulk_store_masked16(dst, tile + start, 0xff00);
                bulk_store(dst + 4, tile + start + 4, (len - 8) * 4);
                bulk_store_masked16(dst + len - 4, tile + start + len - 4, 0x00ff);
            }
            bulk_commit();
            bulk_wait_read();
        } else {
            const uint32_t total = s_start[256];
            for (uint32_t p = tid; p < total; p += 256) {
                // bucket of position p by binary search over 256 starts (stand-in for the digit lookup)
                int lo = 0, hi = 255;
                while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (s_start[mid] <= p) lo = mid; else hi = mid - 1; }
                uint32_t *d2 = out + (size_t)lo * bucket_stride_words + (size_t)blockIdx.x * tiles_per_cta * (mean4 * 8) + (size_t)t * s_len[lo];
                d2[p - s_start[lo]] = tile[p];
            }
        }
        dst += len;
        __syncthreads();
    }
}

int main(int argc, char **argv)
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int tile_words = argc > 1 ? atoi(argv[1]) : 14336;
    const int ctas_per_sm = argc > 2 ? atoi(argv[2]) : 3;
    const int grid = sms * ctas_per_sm;
    const int tiles_per_cta = 64;
    const size_t total_words = (size_t)grid * tiles_per_cta * tile_words;
    const size_t bucket_stride = (total_words / 256 * 2 + 1024) & ~(size_t)3;
    uint32_t *out; cudaMalloc(&out, bucket_stride * 256 * 4);
    cudaFuncSetAttribute(k_store<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, tile_words * 4);
    cudaFuncSetAttribute(k_store<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, tile_words * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaFuncSetAttribute(k_store<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, tile_words * 4);
    for (int mode = 0; mode < 3; mode++) {
        float ms = 0;
        for (int r = 0; r < 3; r++) {
            cudaEventRecord(e0);
            if (mode == 0) k_store<0><<<grid, 256, tile_words * 4>>>(out, bucket_stride, tiles_per_cta, tile_words);
            else if (mode == 2) k_store<2><<<grid, 256, tile_words * 4>>>(out, bucket_stride, tiles_per_cta, tile_words);
            else k_store<1><<<grid, 256, tile_words * 4>>>(out, bucket_stride, tiles_per_cta, tile_words);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
        }
        cudaError_t e = cudaDeviceSynchronize();
        printf("%s tile=%d words, %d CTAs/SM: %.3f ms, %.0f GB/s written, %.1f runs/us/SM (%s)\n", mode == 0 ? "bulk-store " : mode == 2 ? "bulk 3-part" : "lds+stg    ",
               tile_words, ctas_per_sm, ms, total_words * 4.0 / ms / 1e6, (double)grid * tiles_per_cta * 256 / sms / (ms * 1e3), cudaGetErrorString(e));
    }
    return 0;
}
