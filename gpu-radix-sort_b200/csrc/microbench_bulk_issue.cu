// microbench_bulk_issue.cu -- developer tool: how fast can ONE warp hand 256 per-bucket runs of a
// digit-sorted tile to the bulk-copy engine (cp.async.bulk.global.shared::cta, SASS UBLKCP.G.S)?
// The store warp of k_scatter3 (scatter3.cuh) must finish a tile's 256 copies inside one pipeline
// step.  Variants: A = every lane issues its own copy (the compiler serialises the lanes with a
// BRA.U.ANY loop, operands through the same uniform registers); B = operands broadcast with
// shuffles, lane 0 issues, fully unrolled (distinct uniform registers, no loop); C = operands read
// from shared memory at a warp-uniform address, unrolled.  W = warps issuing per CTA.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace b200sort;

__device__ __forceinline__ void bulk_store(void *dst_gmem, uint32_t src_smem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(src_smem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

struct E { unsigned long long dst; uint32_t src; uint32_t bytes; };

template <int MODE>
__global__ void __launch_bounds__(256) k_issue(uint32_t *out, size_t bucket_stride_words, int tiles, int tile_words, int nwarps)
{
    extern __shared__ __align__(128) uint32_t tile[];
    __shared__ __align__(16) E tab[256];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < tile_words; i += 256) tile[i] = i * 2654435761u;
    const uint32_t len = (uint32_t)(tile_words / 256) & ~3u;  // words per run, multiple of 4
    tab[tid].src = smem_u32(tile) + tid * len * 4;
    tab[tid].bytes = len * 4;
    tab[tid].dst = reinterpret_cast<unsigned long long>(out + (size_t)tid * bucket_stride_words + (size_t)blockIdx.x * tiles * len);
    __syncthreads();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp >= nwarps) return;
    const int rounds = 8 / nwarps;  // rounds of 32 buckets per issuing warp
    for (int t = 0; t < tiles; t++) {
        for (int r = 0; r < rounds; r++) {
            const int b0 = (warp * rounds + r) * 32;
            if (MODE == 0) {
                const E e = tab[b0 + lane];
                bulk_store(reinterpret_cast<void *>(e.dst + (size_t)t * len * 4), e.src, e.bytes);
            } else if (MODE == 1) {
                const E e = tab[b0 + lane];
#pragma unroll
                for (int l = 0; l < 32; l++) {
                    const unsigned long long d = __shfl_sync(0xffffffffu, e.dst, l);
                    const uint32_t sa = __shfl_sync(0xffffffffu, e.src, l);
                    const uint32_t by = __shfl_sync(0xffffffffu, e.bytes, l);
                    if (lane == 0 && by) bulk_store(reinterpret_cast<void *>(d + (size_t)t * len * 4), sa, by);
                }
            } else {
#pragma unroll
                for (int l = 0; l < 32; l++) {
                    const E e = tab[b0 + l];
                    if (e.bytes) bulk_store(reinterpret_cast<void *>(e.dst + (size_t)t * len * 4), e.src, e.bytes);
                }
            }
        }
        bulk_commit();
        bulk_wait_read();
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main(int argc, char **argv)
{
    int sms;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int tile_words = argc > 1 ? atoi(argv[1]) : 11264;
    const int ctas_per_sm = argc > 2 ? atoi(argv[2]) : 2;
    const int grid = argc > 3 ? atoi(argv[3]) : sms * ctas_per_sm, tiles = 64;  // a small grid takes DRAM out of the picture
    const size_t total_words = (size_t)grid * tiles * tile_words;
    const size_t bucket_stride = (total_words / 256 * 2 + 1024) & ~(size_t)3;
    uint32_t *out;
    cudaMalloc(&out, bucket_stride * 256 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    void (*kern[3])(uint32_t *, size_t, int, int, int) = {k_issue<0>, k_issue<1>, k_issue<2>};
    const char *names[3] = {"A lanes (compiler loop)", "B shfl + lane 0, unrolled", "C uniform LDS, unrolled"};
    for (int m = 0; m < 3; m++) {
        cudaFuncSetAttribute(kern[m], cudaFuncAttributeMaxDynamicSharedMemorySize, tile_words * 4);
        for (int nw = 1; nw <= 8; nw *= 2) {
            float ms = 0;
            for (int r = 0; r < 3; r++) {
                cudaEventRecord(e0);
                kern[m]<<<grid, 256, tile_words * 4>>>(out, bucket_stride, tiles, tile_words, nw);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                cudaEventElapsedTime(&ms, e0, e1);
            }
            cudaError_t e = cudaDeviceSynchronize();
            const double ops_per_sm = (double)(argc > 3 ? 1 : ctas_per_sm) * tiles * 256;
            printf("%-28s W=%d tile=%d x%d CTAs/SM: %.3f ms, %.0f GB/s, %.1f clk/op/SM, %.0f clk per tile per CTA (%s)\n", names[m], nw,
                   tile_words, ctas_per_sm, ms, total_words * 4.0 / ms / 1e6, ms * 1e-3 * 1.965e9 / ops_per_sm,
                   ms * 1e-3 * 1.965e9 / tiles, cudaGetErrorString(e));
        }
    }
    return 0;
}
