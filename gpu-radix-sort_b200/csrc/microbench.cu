// microbench.cu -- developer tool: B200 rates of the primitives the scatter
// kernel leans on (MATCH.ANY, shared-memory atomics, LDS/STS scatter) and a
// plain copy for the HBM ceiling.  Prints warp-instructions per SM-cycle.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace b200sort;

#define ITER 4096
__device__ __forceinline__ uint32_t rng(uint32_t &s) { s = s * 1664525u + 1013904223u; return s >> 24; }

__global__ void k_match(uint32_t *out, int bits)
{
    uint32_t s = threadIdx.x * 2654435761u + blockIdx.x, acc = 0, m = (1u << bits) - 1;
    long long t0 = clock64();
    for (int i = 0; i < ITER; i++) { uint32_t d = rng(s) & m; acc += __match_any_sync(0xffffffffu, d); }
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x * 2] = (uint32_t)(t1 - t0);
    out[blockIdx.x * 2 + 1] = acc;
}
__global__ void k_rngonly(uint32_t *out, int bits)
{
    uint32_t s = threadIdx.x * 2654435761u + blockIdx.x, acc = 0, m = (1u << bits) - 1;
    long long t0 = clock64();
    for (int i = 0; i < ITER; i++) { uint32_t d = rng(s) & m; acc += d; }
    long long t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x * 2] = (uint32_t)(t1 - t0);
    out[blockIdx.x * 2 + 1] = acc;
}
template <int MODE>  // 0 atomicAdd no return, 1 atomicAdd with return, 2 LDS+STS, 3 STS only
__global__ void k_smem(uint32_t *out, int bits)
{
    __shared__ uint32_t h[32][256];
    for (int i = threadIdx.x; i < 32 * 256; i += blockDim.x) (&h[0][0])[i] = 0;
    __syncthreads();
    uint32_t *my = h[threadIdx.x >> 5];
    uint32_t s = threadIdx.x * 2654435761u + blockIdx.x, acc = 0, m = (1u << bits) - 1;
    long long t0 = clock64();
    for (int i = 0; i < ITER; i++) {
        uint32_t d = rng(s) & m;
        if (MODE == 0) atomicAdd(&my[d], 1u);
        else if (MODE == 1) acc += atomicAdd(&my[d], 1u);
        else if (MODE == 2) { uint32_t v = my[d]; __syncwarp(); my[d] = v + 1; __syncwarp(); acc += v; }
        else my[d] = i;
    }
    long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x * 2] = (uint32_t)(t1 - t0);
    out[blockIdx.x * 2 + 1] = acc + my[threadIdx.x & 255];
}
__global__ void k_copy(const uint4 *in, uint4 *out, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    for (; i + 3 * st < n; i += 4 * st) {
        uint4 a = ld_stream_v4(in + i), b = ld_stream_v4(in + i + st), c = ld_stream_v4(in + i + 2 * st), d = ld_stream_v4(in + i + 3 * st);
        out[i] = a; out[i + st] = b; out[i + 2 * st] = c; out[i + 3 * st] = d;
    }
    for (; i < n; i += st) out[i] = in[i];
}
__global__ void k_read(const uint4 *in, uint32_t *out, size_t n)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
    uint32_t acc = 0;
    for (; i + 3 * st < n; i += 4 * st) {
        uint4 a = ld_stream_v4(in + i), b = ld_stream_v4(in + i + st), c = ld_stream_v4(in + i + 2 * st), d = ld_stream_v4(in + i + 3 * st);
        acc += a.x ^ b.y ^ c.z ^ d.w;
    }
    if (acc == 0x12345) out[0] = acc;
}


// Same-address shared-memory atomics inside ONE warp instruction: in which
// order does the hardware serve the colliding lanes?  Every lane adds 1 to the
// counter of a random digit and gets the old value back; among the lanes that
// named the same counter the returned values are consecutive, and this kernel
// counts how often they are NOT in ascending lane order.  SHARED = 1: all
// warps of the CTA use one counter array (cross-warp contention in flight).
template <int SHARED>
__global__ void k_atoms_order(unsigned long long *out, int bits, int iters)
{
    __shared__ uint32_t h[32][256];
    for (int i = threadIdx.x; i < 32 * 256; i += blockDim.x) (&h[0][0])[i] = 0;
    __syncthreads();
    uint32_t *my = SHARED ? h[0] : h[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31, lt = (1u << lane) - 1, m = (1u << bits) - 1;
    uint32_t s = (threadIdx.x + 1) * 2654435761u + blockIdx.x * 97u;
    unsigned long long bad = 0, pairs = 0;
    for (int i = 0; i < iters; i++) {
        uint32_t d = rng(s) & m;
        if ((i & 7) == 7) d = (lane * 5 + i) & 1;          // heavy collisions now and then
        const uint32_t r = atomicAdd(&my[d], 1u);
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        // the lowest peer lane must hold the smallest value and my value must be its value + my rank
        const int first = __ffs(peers) - 1;
        const uint32_t r0 = __shfl_sync(0xffffffffu, r, first);
        const uint32_t want = r0 + __popc(peers & lt);
        if (SHARED) {
            // other warps interleave: only the relative order inside the warp is checked
            const uint32_t prev_lane = 31 - __clz(peers & lt | 0u);
            const uint32_t rp = __shfl_sync(0xffffffffu, r, (peers & lt) ? prev_lane : lane);
            if ((peers & lt) && !(rp < r)) bad++;
        } else if (r != want) bad++;
        pairs += (peers & lt) ? 1 : 0;
    }
    atomicAdd(&out[0], bad);
    atomicAdd(&out[1], pairs);
}

template <typename F> static void rate(const char *name, F launch, int threads, uint32_t *d_out, int blocks)
{
    launch(); cudaDeviceSynchronize(); launch(); cudaDeviceSynchronize();
    uint32_t h[2]; cudaMemcpy(h, d_out, 8, cudaMemcpyDeviceToHost);
    double cyc = h[0];
    double warp_instr = (double)ITER * (threads / 32);
    printf("%-28s %4d thr/CTA x%d CTA/SM: %8.0f cyc  -> %.3f warp-ops/cyc/CTA, %.2f cyc per warp-op\n", name, threads, blocks, cyc, warp_instr / cyc, cyc / warp_instr);
}

int main()
{
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    uint32_t *d_out; cudaMalloc(&d_out, 1 << 20);
    for (int threads : {256, 1024}) {
        for (int bits : {8, 2}) {
            printf("-- digit bits %d\n", bits);
            rate("rng only", [&] { k_rngonly<<<sms, threads>>>(d_out, bits); }, threads, d_out, 1);
            rate("match_any", [&] { k_match<<<sms, threads>>>(d_out, bits); }, threads, d_out, 1);
            rate("atoms add (no ret)", [&] { k_smem<0><<<sms, threads>>>(d_out, bits); }, threads, d_out, 1);
            rate("atoms add (ret)", [&] { k_smem<1><<<sms, threads>>>(d_out, bits); }, threads, d_out, 1);
            rate("lds+sts rmw", [&] { k_smem<2><<<sms, threads>>>(d_out, bits); }, threads, d_out, 1);
            rate("sts random", [&] { k_smem<3><<<sms, threads>>>(d_out, bits); }, threads, d_out, 1);
        }
    }
    {
        unsigned long long *d_cnt; cudaMalloc(&d_cnt, 16);
        for (int shared = 0; shared < 2; shared++)
            for (int bits : {8, 4, 1}) {
                cudaMemset(d_cnt, 0, 16);
                if (shared) k_atoms_order<1><<<sms * 2, 1024>>>(d_cnt, bits, 20000);
                else k_atoms_order<0><<<sms * 2, 1024>>>(d_cnt, bits, 20000);
                cudaError_t e = cudaDeviceSynchronize();
                unsigned long long h[2]; cudaMemcpy(h, d_cnt, 16, cudaMemcpyDeviceToHost);
                printf("atoms lane order  %s counters, %d-bit digits: %llu colliding lanes checked, %llu out of ascending lane order (%s)\n",
                       shared ? "CTA-shared" : "warp-private", bits, h[1], h[0], cudaGetErrorString(e));
            }
    }
    size_t n = (size_t)1 << 30;  // bytes
    uint4 *a, *b; cudaMalloc(&a, n); cudaMalloc(&b, n); cudaMemset(a, 1, n); cudaMemset(b, 2, n);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int occ : {2, 4, 8}) {
        float ms;
        for (int r = 0; r < 3; r++) { cudaEventRecord(e0); k_copy<<<sms * occ, 256>>>(a, b, n / 16); cudaEventRecord(e1); cudaEventSynchronize(e1); }
        cudaEventElapsedTime(&ms, e0, e1);
        printf("copy 1GiB  grid=%d x256: %.3f ms  %.0f GB/s (r+w)\n", sms * occ, ms, 2.0 * n / ms / 1e6);
        for (int r = 0; r < 3; r++) { cudaEventRecord(e0); k_read<<<sms * occ, 256>>>(a, d_out, n / 16); cudaEventRecord(e1); cudaEventSynchronize(e1); }
        cudaEventElapsedTime(&ms, e0, e1);
        printf("read 1GiB  grid=%d x256: %.3f ms  %.0f GB/s\n", sms * occ, ms, 1.0 * n / ms / 1e6);
    }
    float ms; cudaEventRecord(e0); cudaMemcpyAsync(b, a, n, cudaMemcpyDeviceToDevice); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventRecord(e0); cudaMemcpyAsync(b, a, n, cudaMemcpyDeviceToDevice); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    printf("cudaMemcpy D2D 1GiB: %.3f ms %.0f GB/s (r+w)\n", ms, 2.0 * n / ms / 1e6);
    return 0;
}
