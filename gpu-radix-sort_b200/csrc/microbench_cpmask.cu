#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__global__ void k(uint32_t *out, uint16_t mask)
{
    __shared__ __align__(128) uint32_t t[64];
    for (int i = threadIdx.x; i < 64; i += blockDim.x) t[i] = 100 + i;
    __syncthreads();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (threadIdx.x == 0) {
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group.cp_mask [%0], [%1], %2, %3;" ::"l"(out), "r"(smem_u32(t)), "r"(32), "h"(mask) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}
int main()
{
    uint32_t *d; cudaMalloc(&d, 256); cudaMemset(d, 0, 256);
    k<<<1, 32>>>(d, 0x0FF0);
    cudaError_t e = cudaDeviceSynchronize();
    uint32_t h[16]; cudaMemcpy(h, d, 64, cudaMemcpyDeviceToHost);
    printf("%s:", cudaGetErrorString(e));
    for (int i = 0; i < 12; i++) printf(" %u", h[i]);
    printf("\n");
}
