// scatter3.cuh -- the digit-scatter pass (K3), pipelined and warp-specialised.
//
// Same contract as k_onesweep<RANK_ATOMS, RadixBinner, OUT_INDEX> (kernels.cuh); replaces
// gpu_radix_sort_local + sum_scan_blelloch + gpu_glbl_shuffle (reference libsort/sort.cu:29-213,
// scan.cu:16-250).  Why a third shape: with the bulk-copy write-out (scatter2.cuh) the pass needs
// 13.9 instead of 17 L1/shared wavefronts per 32 keys, but the look-back has to finish BEFORE the
// tile is re-ordered (the shared-memory layout follows the 16-byte phase of every bucket's global
// destination), and that wait, the serial issue of 256 bulk copies and the key load all sat on
// the critical path (profiles/r02a_*: LSU only 58-68 % busy).  Here they leave it:
//   * 8 compute warps keep TWO tiles in registers.  While tile X waits for its look-back, the
//     warps rank and scatter tile Y, then load and count the next tile: every tile gets a whole
//     rank + count phase for its look-back.
//   * warp 8 does the look-back (8 bins per lane, windowed, prompt INCL publication for the
//     successors), warp 9 the write-out: per bucket the ragged head/tail stores and ONE
//     cp.async.bulk.global.shared::cta (SASS UBLKCP.G.S) for the 16-byte-aligned body.  Compute
//     warps and helpers hand over through named barriers (one side arrives, the other waits).
//   * keys come straight from global memory (coalesced 32-bit loads, tile order == rank order);
//     whoever takes ticket t asks L2 for tile t + gridDim.x (cp.async.bulk.prefetch.L2).
// 2 CTAs x 320 threads per SM, 96 registers per thread, 256 x KPT keys per tile.
#pragma once
#include "kernels.cuh"
#include "onesweep_bulk.cuh"
#include "scatter2.cuh"

namespace b200sort {

constexpr int S3_COMPUTE = 256;          // compute threads = bin threads
constexpr int S3_THREADS = S3_COMPUTE + 64;  // + look-back warp + store warp
constexpr uint32_t S3_NONE = 0xffffffffu, S3_QUIT = 0xfffffffeu;

struct S3Bucket {
    unsigned long long dst;  // byte address of the first destination of this tile's run
    uint32_t start4;         // byte offset of the run in keys[]
    uint32_t count;          // keys of the run (padding keys excluded)
};

template <int KPT>
struct Scatter3Smem {
    static constexpr int KEYS_PER_THREAD = KPT;
    static constexpr int TILE = S3_COMPUTE * KPT;
    static constexpr int NWARPS = S3_COMPUTE / 32;
    static constexpr int PAD = RADIX * 6;
    alignas(128) uint32_t keys[TILE + PAD];       // the padded digit-sorted tile
    uint32_t warp_hist[2][NWARPS][RADIX];         // per register set: per-warp digit counters -> running BYTE offsets
    alignas(16) S3Bucket table[2][RADIX];         // per register set: what the helper writes out
    uint32_t lb_cnt[RADIX];                       // counts of the tile whose look-back is requested
    uint32_t lb_excl[RADIX];                      // its result: keys of every bin in all earlier tiles
    uint32_t s_warp[8];
    uint32_t lb_req_tile;    // tile whose look-back is requested (S3_NONE: nothing this step, S3_QUIT: leave)
    uint32_t bulk_req_set;   // register set whose sorted tile is ready for the write-out (S3_NONE / S3_QUIT)
    uint32_t t_new;
    RadixBinner::Arg binner_arg;
    uint32_t heavy_d4;
};

__device__ __forceinline__ void bar_compute() { asm volatile("bar.sync 1, 256;" ::: "memory"); }
__device__ __forceinline__ int bar_compute_or(bool p)
{
    uint32_t r;
    asm volatile(
        "{\n\t"
        ".reg .pred p, q;\n\t"
        "setp.ne.u32 p, %1, 0;\n\t"
        "bar.red.or.pred q, 1, 256, p;\n\t"
        "selp.u32 %0, 1, 0, q;\n\t"
        "}"
        : "=r"(r)
        : "r"((uint32_t)p)
        : "memory");
    return (int)r;
}
// Hand-off barriers between the compute warps and a helper warp (256 + 32 threads): one side
// arrives, the other waits -- no polling of shared memory while the LSU is the bottleneck.
constexpr int BAR_LBREQ = 2, BAR_LBDONE = 3, BAR_FULL = 4, BAR_FREE = 5;
template <int ID>
__device__ __forceinline__ void bar_arrive288()
{
    asm volatile("bar.arrive %0, 288;" ::"n"(ID) : "memory");
}
template <int ID>
__device__ __forceinline__ void bar_sync288()
{
    asm volatile("bar.sync %0, 288;" ::"n"(ID) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_cta(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_cta(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t atoms_add(uint32_t addr, uint32_t v)
{
    uint32_t r;
    asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(r) : "r"(addr), "r"(v) : "memory");
    return r;
}
__device__ __forceinline__ void sts_u32(uint32_t addr, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// 256-thread exclusive scan over the compute warps (named barrier 1)
__device__ __forceinline__ uint32_t compute_excl_scan_256(uint32_t v, uint32_t *s_warp)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    bar_compute();
    uint32_t wbase = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        uint32_t c = s_warp[w];
        if (w < warp) wbase += c;
    }
    return wbase + inc - v;
}

// cold path, see rank_tile_heavy in kernels.cuh (same ranking, explicit counter set)
template <int KPT>
__device__ __noinline__ void rank_tile_heavy3(const uint32_t *keys, Scatter3Smem<KPT> *sp, int set)
{
    Scatter3Smem<KPT> &s = *sp;
    const RadixBinner binner(s.binner_arg);
    const uint32_t hd4 = s.heavy_d4;
    unsigned char *const my_hist = reinterpret_cast<unsigned char *>(s.warp_hist[set][threadIdx.x >> 5]);
    unsigned char *const tile_bytes = reinterpret_cast<unsigned char *>(s.keys);
    const uint32_t lt = lanemask_lt();
    const int lane = threadIdx.x & 31;
#pragma unroll 4
    for (int i = 0; i < KPT; i++) {
        const uint32_t k = keys[i];
        const uint32_t d4 = binner.bin4(k);
        const bool is = d4 == hd4;
        const uint32_t m = __ballot_sync(0xffffffffu, is);  // warp-uniform
        const int top = 31 - __clz(m | 1u);
        const uint32_t before4 = __popc(m & lt) * 4u;
        uint32_t p4 = 0;
        if (!is || lane == top) p4 = atomicAdd(reinterpret_cast<uint32_t *>(my_hist + d4), is ? before4 + 4u : 4u);
        const uint32_t g4 = __shfl_sync(0xffffffffu, p4, top);
        if (is) p4 = g4 + before4;
        *reinterpret_cast<uint32_t *>(tile_bytes + p4) = k;
    }
}

// ---- helper warps -----------------------------------------------------------------------------
// Warp 8: the look-back of the tile named by lb_req_tile.  Nothing else runs in this warp: the
// length of a round trip sets how deep every look-back in the grid has to walk.
__device__ __forceinline__ uint4 ld_relaxed_v4(const uint32_t *p)
{
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_v4(uint32_t *p, uint4 v)
{
    asm volatile("st.relaxed.gpu.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// Lane l owns bins 8l .. 8l+7 (32 contiguous bytes of a tile's descriptor row): a window is LB
// whole rows, two 128-bit loads per lane and row, and the walk steps through tiles, not through
// (tile, bin) pairs.  With tiles published every ~40 cycles and a ~1000-cycle round trip the
// nearest inclusive prefix is ~25 tiles back: rows per round trip is what the latency hangs on.
template <int KPT, int LB>
__device__ __forceinline__ void scatter3_lookback_warp(Scatter3Smem<KPT> &s, uint32_t *__restrict__ desc)
{
    const int lane = threadIdx.x & 31;
    bar_arrive288<BAR_LBDONE>();  // pairs with the compute warps' first wait (nothing was requested yet)
    while (true) {
        bar_sync288<BAR_LBREQ>();
        const uint32_t t = s.lb_req_tile;
        if (t == S3_QUIT) return;
        if (t != S3_NONE) {
            uint32_t excl[8];
#pragma unroll
            for (int q = 0; q < 8; q++) excl[q] = 0;
            uint32_t done_mask = t == 0 ? 0xffu : 0u;  // per bin of this lane
            uint32_t j = t - 1;                         // next row (tile) to account for, common to all bins
            while (!__all_sync(0xffffffffu, done_mask == 0xffu)) {
                uint32_t win[LB][8];
#pragma unroll
                for (int k = 0; k < LB; k++) {
                    if ((uint32_t)k <= j) {
                        const uint32_t *row = desc + (size_t)(j - k) * RADIX + lane * 8;
                        const uint4 a = ld_relaxed_v4(row), b = ld_relaxed_v4(row + 4);
                        win[k][0] = a.x; win[k][1] = a.y; win[k][2] = a.z; win[k][3] = a.w;
                        win[k][4] = b.x; win[k][5] = b.y; win[k][6] = b.z; win[k][7] = b.w;
                    } else {
#pragma unroll
                        for (int q = 0; q < 8; q++) win[k][q] = DESC_INCL;  // before the first tile: nothing
                    }
                }
                bool stop = false;
#pragma unroll
                for (int k = 0; k < LB; k++) {
                    if (!stop) {
                        // a row counts only when every bin that still needs it has been published
                        bool missing = false;
#pragma unroll
                        for (int q = 0; q < 8; q++) missing |= !(done_mask >> q & 1u) && (win[k][q] >> 30) == 0;
                        if (__any_sync(0xffffffffu, missing)) {
                            stop = true;
                        } else {
#pragma unroll
                            for (int q = 0; q < 8; q++) {
                                if (!(done_mask >> q & 1u)) {
                                    excl[q] += win[k][q] & DESC_VAL;
                                    if (win[k][q] & DESC_INCL) done_mask |= 1u << q;
                                }
                            }
                            --j;
                            if (__all_sync(0xffffffffu, done_mask == 0xffu)) stop = true;
                        }
                    }
                }
            }
            if (t != 0) {
                uint32_t inc[8];
#pragma unroll
                for (int q = 0; q < 8; q++) inc[q] = DESC_INCL | (excl[q] + s.lb_cnt[lane * 8 + q]);
                uint32_t *row = desc + (size_t)t * RADIX + lane * 8;
                st_relaxed_v4(row, make_uint4(inc[0], inc[1], inc[2], inc[3]));
                st_relaxed_v4(row + 4, make_uint4(inc[4], inc[5], inc[6], inc[7]));
            }
#pragma unroll
            for (int q = 0; q < 8; q++) s.lb_excl[lane * 8 + q] = excl[q];
        }
        __threadfence_block();
        bar_arrive288<BAR_LBDONE>();
    }
}

// Warp 9: the write-out of the sorted tile of register set bulk_req_set.  Lane l of round r
// owns bucket 32r + l: the <= 3-key ragged head and tail go through the LSU, the 16-byte-aligned
// body is ONE bulk copy (SASS UBLKCP.G.S).
template <int KPT>
__device__ __forceinline__ void scatter3_store_warp(Scatter3Smem<KPT> &s)
{
    const int lane = threadIdx.x & 31;
    bar_arrive288<BAR_FREE>();  // the sorted buffer starts out free
    while (true) {
        bar_sync288<BAR_FULL>();
        const uint32_t set = s.bulk_req_set;
        if (set == S3_QUIT) break;
        if (set != S3_NONE) {
            fence_proxy_async();  // the compute warps' generic-proxy stores come before the engine's reads
#pragma unroll 1
            for (int r = 0; r < 8; r++) {
                const S3Bucket e = s.table[set][r * 32 + lane];
                if (e.count) {
                    const uint32_t *src = reinterpret_cast<const uint32_t *>(reinterpret_cast<const unsigned char *>(s.keys) + e.start4);
                    uint32_t *dst = reinterpret_cast<uint32_t *>(e.dst);
                    const uint32_t phase = (uint32_t)(e.dst >> 2) & 3u;
                    const uint32_t head = min(e.count, (4u - phase) & 3u);
                    const uint32_t body = (e.count - head) & ~3u;
                    const uint32_t tail = e.count - head - body;
                    if (head == ((4u - phase) & 3u)) {
                        // head ends on a 16 B boundary, tail starts on one: 8-byte stores where a pair is whole
                        if (head & 1u) dst[0] = src[0];
                        if (head & 2u) *reinterpret_cast<uint2 *>(dst + (head & 1u)) = *reinterpret_cast<const uint2 *>(src + (head & 1u));
                        const uint32_t t0 = head + body;
                        if (tail & 2u) *reinterpret_cast<uint2 *>(dst + t0) = *reinterpret_cast<const uint2 *>(src + t0);
                        if (tail & 1u) dst[t0 + (tail & 2u)] = src[t0 + (tail & 2u)];
                    } else {
                        for (uint32_t i = 0; i < head; i++) dst[i] = src[i];
                    }
                    if (body) bulk_store(dst + head, src + head, body * 4u);
                }
            }
            bulk_commit();
            bulk_wait_read();  // the engine has read the tile: the sorted buffer is free again
        }
        bar_arrive288<BAR_FREE>();
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ---- the kernel -----------------------------------------------------------------------------
// WOUT: who writes the sorted tile out.  0 -- the store warp (one warp: a bulk copy blocks its
// issuer for ~70 cycles, 256 of them take longer than a pipeline step).  1 -- the bin threads
// themselves, four lanes at a time in between the count atomics of the NEXT tile (the LSU-bound
// count phase has issue slots to spare, and eight warps keep eight copies in flight).
template <int KPT, int LB, int BATCH, bool PREFETCH, int WOUT>
__global__ void __launch_bounds__(S3_THREADS, 2)
    k_scatter3(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, uint32_t n, const RadixBinner::Arg binner_arg,
               const uint32_t *__restrict__ base, uint32_t *__restrict__ base_next, uint32_t *__restrict__ desc,
               uint32_t *__restrict__ ticket, uint32_t ntiles, int in_aligned)
{
    using Smem = Scatter3Smem<KPT>;
    constexpr int TILE = Smem::TILE;
    constexpr int NWARPS = Smem::NWARPS;
    static_assert(KPT % BATCH == 0, "rank batches must tile the keys of a thread");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem &s = *reinterpret_cast<Smem *>(smem_raw);

    if (ticket[SKIP_OFFSET]) {  // constant digit (k_scan_hist): the pass is a copy
        copy_keys(in, out, n);
        return;
    }
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        s.binner_arg = binner_arg;
        s.t_new = atomicAdd(ticket, 1u);
    }
    if (warp < NWARPS) {
#pragma unroll
        for (int j = lane; j < RADIX; j += 32) s.warp_hist[0][warp][j] = s.warp_hist[1][warp][j] = 0;
    }
    __syncthreads();
    if (warp == NWARPS) {
        scatter3_lookback_warp<KPT, LB>(s, desc);
        return;
    }
    if (warp == NWARPS + 1) {
        if (WOUT == 0) scatter3_store_warp<KPT>(s);
        return;
    }

    const RadixBinner binner(binner_arg);
    const uint32_t pad_bin = binner.pad_bin();
    const uint32_t gbase = base[tid];
    const uint32_t tile_addr = smem_u32(s.keys);
    const uint32_t wbase = warp * (KPT * 32) + lane;
    uint32_t kA[KPT], kB[KPT];
    uint32_t my_count[2] = {0, 0}, padded[2] = {0, 0};  // this thread's bin, per register set
    uint32_t tiles[2] = {S3_NONE, S3_NONE};
    bool pending = false;  // WOUT 1: a sorted tile sits in keys[], described by table[pset]
    int pset = 0;

    // WOUT 1: lanes 4*slot .. 4*slot+3 write their buckets of the pending tile out
    auto writeout_slot = [&](int slot) {
        if ((lane >> 2) != slot) return;
        const S3Bucket e = s.table[pset][tid];
        if (!e.count) return;
        const uint32_t *src = reinterpret_cast<const uint32_t *>(reinterpret_cast<const unsigned char *>(s.keys) + e.start4);
        uint32_t *dst = reinterpret_cast<uint32_t *>(e.dst);
        const uint32_t phase = (uint32_t)(e.dst >> 2) & 3u;
        const uint32_t head = min(e.count, (4u - phase) & 3u);
        const uint32_t body = (e.count - head) & ~3u;
        const uint32_t tail = e.count - head - body;
        if (head == ((4u - phase) & 3u)) {
            if (head & 1u) dst[0] = src[0];
            if (head & 2u) *reinterpret_cast<uint2 *>(dst + (head & 1u)) = *reinterpret_cast<const uint2 *>(src + (head & 1u));
            const uint32_t t0 = head + body;
            if (tail & 2u) *reinterpret_cast<uint2 *>(dst + t0) = *reinterpret_cast<const uint2 *>(src + t0);
            if (tail & 1u) dst[t0 + (tail & 2u)] = src[t0 + (tail & 2u)];
        } else {
            for (uint32_t i = 0; i < head; i++) dst[i] = src[i];
        }
        if (body) {
            fence_proxy_async();  // the generic-proxy stores of the rank phase come before the engine's reads
            bulk_store(dst + head, src + head, body * 4u);
            bulk_commit();
        }
    };

    // One pipeline step: take a tile into register set X and count it; then lay out, rank and hand
    // over the tile that has been waiting in register set Y.
    auto step = [&](uint32_t(&keyX)[KPT], const int setX, uint32_t(&keyY)[KPT], const int setY) {
        // (a) acquire + load + count
        const uint32_t tx = s.t_new;
        tiles[setX] = tx;
        const bool haveX = tx < ntiles;
        const uint32_t tile_base = tx * (uint32_t)TILE;
        const uint32_t valid = haveX ? min((uint32_t)TILE, n - tile_base) : 0u;
        if (haveX) {
            if (PREFETCH && tid == 0 && in_aligned && tx + gridDim.x < ntiles) {
                const uint32_t ahead = tx + gridDim.x;
                const uint32_t bytes = (min((uint32_t)TILE, n - ahead * (uint32_t)TILE) * 4u) & ~15u;
                if (bytes) prefetch_l2_bulk(in + (size_t)ahead * TILE, bytes);
            }
            const uint32_t *src = in + (size_t)tile_base + wbase;
            if (valid == TILE) {
#pragma unroll
                for (int i = 0; i < KPT; i++) keyX[i] = ld_stream_u32(src + i * 32);
            } else {
#pragma unroll
                for (int i = 0; i < KPT; i++) keyX[i] = wbase + i * 32 < valid ? ld_stream_u32(src + i * 32) : 0xffffffffu;
            }
#pragma unroll
            for (int i = 0; i < KPT; i++) {
                atomicAdd(&s.warp_hist[setX][warp][binner.bin(keyX[i])], 1u);
                if (WOUT == 1 && i % 5 == 0 && i / 5 < 8 && pending) writeout_slot(i / 5);
            }
            if (WOUT == 1 && KPT < 36 && pending)
                for (int slot = (KPT + 4) / 5; slot < 8; slot++) writeout_slot(slot);
        } else if (WOUT == 1 && pending) {
            for (int slot = 0; slot < 8; slot++) writeout_slot(slot);
        }
        pending = false;
        bar_compute();
        if (haveX) {
            uint32_t p = 0;
#pragma unroll
            for (int w = 0; w < NWARPS; w++) p += s.warp_hist[setX][w][tid];
            padded[setX] = p;
            uint32_t c = p;
            if ((uint32_t)tid == pad_bin) c -= (uint32_t)TILE - valid;  // padding keys carry the top digit
            my_count[setX] = c;
            st_relaxed(&desc[(size_t)tx * RADIX + tid], (tx == 0 ? DESC_INCL : DESC_AGG) | c);
        }
        // the waiting tile's look-back result (requested one step ago)
        const uint32_t ty = tiles[setY];
        const bool haveY = ty < ntiles;
        bar_sync288<BAR_LBDONE>();  // the look-back warp has finished the previous request
        const uint32_t excl = haveY ? s.lb_excl[tid] : 0u;
        if (haveX) s.lb_cnt[tid] = my_count[setX];
        if (tid == 0) s.lb_req_tile = haveX ? tx : S3_NONE;
        __threadfence_block();
        bar_arrive288<BAR_LBREQ>();  // every bin thread has read lb_excl and written lb_cnt
        if (!haveY) {
            // nothing is waiting (first step): pass the sorted buffer on unused, take the next ticket
            if (WOUT == 0) bar_sync288<BAR_FREE>();
            if (tid == 0) {
                s.bulk_req_set = S3_NONE;
                if (haveX) s.t_new = atomicAdd(ticket, 1u);
            }
            if (WOUT == 0) {
                __threadfence_block();
                bar_arrive288<BAR_FULL>();
            }
            bar_compute();
            return;
        }

        // (b) lay the waiting tile out with the 16-byte phase of every bucket's destination
        const unsigned long long dst_addr = reinterpret_cast<unsigned long long>(out) + 4ull * ((unsigned long long)gbase + excl);
        if (base_next && ty == ntiles - 1) base_next[tid] = gbase + excl + my_count[setY];
        const uint32_t phase = (uint32_t)(dst_addr >> 2) & 3u;
        const uint32_t pY = padded[setY];
        const uint32_t span4 = pY ? ((phase + pY + 3u) & ~3u) * 4u : 0u;
        const uint32_t slot4 = compute_excl_scan_256(span4, s.s_warp);
        const uint32_t start4 = slot4 + phase * 4u;
        const bool is_heavy = pY * 4u > (uint32_t)TILE;  // one digit holds > 1/4 of the tile
        {
            uint32_t run = start4;
#pragma unroll
            for (int w = 0; w < NWARPS; w++) {
                const uint32_t c = s.warp_hist[setY][w][tid];
                s.warp_hist[setY][w][tid] = run;
                run += c << 2;
            }
            if (is_heavy) s.heavy_d4 = (uint32_t)tid * 4u;
            S3Bucket e;
            e.dst = dst_addr;
            e.start4 = start4;
            e.count = my_count[setY];
            s.table[setY][tid] = e;
        }
        // the sorted buffer must be free: the copies of the previous tile have read it
        if (WOUT == 0) bar_sync288<BAR_FREE>();
        else bulk_wait_read();
        uint32_t t_next = tx;
        if (tid == 0 && haveX) t_next = atomicAdd(ticket, 1u);  // in flight during the rank phase
        const int heavy = bar_compute_or(is_heavy);

        // (c) rank every key of the waiting tile and move it to its place in the padded sorted tile
        if (!heavy) {
            const uint32_t hist_addr = smem_u32(s.warp_hist[setY][warp]);
#pragma unroll
            for (int i0 = 0; i0 < KPT; i0 += BATCH) {
                uint32_t p4[BATCH];
#pragma unroll
                for (int b = 0; b < BATCH; b++) p4[b] = atoms_add(hist_addr + binner.bin4(keyY[i0 + b]), 4u);
#pragma unroll
                for (int b = 0; b < BATCH; b++) sts_u32(tile_addr + p4[b], keyY[i0 + b]);
            }
        } else {
            uint32_t spilled[KPT];
#pragma unroll
            for (int i = 0; i < KPT; i++) spilled[i] = keyY[i];
            rank_tile_heavy3<KPT>(spilled, &s, setY);
        }
        __syncwarp();
#pragma unroll
        for (int j = lane; j < RADIX; j += 32) s.warp_hist[setY][warp][j] = 0;
        if (tid == 0) {
            s.t_new = t_next;
            s.bulk_req_set = (uint32_t)setY;
        }
        if (WOUT == 0) {
            __threadfence_block();
            bar_arrive288<BAR_FULL>();  // this thread's keys are in place
        }
        bar_compute();  // t_new is visible, every key is in place
        pending = true;
        pset = setY;
        tiles[setY] = S3_NONE;
    };

    while (true) {
        step(kA, 0, kB, 1);
        if (tiles[0] >= ntiles) break;
        step(kB, 1, kA, 0);
        if (tiles[1] >= ntiles) break;
    }
    if (WOUT == 1 && pending) {
        for (int slot = 0; slot < 8; slot++) writeout_slot(slot);
    }
    // consume the helpers' last hand-backs, then send them home
    bar_sync288<BAR_LBDONE>();
    if (WOUT == 0) bar_sync288<BAR_FREE>();
    if (tid == 0) {
        s.lb_req_tile = S3_QUIT;
        s.bulk_req_set = S3_QUIT;
    }
    __threadfence_block();
    bar_arrive288<BAR_LBREQ>();
    if (WOUT == 0) bar_arrive288<BAR_FULL>();
    else asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace b200sort
