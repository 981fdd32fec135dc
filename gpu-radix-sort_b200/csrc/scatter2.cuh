// scatter2.cuh -- the digit-scatter pass (K3), second generation.
//
// Same contract as k_onesweep<RANK_ATOMS, RadixBinner, OUT_INDEX> (kernels.cuh): one stable
// 8-bit-digit scatter of `in` into `out`, decoupled look-back over per-tile bucket counts
// (replaces gpu_radix_sort_local + sum_scan_blelloch + gpu_glbl_shuffle, reference
// libsort/sort.cu:29-213, scan.cu:16-250).  What changed, and why: the round-1 kernel is
// bound by L1/shared-memory LSU wavefronts (17 per 32 keys, profiles/r01c_k_onesweep_lsu_
// wavefronts.txt), 4.5 of which are its write-out (sorted tile -> registers -> per-key
// destination lookup -> STG at 2.5 lines per warp row).  Here the digit-sorted tile is laid
// out with the 16-byte phase of every bucket's global destination and the bulk-copy engine
// (cp.async.bulk.global.shared::cta, SASS UBLKCP.G.S) writes each bucket's run; only the
// <= 3-key ragged ends go through the LSU (or, RAG_MASK, through byte-masked 16 B bulk
// copies).  The sorted tile has its own buffer, so the next tile's load no longer waits for
// the copy engine to drain this one:
//   LOAD_TMA_SEP      1-D bulk TMA into a separate landing buffer (issued as soon as the
//                     keys of the current tile are in registers)
//   LOAD_TMA_INPLACE  landing buffer == sorted buffer (round-1 bulk kernel's scheme)
//   LOAD_LDG          no landing buffer: coalesced 32-bit loads of a tile that the CTA which
//                     took ticket t - gridDim.x prefetched into L2
#pragma once
#include "kernels.cuh"
#include "onesweep_bulk.cuh"

namespace b200sort {

constexpr int LOAD_TMA_SEP = 0, LOAD_TMA_INPLACE = 1, LOAD_LDG = 2;
constexpr int RAG_SCALAR = 0, RAG_PAIR = 1, RAG_MASK = 2;

// 16-byte bulk copy shared -> global of the bytes selected by `mask` (bit i = byte i)
__device__ __forceinline__ void bulk_store_masked16(void *dst_gmem, const void *src_smem, uint32_t mask)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.cp_mask [%0], [%1], 16, %2;" ::"l"(dst_gmem),
                 "r"(smem_u32(src_smem)), "h"((unsigned short)mask)
                 : "memory");
}
__device__ __forceinline__ void prefetch_l2_bulk(const void *gmem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t ld_stream_u32(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

template <int THREADS, int KPT, int LOAD, class BinnerArg>
struct Scatter2Smem {
    static constexpr int KEYS_PER_THREAD = KPT;
    static constexpr int TILE = THREADS * KPT;
    static constexpr int NWARPS = THREADS / 32;
    static constexpr int PAD = RADIX * 6;        // <= 3 keys of head and of tail padding per bucket
    alignas(128) uint32_t keys[TILE + PAD];      // the padded digit-sorted tile (LOAD_TMA_INPLACE: first the landing buffer)
    alignas(128) uint32_t landing[LOAD == LOAD_TMA_SEP ? TILE : 4];
    uint32_t warp_hist[NWARPS][RADIX];           // per-warp digit counters -> running BYTE offsets into keys[]
    uint32_t s_warp[8];
    alignas(8) uint64_t mbar;
    uint32_t next_tile;
    BinnerArg binner_arg;                        // copy for the out-of-line heavy-tile path
    uint32_t heavy_d4;
};

// EARLY (LOAD_TMA_SEP only): take the next ticket and start its load right after the count
// phase (the landing buffer is free from there on) instead of after the rank phase.
// PREFETCH: whoever takes ticket t also asks L2 for tile t + gridDim.x.
template <int THREADS, int KPT, int MINB, int LB, int LOAD, int RAG, bool EARLY, bool PREFETCH>
__global__ void __launch_bounds__(THREADS, MINB)
    k_scatter2(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, uint32_t n, const RadixBinner::Arg binner_arg,
               const uint32_t *__restrict__ base, uint32_t *__restrict__ base_next, uint32_t *__restrict__ desc,
               uint32_t *__restrict__ ticket, uint32_t ntiles, int in_aligned)
{
    using Binner = RadixBinner;
    using Smem = Scatter2Smem<THREADS, KPT, LOAD, RadixBinner::Arg>;
    constexpr int TILE = Smem::TILE;
    constexpr int NWARPS = Smem::NWARPS;
    static_assert(THREADS >= RADIX, "need one thread per bin");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem &s = *reinterpret_cast<Smem *>(smem_raw);

    if (ticket[SKIP_OFFSET]) {  // constant digit (k_scan_hist): the pass is a copy
        copy_keys(in, out, n);
        return;
    }
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const Binner binner(binner_arg);
    const uint32_t pad_bin = binner.pad_bin();
    unsigned char *const my_hist = reinterpret_cast<unsigned char *>(s.warp_hist[warp]);
    unsigned char *const tile_bytes = reinterpret_cast<unsigned char *>(s.keys);
    const bool use_tma = LOAD != LOAD_LDG && in_aligned;
    uint32_t *const land = LOAD == LOAD_TMA_SEP ? s.landing : s.keys;

    auto issue_load = [&](uint32_t tile) {
        // elected thread: arm the barrier and start the bulk copy of the 16 B-multiple part of
        // the tile; a ragged tail (<= 3 keys, last tile only) is read from global by its consumer
        const uint32_t valid = min((uint32_t)TILE, n - tile * (uint32_t)TILE);
        const uint32_t bytes = (valid * 4u) & ~15u;
        if (bytes) {
            mbar_arrive_expect_tx(&s.mbar, bytes);
            tma_load_1d(land, in + (size_t)tile * TILE, bytes, &s.mbar);
        } else {
            mbar_arrive_expect_tx(&s.mbar, 0);
        }
    };
    auto prefetch_ahead = [&](uint32_t tile) {
        const uint32_t ahead = tile + gridDim.x;
        if (PREFETCH && in_aligned && ahead < ntiles) {
            const uint32_t bytes = (min((uint32_t)TILE, n - ahead * (uint32_t)TILE) * 4u) & ~15u;
            if (bytes) prefetch_l2_bulk(in + (size_t)ahead * TILE, bytes);
        }
    };

    if (tid == 0) {
        mbar_init(&s.mbar, 1);
        fence_mbar_init();
        const uint32_t t = atomicAdd(ticket, 1u);
        s.next_tile = t;
        s.binner_arg = binner_arg;
        if (t < ntiles) {
            if (use_tma) issue_load(t);
            prefetch_ahead(t);
        }
    }
#pragma unroll
    for (int j = lane; j < RADIX; j += 32) s.warp_hist[warp][j] = 0;
    __syncthreads();
    uint32_t cur = s.next_tile;

    for (uint32_t it = 0; cur < ntiles; ++it) {
        const uint32_t tile_base = cur * (uint32_t)TILE;
        const uint32_t valid = min((uint32_t)TILE, n - tile_base);

        // (1) keys -> registers, warp-striped so that tile order == rank order
        uint32_t key[KPT];
        const uint32_t wbase = warp * (KPT * 32) + lane;
        if (use_tma) {
            mbar_wait(&s.mbar, it & 1);
            if (valid == TILE) {
#pragma unroll
                for (int i = 0; i < KPT; i++) key[i] = land[wbase + i * 32];
            } else {
                const uint32_t staged = ((valid * 4u) & ~15u) >> 2;
#pragma unroll
                for (int i = 0; i < KPT; i++) {
                    const uint32_t idx = wbase + i * 32;
                    key[i] = idx < staged ? land[idx] : (idx < valid ? in[(size_t)tile_base + idx] : 0xffffffffu);
                }
            }
        } else {
            const uint32_t *src = in + (size_t)tile_base + wbase;
            if (valid == TILE) {
#pragma unroll
                for (int i = 0; i < KPT; i++) key[i] = ld_stream_u32(src + i * 32);
            } else {
#pragma unroll
                for (int i = 0; i < KPT; i++) key[i] = wbase + i * 32 < valid ? ld_stream_u32(src + i * 32) : 0xffffffffu;
            }
        }
        __syncwarp();  // this warp's counters were zeroed by other lanes

        // (2) early counts: per-warp digit histogram
#pragma unroll
        for (int i = 0; i < KPT; i++) atomicAdd(&s.warp_hist[warp][binner.bin(key[i])], 1u);
        __syncthreads();

        uint32_t t_next = 0;
        if (LOAD == LOAD_TMA_SEP && EARLY && tid == 0) {
            // every warp holds its keys in registers: the landing buffer is free for the next tile
            t_next = atomicAdd(ticket, 1u);
            s.next_tile = t_next;
            if (t_next < ntiles) {
                if (use_tma) {
                    fence_proxy_async();
                    issue_load(t_next);
                }
                prefetch_ahead(t_next);
            }
        }

        // (3) bin threads: reduce over warps, publish the tile aggregate, finish the look-back,
        //     then lay the buckets out with the 16-byte phase of their destinations
        uint32_t my_count = 0, padded = 0, phase = 0, span4 = 0;
        unsigned long long dst_addr = 0;  // byte address of the first destination of this tile's run
        if (tid < RADIX) {
#pragma unroll
            for (int w = 0; w < NWARPS; w++) padded += s.warp_hist[w][tid];
            my_count = padded;
            if ((uint32_t)tid == pad_bin) my_count -= (uint32_t)TILE - valid;  // padding keys carry the top digit
            st_relaxed(&desc[(size_t)cur * RADIX + tid], (cur == 0 ? DESC_INCL : DESC_AGG) | my_count);

            uint32_t excl = 0;
            if (cur != 0) {
                // windowed decoupled look-back: LB descriptors per round trip
                uint32_t j = cur - 1;
                bool done = false;
                while (true) {
                    uint32_t lbv[LB];
#pragma unroll
                    for (int k = 0; k < LB; k++)
                        lbv[k] = (uint32_t)k <= j ? ld_relaxed(&desc[(size_t)(j - k) * RADIX + tid]) : DESC_INCL;
#pragma unroll
                    for (int k = 0; k < LB; k++) {
                        if (!done) {
                            const uint32_t v = lbv[k];
                            if ((v >> 30) == 0) break;
                            excl += v & DESC_VAL;
                            if (v & DESC_INCL) done = true;
                            else --j;
                        }
                    }
                    if (done) break;
                }
                st_relaxed(&desc[(size_t)cur * RADIX + tid], DESC_INCL | (excl + my_count));
            }
            const uint32_t gbase = base[tid];
            dst_addr = reinterpret_cast<unsigned long long>(out) + 4ull * ((unsigned long long)gbase + excl);
            if (base_next && cur == ntiles - 1) base_next[tid] = gbase + excl + my_count;
            phase = (uint32_t)(dst_addr >> 2) & 3u;                      // keys before it in its 16-byte chunk
            span4 = padded ? ((phase + padded + 3u) & ~3u) * 4u : 0u;    // bytes taken in the padded layout
            // the sorted buffer is about to be rewritten: this thread's bulk copies of the previous
            // tile must have finished reading it (every bin thread waits before the barriers below)
            if (LOAD != LOAD_TMA_INPLACE) bulk_wait_read();
        }
        const uint32_t slot4 = block_excl_scan_256(span4, s.s_warp, nullptr);
        const uint32_t start4 = slot4 + phase * 4u;  // byte offset of the run in keys[]
        const bool is_heavy = padded * 4u > (uint32_t)TILE;  // one digit holds > 1/4 of the tile
        if (tid < RADIX) {
            uint32_t run = start4;
#pragma unroll
            for (int w = 0; w < NWARPS; w++) {
                const uint32_t c = s.warp_hist[w][tid];
                s.warp_hist[w][tid] = run;
                run += c << 2;
            }
            if (is_heavy) s.heavy_d4 = (uint32_t)tid * 4u;
        }
        const int heavy = __syncthreads_or(is_heavy);

        // (4) rank every key and move it to its place in the padded digit-sorted tile
        if (!heavy) {
#pragma unroll
            for (int i = 0; i < KPT; i++) {
                const uint32_t p4 = atomicAdd(reinterpret_cast<uint32_t *>(my_hist + binner.bin4(key[i])), 4u);
                *reinterpret_cast<uint32_t *>(tile_bytes + p4) = key[i];
            }
        } else {
            uint32_t spilled[KPT];
#pragma unroll
            for (int i = 0; i < KPT; i++) spilled[i] = key[i];
            rank_tile_heavy<Binner, Smem>(spilled, &s);
        }
        __syncwarp();
#pragma unroll
        for (int j = lane; j < RADIX; j += 32) s.warp_hist[warp][j] = 0;
        if (!(LOAD == LOAD_TMA_SEP && EARLY) && tid == 0) {
            t_next = atomicAdd(ticket, 1u);  // in flight during the barrier
            if (LOAD != LOAD_TMA_INPLACE) s.next_tile = t_next;
        }
        __syncthreads();

        if (LOAD == LOAD_TMA_SEP && !EARLY && tid == 0 && t_next < ntiles) {
            if (use_tma) {
                fence_proxy_async();
                issue_load(t_next);
            }
            prefetch_ahead(t_next);
        }
        if (LOAD == LOAD_LDG && tid == 0 && t_next < ntiles) prefetch_ahead(t_next);

        // (5) write-out: thread d owns bucket d.  The whole 16-byte chunks of the destination go
        //     to the bulk-copy engine, the keys outside them (head and tail, <= 3 each) either
        //     through the LSU or as byte-masked 16-byte bulk copies.
        if (tid < RADIX && my_count) {
            const uint32_t *src = reinterpret_cast<const uint32_t *>(tile_bytes + start4);
            uint32_t *dst = reinterpret_cast<uint32_t *>(dst_addr);
            const uint32_t head = min(my_count, (4u - phase) & 3u);
            const uint32_t body = (my_count - head) & ~3u;
            const uint32_t tail = my_count - head - body;
            if (RAG == RAG_MASK) {
                fence_proxy_async();  // the generic-proxy stores of (4) come before the engine's reads
                if (head) {
                    uint32_t m = (0xffffu << (4u * phase)) & 0xffffu;
                    if (head != ((4u - phase) & 3u)) m &= (1u << (4u * (phase + head))) - 1u;  // run ends inside the chunk
                    bulk_store_masked16(dst - phase, src - phase, m);
                }
                if (body) bulk_store(dst + head, src + head, body * 4u);
                if (tail) bulk_store_masked16(dst + head + body, src + head + body, (1u << (4u * tail)) - 1u);
                bulk_commit();
            } else {
                if (RAG == RAG_PAIR && head == ((4u - phase) & 3u)) {
                    // head ends on a 16 B boundary, tail starts on one: 8-byte stores where a pair is whole
                    if (head & 1u) dst[0] = src[0];
                    if (head & 2u) *reinterpret_cast<uint2 *>(dst + (head & 1u)) = *reinterpret_cast<const uint2 *>(src + (head & 1u));
                    const uint32_t t0 = head + body;
                    if (tail & 2u) *reinterpret_cast<uint2 *>(dst + t0) = *reinterpret_cast<const uint2 *>(src + t0);
                    if (tail & 1u) dst[t0 + (tail & 2u)] = src[t0 + (tail & 2u)];
                } else {
                    for (uint32_t j = 0; j < head; j++) dst[j] = src[j];
                    for (uint32_t j = head + body; j < my_count; j++) dst[j] = src[j];
                }
                if (body) {
                    fence_proxy_async();
                    bulk_store(dst + head, src + head, body * 4u);
                    bulk_commit();
                }
            }
            if (LOAD == LOAD_TMA_INPLACE) bulk_wait_read();
        }
        if (LOAD == LOAD_TMA_INPLACE) {
            if (tid == 0) s.next_tile = t_next;
            __syncthreads();  // the tile buffer is free again
            cur = s.next_tile;
            if (tid == 0 && cur < ntiles) {
                if (use_tma) {
                    fence_proxy_async();
                    issue_load(cur);
                }
                prefetch_ahead(cur);
            }
        } else {
            cur = s.next_tile;
        }
    }
    // bulk stores still in flight belong to this thread's bulk async-group: drain them
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace b200sort
