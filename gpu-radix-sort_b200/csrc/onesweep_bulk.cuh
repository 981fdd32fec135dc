// onesweep_bulk.cuh -- the scatter pass with a bulk-copy write-out.
//
// Same contract and the same phases (1)-(5) as k_onesweep<RANK_ATOMS, RadixBinner>
// in kernels.cuh, but the digit-sorted tile is laid out in shared memory so that
// every bucket's run has the same 16-byte phase as its destination in global
// memory.  One thread per bucket then hands the 16 B-aligned body of its run to
// the bulk-copy engine (cp.async.bulk.global.shared::cta, SASS UBLKCP.G.S) and
// stores the <= 3 keys of the ragged head and tail itself.  The write-out costs
// no LDS/STG wavefronts for the body, which matters because the pass is bound by
// the L1/shared-memory pipe (profiles/r01b_k_onesweep_lsu_wavefronts.txt), not
// by HBM.  The price: the look-back must finish before the tile is re-ordered,
// because the layout depends on the global offsets modulo 4 keys.
#pragma once
#include "kernels.cuh"

namespace b200sort {

__device__ __forceinline__ void bulk_store(void *dst_gmem, const void *src_smem, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// wait until the bulk copies of this thread have finished READING shared memory
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

template <int THREADS, int KPT, class BinnerArg>
struct BulkSmem {
    static constexpr int KEYS_PER_THREAD = KPT;
    static constexpr int TILE = THREADS * KPT;
    static constexpr int NWARPS = THREADS / 32;
    static constexpr int PAD = RADIX * 6;      // <= 3 keys of head and of tail padding per bucket
    alignas(128) uint32_t keys[TILE + PAD];    // TMA landing buffer [0, TILE), then the padded digit-sorted tile
    uint32_t warp_hist[NWARPS][RADIX];         // per-warp digit counters -> running BYTE offsets into keys[]
    uint32_t s_warp[8];
    alignas(8) uint64_t mbar;
    uint32_t next_tile;
    BinnerArg binner_arg;                      // copy for the out-of-line heavy-tile path
    uint32_t heavy_d4;
};

// RANK / Binner / PTR_OUT as in k_onesweep: with PTR_OUT `base_v` holds one 64-bit
// destination address per bin (local or NVLink-mapped peer memory) -- the splitter
// scatter of the distributed sort, where a tile has at most 16 long runs and the
// bulk-copy engine, not the SM's store path, rides out the NVLink latency.
// UNORDERED (PTR_OUT only): the tiles of this launch may land in any order inside a destination's
// range -- the distributed sort sorts what arrived anyway -- so a tile reserves its place with ONE
// atomicAdd per destination on a cursor array (`desc`, 16 zeroed words) instead of waiting, fully
// exposed here, for the look-back over all earlier tiles.
// UNORDERED == EXCH_REMOTE_CURSOR: the cursors are not a local array but one word PER DESTINATION BUFFER,
// living on the destination GPU (base_v = CursorPlan): every source reserves room for a tile's run
// with one system-scope atomicAdd over NVLink.  Nobody needs to know in advance how many keys go
// where -- no counting pass, no count exchange -- and when all sources are done the word holds the
// number of keys received.  A reservation that would pass the buffer's capacity is dropped and
// raises *status (the word still counts it, so the receiver sees cursor > capacity).
// (Measured and dropped: reserving whole 16-byte chunks, runs padded with 0xFFFFFFFF, so that the
// layout does not wait for the atomic -- the exchange was no faster, 0.948 vs 0.956 ms at 2 GPUs, and
// the padding keys spoil the constant-digit and heavy-tile shortcuts of the local sort on skewed
// inputs: all-equal keys 5.07 ms instead of 3.55 ms per sort.)
constexpr int EXCH_ORDERED = 0, EXCH_LOCAL_CURSOR = 1, EXCH_REMOTE_CURSOR = 2;
struct CursorPlan {
    unsigned long long dest[MAX_SPLITTERS + 1];    // receive buffer of every rank as mapped here
    unsigned long long cursor[MAX_SPLITTERS + 1];  // its cursor word
    uint32_t cap;                                  // keys a receive buffer holds
    uint32_t unused;
    uint32_t *status;                              // local overflow flag
};
__global__ void k_store_cursor_plan(const unsigned long long *__restrict__ peer_base, int nranks, unsigned long long cap,
                                    uint32_t *status, CursorPlan *__restrict__ plan)
{
    const int d = threadIdx.x;
    if (d <= MAX_SPLITTERS) {
        plan->dest[d] = d < nranks ? peer_base[d] : 0ull;
        plan->cursor[d] = d < nranks ? peer_base[d] + 4ull * cap : 0ull;
    }
    if (d == 0) {
        plan->cap = (uint32_t)cap;
        plan->unused = 0;
        plan->status = status;
    }
}

// SLICED (step-wise distributed sort): base_v is a SliceTable -- bucket bases are positions in ONE array cut
// into slices that live on different GPUs (k_onesweep<OUT_SLICED> is the per-key-store version).  A
// bucket's run in a tile starts at a global position inside slice j; the part that fits goes to slice
// j's owner through the bulk-copy engine like any run, the rare rest (a run crosses at most one slice
// boundary, and only the one tile whose run covers the boundary does) is stored key by key at the start
// of slice j + 1.
template <int THREADS, int KPT, int MINB, int LB, int RANK, class Binner, bool PTR_OUT, int UNORDERED = EXCH_ORDERED,
          bool SLICED = false>
__global__ void __launch_bounds__(THREADS, MINB)
    k_onesweep_bulk(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, uint32_t n, const typename Binner::Arg binner_arg,
                    const void *__restrict__ base_v, uint32_t *__restrict__ base_next, uint32_t *__restrict__ desc,
                    uint32_t *__restrict__ ticket, uint32_t ntiles, int use_tma)
{
    using Smem = BulkSmem<THREADS, KPT, typename Binner::Arg>;
    static_assert(!(RANK == RANK_ATOMS && Binner::kByPosition), "the heavy-tile path bins by key alone");
    constexpr int TILE = Smem::TILE;
    constexpr int NWARPS = Smem::NWARPS;
    static_assert(THREADS >= RADIX, "need one thread per bin");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem &s = *reinterpret_cast<Smem *>(smem_raw);

    static_assert(!SLICED || (!PTR_OUT && UNORDERED == EXCH_ORDERED), "sliced output: ordered radix pass");
    if (!PTR_OUT && !SLICED && ticket[SKIP_OFFSET]) {  // constant digit (k_scan_hist): the pass is a copy
        copy_keys(in, out, n);
        return;
    }
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const Binner binner(binner_arg);
    const uint32_t pad_bin = binner.pad_bin();
    unsigned char *const my_hist = reinterpret_cast<unsigned char *>(s.warp_hist[warp]);
    unsigned char *const tile_bytes = reinterpret_cast<unsigned char *>(s.keys);

    auto issue_load = [&](uint32_t tile) {
        const uint32_t valid = min((uint32_t)TILE, n - tile * (uint32_t)TILE);
        const uint32_t bytes = (valid * 4u) & ~15u;
        if (bytes) {
            mbar_arrive_expect_tx(&s.mbar, bytes);
            tma_load_1d(s.keys, in + (size_t)tile * TILE, bytes, &s.mbar);
        } else {
            mbar_arrive_expect_tx(&s.mbar, 0);
        }
    };

    if (tid == 0) {
        mbar_init(&s.mbar, 1);
        fence_mbar_init();
        const uint32_t t = atomicAdd(ticket, 1u);
        s.next_tile = t;
        s.binner_arg = binner_arg;
        if (use_tma && t < ntiles) issue_load(t);
    }
#pragma unroll
    for (int j = lane; j < RADIX; j += 32) s.warp_hist[warp][j] = 0;
    __syncthreads();
    uint32_t cur = s.next_tile;

    for (uint32_t it = 0; cur < ntiles; ++it) {
        const uint32_t tile_base = cur * (uint32_t)TILE;
        const uint32_t valid = min((uint32_t)TILE, n - tile_base);

        // (1) keys -> registers, (2) early per-warp digit counts (shared with k_onesweep, kernels.cuh)
        uint32_t key[KPT];
        const uint32_t wbase = warp * (KPT * 32) + lane;
        tile_keys_to_registers<KPT, TILE>(key, s, in, tile_base, valid, wbase, use_tma, it & 1);
        uint32_t dig[Binner::kByPosition ? (KPT + 7) / 8 : 1] = {};
        tile_count_digits(key, dig, binner, s.warp_hist[warp], tile_base + wbase);
        __syncthreads();

        // (3) bin threads: reduce over warps, publish the tile aggregate, finish the
        //     look-back, then lay the buckets out with the 16-byte phase of their destinations
        uint32_t my_count = 0, padded = 0, phase = 0, span4 = 0;
        bool drop = false;  // EXCH_REMOTE_CURSOR: the destination is full, this run is not stored
        uint32_t lo_len = 0;               // SLICED: keys of this run that belong to the slice it starts in
        unsigned long long hi_addr = 0;    // SLICED: first key of the next slice (where the rest goes)
        unsigned long long dst_addr = 0;  // byte address of the first destination of this tile's run
        if (tid < RADIX) {
#pragma unroll
            for (int w = 0; w < NWARPS; w++) padded += s.warp_hist[w][tid];
            my_count = padded;
            if ((uint32_t)tid == pad_bin) my_count -= (uint32_t)TILE - valid;  // padding keys carry the top digit
            uint32_t excl = 0;
            if (UNORDERED == EXCH_REMOTE_CURSOR) {
                const CursorPlan *cp = static_cast<const CursorPlan *>(base_v);
                if (my_count) {
                    excl = atomicAdd_system(reinterpret_cast<uint32_t *>(cp->cursor[tid]), my_count);
                    if (excl + my_count > cp->cap || excl + my_count < excl) {
                        drop = true;
                        atomicOr(cp->status, 1u);
                    }
                }
            } else if (UNORDERED == EXCH_LOCAL_CURSOR) {
                if (my_count) excl = atomicAdd(&desc[tid], my_count);
            } else
                st_relaxed(&desc[(size_t)cur * RADIX + tid], (cur == 0 ? DESC_INCL : DESC_AGG) | my_count);
            if (UNORDERED == EXCH_ORDERED && cur != 0) {
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
            if (PTR_OUT && UNORDERED == EXCH_REMOTE_CURSOR) {
                dst_addr = my_count ? static_cast<const CursorPlan *>(base_v)->dest[tid] + 4ull * excl : 0ull;
            } else if (SLICED) {
                const SliceTable *tab = static_cast<const SliceTable *>(base_v);
                const uint32_t g = tab->gbase[tid] + excl, ns = tab->nslice;
                uint32_t j = 0;
                for (uint32_t k = 1; k < ns; k++) j += g >= tab->slice_start[k] ? 1u : 0u;
                dst_addr = tab->peer[j] + 4ull * (unsigned long long)(g - tab->slice_start[j]);
                lo_len = min(my_count, tab->slice_start[j + 1] - g);
                hi_addr = j + 1 < ns ? tab->peer[j + 1] : 0ull;
            } else if (PTR_OUT) {
                dst_addr = static_cast<const unsigned long long *>(base_v)[tid] + 4ull * excl;
            } else {
                const uint32_t gbase = static_cast<const uint32_t *>(base_v)[tid];
                dst_addr = reinterpret_cast<unsigned long long>(out) + 4ull * ((unsigned long long)gbase + excl);
                if (base_next && cur == ntiles - 1) base_next[tid] = gbase + excl + my_count;
            }
            phase = (uint32_t)(dst_addr >> 2) & 3u;       // keys before it in its 16-byte chunk
            span4 = padded ? ((phase + padded + 3u) & ~3u) * 4u : 0u;  // bytes taken in the padded layout
        }
        const uint32_t slot4 = block_excl_scan_256(span4, s.s_warp, nullptr);
        const uint32_t start4 = slot4 + phase * 4u;       // byte offset of the run in keys[]
        const bool is_heavy = RANK == RANK_ATOMS && padded * 4u > (uint32_t)TILE;  // one digit holds > 1/4 of the tile
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

        // (4) rank every key and move it to its place in the padded digit-sorted tile (shared with k_onesweep)
        tile_rank_and_place<RANK, 0, Binner, KPT, Smem>(key, dig, binner, s, my_hist, tile_bytes, heavy, tile_base + wbase);
        __syncwarp();
#pragma unroll
        for (int j = lane; j < RADIX; j += 32) s.warp_hist[warp][j] = 0;
        uint32_t t_next = 0;
        if (tid == 0) t_next = atomicAdd(ticket, 1u);  // in flight during the write-out
        __syncthreads();

        // (5) write-out: thread d owns bucket d.  Head and tail (the keys outside whole
        //     16-byte chunks of the destination) are stored one by one, the body goes to
        //     the bulk-copy engine.
        if (tid < RADIX && my_count && !drop) {
            const uint32_t *src = reinterpret_cast<const uint32_t *>(tile_bytes + start4);
            uint32_t *dst = reinterpret_cast<uint32_t *>(dst_addr);
            const uint32_t here = SLICED ? lo_len : my_count;  // keys for the destination the run starts in
            const uint32_t head = min(here, (4u - phase) & 3u);
            const uint32_t body = (here - head) & ~3u;
            for (uint32_t j = 0; j < head; j++) dst[j] = src[j];
            for (uint32_t j = head + body; j < here; j++) dst[j] = src[j];
            if (SLICED)  // the run crosses into the next slice (one tile per slice boundary and bucket)
                for (uint32_t j = here; j < my_count; j++) reinterpret_cast<uint32_t *>(hi_addr)[j - here] = src[j];
            if (body) {
                fence_proxy_async();  // the generic-proxy stores of (4) come before the engine's reads
                bulk_store(dst + head, src + head, body * 4u);
                bulk_commit();
                bulk_wait_read();
            }
        }
        if (tid == 0) s.next_tile = t_next;
        __syncthreads();  // the tile buffer is free again
        cur = s.next_tile;
        if (tid == 0 && use_tma && cur < ntiles) {
            fence_proxy_async();
            issue_load(cur);
        }
    }
    // bulk stores still in flight belong to this thread's bulk async-group: drain them
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace b200sort
