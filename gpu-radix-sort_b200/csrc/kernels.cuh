// kernels.cuh -- the sm_100a kernels of the sort hot path.
//
//   K1  k_histogram      one upsweep over the keys, 128-bit streaming loads,
//                        digit histograms for ALL passes at once (4 B/key).
//       k_scan_hist      256-bin exclusive scans -> global bucket bases
//                        (and, for width <= 8, the exact boundaries).
//   K3  k_onesweep       one stable digit-scatter pass (8 B/key): 1-D bulk TMA
//                        stages each key tile in shared memory (ONE tile buffer,
//                        persistent CTAs, late dynamic tile tickets; the next
//                        tile's load runs under this tile's stores).  Keys are
//                        ranked with returning shared-memory atomics (default,
//                        self-tested and audited per tile) or with one ballot per
//                        digit bit (RANK_BALLOT) -- MATCH.ANY measured 62 cycles
//                        per warp-op on B200 (profiles/r01_microbench_primitives.log)
//                        and is only used for <= 2-bit digits -- re-ordered in
//                        shared memory and written coalesced per bucket.
//   K2  (inside K3)      chained scan with decoupled look-back over per-tile
//                        bucket counts; replaces the reference's multi-kernel
//                        Blelloch scan (libsort/scan.cu:16-250).
//       k_bounds_*       boundaries for wide groups / reference quirk mode.
//
// Together K1+K3 replace gpu_radix_sort_local, sum_scan_blelloch and
// gpu_glbl_shuffle (reference libsort/sort.cu:29-213, scan.cu) and k_bounds_*
// replace gpu_groups + the host back-fill (sort.cu:14-27, 367-394).
#pragma once
#include <type_traits>
#include "ptx.cuh"

namespace b200sort {

constexpr int RADIX = 256;       // bins per pass (8-bit digits)
constexpr int MAX_PASSES = 4;    // 32 bits / 8
constexpr uint32_t DESC_AGG = 1u << 30;   // tile aggregate published
constexpr uint32_t DESC_INCL = 2u << 30;  // inclusive prefix published
constexpr uint32_t DESC_VAL = (1u << 30) - 1;

struct PassPlan {
    int npass;
    uint32_t shift[MAX_PASSES];
    uint32_t mask[MAX_PASSES];
};

__device__ __forceinline__ uint32_t digit_of(uint32_t key, uint32_t shift, uint32_t mask)
{
    return (key >> shift) & mask;
}

// ---------------------------------------------------------------- K1 ------
// Grid: persistent, a multiple of the SM count.  Each CTA keeps NPASS x 256
// counters per replica in shared memory; warps are spread over REPL replicas
// to thin out same-address atomic collisions.  Flushed once per CTA.
template <int NPASS, int THREADS, int REPL>
__global__ void __launch_bounds__(THREADS) k_histogram(const uint32_t *__restrict__ in, size_t n, PassPlan plan,
                                                      uint32_t *__restrict__ ghist, const uint32_t *__restrict__ dyn_n)
{
    __shared__ uint32_t s_hist[REPL][NPASS][RADIX];
    if (dyn_n) n = min((size_t)*dyn_n, n);  // key count decided on the device (the launch was sized for the capacity = n)
    for (int i = threadIdx.x; i < REPL * NPASS * RADIX; i += THREADS) (&s_hist[0][0][0])[i] = 0;
    __syncthreads();

    uint32_t(*my)[RADIX] = s_hist[(threadIdx.x >> 5) % REPL];
    uint32_t shift[NPASS], mask[NPASS];
#pragma unroll
    for (int p = 0; p < NPASS; p++) { shift[p] = plan.shift[p]; mask[p] = plan.mask[p]; }

    auto count = [&](uint32_t k) {
#pragma unroll
        for (int p = 0; p < NPASS; p++) atomicAdd(&my[p][digit_of(k, shift[p], mask[p])], 1u);
    };

    // scalar head up to the first 16 B boundary, vector body, scalar tail
    size_t head = ((16u - (reinterpret_cast<uintptr_t>(in) & 15u)) & 15u) >> 2;
    if (head > n) head = n;
    const size_t nvec = (n - head) >> 2;
    const size_t tail0 = head + (nvec << 2);
    const size_t gtid = (size_t)blockIdx.x * THREADS + threadIdx.x;
    const size_t gstride = (size_t)gridDim.x * THREADS;
    if (gtid < head) count(in[gtid]);
    if (gtid < n - tail0) count(in[tail0 + gtid]);

    const uint4 *vin = reinterpret_cast<const uint4 *>(in + head);
    constexpr int U = 4;
    size_t i = gtid;
    for (; i + (U - 1) * gstride < nvec; i += U * gstride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = ld_stream_v4(vin + i + u * gstride);
#pragma unroll
        for (int u = 0; u < U; u++) { count(v[u].x); count(v[u].y); count(v[u].z); count(v[u].w); }
    }
    for (; i < nvec; i += gstride) {
        uint4 v = ld_stream_v4(vin + i);
        count(v.x); count(v.y); count(v.z); count(v.w);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < NPASS * RADIX; j += THREADS) {
        uint32_t c = 0;
#pragma unroll
        for (int r = 0; r < REPL; r++) c += (&s_hist[r][0][0])[j];
        if (c) atomicAdd(&ghist[j], c);
    }
}

// K1, bank-private layout: counter (pass p, digit d) has 32 copies, copy l in
// bank l, and lane l only ever touches its own bank, so a warp's 32 shared
// atomics never collide (random 8-bit digits otherwise cost ~2.35 cycles per
// warp-op in bank conflicts, which caps a 4-digit histogram below HBM speed).
// NPASS * 32 KB of dynamic shared memory, one CTA per SM.
template <int NPASS, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_histogram_lanes(const uint32_t *__restrict__ in, size_t n, PassPlan plan,
                                                               uint32_t *__restrict__ ghist, const uint32_t *__restrict__ dyn_n)
{
    extern __shared__ __align__(16) uint32_t s_lane_hist[];  // [NPASS][RADIX][32]
    if (dyn_n) n = min((size_t)*dyn_n, n);
    for (int i = threadIdx.x; i < NPASS * RADIX * 32; i += THREADS) s_lane_hist[i] = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    uint32_t shift[NPASS], mask[NPASS];
#pragma unroll
    for (int p = 0; p < NPASS; p++) { shift[p] = plan.shift[p]; mask[p] = plan.mask[p]; }
    auto count = [&](uint32_t k) {
#pragma unroll
        for (int p = 0; p < NPASS; p++)
            atomicAdd(&s_lane_hist[(p * RADIX + digit_of(k, shift[p], mask[p])) * 32 + lane], 1u);
    };
    size_t head = ((16u - (reinterpret_cast<uintptr_t>(in) & 15u)) & 15u) >> 2;
    if (head > n) head = n;
    const size_t nvec = (n - head) >> 2;
    const size_t tail0 = head + (nvec << 2);
    const size_t gtid = (size_t)blockIdx.x * THREADS + threadIdx.x;
    const size_t gstride = (size_t)gridDim.x * THREADS;
    if (gtid < head) count(in[gtid]);
    if (gtid < n - tail0) count(in[tail0 + gtid]);
    const uint4 *vin = reinterpret_cast<const uint4 *>(in + head);
    constexpr int U = 4;
    size_t i = gtid;
    for (; i + (U - 1) * gstride < nvec; i += U * gstride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = ld_stream_v4(vin + i + u * gstride);
#pragma unroll
        for (int u = 0; u < U; u++) { count(v[u].x); count(v[u].y); count(v[u].z); count(v[u].w); }
    }
    for (; i < nvec; i += gstride) {
        uint4 v = ld_stream_v4(vin + i);
        count(v.x); count(v.y); count(v.z); count(v.w);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < NPASS * RADIX; j += THREADS) {
        uint32_t c = 0;
#pragma unroll
        for (int l = 0; l < 32; l++) c += s_lane_hist[j * 32 + ((l + lane) & 31)];
        if (c) atomicAdd(&ghist[j], c);
    }
}

// 256-thread block-wide exclusive scan helper (threads >= 256 pass 0 / ignore).
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t *s_warp /*>=8*/, uint32_t *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31 && warp < 8) s_warp[warp] = inc;
    __syncthreads();
    uint32_t wbase = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        uint32_t c = s_warp[w];
        if (w < warp) wbase += c;
        tot += c;
    }
    if (total) *total = tot;
    return wbase + inc - v;
}

// One 256-thread CTA per pass: base[p][b] = #keys whose digit p is < b.
// When d_exact != nullptr (single-pass partial sort, width <= 8) the scan of
// pass 0 IS the exact boundary array (localTest/tests.cpp:44-54).
// skip (nullable): skip[p * SKIP_STRIDE] = 1 when every key has the same digit in pass p
// (one bin holds all n keys) -- that pass is the identity permutation and the scatter
// kernel just copies.
constexpr int SKIP_STRIDE = 8;   // one flag per (pass, portion) slot, beside the tile tickets
constexpr int SKIP_OFFSET = 32;  // flag of a pass = its ticket word + SKIP_OFFSET
__global__ void __launch_bounds__(256) k_scan_hist(const uint32_t *__restrict__ ghist, uint32_t *__restrict__ base,
                                                  uint32_t *__restrict__ d_exact, uint32_t ngroups, uint32_t n,
                                                  uint32_t *__restrict__ skip, const uint32_t *__restrict__ dyn_n)
{
    __shared__ uint32_t s_warp[8];
    pdl_wait();  // K1's counts (programmatic dependent launch, engine.cu launch_after)
    pdl_trigger();
    if (dyn_n) n = min(*dyn_n, n);  // never past the capacity the launch was sized for
    const int p = blockIdx.x;
    uint32_t c = ghist[p * RADIX + threadIdx.x];
    uint32_t ex = block_excl_scan_256(c, s_warp, nullptr);
    base[p * RADIX + threadIdx.x] = ex;
    if (d_exact && p == 0 && threadIdx.x < ngroups) d_exact[threadIdx.x] = ex;
    if (skip && c == n) skip[p * SKIP_STRIDE] = 1;
}

// The copy a scatter pass degenerates to when its digit is constant over the whole input.
__device__ __forceinline__ void copy_keys(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, uint32_t n)
{
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n; i += 4 * stride) {
        const uint32_t a = in[i], b = in[i + stride], c = in[i + 2 * stride], d = in[i + 3 * stride];
        out[i] = a; out[i + stride] = b; out[i + 2 * stride] = c; out[i + 3 * stride] = d;
    }
    for (; i < n; i += stride) out[i] = in[i];
}

// ---------------------------------------------------------------- K3 ------
// Lanes of the warp whose digit equals this lane's.  MATCH.ANY costs about two
// cycles per DISTINCT value on B200 (measured: 62 cycles per warp-op for random
// 8-bit digits, SM-wide, not per scheduler), so the peer mask is built from one
// ballot per digit bit instead (VOTE runs at two warp-ops per cycle per SM).
// MATCH.ANY is kept for the narrow (<= 2-bit) passes where it is cheaper.
// One digit bit of the ballot match, in PTX so that it stays at four SASS
// instructions (LOP3 -> predicate, VOTE, SEL, LOP3): peers &= ballot(bit) ^ (bit ? 0 : ~0).
__device__ __forceinline__ uint32_t match_bit(uint32_t peers, uint32_t d, uint32_t bitmask)
{
    asm("{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b32 t, m, x;\n\t"
        "and.b32 t, %1, %2;\n\t"
        "setp.ne.u32 p, t, 0;\n\t"
        "vote.sync.ballot.b32 m, p, 0xffffffff;\n\t"
        "selp.b32 x, 0, -1, p;\n\t"
        "lop3.b32 %0, %0, m, x, 0x60;\n\t"
        "}"
        : "+r"(peers)
        : "r"(d), "r"(bitmask));
    return peers;
}

template <int NBITS>
__device__ __forceinline__ uint32_t match_digit(uint32_t d, int nbits_rt)
{
    if (NBITS == 0 && nbits_rt <= 2) return __match_any_sync(0xffffffffu, d);
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; b++)
        if (NBITS ? (b < NBITS) : (b < nbits_rt)) peers = match_bit(peers, d, 1u << b);
    return peers;
}

// Leader-only shared atomic add without a branch: the highest peer lane adds
// the group's size to the warp's running digit counter and gets the old value.
__device__ __forceinline__ uint32_t leader_atom_add(uint32_t smem_addr, uint32_t val, bool leader)
{
    uint32_t old = 0;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.u32 p, %3, 0;\n\t"
        "@p atom.shared.add.u32 %0, [%1], %2;\n\t"
        "}"
        : "+r"(old)
        : "r"(smem_addr), "r"(val), "r"((uint32_t)leader)
        : "memory");
    return old;
}

// Which bin a key falls into.  RadixBinner: an (up to) 8-bit digit, the LSD/MSD
// pass.  SplitBinner: destination rank of the distributed sort = number of
// splitters <= key (splitters ascending, at most 15 -> up to 16 ranks).
struct RadixBinner {
    struct Arg { uint32_t shift, mask; };
    uint32_t shift, mask;
    __device__ __forceinline__ explicit RadixBinner(const Arg &a) : shift(a.shift), mask(a.mask) {}
    __device__ __forceinline__ uint32_t bin(uint32_t key) const { return (key >> shift) & mask; }
    // 4 * bin(key) in two ALU ops: rotate the digit to bits [2, 10) and mask
    __device__ __forceinline__ uint32_t bin4(uint32_t key) const
    {
        return __funnelshift_r(key, key, (shift + 30u) & 31u) & (mask << 2);
    }
    // position-aware forms (SplitBinner splits ties by position; a digit does not care)
    __device__ __forceinline__ uint32_t bin(uint32_t key, uint32_t) const { return bin(key); }
    __device__ __forceinline__ uint32_t bin4(uint32_t key, uint32_t) const { return bin4(key); }
    static constexpr bool kByPosition = false;  // bin is a function of the key alone
    static constexpr int kBits = 0;             // digit width known at run time only
    __device__ __forceinline__ uint32_t pad_bin() const { return mask; }  // bin of the 0xffffffff padding keys
    __device__ __forceinline__ int nbits() const { return __popc(mask); }
};
constexpr int MAX_SPLITTERS = 15;
constexpr int SPLIT_WORDS = 32;    // device splitter table: [0,16) values, [16,32) tie thresholds
constexpr int SPLIT_TIE_OFFSET = 16;
struct SplitArg { const uint32_t *d_splitters; int nsplit; };  // splitters live in device memory (no host round trip)
// Position hash used to split ties: consecutive positions spread evenly over [0, 2^32).
__device__ __forceinline__ uint32_t tie_hash(uint32_t idx) { return idx * 0x9E3779B1u; }
// NS = compiled-in capacity (1, 3, 7 or 15 splitters for up to 2, 4, 8, 16 ranks):
// only NS compares per key and NS registers.
//
// Ties: a key value that fills more than one rank's share appears as several equal
// splitters.  Splitter j then carries a threshold T_j (ascending within the run of equal
// splitters, 0 = "no tie splitting here"): a key EQUAL to splitter j counts as past it only
// if tie_hash(position) >= T_j.  Keys-only sort, so which copy goes where does not matter;
// the count kernel and the scatter kernel evaluate the same function of (key, position).
template <int NS>
struct SplitBinner {
    using Arg = SplitArg;
    uint32_t v[NS];
    int nsplit;
    const uint32_t *table;
    bool ties;
    __device__ __forceinline__ explicit SplitBinner(const Arg &a) : nsplit(a.nsplit), table(a.d_splitters)
    {
        uint32_t any = 0;
#pragma unroll
        for (int j = 0; j < NS; j++) {
            v[j] = j < a.nsplit ? a.d_splitters[j] : 0xffffffffu;
            any |= j < a.nsplit ? a.d_splitters[SPLIT_TIE_OFFSET + j] : 0u;
        }
        ties = any != 0;
    }
    // number of splitters <= key.  Unused slots hold 0xffffffff; the min() undoes what a key of
    // 0xffffffff would count against them.  NS >= 7: branch-free binary search (3 or 4 compares
    // and a few selects instead of NS compares).
    __device__ __forceinline__ uint32_t bin(uint32_t key) const
    {
        uint32_t d = 0;
        if (NS == 7) {
            const bool c2 = key >= v[3];
            const uint32_t m1 = c2 ? v[5] : v[1];
            const bool c1 = key >= m1;
            const uint32_t lo = c1 ? v[2] : v[0], hi = c1 ? v[6] : v[4];
            const bool c0 = key >= (c2 ? hi : lo);
            d = (c2 ? 4u : 0u) + (c1 ? 2u : 0u) + (c0 ? 1u : 0u);
        } else if (NS == 15) {
            const bool c3 = key >= v[7];
            const uint32_t m2 = c3 ? v[11] : v[3];
            const bool c2 = key >= m2;
            const uint32_t a = c2 ? v[5] : v[1], b = c2 ? v[13] : v[9];
            const bool c1 = key >= (c3 ? b : a);
            const uint32_t q0 = c1 ? v[2] : v[0], q1 = c1 ? v[6] : v[4], q2 = c1 ? v[10] : v[8], q3 = c1 ? v[14] : v[12];
            const uint32_t r0 = c2 ? q1 : q0, r1 = c2 ? q3 : q2;
            const bool c0 = key >= (c3 ? r1 : r0);
            d = (c3 ? 8u : 0u) + (c2 ? 4u : 0u) + (c1 ? 2u : 0u) + (c0 ? 1u : 0u);
        } else {
#pragma unroll
            for (int j = 0; j < NS; j++) d += key >= v[j] ? 1u : 0u;
        }
        return min(d, (uint32_t)nsplit);
    }
    __device__ __forceinline__ uint32_t bin4(uint32_t key) const { return bin(key) << 2; }
    // destination of the key at position idx of the shard.  The splitters equal to the key are
    // [lo, d) with lo = number of splitters < key; the key passes those whose threshold the
    // position's hash reaches (thresholds ascend inside a run, zeros first).  Uniform test: most
    // inputs have no ties; divergent branch: only keys that equal a splitter.  (A splitter of
    // 0xffffffff never has a threshold, so the padding keys stay in the last bin.)
    __device__ __forceinline__ uint32_t bin(uint32_t key, uint32_t idx) const
    {
        uint32_t d = bin(key);
        if (ties && d > 0 && __ldg(table + d - 1) == key) {  // one load tells whether the key equals a splitter at all
            const uint32_t lo = key ? bin(key - 1u) : 0u;
            {
                const uint32_t h = tie_hash(idx);
                uint32_t c = 0;
#pragma unroll
                for (int j = 0; j < NS; j++)  // independent, warp-uniform loads
                    c += ((uint32_t)j >= lo && (uint32_t)j < d && h >= __ldg(table + SPLIT_TIE_OFFSET + j)) ? 1u : 0u;
                d = lo + c;
            }
        }
        return d;
    }
    __device__ __forceinline__ uint32_t bin4(uint32_t key, uint32_t idx) const { return bin(key, idx) << 2; }
    static constexpr bool kByPosition = true;  // equal keys may sit in different bins
    static constexpr int kBits = NS == 1 ? 1 : NS == 3 ? 2 : NS == 7 ? 3 : 4;  // bits of a bin index, at compile time
    __device__ __forceinline__ uint32_t pad_bin() const { return (uint32_t)nsplit; }
    __device__ __forceinline__ int nbits() const { return 32 - __clz(nsplit); }  // bins 0..nsplit
};

// Out-of-line ranking for a tile dominated by one digit (cold path of the
// atomic-ranking kernel; keys come through local memory so that the hot path keeps
// its registers).  The hardware serialises the lanes of a row that add to the same
// counter, so the lanes holding the dominant digit hd4/4 are ranked with one ballot
// and a single atomic by their highest lane; all other lanes use their own atomic.
template <class Binner, class Smem>
__device__ __noinline__ void rank_tile_heavy(const uint32_t *keys, Smem *sp)
{
    Smem &s = *sp;
    const Binner binner(s.binner_arg);
    const uint32_t hd4 = s.heavy_d4;
    unsigned char *const my_hist = reinterpret_cast<unsigned char *>(s.warp_hist[threadIdx.x >> 5]);
    unsigned char *const tile_bytes = reinterpret_cast<unsigned char *>(s.keys);
    const uint32_t lt = lanemask_lt();
    const int lane = threadIdx.x & 31;
#pragma unroll 4
    for (int i = 0; i < Smem::KEYS_PER_THREAD; i++) {
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

// Where a scatter pass writes: OUT_INDEX -- bucket bases are indices into `out`; OUT_PTR -- one
// 64-bit destination address per bin (local or NVLink-mapped peer memory; the splitter exchange);
// OUT_SLICED -- bucket bases are positions in ONE array that is cut into contiguous slices living
// on different GPUs (the step-wise distributed LSD sort): a SliceTable maps a position to its owner.
constexpr int OUT_INDEX = 0, OUT_PTR = 1, OUT_SLICED = 2;
constexpr int MAX_SLICES = 16;
struct SliceTable {                        // device memory, one per rank
    unsigned long long peer[MAX_SLICES];   // where slice j lives (its first key), as seen from this GPU
    uint32_t slice_start[MAX_SLICES + 1];  // global position of slice j's first key; [nslice] = total keys
    uint32_t nslice;
    uint32_t pad_[2];
    uint32_t gbase[RADIX];                 // global position of this rank's first key of every digit (k_step_bases)
};

template <int THREADS, int KPT, int PTR_OUT, class BinnerArg>
struct OnesweepSmem {
    static constexpr int KEYS_PER_THREAD = KPT;
    static constexpr int TILE = THREADS * KPT;
    static constexpr int NWARPS = THREADS / 32;
    alignas(128) uint32_t keys[TILE];          // TMA landing buffer, then the digit-sorted tile
    uint32_t warp_hist[NWARPS][RADIX];         // per-warp digit counters -> running BYTE offsets into keys[]
    // global destination of bucket b minus its start in the tile: an index into `out`,
    // or (PTR_OUT) a byte address, possibly in peer memory
    typename std::conditional<PTR_OUT != OUT_INDEX, unsigned long long, uint32_t>::type tile_dst[RADIX];
    uint32_t s_warp[8];
    alignas(8) uint64_t mbar;
    uint32_t next_tile;
    BinnerArg binner_arg;  // copy for the out-of-line heavy-tile path
    uint32_t heavy_d4;  // 4 * (a digit holding more than a quarter of the current tile), valid when the tile votes heavy
    uint32_t bucket_start[PTR_OUT == OUT_PTR ? MAX_SPLITTERS + 2 : 1];  // position-binned write-out (SplitBinner): bucket b starts here
    // OUT_SLICED: a bucket's run in this tile may straddle one slice boundary -- keys at tile positions
    // >= split[b] use tile_dst_hi[b] (the next slice's owner) instead of tile_dst[b]
    unsigned long long tile_dst_hi[PTR_OUT == OUT_SLICED ? RADIX : 1];
    uint32_t split[PTR_OUT == OUT_SLICED ? RADIX : 1];
    unsigned long long peer[PTR_OUT == OUT_SLICED ? MAX_SLICES : 1];
    uint32_t slice_start[PTR_OUT == OUT_SLICED ? MAX_SLICES + 1 : 1];
    uint32_t nslice;
};

// Raised by the atomic-ranking kernels when the per-tile audit sees colliding lanes served out of
// lane order (never observed on sm_100; see DESIGN.md "atomic ranking").  One word per device.
static __device__ unsigned int g_rank_violation = 0;

constexpr int RANK_BALLOT = 0;  // one vote.ballot per digit bit + a leader atomic: architecturally ordered
constexpr int RANK_ATOMS = 1;   // rank = return value of the warp's shared atomic (see DESIGN.md "atomic ranking")

// ---- phases shared by k_onesweep and k_onesweep_bulk (onesweep_bulk.cuh) ----------------------
// (1) the tile's keys -> registers, warp-striped so that tile order == rank order.  With use_tma the
// 16-byte-multiple part of the tile has landed in s.keys (mbarrier phase `parity`); a ragged tail
// (<= 3 keys, last tile only) is read straight from global; positions past `valid` become 0xFFFFFFFF.
template <int KPT, int TILE, class Smem>
__device__ __forceinline__ void tile_keys_to_registers(uint32_t (&key)[KPT], Smem &s, const uint32_t *__restrict__ in,
                                                       uint32_t tile_base, uint32_t valid, uint32_t wbase, int use_tma,
                                                       uint32_t parity)
{
    if (use_tma) {
        mbar_wait(&s.mbar, parity);
        if (valid == TILE) {
#pragma unroll
            for (int i = 0; i < KPT; i++) key[i] = s.keys[wbase + i * 32];
        } else {
            const uint32_t staged = ((valid * 4u) & ~15u) >> 2;
#pragma unroll
            for (int i = 0; i < KPT; i++) {
                const uint32_t idx = wbase + i * 32;
                key[i] = idx < staged ? s.keys[idx] : (idx < valid ? in[(size_t)tile_base + idx] : 0xffffffffu);
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < KPT; i++) {
            const uint32_t idx = wbase + i * 32;
            key[i] = idx < valid ? in[(size_t)tile_base + idx] : 0xffffffffu;
        }
    }
    __syncwarp();  // this warp's counters were zeroed by other lanes
}

// (2) early counts: per-warp digit histogram (adding 1 lets same-digit lanes be merged).  Indexed with
// bin(), not bin4(): a different expression than (4) uses, so that the optimiser recomputes the address
// there instead of spilling KPT of them across barriers.  A splitter bin costs a binary search and
// depends on the position: it is kept, 4 bits per key, in dig[].
template <class Binner, int KPT, int NDIG>
__device__ __forceinline__ void tile_count_digits(const uint32_t (&key)[KPT], uint32_t (&dig)[NDIG], const Binner &binner,
                                                  uint32_t *warp_hist, uint32_t first_pos)
{
#pragma unroll
    for (int i = 0; i < KPT; i++) {
        const uint32_t d = binner.bin(key[i], first_pos + i * 32);
        atomicAdd(&warp_hist[d], 1u);
        if (Binner::kByPosition) dig[i >> 3] |= d << (4 * (i & 7));
    }
}

// (4) rank every key of the tile and move it to its place in the digit-sorted tile.  warp counters hold
// running BYTE offsets into the tile buffer.  RANK_ATOMS: a key's place is the value its warp's shared
// atomicAdd returns (see DESIGN.md "atomic ranking"); one row of every tile is re-checked with ballots
// (the audit: lanes of one instruction with the same digit must have been served in ascending lane
// order; a violation raises g_rank_violation and the host side repeats the call with the ballot
// kernels, engine.cu rank_audit).  A tile in which one digit dominates (`heavy`) takes the out-of-line
// path.  RANK_BALLOT: one vote per digit bit + a leader atomic.  first_pos = position of key[0].
template <int RANK, int NBITS, class Binner, int KPT, class Smem, int NDIG>
__device__ __forceinline__ void tile_rank_and_place(const uint32_t (&key)[KPT], const uint32_t (&dig)[NDIG], const Binner &binner,
                                                    Smem &s, unsigned char *my_hist, unsigned char *tile_bytes, int heavy,
                                                    uint32_t first_pos)
{
    if (RANK == RANK_ATOMS) {
        if (!heavy) {
#pragma unroll
            for (int i = 0; i < KPT; i++) {
                const uint32_t p4 = atomicAdd(reinterpret_cast<uint32_t *>(my_hist + binner.bin4(key[i], first_pos + i * 32)), 4u);
                *reinterpret_cast<uint32_t *>(tile_bytes + p4) = key[i];
                if (i == 0) {  // the audit row (1/KPT of the keys)
                    const uint32_t d4 = binner.bin4(key[0], first_pos);
                    const uint32_t peers = match_digit<NBITS>(d4 >> 2, binner.nbits());
                    const uint32_t first = __shfl_sync(0xffffffffu, p4, __ffs(peers) - 1);
                    if (p4 != first + 4u * __popc(peers & lanemask_lt())) atomicOr(&g_rank_violation, 1u);
                }
            }
        } else {
            // one digit dominates the tile (skewed input): out-of-line path, see rank_tile_heavy
            uint32_t spilled[KPT];
#pragma unroll
            for (int i = 0; i < KPT; i++) spilled[i] = key[i];
            rank_tile_heavy<Binner, Smem>(spilled, &s);
        }
    } else {
        const uint32_t lt = lanemask_lt();
        const int lane = threadIdx.x & 31;
        const int nbits = binner.nbits();
        const uint32_t hist_addr = smem_u32(my_hist);
#pragma unroll
        for (int i = 0; i < KPT; i++) {
            const uint32_t d4 = Binner::kByPosition ? ((dig[i >> 3] >> (4 * (i & 7))) & 15u) << 2
                                                    : binner.bin4(key[i], first_pos + i * 32);
            const uint32_t peers = match_digit<(Binner::kBits ? Binner::kBits : NBITS)>(d4 >> 2, nbits);
            const uint32_t before = __popc(peers & lt);
            const int top = 31 - __clz(peers);  // highest peer lane leads; its `before` + 1 is the group size
            uint32_t off4 = leader_atom_add(hist_addr + d4, (before + 1u) * 4u, lane == top);
            off4 = __shfl_sync(0xffffffffu, off4, top);
            *reinterpret_cast<uint32_t *>(tile_bytes + off4 + before * 4u) = key[i];
        }
    }
}

// in/out: this portion's keys (n < 2^30 so that counts fit the 30-bit
// descriptor value).  base: global bucket bases for this portion;
// base_next (nullable): receives base + portion bucket totals for the next
// portion.  desc: ntiles x 256 zero-initialised descriptors; ticket: zeroed.
// PTR_OUT (an OUT_* mode): OUT_PTR -- `base` holds one 64-bit destination pointer per
// bin (local or NVLink-mapped peer memory) instead of an offset into `out`;
// OUT_SLICED -- `base` is a SliceTable (slices >= one tile long, single portion).
//
// One tile per loop trip, one 1-D bulk TMA per tile.  The ticket for a tile is
// taken only when the CTA is about to load it, so tiles publish their counts
// in (nearly) ticket order and a look-back rarely meets an unpublished
// predecessor; load latency is covered by the other CTAs resident on the SM.
template <int THREADS, int KPT, int MINB, int RANK, int NBITS, int LB, class Binner, int PTR_OUT>
__global__ void __launch_bounds__(THREADS, MINB)
    k_onesweep(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, uint32_t n, const typename Binner::Arg binner_arg,
               const void *__restrict__ base_v, uint32_t *__restrict__ base_next, uint32_t *__restrict__ desc,
               uint32_t *__restrict__ ticket, uint32_t ntiles, int use_tma, const uint32_t *__restrict__ dyn_n)
{
    using Smem = OnesweepSmem<THREADS, KPT, PTR_OUT, typename Binner::Arg>;
    static_assert(!(RANK == RANK_ATOMS && Binner::kByPosition), "the heavy-tile path bins by key alone");
    constexpr int TILE = Smem::TILE;
    constexpr int NWARPS = Smem::NWARPS;
    static_assert(THREADS >= RADIX, "need one thread per bin");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem &s = *reinterpret_cast<Smem *>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // Prologue that touches no global memory (barrier init, counters zeroed): under a programmatic
    // dependent launch it runs while the previous pass drains.  Everything after pdl_wait() sees the
    // previous kernel's keys, bases, tickets and descriptors.
    if (tid == 0) {
        mbar_init(&s.mbar, 1);
        fence_mbar_init();
    }
#pragma unroll
    for (int j = lane; j < RADIX; j += 32) s.warp_hist[warp][j] = 0;
    pdl_wait();
    pdl_trigger();
    if (dyn_n) {  // key count decided on the device: the grid and the descriptors were sized for the capacity
        n = min(*dyn_n, n);  // never past the capacity the launch was sized for
        ntiles = (n + (uint32_t)(THREADS * KPT) - 1u) / (uint32_t)(THREADS * KPT);
    }
    if (PTR_OUT == OUT_INDEX && ticket[SKIP_OFFSET]) {  // constant digit (k_scan_hist): the pass is a copy
        copy_keys(in, out, n);
        return;
    }
    const Binner binner(binner_arg);
    const uint32_t pad_bin = binner.pad_bin();
    unsigned char *const my_hist = reinterpret_cast<unsigned char *>(s.warp_hist[warp]);
    unsigned char *const tile_bytes = reinterpret_cast<unsigned char *>(s.keys);

    auto issue_load = [&](uint32_t tile) {
        // elected thread: arm the barrier and start the bulk copy of the
        // 16 B-multiple part of the tile; a ragged tail (<= 3 keys, last tile
        // only) is read straight from global by its consumer.
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
        const uint32_t t = atomicAdd(ticket, 1u);
        s.next_tile = t;
        s.binner_arg = binner_arg;
        if (use_tma && t < ntiles) issue_load(t);
    }
    if (PTR_OUT == OUT_SLICED) {  // the slice map, once per CTA
        const SliceTable *tab = static_cast<const SliceTable *>(base_v);
        if (tid < MAX_SLICES) s.peer[tid] = tab->peer[tid];
        if (tid <= MAX_SLICES) s.slice_start[tid] = tab->slice_start[tid];
        if (tid == 0) s.nslice = tab->nslice;
    }
    __syncthreads();
    uint32_t cur = s.next_tile;

    for (uint32_t it = 0; cur < ntiles; ++it) {
        const uint32_t tile_base = cur * (uint32_t)TILE;
        const uint32_t valid = min((uint32_t)TILE, n - tile_base);

        // (1) keys -> registers, (2) early per-warp digit counts (helpers above)
        uint32_t key[KPT];
        const uint32_t wbase = warp * (KPT * 32) + lane;
        tile_keys_to_registers<KPT, TILE>(key, s, in, tile_base, valid, wbase, use_tma, it & 1);
        uint32_t dig[Binner::kByPosition ? (KPT + 7) / 8 : 1] = {};
        tile_count_digits(key, dig, binner, s.warp_hist[warp], tile_base + wbase);
        __syncthreads();

        // (3) bin threads: reduce over warps, publish the tile aggregate,
        //     scan bins, turn warp counters into running offsets in the tile
        uint32_t my_count = 0, my_start4 = 0, padded4 = 0;
        if (tid < RADIX) {
#pragma unroll
            for (int w = 0; w < NWARPS; w++) padded4 += s.warp_hist[w][tid];
            my_count = padded4;
            padded4 <<= 2;  // bytes from here on
            if ((uint32_t)tid == pad_bin) my_count -= (uint32_t)TILE - valid;  // padding keys carry the top digit
            st_relaxed(&desc[(size_t)cur * RADIX + tid], (cur == 0 ? DESC_INCL : DESC_AGG) | my_count);
        }
        // first look-back loads go out now and land while the tile is ranked
        uint32_t lbv[LB];
        if (tid < RADIX) {
#pragma unroll
            for (int k = 0; k < LB; k++)
                lbv[k] = (uint32_t)k < cur ? ld_relaxed(&desc[(size_t)(cur - 1 - k) * RADIX + tid]) : DESC_INCL;
        }
        my_start4 = block_excl_scan_256(padded4, s.s_warp, nullptr);
        if (tid < RADIX) {
            uint32_t run = my_start4;
#pragma unroll
            for (int w = 0; w < NWARPS; w++) {
                const uint32_t c = s.warp_hist[w][tid];
                s.warp_hist[w][tid] = run;
                run += c << 2;
            }
        }
        // barrier + vote: does one digit hold more than a quarter of the tile?  (padded4 counts bytes)
        const bool is_heavy = RANK == RANK_ATOMS && padded4 > (uint32_t)TILE;
        if (is_heavy) s.heavy_d4 = (uint32_t)tid * 4u;  // up to three candidates: any one will do
        const int heavy = __syncthreads_or(is_heavy);

        // (4) rank every key and move it to its place in the digit-sorted tile (all warps hold their
        //     keys in registers since (2), so the landing buffer can be overwritten)
        tile_rank_and_place<RANK, NBITS, Binner, KPT, Smem>(key, dig, binner, s, my_hist, tile_bytes, heavy, tile_base + wbase);

        // (5) decoupled look-back for this thread's bin
        if (tid < RADIX) {
            uint32_t excl = 0;
            if (cur != 0) {
                // walk back LB predecessors per round trip: the loads of a window are
                // in flight together; an unpublished descriptor restarts the window there
                uint32_t j = cur - 1;  // next predecessor to account for
                bool done = false;
                while (true) {
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
#pragma unroll
                    for (int k = 0; k < LB; k++)
                        lbv[k] = (uint32_t)k <= j ? ld_relaxed(&desc[(size_t)(j - k) * RADIX + tid]) : DESC_INCL;
                }
                st_relaxed(&desc[(size_t)cur * RADIX + tid], DESC_INCL | (excl + my_count));
            }
            const uint32_t my_start = my_start4 >> 2;
            if (PTR_OUT == OUT_PTR && Binner::kByPosition && tid <= MAX_SPLITTERS + 1) s.bucket_start[tid] = my_start;
            if (PTR_OUT == OUT_SLICED) {
                // this tile's run of bucket `tid` starts at global position g: find the slice it falls
                // into; the part of the run past that slice's end goes to the next slice's owner
                const uint32_t g = static_cast<const SliceTable *>(base_v)->gbase[tid] + excl;
                uint32_t j = 0;
                for (uint32_t k = 1; k < s.nslice; k++) j += g >= s.slice_start[k] ? 1u : 0u;
                const uint32_t sp = my_start + (s.slice_start[j + 1] - g);  // tile position where slice j+1 begins
                s.tile_dst[tid] = s.peer[j] + 4ull * (unsigned long long)(g - s.slice_start[j]) - 4ull * (unsigned long long)my_start;
                s.tile_dst_hi[tid] = (j + 1 < s.nslice ? s.peer[j + 1] : 0ull) - 4ull * (unsigned long long)sp;
                s.split[tid] = sp;
            } else if (PTR_OUT == OUT_PTR) {
                const unsigned long long gptr = static_cast<const unsigned long long *>(base_v)[tid];
                s.tile_dst[tid] = gptr + 4ull * (unsigned long long)excl - 4ull * (unsigned long long)my_start;
            } else {
                const uint32_t gbase = static_cast<const uint32_t *>(base_v)[tid];
                s.tile_dst[tid] = gbase + excl - my_start;
                if (base_next && cur == ntiles - 1) base_next[tid] = gbase + excl + my_count;
            }
        }
        // the counters are dead until the next tile's (2): clear this warp's own
#pragma unroll
        for (int j = lane; j < RADIX; j += 32) s.warp_hist[warp][j] = 0;
        uint32_t t_next = 0;
        if (tid == 0) t_next = atomicAdd(ticket, 1u);  // in flight during the write-out
        __syncthreads();

        // (6) coalesced write-out: consecutive threads, consecutive sorted keys.  The sorted
        //     keys go to registers first; from the barrier on the tile buffer is free, so the
        //     next tile's bulk load runs under this tile's stores.
#pragma unroll
        for (int i = 0; i < KPT; i++) key[i] = s.keys[tid + i * THREADS];
        if (tid == 0) s.next_tile = t_next;
        __syncthreads();
        const uint32_t nxt = s.next_tile;
        if (tid == 0 && use_tma && nxt < ntiles) {
            fence_proxy_async();  // generic-proxy accesses to keys[] come before the bulk copy's writes
            issue_load(nxt);
        }
#pragma unroll
        for (int i = 0; i < KPT; i++) {
            const uint32_t p = tid + i * THREADS;
            uint32_t d4;
            if (PTR_OUT == OUT_PTR && Binner::kByPosition) {
                // equal keys may sit in different buckets: the bucket follows from the position
                uint32_t d = 0;
                for (uint32_t b = 1; b <= pad_bin; b++) d += s.bucket_start[b] <= p ? 1u : 0u;
                d4 = d << 2;
            } else {
                d4 = binner.bin4(key[i]);
            }
            if (valid == TILE || p < valid) {
                if (PTR_OUT == OUT_SLICED) {
                    const uint32_t sp = *reinterpret_cast<const uint32_t *>(reinterpret_cast<const unsigned char *>(s.split) + d4);
                    const unsigned char *tab = p < sp ? reinterpret_cast<const unsigned char *>(s.tile_dst)
                                                      : reinterpret_cast<const unsigned char *>(s.tile_dst_hi);
                    *reinterpret_cast<uint32_t *>(*reinterpret_cast<const unsigned long long *>(tab + 2 * d4) + 4ull * p) = key[i];
                } else if (PTR_OUT == OUT_PTR)
                    *reinterpret_cast<uint32_t *>(*reinterpret_cast<const unsigned long long *>(
                                                      reinterpret_cast<const unsigned char *>(s.tile_dst) + 2 * d4) +
                                                  4ull * p) = key[i];
                else
                    out[*reinterpret_cast<const uint32_t *>(reinterpret_cast<const unsigned char *>(s.tile_dst) + d4) + p] = key[i];
            }
        }
        cur = nxt;
    }
}

// Device self-test behind LIBSORT_RANK=auto (engine.cu choose_rank_mode): every
// lane adds to a pseudo-random counter of its warp and gets the old value; the
// lanes of one instruction that named the same counter must have been served
// in ascending lane order.  out[0] += violations, out[1] += colliding lanes seen.
__global__ void __launch_bounds__(1024) k_selftest_atoms_order(unsigned long long *out, int iters)
{
    __shared__ uint32_t h[32][RADIX];
    for (int i = threadIdx.x; i < 32 * RADIX; i += blockDim.x) (&h[0][0])[i] = 0;
    __syncthreads();
    uint32_t *my = h[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31, lt = lanemask_lt();
    uint32_t s = (threadIdx.x + 1) * 2654435761u + blockIdx.x * 97u;
    uint32_t bad = 0, seen = 0;
    for (int i = 0; i < iters; i++) {
        s = s * 1664525u + 1013904223u;
        uint32_t d = s >> 24;
        if ((i & 3) == 1) d &= 15;               // denser collisions
        if ((i & 3) == 3) d = (lane * 5 + i) & 1;  // two 16-lane pile-ups
        const uint32_t r = atomicAdd(&my[d], 1u);
        const uint32_t peers = match_digit<8>(d, 8);
        const uint32_t r0 = __shfl_sync(0xffffffffu, r, __ffs(peers) - 1);
        if (r != r0 + __popc(peers & lt)) bad++;
        seen += (peers & lt) ? 1 : 0;
    }
    if (bad) atomicAdd(&out[0], (unsigned long long)bad);
    atomicAdd(&out[1], (unsigned long long)seen);
}

// ------------------------------------------------------- distributed sort --
// Pseudo-random sample of the local shard (splitter selection).
__global__ void k_sample(const uint32_t *__restrict__ in, uint32_t n, uint32_t *__restrict__ out, uint32_t nsamples,
                         uint32_t seed)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nsamples) return;
    // stratified: one draw from each of nsamples equal strata (a multiply-xorshift hash picks the slot)
    uint32_t h = (i + seed) * 0x9E3779B1u;
    h ^= h >> 15; h *= 0x85EBCA77u; h ^= h >> 13;
    const unsigned long long lo = (unsigned long long)i * n / nsamples, hi = (unsigned long long)(i + 1) * n / nsamples;
    const unsigned long long span = hi > lo ? hi - lo : 1;
    out[i] = in[lo + h % span];
}

// Keys per destination rank (bank-private counters: lane l owns column l).
struct PtrTable { unsigned long long p[MAX_SPLITTERS + 1]; };
template <int NS>
__global__ void __launch_bounds__(1024, 1) k_dest_hist(const uint32_t *__restrict__ in, size_t n, const SplitArg arg,
                                                      uint32_t *__restrict__ counts)
{
    const SplitBinner<NS> binner(arg);
    __shared__ uint32_t cnt[MAX_SPLITTERS + 1][32];
    for (int i = threadIdx.x; i < (MAX_SPLITTERS + 1) * 32; i += blockDim.x) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    size_t head = ((16u - (reinterpret_cast<uintptr_t>(in) & 15u)) & 15u) >> 2;
    if (head > n) head = n;
    const size_t nvec = (n - head) >> 2, tail0 = head + (nvec << 2);
    const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, gstride = (size_t)gridDim.x * blockDim.x;
    const uint4 *vin = reinterpret_cast<const uint4 *>(in + head);
    if (!binner.ties) {  // the usual case, kept free of the tie walk so that the streaming loop stays tight
        if (gtid < head) atomicAdd(&cnt[binner.bin(in[gtid])][lane], 1u);
        if (gtid < n - tail0) atomicAdd(&cnt[binner.bin(in[tail0 + gtid])][lane], 1u);
        for (size_t i = gtid; i < nvec; i += gstride) {
            const uint4 v = ld_stream_v4(vin + i);
            atomicAdd(&cnt[binner.bin(v.x)][lane], 1u);
            atomicAdd(&cnt[binner.bin(v.y)][lane], 1u);
            atomicAdd(&cnt[binner.bin(v.z)][lane], 1u);
            atomicAdd(&cnt[binner.bin(v.w)][lane], 1u);
        }
    } else {  // ties are split by position (SplitBinner), exactly as the scatter will
        if (gtid < head) atomicAdd(&cnt[binner.bin(in[gtid], (uint32_t)gtid)][lane], 1u);
        if (gtid < n - tail0) atomicAdd(&cnt[binner.bin(in[tail0 + gtid], (uint32_t)(tail0 + gtid))][lane], 1u);
        for (size_t i = gtid; i < nvec; i += gstride) {
            const uint4 v = ld_stream_v4(vin + i);
            const uint32_t p = (uint32_t)(head + (i << 2));
            atomicAdd(&cnt[binner.bin(v.x, p)][lane], 1u);
            atomicAdd(&cnt[binner.bin(v.y, p + 1)][lane], 1u);
            atomicAdd(&cnt[binner.bin(v.z, p + 2)][lane], 1u);
            atomicAdd(&cnt[binner.bin(v.w, p + 3)][lane], 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x <= (unsigned)binner.nsplit) {
        uint32_t c = 0;
        for (int l = 0; l < 32; l++) c += cnt[threadIdx.x][(l + lane) & 31];
        if (c) atomicAdd(&counts[threadIdx.x], c);
    }
}

// splitters[g-1] = the g/P quantile of the sorted all-rank sample; splitters[16 + g-1] = its tie
// threshold (see SplitBinner).  When the quantile falls inside a run [r0, r1) of >= TIE_MIN_RUN equal
// samples -- a value heavy enough to matter -- the copies of that value are cut at the quantile too:
// the fraction (q - r0) / (r1 - r0) of them stays below this splitter.
constexpr uint32_t TIE_MIN_RUN = 8;
__global__ void k_pick_splitters(const uint32_t *__restrict__ sorted, uint32_t nsorted, int nranks,
                                 uint32_t *__restrict__ splitters)
{
    const int g = threadIdx.x + 1;
    if (threadIdx.x < SPLIT_WORDS - SPLIT_TIE_OFFSET) splitters[SPLIT_TIE_OFFSET + threadIdx.x] = 0;
    if (g >= nranks) return;
    const uint32_t q = (uint32_t)((unsigned long long)g * nsorted / nranks);
    const uint32_t v = sorted[q];
    splitters[g - 1] = v;
    uint32_t lo = 0, hi = q;  // r0 = first index holding v
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (sorted[mid] < v) lo = mid + 1; else hi = mid;
    }
    const uint32_t r0 = lo;
    lo = q + 1, hi = nsorted;  // r1 = first index holding more than v
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (sorted[mid] <= v) lo = mid + 1; else hi = mid;
    }
    const uint32_t r1 = lo;
    if (r1 - r0 >= TIE_MIN_RUN && q > r0 && v != 0xffffffffu)  // 0xffffffff pads partial tiles: never split
        splitters[SPLIT_TIE_OFFSET + g - 1] = (uint32_t)((((unsigned long long)(q - r0)) << 32) / (r1 - r0));
}

// Step-wise distributed LSD sort: gathered[k][d] = keys of digit d on rank k.  The global array
// after the step is bucket-major, rank-minor (the reference's regather order, localTest/
// benchmarks.cpp:120-136, benchmark/pkg/sort/helpers.go:67-78), so rank `me`'s first key of digit
// d lands at  sum over digits < d of all ranks' counts  +  sum over ranks < me of their digit-d counts.
__global__ void __launch_bounds__(RADIX) k_step_bases(const uint32_t *__restrict__ gathered, int nranks, int me,
                                                      SliceTable *__restrict__ tab)
{
    __shared__ uint32_t s_warp[8];
    const int d = threadIdx.x;
    uint32_t tot = 0, before = 0;
    for (int k = 0; k < nranks; k++) {
        const uint32_t c = gathered[k * RADIX + d];
        tot += c;
        if (k < me) before += c;
    }
    tab->gbase[d] = block_excl_scan_256(tot, s_warp, nullptr) + before;
}

__global__ void k_store_ptrs(const PtrTable t, unsigned long long *__restrict__ dst)
{
    if (threadIdx.x <= MAX_SPLITTERS) dst[threadIdx.x] = t.p[threadIdx.x];
}

// Receive layout of the exchange, computed where the count matrix already is (no host round
// trip): cm[s][d] = keys rank s sends to rank d (nranks x 16 words, all-gathered).  Rank `me`
// writes its keys for rank d behind those of the ranks before it: dest_ptrs[d] = peer_base[d] +
// 4 * sum over s < me of cm[s][d]; what arrives here is nrecv = sum over s of cm[s][me].  If any
// rank would receive more than `cap` keys, status is raised and the scatter that follows is
// switched off (its tile ticket starts past the last tile) -- every rank sees the same matrix and
// takes the same decision.
__global__ void k_plan_exchange(const uint32_t *__restrict__ cm, int nranks, int me,
                                const unsigned long long *__restrict__ peer_base, unsigned long long cap,
                                unsigned long long *__restrict__ dest_ptrs, uint32_t *__restrict__ nrecv,
                                uint32_t *__restrict__ status, uint32_t *__restrict__ ticket)
{
    const int d = threadIdx.x;
    unsigned long long off = 0, tot = 0;
    if (d < nranks) {
        for (int s = 0; s < nranks; s++) {
            const uint32_t c = cm[s * (MAX_SPLITTERS + 1) + d];
            if (s < me) off += c;
            tot += c;
        }
    }
    const bool any_over = __any_sync(0xffffffffu, d < nranks && tot > cap);
    if (d < nranks) dest_ptrs[d] = peer_base[d] + 4ull * off;
    if (d == me) *nrecv = any_over ? 0u : (uint32_t)tot;
    if (d == 0 && any_over) {
        *status = 1u;
        *ticket = 0x7fffffffu;
    }
}

// ---- splitters straight from a histogram (distributed sort, plan = "hist") -------------------
// For inputs whose top 12 bits are spread well enough, neither a sample nor a counting pass is
// needed: one 4096-bin histogram of key >> 20 per shard (k_hist_top12, one read of the shard),
// all-gathered, gives (a) splitters on bin boundaries that cut the union into nranks near-equal
// ranges and (b) the exact nranks x nranks count matrix for those splitters -- what the sample sort
// + k_dest_hist pass + count all-gather produce otherwise.  k_splitters_from_hist also reports the
// resulting imbalance (largest range * 1000 / mean) so that the host side can fall back to sampled
// splitters with tie thresholds for inputs this cannot balance (few distinct top bits, heavy ties).
constexpr int TOP_BITS = 12, TOP_BINS = 1 << TOP_BITS, TOP_SHIFT = 32 - TOP_BITS;

template <int THREADS>
__global__ void __launch_bounds__(THREADS) k_hist_top12(const uint32_t *__restrict__ in, size_t n, uint32_t *__restrict__ ghist)
{
    __shared__ uint32_t s_hist[TOP_BINS];
    for (int i = threadIdx.x; i < TOP_BINS; i += THREADS) s_hist[i] = 0;
    __syncthreads();
    auto count = [&](uint32_t k) { atomicAdd(&s_hist[k >> TOP_SHIFT], 1u); };
    size_t head = ((16u - (reinterpret_cast<uintptr_t>(in) & 15u)) & 15u) >> 2;
    if (head > n) head = n;
    const size_t nvec = (n - head) >> 2, tail0 = head + (nvec << 2);
    const size_t gtid = (size_t)blockIdx.x * THREADS + threadIdx.x, gstride = (size_t)gridDim.x * THREADS;
    if (gtid < head) count(in[gtid]);
    if (gtid < n - tail0) count(in[tail0 + gtid]);
    const uint4 *vin = reinterpret_cast<const uint4 *>(in + head);
    constexpr int U = 4;
    size_t i = gtid;
    for (; i + (U - 1) * gstride < nvec; i += U * gstride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; u++) v[u] = ld_stream_v4(vin + i + u * gstride);
#pragma unroll
        for (int u = 0; u < U; u++) { count(v[u].x); count(v[u].y); count(v[u].z); count(v[u].w); }
    }
    for (; i < nvec; i += gstride) {
        const uint4 v = ld_stream_v4(vin + i);
        count(v.x); count(v.y); count(v.z); count(v.w);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < TOP_BINS; j += THREADS) {
        const uint32_t c = s_hist[j];
        if (c) atomicAdd(&ghist[j], c);
    }
}

// One CTA of 1024 threads, 4 bins each.  hists: nranks x 4096 (all-gathered).  Writes the 32-word
// splitter table (values on bin boundaries, tie thresholds 0), the count matrix in the layout of the
// all-gathered k_dest_hist counts (nranks x 16) and *imbalance_permille.
__global__ void __launch_bounds__(1024) k_splitters_from_hist(const uint32_t *__restrict__ hists, int nranks,
                                                              uint32_t *__restrict__ splitters, uint32_t *__restrict__ cm,
                                                              uint32_t *__restrict__ imbalance_permille)
{
    __shared__ unsigned long long s_warp[32];
    __shared__ uint32_t s_e[MAX_SPLITTERS + 2];
    __shared__ unsigned long long s_cm[MAX_SPLITTERS + 1][MAX_SPLITTERS + 1];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    uint32_t cnt[4];
    unsigned long long mine = 0;
    cnt[0] = cnt[1] = cnt[2] = cnt[3] = 0;  // a bin of the union holds < 2^32 keys (shards are < 2^28 keys in practice)
    for (int r = 0; r < nranks; r++) {
        const uint4 h = *reinterpret_cast<const uint4 *>(hists + (size_t)r * TOP_BINS + 4 * t);
        cnt[0] += h.x; cnt[1] += h.y; cnt[2] += h.z; cnt[3] += h.w;
    }
    mine = (unsigned long long)cnt[0] + cnt[1] + cnt[2] + cnt[3];
    if (t < (MAX_SPLITTERS + 1) * (MAX_SPLITTERS + 1)) (&s_cm[0][0])[t] = 0;
    if (t < MAX_SPLITTERS + 2) s_e[t] = 0;
    // block exclusive scan of `mine`
    unsigned long long inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    unsigned long long wbase = 0, total = 0;
    for (int w = 0; w < 32; w++) {
        const unsigned long long c = s_warp[w];
        if (w < warp) wbase += c;
        total += c;
    }
    unsigned long long excl = wbase + inc - mine;
    __shared__ unsigned long long s_target[MAX_SPLITTERS + 1];
    if (t >= 1 && t < nranks) s_target[t] = total * (unsigned long long)t / (unsigned long long)nranks;
    __syncthreads();
    // boundary g (between rank g-1 and g) = the bin edge closest to the g/nranks quantile
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t b = 4 * t + k;
        if (cnt[k]) {
            for (int g = 1; g < nranks; g++) {
                const unsigned long long target = s_target[g];
                if (excl <= target && target < excl + cnt[k]) {
                    uint32_t e = (target - excl <= cnt[k] / 2) ? b : b + 1;
                    s_e[g] = min(e, (uint32_t)TOP_BINS - 1u);  // the last bin always stays right of every splitter
                }
            }
        }
        excl += cnt[k];
    }
    if (t == 0) s_e[nranks] = TOP_BINS;
    __syncthreads();
    // count matrix for those boundaries: destinations are monotone in the bin index, so a warp (128
    // consecutive bins) meets only a few of them -- reduce inside the warp, one atomic per (warp, source, destination)
    int dk[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        int d = 0;
        for (int g = 1; g < nranks; g++) d += s_e[g] <= (uint32_t)(4 * t + k) ? 1 : 0;
        dk[k] = d;
    }
    const int dlo = __shfl_sync(0xffffffffu, dk[0], 0), dhi = __shfl_sync(0xffffffffu, dk[3], 31);
    for (int r = 0; r < nranks; r++) {
        const uint4 h = *reinterpret_cast<const uint4 *>(hists + (size_t)r * TOP_BINS + 4 * t);
        const uint32_t hv[4] = {h.x, h.y, h.z, h.w};
        for (int d = dlo; d <= dhi; d++) {
            uint32_t v = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) v += dk[k] == d ? hv[k] : 0u;
            unsigned long long sum = v;  // <= 128 bins of < 2^30 keys each
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            if (lane == 0 && sum) atomicAdd(&s_cm[r][d], sum);
        }
    }
    __syncthreads();
    if (t < (MAX_SPLITTERS + 1) * (MAX_SPLITTERS + 1)) {
        const int r = t / (MAX_SPLITTERS + 1), d = t % (MAX_SPLITTERS + 1);
        if (r < nranks) cm[r * (MAX_SPLITTERS + 1) + d] = (uint32_t)min(s_cm[r][d], 0xffffffffull);
    }
    if (t < 2 * (MAX_SPLITTERS + 1)) {
        uint32_t v = 0;
        if (t < nranks - 1) v = s_e[t + 1] << TOP_SHIFT;
        splitters[t] = v;  // [0,16): values, [16,32): tie thresholds (none)
    }
    if (t == 0) {
        unsigned long long worst = 0;
        for (int d = 0; d < nranks; d++) {
            unsigned long long tot = 0;
            for (int r = 0; r < nranks; r++) tot += s_cm[r][d];
            worst = max(worst, tot);
        }
        *imbalance_permille = total ? (uint32_t)min(worst * 1000ull * (unsigned long long)nranks / total, 0xffffffffull) : 1000u;
    }
}

// ------------------------------------------------------------ boundaries --
// Exact boundaries for wide groups: thread g finds the first index whose
// group is >= g in the array already sorted by those bits.
__global__ void k_bounds_search(const uint32_t *__restrict__ sorted, uint32_t n, uint32_t offset, uint32_t gmask,
                                uint32_t *__restrict__ exact, size_t ngroups)
{
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < ngroups; g += (size_t)gridDim.x * blockDim.x) {
        uint32_t lo = 0, hi = n;
        while (lo < hi) {
            uint32_t mid = lo + ((hi - lo) >> 1);
            if (((sorted[mid] >> offset) & gmask) < (uint32_t)g) lo = mid + 1; else hi = mid;
        }
        exact[g] = lo;
    }
}

// In place: exact boundaries -> what the reference's gpu_groups + host
// back-fill produce (libsort/sort.cu:14-27, 381-387).  With g0 the first
// non-empty group:  R[0] = 0;  R[1] = E[1] iff groups 0 and 1 are both
// non-empty, else 0;  R[g>=2] = E[g] for g > g0, else E[g0+1] (E[G] = n).
// Single CTA: the reads of E[g0+1], E[1], E[2] complete before any write.
__global__ void __launch_bounds__(1024) k_bounds_refquirk(uint32_t *__restrict__ b, size_t ngroups, uint32_t n)
{
    __shared__ unsigned long long s_g0;
    __shared__ uint32_t s_fill, s_r1;
    if (threadIdx.x == 0) s_g0 = ngroups;
    __syncthreads();
    unsigned long long mine = ngroups;
    for (size_t g = threadIdx.x; g < ngroups; g += blockDim.x) {
        uint32_t nxt = (g + 1 < ngroups) ? b[g + 1] : n;
        if (nxt != b[g]) { mine = g; break; }
    }
    if (mine < ngroups) atomicMin(&s_g0, mine);
    __syncthreads();
    if (threadIdx.x == 0) {
        const size_t g0 = s_g0;
        s_fill = (g0 + 1 < ngroups) ? b[g0 + 1] : n;
        uint32_t r1 = 0;
        if (ngroups > 1) {
            const uint32_t e1 = b[1], e2 = ngroups > 2 ? b[2] : n;
            if (e1 != 0 && e2 != e1) r1 = e1;  // count[0] > 0 and count[1] > 0
        }
        s_r1 = r1;
    }
    __syncthreads();
    const size_t g0 = s_g0;
    for (size_t g = threadIdx.x; g < ngroups; g += blockDim.x) {
        if (g == 0) b[0] = 0;
        else if (g == 1) b[1] = s_r1;
        else if (g <= g0) b[g] = s_fill;
    }
}

}  // namespace b200sort
