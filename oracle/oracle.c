/*
 * oracle.c -- CPU restatement of the jssmith/gpu-radix-sort hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product (gpu-radix-sort_b200/)
 * links, loads or calls this file.  It is used by tests/, by
 * __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference
 * legs, and only as the checker or as the reported CPU baseline.
 *
 * Parity status: PINNED.  The restatement is checked against
 *   (a) the reference's own host code compiled from /root/reference
 *       (oracle/_ref/libsort_ref.so: populateInput, providedCpu) and
 *   (b) outputs of the reference's own GPU path (providedGpu / gpuPartial)
 *       captured on a B200 from that same unmodified build and committed
 *       under tests/golden/ (see tests/golden/make_golden.py).
 *
 * Every function cites the reference file:line (relative to the reference
 * repo root) whose behaviour it restates.  All arithmetic is uint32.
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

/* libsort/sort.cu:9 and localTest/local.h:10 -- group_bits(val,pos,width).
 * A shift count >= 32 yields 0 on the GPU (PTX shr.u32 clamps), which is what
 * the reference kernels see when offset+width runs past bit 31. */
static inline uint32_t group_bits(uint32_t v, uint32_t pos, uint32_t width)
{
    uint32_t s = pos >= 32 ? 0u : (v >> pos);
    uint32_t m = width >= 32 ? 0xffffffffu : ((1u << width) - 1u);
    return s & m;
}

/* libsort/utils.cu:63-79 -- populateInput: PCG-XSH-RR 64->32.  The reference
 * keeps the state in a function-static; here the caller owns it so tests can
 * replay "fresh process" streams.  Initial state: 0x4d595df4d0f33173. */
void oracle_pcg_fill(uint64_t *state, uint32_t *arr, size_t nelem)
{
    const uint64_t mult = 6364136223846793005ull;
    const uint64_t inc = 1442695040888963407ull;
    uint64_t s = *state;
    for (size_t i = 0; i < nelem; i++) {
        uint64_t x = s;
        unsigned rot = (unsigned)(x >> 59);
        s = x * mult + inc;
        x ^= x >> 18;
        uint32_t xs = (uint32_t)(x >> 27);
        arr[i] = (xs >> rot) | (xs << ((32 - rot) & 31));
    }
    *state = s;
}

uint64_t oracle_pcg_initial_state(void) { return 0x4d595df4d0f33173ull; }

/* One 2-bit pass of SortState::Step (libsort/sort.cu:322-346).  The three
 * launches of a pass -- gpu_radix_sort_local (sort.cu:29-184: block-local
 * stable 4-way split + per-(digit,block) counts), sum_scan_blelloch over the
 * digit-major count table (scan.cu:165-250) and gpu_glbl_shuffle
 * (sort.cu:186-213: dst = scan[digit*grid+blk] + rank) -- compose to exactly
 * one stable counting sort of the whole array on the 2-bit digit. */
static void pass_2bit(const uint32_t *in, uint32_t *out, size_t n, uint32_t shift)
{
    size_t cnt[4] = {0, 0, 0, 0};
    for (size_t i = 0; i < n; i++) cnt[group_bits(in[i], shift, 2)]++;
    size_t pos[4];
    size_t run = 0;
    for (int d = 0; d < 4; d++) { pos[d] = run; run += cnt[d]; }
    for (size_t i = 0; i < n; i++) out[pos[group_bits(in[i], shift, 2)]++] = in[i];
}

/* SortState::Step(offset,width) (libsort/sort.cu:322-346): 2-bit LSD passes
 * for shift = offset, offset+2, ... < offset+width.  Result is left in `keys`
 * (the reference ping-pongs d_in -> d_out -> d_in inside every pass).
 * Returns 0 on success, -1 on allocation failure. */
int oracle_step(uint32_t *keys, size_t n, uint32_t offset, uint32_t width)
{
    if (n == 0) return 0;
    uint32_t *tmp = (uint32_t *)malloc(n * sizeof(uint32_t));
    if (!tmp) return -1;
    uint32_t *src = keys, *dst = tmp;
    for (uint32_t shift = offset; shift < offset + width; shift += 2) {
        pass_2bit(src, dst, n, shift);
        uint32_t *t = src; src = dst; dst = t;
    }
    if (src != keys) memcpy(keys, src, n * sizeof(uint32_t));
    free(tmp);
    return 0;
}

/* providedGpu (libsort/invokers.cu:45-64): Step(0,32) then copy back. */
int oracle_full_sort(uint32_t *keys, size_t n) { return oracle_step(keys, n, 0, 32); }

/* gpu_groups + host back-fill, literally (libsort/sort.cu:14-27 and
 * sort.cu:367-394).  `sorted` must already be ordered by the group bits.
 * The GPU threads race only when it is not; iterating i upwards reproduces
 * the in-contract result.  Thread 0 compares against itself and never writes
 * (sort.cu:19-23); the back-fill loop stops at group 2 (sort.cu:381-387). */
void oracle_boundaries_ref(const uint32_t *sorted, size_t n, uint32_t offset, uint32_t width,
                           uint32_t *boundaries)
{
    size_t nb = (size_t)1 << width;
    memset(boundaries, 0, nb * sizeof(uint32_t));
    for (size_t i = 1; i < n; i++) {
        uint32_t g = group_bits(sorted[i], offset, width);
        uint32_t pg = group_bits(sorted[i - 1], offset, width);
        if (g != pg) boundaries[g] = (uint32_t)i;
    }
    /* `int prev = data_len;` in the reference: the int<->unsigned round trips
     * are value preserving mod 2^32, so uint32 arithmetic is equivalent. */
    uint32_t prev = (uint32_t)n;
    for (size_t g = nb - 1; g > 1; g--) {
        if (boundaries[g] == 0) boundaries[g] = prev;
        prev = boundaries[g];
    }
}

/* The boundary contract the reference's own tests assert
 * (localTest/tests.cpp:44-68, benchmark/pkg/sort/testHelpers.go:430-439):
 * boundaries[g] = number of keys whose group is < g. */
void oracle_boundaries_exact(const uint32_t *keys, size_t n, uint32_t offset, uint32_t width,
                             uint32_t *boundaries)
{
    size_t nb = (size_t)1 << width;
    uint32_t *cnt = (uint32_t *)calloc(nb, sizeof(uint32_t));
    for (size_t i = 0; i < n; i++) cnt[group_bits(keys[i], offset, width)]++;
    uint32_t run = 0;
    for (size_t g = 0; g < nb; g++) { boundaries[g] = run; run += cnt[g]; }
    free(cnt);
}

/* gpuPartial (libsort/invokers.cu:15-41): Step(offset,width), result back to
 * the caller's array, then GetBoundaries. */
int oracle_gpu_partial(uint32_t *keys, uint32_t *boundaries, size_t n, uint32_t offset, uint32_t width)
{
    if (n > 0xffffffffull) return -1; /* invokers.cu:30-33 */
    if (oracle_step(keys, n, offset, width)) return -1;
    oracle_boundaries_ref(keys, n, offset, width, boundaries);
    return 0;
}

/* distribSort (localTest/benchmarks.cpp:70-160), generalised from 2 to
 * `nworker` contiguous slices the way the Go driver does it
 * (benchmark/pkg/sort/distrib.go:107-158; slice g = [g*per, min(n,(g+1)*per))
 * with per = ceil(n/nworker); the C++ simulation's len/2 split is the
 * nworker==2, even-n case).  32/width steps, LSD first.  Each step: every
 * worker stably partial-sorts its slice on bits [step*width, +width) and the
 * next global order is bucket-major, worker-minor (BucketReader STRIDED,
 * benchmark/pkg/sort/helpers.go:67-78; memcpy loop benchmarks.cpp:120-136).
 * Bucket sizes come from exact boundaries, i.e. what the reference's
 * boundaries are whenever its own shuffle is well defined. */
int oracle_distrib_sort(uint32_t *data, size_t n, int nworker, uint32_t width, int cxx_split)
{
    if (n == 0) return 0;
    size_t nb = (size_t)1 << width;
    uint32_t *tmp = (uint32_t *)malloc(n * sizeof(uint32_t));
    uint32_t *bounds = (uint32_t *)malloc((size_t)nworker * nb * sizeof(uint32_t));
    size_t *start = (size_t *)malloc(((size_t)nworker + 1) * sizeof(size_t));
    if (!tmp || !bounds || !start) return -1;
    if (cxx_split && nworker == 2) { /* benchmarks.cpp:79-80 */
        start[0] = 0; start[1] = n / 2; start[2] = n;
    } else {
        size_t per = (n + (size_t)nworker - 1) / (size_t)nworker;
        for (int w = 0; w <= nworker; w++) {
            size_t s = (size_t)w * per;
            start[w] = s > n ? n : s;
        }
    }
    uint32_t *cur = data, *next = tmp;
    for (uint32_t step = 0; step < 32 / width; step++) {
        for (int w = 0; w < nworker; w++) {
            size_t len = start[w + 1] - start[w];
            if (oracle_step(cur + start[w], len, step * width, width)) return -1;
            oracle_boundaries_exact(cur + start[w], len, step * width, width, bounds + (size_t)w * nb);
        }
        size_t slot = 0;
        for (size_t b = 0; b < nb; b++) {
            for (int w = 0; w < nworker; w++) {
                size_t len = start[w + 1] - start[w];
                size_t lo = bounds[(size_t)w * nb + b];
                size_t hi = (b == nb - 1) ? len : bounds[(size_t)w * nb + b + 1];
                memcpy(next + slot, cur + start[w] + lo, (hi - lo) * sizeof(uint32_t));
                slot += hi - lo;
            }
        }
        uint32_t *t = cur; cur = next; next = t;
    }
    if (cur != data) memcpy(data, cur, n * sizeof(uint32_t));
    free(tmp); free(bounds); free(start);
    return 0;
}

/* Order-independent multiset fingerprint used by the full-size property
 * tests: (sum, xor, sum of k*k mod 2^64). */
void oracle_fingerprint(const uint32_t *keys, size_t n, uint64_t out[3])
{
    uint64_t s = 0, x = 0, q = 0;
    for (size_t i = 0; i < n; i++) {
        uint64_t k = keys[i];
        s += k; x ^= k * 0x9e3779b97f4a7c15ull; q += k * k;
    }
    out[0] = s; out[1] = x; out[2] = q;
}
