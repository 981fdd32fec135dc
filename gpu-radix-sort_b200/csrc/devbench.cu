// devbench.cu -- developer tool (not part of libsort.so): device-resident
// correctness sweep and per-kernel timing through the additive C ABI.
//   devbench check            parity sweep vs. a host stable counting sort
//   devbench time N [reps]    full sort + 8/16-bit partial timings at N keys
#include <algorithm>
#include <cmath>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/libsort_b200.h"

#define CHECK(x)                                                                        \
    do {                                                                                \
        cudaError_t e_ = (x);                                                           \
        if (e_ != cudaSuccess) {                                                        \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            exit(2);                                                                    \
        }                                                                               \
    } while (0)

static std::vector<uint32_t> host_step(const std::vector<uint32_t> &in, uint32_t off, uint32_t width)
{
    // stable sort by bits [off, off+width) (width even, off+width <= 32)
    std::vector<uint32_t> a = in, b(in.size());
    for (uint32_t s = off; s < off + width; s += 2) {
        size_t cnt[4] = {0, 0, 0, 0}, pos[4];
        for (uint32_t k : a) cnt[(k >> s) & 3]++;
        size_t run = 0;
        for (int d = 0; d < 4; d++) { pos[d] = run; run += cnt[d]; }
        for (uint32_t k : a) b[pos[(k >> s) & 3]++] = k;
        a.swap(b);
    }
    return a;
}

static std::vector<uint32_t> host_bounds(const std::vector<uint32_t> &in, uint32_t off, uint32_t width)
{
    std::vector<uint32_t> c((size_t)1 << width, 0), b((size_t)1 << width);
    uint32_t m = (uint32_t)(((uint64_t)1 << width) - 1);
    for (uint32_t k : in) c[(k >> off) & m]++;
    uint32_t run = 0;
    for (size_t g = 0; g < c.size(); g++) { b[g] = run; run += c[g]; }
    return b;
}

struct Dev {
    uint32_t *in = nullptr, *out = nullptr, *tmp = nullptr, *bounds = nullptr;
    void *ws = nullptr;
    size_t cap = 0, wsb = 0;
    void ensure(size_t n)
    {
        if (n <= cap) return;
        if (in) { cudaFree(in); cudaFree(out); cudaFree(tmp); cudaFree(ws); }
        cap = n;
        CHECK(cudaMalloc(&in, (n + 4) * 4));
        CHECK(cudaMalloc(&out, (n + 4) * 4));
        CHECK(cudaMalloc(&tmp, (n + 4) * 4));
        wsb = libsortWorkspaceBytes(n);
        CHECK(cudaMalloc(&ws, wsb));
        if (!bounds) CHECK(cudaMalloc(&bounds, (1u << 16) * 4));
    }
};

static int check_one(Dev &d, const std::vector<uint32_t> &h, uint32_t off, uint32_t width, int misalign, const char *tag)
{
    size_t n = h.size();
    d.ensure(n + 4);
    uint32_t *din = d.in + misalign;
    CHECK(cudaMemcpy(din, h.data(), n * 4, cudaMemcpyHostToDevice));
    CHECK(cudaMemset(d.out, 0xEE, n * 4));
    uint32_t *bptr = width <= 16 ? d.bounds : nullptr;
    int rc = libsortDeviceSort(din, d.out, d.tmp, n, off, width, bptr, LIBSORT_BOUNDS_EXACT, d.ws, d.wsb, 0, nullptr);
    if (rc) { printf("FAIL %s n=%zu off=%u w=%u rc=%d (%s)\n", tag, n, off, width, rc, libsortErrorString(rc)); return 1; }
    CHECK(cudaDeviceSynchronize());
    std::vector<uint32_t> got(n), ref = host_step(h, off, width);
    CHECK(cudaMemcpy(got.data(), d.out, n * 4, cudaMemcpyDeviceToHost));
    if (got != ref) {
        size_t i = 0;
        while (i < n && got[i] == ref[i]) i++;
        printf("FAIL %s n=%zu off=%u w=%u mis=%d first diff at %zu: got %08x want %08x\n", tag, n, off, width, misalign, i,
               got[i], ref[i]);
        return 1;
    }
    if (bptr) {
        std::vector<uint32_t> gb((size_t)1 << width), rb = host_bounds(h, off, width);
        CHECK(cudaMemcpy(gb.data(), d.bounds, gb.size() * 4, cudaMemcpyDeviceToHost));
        if (gb != rb) { printf("FAIL %s n=%zu off=%u w=%u boundaries differ\n", tag, n, off, width); return 1; }
    }
    return 0;
}

static int run_check()
{
    Dev d;
    int fails = 0, cases = 0;
    const size_t sizes[] = {1, 2, 3, 4, 5, 127, 128, 129, 255, 1021, 1111, 4099, 8191, 8192, 8193, 16385, 100003, (1u << 20) + 5, 3000001};
    const uint32_t ow[][2] = {{0, 32}, {0, 8}, {0, 16}, {4, 8}, {0, 4}, {24, 8}, {16, 16}, {8, 8}, {0, 2}, {30, 2}, {6, 10}, {0, 24}};
    for (size_t n : sizes) {
        std::vector<uint32_t> h(n);
        populateInput(h.data(), n);
        for (auto &p : ow) { fails += check_one(d, h, p[0], p[1], 0, "uniform"); cases++; }
        fails += check_one(d, h, 0, 32, 1, "misaligned"); cases++;
        fails += check_one(d, h, 0, 8, 3, "misaligned"); cases++;
        std::vector<uint32_t> s(n);
        for (size_t i = 0; i < n; i++) s[i] = 0xAB000000u | (h[i] & 0xffff);
        fails += check_one(d, s, 0, 32, 0, "consthigh"); cases++;
        fails += check_one(d, s, 16, 16, 0, "consthigh"); cases++;
        for (size_t i = 0; i < n; i++) s[i] = 0xffffffffu;
        fails += check_one(d, s, 0, 32, 0, "allones"); cases++;
        for (size_t i = 0; i < n; i++) s[i] = (uint32_t)(n - i);
        fails += check_one(d, s, 0, 32, 0, "descending"); cases++;
    }
    // multi-portion look-back carry
    libsortDebugSetPortionTiles(3);
    {
        size_t n = (size_t)libsortTileKeys() * 17 + 123;
        std::vector<uint32_t> h(n);
        populateInput(h.data(), n);
        fails += check_one(d, h, 0, 32, 0, "portions"); cases++;
        fails += check_one(d, h, 8, 16, 0, "portions"); cases++;
    }
    libsortDebugSetPortionTiles(0);
    printf("devbench check: %d cases, %d failures\n", cases, fails);
    return fails ? 1 : 0;
}

static int run_time(size_t n, int reps, const char *dist)
{
    Dev d;
    d.ensure(n);
    std::vector<uint32_t> h(n);
    populateInput(h.data(), n);
    if (!strcmp(dist, "sorted")) std::sort(h.begin(), h.end());
    else if (!strcmp(dist, "consthigh")) for (auto &k : h) k = 0xAB000000u | (k & 0xffffffu);
    else if (!strcmp(dist, "zipf")) for (auto &k : h) k = (uint32_t)(exp2(32.0 * k / 4294967296.0)) - 1u;  // log-uniform
    else if (!strcmp(dist, "allsame")) for (auto &k : h) k = 0x12345678u;
    printf("-- input: %s\n", dist);
    CHECK(cudaMemcpy(d.in, h.data(), n * 4, cudaMemcpyHostToDevice));
    struct Cfg { uint32_t off, width; const char *name; double bytes_per_key; } cfgs[] = {
        {0, 32, "full32", 36}, {0, 8, "partial8", 12}, {0, 16, "partial16", 20}};
    for (auto &c : cfgs) {
        LibsortKernelTimes t{};
        for (int w = 0; w < 3; w++)
            libsortDeviceSort(d.in, d.out, d.tmp, n, c.off, c.width, c.width <= 16 ? d.bounds : nullptr, 0, d.ws, d.wsb, 0, nullptr);
        CHECK(cudaDeviceSynchronize());
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        for (int r = 0; r < reps; r++)
            libsortDeviceSort(d.in, d.out, d.tmp, n, c.off, c.width, c.width <= 16 ? d.bounds : nullptr, 0, d.ws, d.wsb, 0, nullptr);
        cudaEventRecord(e1);
        CHECK(cudaDeviceSynchronize());
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        ms /= reps;
        libsortDeviceSort(d.in, d.out, d.tmp, n, c.off, c.width, c.width <= 16 ? d.bounds : nullptr, 0, d.ws, d.wsb, 0, &t);
        printf("%-9s n=%zu  %.3f ms  %.2f Gkeys/s  %.0f GB/s algorithmic | memset %.3f hist %.3f scan %.3f passes", c.name, n, ms,
               n / ms / 1e6, c.bytes_per_key * n / ms / 1e6, t.memset_ms, t.histogram_ms, t.scan_ms);
        for (int p = 0; p < t.npass; p++) printf(" %.3f", t.pass_ms[p]);
        printf(" (pass0 %.0f GB/s) bounds %.3f launches %d\n", 8.0 * n / t.pass_ms[0] / 1e6, t.boundaries_ms, t.launches);
        if (c.width == 32) {
            // cheap full-size check of the timed configuration: ascending + same multiset fingerprint
            std::vector<uint32_t> got(n);
            CHECK(cudaMemcpy(got.data(), d.out, n * 4, cudaMemcpyDeviceToHost));
            uint64_t s0 = 0, s1 = 0, x0 = 0, x1 = 0;
            size_t bad = 0;
            for (size_t i = 0; i < n; i++) {
                s0 += h[i]; x0 ^= (uint64_t)h[i] * 0x9E3779B97F4A7C15ull;
                s1 += got[i]; x1 ^= (uint64_t)got[i] * 0x9E3779B97F4A7C15ull;
                if (i && got[i - 1] > got[i]) bad++;
            }
            printf("verify full32: %s (descents %zu, sum %s, xor %s)\n", (!bad && s0 == s1 && x0 == x1) ? "OK" : "FAIL", bad,
                   s0 == s1 ? "ok" : "DIFF", x0 == x1 ? "ok" : "DIFF");
        }
    }
    return 0;
}

int main(int argc, char **argv)
{
    if (argc >= 2 && !strcmp(argv[1], "check")) return run_check();
    if (argc >= 3 && !strcmp(argv[1], "time")) return run_time(strtoull(argv[2], 0, 0), argc >= 4 ? atoi(argv[3]) : 10, argc >= 5 ? argv[4] : "uniform");
    fprintf(stderr, "usage: devbench check | time N [reps] [uniform|sorted|consthigh|zipf|allsame]\n");
    return 2;
}
