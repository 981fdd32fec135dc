// radix_sort.cpp -- C++ consumer of the drop-in ABI, the counterpart of the
// reference's localTest/radix_sort (localTest/main.cpp, tests.cpp, benchmarks.cpp).
// It binds to libsort.so exactly as the reference harness does (libsort.h, -lsort)
// and runs BASELINE.json configs[0]: a 1M-key full sort checked against the
// library's CPU local sort, plus the reference's own test plan
// (tests.cpp:88-161): partial sort boundaries == CPU prefix sums, GPU == CPU
// element for element, two-partition distributed sort == single GPU.
//   radix_sort [nelem]        tests at nelem keys (default 1<<20)
//   radix_sort bench [nelem]  timings of full / 8-bit / 16-bit sorts (default 1<<28)
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <vector>

#include "libsort.h"
#include "libsort_b200.h"  // additive: libsortDistribSort (every GPU of the box, exchange over NVLink)

static inline uint32_t group_of(uint32_t v, int pos, int width) { return (v >> pos) & ((1u << width) - 1); }

static bool ascending(const std::vector<uint32_t> &a)
{
    for (size_t i = 1; i < a.size(); i++)
        if (a[i] < a[i - 1]) return false;
    return true;
}

// reference tests.cpp:38-86: boundaries are the exclusive prefix sums of the
// group histogram and every key sits in its bucket
static bool check_partial(std::vector<uint32_t> data, int offset, int width)
{
    const size_t nb = (size_t)1 << width;
    std::vector<uint32_t> counts(nb, 0), want(nb, 0), got(nb, 0);
    for (uint32_t k : data) counts[group_of(k, offset, width)]++;
    for (size_t g = 1; g < nb; g++) want[g] = want[g - 1] + counts[g - 1];
    if (!gpuPartial(data.data(), got.data(), data.size(), offset, width)) return false;
    for (size_t g = 0; g < nb; g++)
        if (want[g] != got[g]) {
            fprintf(stderr, "boundary %zu: expected %u got %u\n", g, want[g], got[g]);
            return false;
        }
    size_t g = 0;
    for (size_t i = 0; i < data.size(); i++) {
        while (g + 1 < nb && i >= want[g + 1]) g++;
        if (group_of(data[i], offset, width) != g) {
            fprintf(stderr, "data[%zu] in wrong bucket\n", i);
            return false;
        }
    }
    return true;
}

// reference benchmarks.cpp:70-160: 4 LSD steps of 8 bits over two halves sorted
// concurrently (one GPU each through the device pool), bucket-major regather
static bool distrib_sort(std::vector<uint32_t> &data, double *worker_ms, double *shuffle_ms)
{
    const int width = 8, nb = 1 << width;
    std::vector<uint32_t> other(data.size()), b1(nb), b2(nb);
    uint32_t *cur = data.data(), *next = other.data();
    const size_t len1 = data.size() / 2, len2 = data.size() - len1;
    for (int step = 0; step < 32 / width; step++) {
        auto t0 = std::chrono::steady_clock::now();
        auto f1 = std::async(std::launch::async, gpuPartial, cur, b1.data(), len1, step * width, width);
        auto f2 = std::async(std::launch::async, gpuPartial, cur + len1, b2.data(), len2, step * width, width);
        const bool ok1 = f1.get(), ok2 = f2.get();
        if (!ok1 || !ok2) return false;
        auto t1 = std::chrono::steady_clock::now();
        size_t slot = 0;
        for (int b = 0; b < nb; b++) {
            const size_t n1 = (b == nb - 1 ? len1 : b1[b + 1]) - b1[b], n2 = (b == nb - 1 ? len2 : b2[b + 1]) - b2[b];
            memcpy(next + slot, cur + b1[b], n1 * 4);
            slot += n1;
            memcpy(next + slot, cur + len1 + b2[b], n2 * 4);
            slot += n2;
        }
        auto t2 = std::chrono::steady_clock::now();
        if (worker_ms) *worker_ms += std::chrono::duration<double, std::milli>(t1 - t0).count();
        if (shuffle_ms) *shuffle_ms += std::chrono::duration<double, std::milli>(t2 - t1).count();
        std::swap(cur, next);
    }
    if (cur != data.data()) memcpy(data.data(), cur, data.size() * 4);
    return true;
}

static int run_tests(size_t n)
{
    std::vector<uint32_t> in(n);
    populateInput(in.data(), n);
    if (!check_partial(in, 0, 8)) { fprintf(stderr, "GPU partial test failed\n"); return 1; }
    std::vector<uint32_t> gpu = in, cpu = in, dis = in, nvl = in;
    if (!providedGpu(gpu.data(), n) || !ascending(gpu)) { fprintf(stderr, "GPU sort failed\n"); return 1; }
    if (!providedCpu(cpu.data(), n) || !ascending(cpu)) { fprintf(stderr, "CPU sort failed\n"); return 1; }
    if (!distrib_sort(dis, nullptr, nullptr) || !ascending(dis)) { fprintf(stderr, "distributed sort failed\n"); return 1; }
    if (gpu != cpu) { fprintf(stderr, "CPU and GPU results disagree\n"); return 1; }
    if (gpu != dis) { fprintf(stderr, "distributed and single-GPU results disagree\n"); return 1; }
    if (!libsortDistribSort(nvl.data(), n, 0)) { fprintf(stderr, "NVLink distributed sort failed\n"); return 1; }
    if (gpu != nvl) { fprintf(stderr, "NVLink distributed and single-GPU results disagree\n"); return 1; }
    printf("Success! (%zu keys: gpuPartial boundaries, providedGpu == providedCpu == distribSort == libsortDistribSort)\n", n);
    return 0;
}

template <class F> static double ms_of(F f)
{
    auto t0 = std::chrono::steady_clock::now();
    f();
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

static int run_bench(size_t n)
{
    std::vector<uint32_t> orig(n), work(n), bounds(1 << 16);
    populateInput(orig.data(), n);
    for (int rep = 0; rep < 2; rep++) {  // first repetition warms the per-device caches
        work = orig;
        double full = ms_of([&] { providedGpu(work.data(), n); });
        work = orig;
        double p8 = ms_of([&] { gpuPartial(work.data(), bounds.data(), n, 0, 8); });
        work = orig;
        double p16 = ms_of([&] { gpuPartialProfile(work.data(), bounds.data(), n, 0, 16); });
        work = orig;
        double wk = 0, sh = 0, dt = ms_of([&] { distrib_sort(work, &wk, &sh); });
        work = orig;
        double nv = ms_of([&] { libsortDistribSort(work.data(), n, 0); });
        if (rep) printf("libsortDistribSort (all GPUs, NVLink exchange): %.2f ms\n", nv);
        if (rep)
            printf("%.1f MiB through the host-pointer ABI (pageable memory):\n\tfull sort %.2f ms\n\tpartial 8b %.2f ms\n"
                   "\tpartial 16b %.2f ms\n\tdistribSort (2 partitions, 4x8b) %.2f ms (workers %.2f, shuffle %.2f)\n",
                   n * 4.0 / (1 << 20), full, p8, p16, dt, wk, sh);
    }
    size_t cpu_n = n < ((size_t)1 << 24) ? n : ((size_t)1 << 24);
    work.assign(orig.begin(), orig.begin() + cpu_n);
    double c = ms_of([&] { providedCpu(work.data(), cpu_n); });
    printf("\tprovidedCpu (std::sort, 1 thread) on %zu keys: %.2f ms\n", cpu_n, c);
    return 0;
}

int main(int argc, char **argv)
{
    if (!initLibSort()) { fprintf(stderr, "Failed to initialize libsort\n"); return EXIT_FAILURE; }
    if (argc > 1 && !strcmp(argv[1], "bench")) return run_bench(argc > 2 ? strtoull(argv[2], 0, 0) : (size_t)1 << 28);
    return run_tests(argc > 1 ? strtoull(argv[1], 0, 0) : (size_t)1 << 20);
}
