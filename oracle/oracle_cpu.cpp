// oracle_cpu.cpp -- TEST INFRASTRUCTURE ONLY (see oracle.c header).
// Restatement of the reference's CPU local sort, providedCpu
// (libsort/invokers.cu:68-71): a single-threaded std::sort over the keys.
// Used as the "port" CPU baseline when oracle/_ref is not available.
#include <algorithm>
#include <cstddef>
#include <cstdint>

extern "C" int oracle_cpu_sort(uint32_t *keys, size_t n)
{
    std::sort(keys, keys + n);
    return 0;
}
