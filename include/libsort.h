/*
 * libsort.h -- the drop-in C ABI of the B200-native radix-sort library.
 *
 * These seven symbols are exactly the ones the reference's libsort.so exports
 * (reference libsort/libsort.h:12-47) with the same argument meaning, units
 * and error behaviour, so the reference's three consumers bind unchanged:
 *   - C++  localTest/local.h:8            (#include "../libsort/libsort.h", -lsort)
 *   - Go   benchmark/pkg/sort/libsort.go:11-14   (cgo, --std=gnu99)
 *   - Py   faasTest/pylibsort/__init__.py:13-20  (ctypes, LD_LIBRARY_PATH)
 * The header compiles as C99 and as C++14 and includes only the three
 * freestanding headers the reference's does.
 *
 * Units: lengths are in ELEMENTS (uint32 keys); boundaries are ELEMENT
 * indices (callers multiply by 4 themselves: libsort.go:58-60).
 * Ownership: every buffer is caller-owned ordinary host memory (pageable or
 * pinned); keys are sorted in place; no pointer is retained after return.
 * Threading: every entry point except populateInput may be called from many
 * host threads at once; each call owns one GPU of the pool for its duration
 * and blocks while all GPUs are busy (reference libsort/utils.cu:42-61).
 * There is no CPU fallback: without a usable GPU the GPU entry points fail.
 */
#ifndef LIBSORT_LIBSORT_H
#define LIBSORT_LIBSORT_H
#include <stdint.h>
#include <stdbool.h>
#include <stddef.h>

#ifdef __cplusplus
#define LIBSORT_API extern "C"
#else
#define LIBSORT_API
#endif

/* Call exactly once per process before any GPU entry point.  A second call
 * prints a message and returns false; returns false if no CUDA device can be
 * enumerated.  (reference libsort/utils.cu:10-32) */
LIBSORT_API bool initLibSort(void);

/* Stable partial sort of h_in by bits [offset, offset+width) and the start
 * index of every group value in `boundaries` (caller-allocated, 1<<width
 * entries, fully overwritten).  offset and width are multiples of 2
 * (reference libsort/sort.h:14-17).  h_in_len > UINT32_MAX fails with a
 * message on stderr.  (reference libsort/invokers.cu:15-41)
 * Boundaries reproduce the reference's output bit for bit, including its two
 * empty-leading-group deviations from the exclusive-prefix-sum contract
 * (see DESIGN.md "boundary semantics"); set LIBSORT_BOUNDARIES=exact to get
 * the pure exclusive prefix sums instead. */
LIBSORT_API bool gpuPartial(uint32_t *h_in, uint32_t *boundaries, size_t h_in_len, uint32_t offset,
                            uint32_t width);

/* Full ascending sort of h_in in place on a GPU.
 * (reference libsort/invokers.cu:45-64) */
LIBSORT_API bool providedGpu(unsigned int *h_in, size_t len);

/* The reference's CPU local sort: single-threaded std::sort; always true.
 * Kept because localTest compares against it; it is never used as a
 * fallback for the GPU path.  (reference libsort/invokers.cu:68-71) */
LIBSORT_API bool providedCpu(unsigned int *in, size_t len);

/* Deterministic PCG-XSH-RR generator whose state persists across calls
 * within a process; not thread safe.  (reference libsort/utils.cu:63-79) */
LIBSORT_API void populateInput(uint32_t *arr, size_t nelem);

/* The same two sorts bracketed by cudaProfilerStart/Stop so that
 * `ncu --profile-from-start off` captures only the sort.
 * (reference libsort/invokers.cu:73-85) */
LIBSORT_API bool gpuPartialProfile(uint32_t *h_in, uint32_t *boundaries, size_t h_in_len,
                                   uint32_t offset, uint32_t width);
LIBSORT_API bool providedGpuProfile(unsigned int *h_in, size_t h_in_len);

#endif /* LIBSORT_LIBSORT_H */
