/*
 * libsort_b200.h -- additive C ABI of the B200-native libsort.so.
 *
 * Nothing here exists in the reference; these entry points expose the same
 * sort on DEVICE-resident keys (for roofline measurement, the PyTorch-facing
 * wrapper and the multi-GPU sort, SURVEY.md 8(b) "Additive") and never
 * collide with the seven legacy names in libsort.h.  Plain pointers and
 * sizes only; `stream` is a cudaStream_t passed as void*.  Unless stated
 * otherwise a function returns 0 on success, a positive cudaError_t value on
 * a CUDA failure, or a negative LIBSORT_E* code; nothing synchronises the
 * stream unless stated.
 */
#ifndef LIBSORT_B200_H
#define LIBSORT_B200_H
#include "libsort.h"

#define LIBSORT_EINVAL (-1)   /* bad argument */
#define LIBSORT_ENOINIT (-2)  /* initLibSort has not succeeded */
#define LIBSORT_ENOMEM (-3)

/* boundary_mode */
#define LIBSORT_BOUNDS_REFERENCE 0 /* bit-exact with reference libsort/sort.cu:14-27,367-394 */
#define LIBSORT_BOUNDS_EXACT 1     /* exclusive prefix sums (localTest/tests.cpp:44-54) */

typedef struct LibsortKernelTimes {
    float memset_ms;     /* workspace clear */
    float histogram_ms;  /* K1: upsweep histogram, all digits */
    float scan_ms;       /* bucket-base scan */
    float pass_ms[4];    /* K3: one scatter pass each */
    float boundaries_ms;
    int npass;
    int launches;        /* kernels launched by the call */
} LibsortKernelTimes;

/* Keys per scatter tile of the built kernel (informational). */
LIBSORT_API int libsortTileKeys(void);

/* Device scratch bytes libsortDeviceSort needs for n keys. */
LIBSORT_API size_t libsortWorkspaceBytes(size_t n);

/*
 * Device-resident equivalent of SortState::Step (+ GetBoundaries)
 * (reference libsort/sort.cu:322-394): stable sort of d_in[0..n) by bits
 * [offset, offset+width) into d_out, d_tmp (n keys) as the ping-pong buffer.
 * d_in may equal d_out (in place) and is otherwise left untouched.
 * d_boundaries: NULL or 1<<width device words.  times: NULL, or receives the
 * per-kernel CUDA-event durations (this variant synchronises the stream).
 * Runs on the calling thread's current device; does not use the device pool.
 */
LIBSORT_API int libsortDeviceSort(const uint32_t *d_in, uint32_t *d_out, uint32_t *d_tmp, size_t n, uint32_t offset,
                                  uint32_t width, uint32_t *d_boundaries, int boundary_mode, void *d_workspace,
                                  size_t workspace_bytes, void *stream, LibsortKernelTimes *times);

/* Number of GPUs in the pool (0 before initLibSort). */
LIBSORT_API int libsortDeviceCount(void);

/* Free the per-device buffers the host-pointer entry points cache. */
LIBSORT_API void libsortReleaseCaches(void);

/* Message for a code returned by this header's functions. */
LIBSORT_API const char *libsortErrorString(int code);

/* Test hook: cap tiles per look-back portion (0 restores the default). */
LIBSORT_API void libsortDebugSetPortionTiles(uint32_t tiles);

#endif /* LIBSORT_B200_H */
