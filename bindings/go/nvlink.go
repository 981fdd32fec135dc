package sort

// NVLink backend of the distributed sort for the Go benchmark driver (drop this file into
// benchmark/pkg/sort/ next to libsort.go; SURVEY.md 8(f) N1).
//
// The reference sorts a large array across workers in 32/width steps, moving every bucket
// through DistribArrays between steps (SortDistribFromArr, distrib.go:90-179).  On one box
// with several B200s the whole thing is ONE library call: libsortDistribSort cuts the array
// into one shard per GPU, exchanges the keys GPU-to-GPU over NVLink in a single fused
// partition+exchange kernel and sorts each value range locally (include/libsort_b200.h).
//
// Source only: there is no Go toolchain in the build image.  The same entry point is
// exercised from C++ (gpu-radix-sort_b200/localTest/radix_sort.cpp) and Python
// (tests/distrib_host_worker.py).
//
// By the package's convention 'len' counts uint32s and 'size' counts bytes.

// #cgo CFLAGS: -O3 -I${SRCDIR}/../../../../repo/gpu-radix-sort_b200/libsort --std=gnu99
// #cgo LDFLAGS: -L${SRCDIR}/../../../../repo/gpu-radix-sort_b200/libsort -lsort
// #include "libsort.h"
// #include "libsort_b200.h"
import "C"
import (
	"errors"
	"fmt"
	"unsafe"
)

// Number of GPUs SortDistribNVLink spreads the array over; 0 = every GPU of libsort's pool.
var nvlinkGpus int = 0

func SetNVLinkGpus(n int) {
	nvlinkGpus = n
}

// How many GPUs libsort's device pool holds (0 before InitLibSort).
func NVLinkDeviceCount() int {
	return (int)(C.libsortDeviceCount())
}

// Sort a native byte array of uint32s across the GPUs of this box. Same contract as
// SortDistribFromRaw (distrib.go:183-248) minus the array factory and the worker: the input
// is left untouched and the sorted copy is returned.
func SortDistribNVLink(inRaw []byte) ([]byte, error) {
	if err := InitLibSort(); err != nil {
		return nil, fmt.Errorf("Failed to initialize libsort: %v", err)
	}
	if len(inRaw)%4 != 0 {
		return nil, errors.New("input size must be a multiple of 4 bytes")
	}

	outRaw := make([]byte, len(inRaw))
	copy(outRaw, inRaw)
	if len(outRaw) == 0 {
		return outRaw, nil
	}

	if err := SortDistribNVLinkInPlace(outRaw); err != nil {
		return nil, err
	}
	return outRaw, nil
}

// In-place variant for callers that own the buffer (no copy; pinned memory is DMA'd directly).
func SortDistribNVLinkInPlace(raw []byte) error {
	if len(raw) == 0 {
		return nil
	}
	cints := (*C.uint32_t)(unsafe.Pointer(&raw[0]))
	ok := C.libsortDistribSort(cints, (C.size_t)(len(raw)/4), (C.int)(nvlinkGpus))
	if ok == 0 {
		return errors.New("libsort libsortDistribSort failed\n")
	}
	return nil
}
