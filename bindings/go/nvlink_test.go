package sort

// Tests for nvlink.go, in the style of libsort_test.go (TestLocal) and distrib_test.go
// (SortDistribTest): PCG inputs from the library, result checked with CheckSort.

import (
	"testing"

	"github.com/stretchr/testify/require"
)

func testSortDistribNVLink(t *testing.T, nElem uint64) {
	err := InitLibSort()
	require.Nil(t, err, "Failed to initialize libsort")

	origRaw, err := GenerateInputs(nElem)
	require.Nil(t, err, "Failed to generate inputs")

	outRaw, err := SortDistribNVLink(origRaw)
	require.Nilf(t, err, "Failed to sort: %v", err)

	err = CheckSort(origRaw, outRaw)
	require.Nilf(t, err, "Did not sort correctly: %v", err)
}

func TestSortNVLinkSmall(t *testing.T) {
	testSortDistribNVLink(t, 1111)
}

func TestSortNVLinkRagged(t *testing.T) {
	// not a multiple of the GPU count, the tile size or the 16-byte shard alignment
	testSortDistribNVLink(t, (1<<22)+12345)
}

func TestSortNVLinkEmpty(t *testing.T) {
	out, err := SortDistribNVLink([]byte{})
	require.Nil(t, err)
	require.Equal(t, 0, len(out))
}

func TestSortNVLinkAllEqual(t *testing.T) {
	// every key the same: the library cuts the ties between the GPUs
	raw := make([]byte, 4*(1<<20))
	for i := range raw {
		raw[i] = 0x5a
	}
	ref := make([]byte, len(raw))
	copy(ref, raw)
	require.Nil(t, SortDistribNVLinkInPlace(raw))
	require.Nil(t, CheckSort(ref, raw))
}
