"""The reference's distributed sort over file arrays with the real worker (faas.f -> gpuPartial on the GPU).
Kept in a file that sorts last: it is the widest end-to-end path (files, requests, threads, GPU), run after
the parity suites.  The orchestration itself is covered on CPU in test_distrib_array.py."""
import importlib

import numpy as np
import pytest


@pytest.fixture()
def da(pkg):
    return importlib.import_module(pkg.__name__ + ".distrib_array")


@pytest.fixture()
def dfiles(pkg):
    return importlib.import_module(pkg.__name__ + ".distrib_files")


@pytest.mark.gpu
@pytest.mark.parametrize("nworker,width", [(2, 8), (3, 16)])
def test_sort_distrib_over_files_gpu_worker(pkg, da, dfiles, tmp_path, nworker, width):
    """The same path with the real worker (faas.f -> gpuPartial), workers on concurrent threads."""
    da.SetDistribMount(tmp_path)
    keys = np.random.default_rng(3).integers(0, 1 << 32, 200003, dtype=np.uint64).astype(np.uint32)
    out = dfiles.SortDistribFromRaw(keys.tobytes(), "g", nworker=nworker, width=width)
    assert np.array_equal(np.frombuffer(bytes(out), np.uint32), np.sort(keys))
