"""File distributed arrays and the sort-request plumbing.  The CPU tests follow the
reference's own plan for this layer (faasTest/testPyLibSort.py:35-199: partitions, whole-array I/O,
reopen/destroy, partRefs, request parsing) plus the on-disk format itself; the GPU test runs the FaaS
worker end to end (faasTest/f.py selfTest)."""
import importlib
import json
import os
import random

import numpy as np
import pytest


@pytest.fixture()
def da(pkg):
    return importlib.import_module(pkg.__name__ + ".distrib_array")


def fill(arr, szs=None):
    szs = arr.shape.caps if szs is None else szs
    bufs = []
    for p in range(arr.shape.npart):
        b = bytes(random.getrandbits(8) for _ in range(szs[p]))
        arr.WritePart(p, b)
        bufs.append(b)
    return bufs


def test_shape(da):
    s = da.ArrayShape.fromUniform(40, 3)
    assert (s.caps, s.lens, s.npart, s.starts) == ([40, 40, 40], [0, 0, 0], 3, [0, 40, 80, 120])
    s = da.ArrayShape.fromCaps([8, 0, 12])
    assert s.starts == [0, 8, 8, 20] and s.lens == [0, 0, 0]


def test_partitions_and_whole_array(da, tmp_path):
    shape = da.ArrayShape.fromUniform(40, 2)
    arr = da.fileDistribArray.Create(tmp_path / "partTest0", shape)
    first = fill(arr, [20, 20])
    assert arr.ReadPart(0) == first[0]
    assert arr.ReadPart(0, start=4, nbyte=4) == first[0][4:8]
    more = fill(arr, [20, 20])                                     # appends
    assert arr.ReadPart(0) == first[0] + more[0]
    whole = first[0] + more[0] + first[1] + more[1]
    assert arr.ReadAll() == whole
    with pytest.raises(da.DistribArrayError, match="Wrote beyond end of partition"):
        arr.WritePart(0, b"x")
    with pytest.raises(da.DistribArrayError, match="Read beyond end of partition"):
        arr.ReadPart(1, start=30, nbyte=20)
    other = da.fileDistribArray.Create(tmp_path / "partTest1", shape)
    other.WriteAll(whole)
    assert other.ReadAll() == whole and other.shape.lens == [40, 40]
    with pytest.raises(da.DistribArrayError, match="does not match array capacity"):
        other.WriteAll(whole[:-1])
    dest = bytearray(12)                                            # read into a caller's buffer
    assert other.ReadPart(1, start=8, nbyte=12, dest=dest) is None and dest == whole[48:60]


def test_on_disk_format(da, tmp_path):
    """meta.json = {"Lens", "Caps"} in bytes; data.dat = partitions at their capacity offsets (file.go:36-41)."""
    arr = da.fileDistribArray.Create(tmp_path / "fmt", da.ArrayShape.fromCaps([16, 8, 24]))
    arr.WritePart(0, b"A" * 10)
    arr.WritePart(2, b"C" * 24)
    arr.WritePart(1, b"B" * 3)
    arr.Close()
    arr.Close()                                                     # idempotent
    assert json.load(open(tmp_path / "fmt" / "meta.json")) == {"Lens": [10, 3, 24], "Caps": [16, 8, 24]}
    raw = open(tmp_path / "fmt" / "data.dat", "rb").read()
    assert len(raw) == 48 and raw[:10] == b"A" * 10 and raw[16:19] == b"B" * 3 and raw[24:] == b"C" * 24
    # an array written by "the other side" (hand-made files) opens here; a short data file reads as zeros
    os.mkdir(tmp_path / "foreign")
    json.dump({"Lens": [4, 4], "Caps": [8, 8]}, open(tmp_path / "foreign" / "meta.json", "w"))
    open(tmp_path / "foreign" / "data.dat", "wb").write(b"abcd\0\0\0\0wxyz")
    f = da.fileDistribArray.Open(tmp_path / "foreign")
    assert f.ReadPart(0) == b"abcd" and f.ReadPart(1) == b"wxyz" and bytes(f.ReadAll()) == b"abcd\0\0\0\0wxyz\0\0\0\0"


def test_reopen_and_destroy(da, tmp_path):
    shape = da.ArrayShape.fromUniform(10, 2)
    arr = da.fileDistribArray.Create(tmp_path / "distribArrayTest", shape)
    assert arr.shape.npart == 2
    bufs = fill(arr)
    arr.Close()
    again = da.fileDistribArray.Open(tmp_path / "distribArrayTest")
    assert again.shape.caps == shape.caps and [again.ReadPart(p) for p in range(2)] == bufs
    with pytest.raises(FileExistsError):
        da.fileDistribArray.Create(tmp_path / "distribArrayTest", shape)
    again.Destroy()
    with pytest.raises(da.DistribArrayError, match="does not exist"):
        da.fileDistribArray.Open(tmp_path / "distribArrayTest")


def test_part_refs_and_requests(da, tmp_path):
    da.SetDistribMount(tmp_path)
    arr = da.fileDistribArray.Create(tmp_path / "distribArrayTest", da.ArrayShape.fromUniform(10, 2))
    bufs = fill(arr)
    assert da.partRef(arr, partID=0, start=0, nbyte=5).read() == bufs[0][:5]
    assert da.partRef(arr, partID=1, start=2, nbyte=6).read() == bufs[1][2:8]
    arr.Close()
    req = {"offset": 0, "width": 2, "arrType": "file",
           "input": [{"arrayName": "distribArrayTest", "partID": 0, "start": 0, "nbyte": 5},
                     {"arrayName": "distribArrayTest", "partID": 1, "start": 2, "nbyte": 6},
                     {"arrayName": "distribArrayTest", "partID": 1, "start": 4, "nbyte": -1}]}
    refs = da.getPartRefs(req)
    assert [r.read() for r in refs] == [bufs[0][:5], bufs[1][2:8], bufs[1][4:]]
    assert bytes(da.readPartRefs(refs)) == bufs[0][:5] + bufs[1][2:8] + bufs[1][4:]
    pinned_like = bytearray(32)
    view = da.readPartRefs(refs, dest=pinned_like)
    assert len(view) == 17 and bytes(pinned_like[:17]) == bytes(view)
    with pytest.raises(ValueError, match="Invalid request type"):
        da.getPartRefs(dict(req, arrType="redis"))
    da.closeOpenArrays()


def test_request_cache_sees_recreated_arrays_and_is_thread_safe(da, tmp_path):
    """getPartRefs caches one open handle per array name for the life of a worker: a name that is
    destroyed and created again (new inode, new lengths) must not be served from the stale handle,
    and concurrent requests share one handle instead of leaking descriptors."""
    from concurrent.futures import ThreadPoolExecutor
    da.SetDistribMount(tmp_path)
    a = da.fileDistribArray.Create(tmp_path / "reused", da.ArrayShape.fromUniform(8, 1))
    a.WritePart(0, b"AAAAAAAA")
    a.Close()
    req = {"offset": 0, "width": 2, "arrType": "file", "input": [{"arrayName": "reused", "partID": 0, "start": 0, "nbyte": -1}]}
    assert da.getPartRefs(req)[0].read() == b"AAAAAAAA"
    first = da.openArrs["reused"]
    da.fileDistribArray.Open(tmp_path / "reused").Destroy()
    b = da.fileDistribArray.Create(tmp_path / "reused", da.ArrayShape.fromUniform(12, 1))
    b.WritePart(0, b"BBBBBBBBBBBB")
    b.Close()
    assert da.getPartRefs(req)[0].read() == b"BBBBBBBBBBBB"        # not the unlinked inode, not the old length
    assert da.openArrs["reused"] is not first and first.closed
    da.closeOpenArrays()
    with ThreadPoolExecutor(8) as ex:
        got = list(ex.map(lambda _: (da.getPartRefs(req)[0].read(), id(da.openArrs["reused"])), range(64)))
    assert all(g[0] == b"BBBBBBBBBBBB" for g in got) and len({g[1] for g in got}) == 1
    da.closeOpenArrays()


def test_write_output_buckets(da, tmp_path):
    """One output partition per bucket, sized by the boundaries (element indices -> bytes; data.py:289-298)."""
    da.SetDistribMount(tmp_path)
    keys = np.array([0, 0, 1, 3, 3, 3], dtype=np.uint32)            # sorted by a 2-bit group; group 2 is empty
    da.writeOutput({"arrType": "file", "output": "out"}, keys.tobytes(), [0, 2, 3, 3])
    out = da.fileDistribArray.Open(tmp_path / "out")
    assert out.shape.caps == [8, 4, 0, 12] and out.shape.lens == out.shape.caps
    assert [np.frombuffer(bytes(out.ReadPart(p)), np.uint32).tolist() for p in range(4)] == [[0, 0], [1], [], [3, 3, 3]]
    out.Destroy()


@pytest.mark.gpu
def test_faas_worker_end_to_end(pkg, tmp_path):
    """The reference's f.py selfTest flow: arrays on disk -> JSON request -> GPU partial sort -> bucket
    partitions on disk, checked with checkPartial; then the error replies of the protocol."""
    faas = importlib.import_module(pkg.__name__ + ".faas")
    for n, off, w in ((1 << 20, 0, 8), (4096, 8, 4), (1 << 18, 16, 16)):
        assert faas.selfTest(n, offset=off, width=w, mount=str(tmp_path)) == {"success": True, "err": ""}
    assert faas.f({"arrType": "redis"})["success"] is False
    pls = importlib.import_module(pkg.__name__ + ".pylibsort")
    pls.SetDistribMount(tmp_path)
    r = faas.f({"offset": 0, "width": 8, "arrType": "file", "output": "o",
                "input": [{"arrayName": "missing", "partID": 0, "start": 0, "nbyte": -1}]})
    assert r["success"] is False and "does not exist" in r["err"]


def _invoke(pkg, stdin, env_extra):
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(pkg.__file__)))
    env = dict(os.environ)
    env.pop("OL_SHARED_VOLUME", None)
    env.update(env_extra)
    return subprocess.run([sys.executable, "-m", pkg.__name__ + ".faas"], input=stdin, capture_output=True, text=True,
                          cwd=root, env=env, timeout=300)


def test_direct_invoke_argument_errors(pkg, tmp_path):
    """directInvoke() replies in JSON and exits 1 on a missing mount or a malformed request (f.py:162-174)."""
    r = _invoke(pkg, "{}", {})
    assert r.returncode == 1 and "OL_SHARED_VOLUME not set" in json.loads(r.stdout.strip().splitlines()[-1])["err"]
    r = _invoke(pkg, "not json", {"OL_SHARED_VOLUME": str(tmp_path)})
    assert r.returncode == 1 and "Argument parsing error" in json.loads(r.stdout.strip().splitlines()[-1])["err"]


@pytest.mark.gpu
def test_direct_invoke_sorts_request(pkg, da, tmp_path):
    """The worker as a process: request on stdin, arrays under OL_SHARED_VOLUME, reply on stdout, exit 0."""
    keys = np.random.default_rng(5).integers(0, 1 << 32, 50000, dtype=np.uint64).astype(np.uint32)
    arr = da.fileDistribArray.Create(tmp_path / "in0", da.ArrayShape.fromUniform(100000, 2))
    arr.WriteAll(keys.tobytes())
    arr.Close()
    req = {"offset": 8, "width": 8, "arrType": "file", "output": "out0",
           "input": [{"arrayName": "in0", "partID": 1, "start": 0, "nbyte": -1},
                     {"arrayName": "in0", "partID": 0, "start": 40, "nbyte": 60000}]}
    r = _invoke(pkg, json.dumps(req), {"OL_SHARED_VOLUME": str(tmp_path)})
    assert r.returncode == 0 and json.loads(r.stdout.strip().splitlines()[-1]) == {"success": True, "err": ""}, r.stdout + r.stderr
    fed = np.concatenate([keys[25000:], keys[10:15010]])
    out = da.fileDistribArray.Open(tmp_path / "out0")
    got = np.frombuffer(bytes(out.ReadAll()), np.uint32)
    order = np.argsort((fed >> 8) & 0xFF, kind="stable")
    assert np.array_equal(got, fed[order])                            # stable partial sort of exactly the requested slices
    assert out.shape.caps == (np.bincount((fed >> 8) & 0xFF, minlength=256) * 4).tolist()


# ---- the reference's distributed sort over file arrays (benchmark/pkg/sort: BucketReader, SortDistribFromArr) ----
@pytest.fixture()
def dfiles(pkg):
    return importlib.import_module(pkg.__name__ + ".distrib_files")


def numpy_worker(da):
    """Stand-in for faas.f on machines without a GPU: the same request handling, numpy as the sorter."""
    def work(req):
        try:
            raw = da.readPartRefs(da.getPartRefs(req))
            keys = np.frombuffer(bytes(raw), np.uint32)
            digit = (keys >> np.uint32(req["offset"])) & np.uint32((1 << req["width"]) - 1)
            counts = np.bincount(digit, minlength=1 << req["width"])
            da.writeOutput(req, keys[np.argsort(digit, kind="stable")].tobytes(), (np.cumsum(counts) - counts).tolist())
        except Exception as e:  # noqa: BLE001
            return {"success": False, "err": str(e)}
        return {"success": True, "err": ""}
    return work


def test_bucket_reader_orders(da, dfiles, tmp_path):
    """helpers.go:30-140 / distrib_test.go:54-149: STRIDED = bucket-major, INORDER = array-major; refs cut at any byte."""
    arrs, bufs = [], []
    for a in range(3):
        arr = da.fileDistribArray.Create(tmp_path / ("a%d" % a), da.ArrayShape.fromCaps([8, 0, 12, 4]))
        parts = [bytes([16 * a + p] * n) for p, n in enumerate([8, 0, 10, 4])]      # partition 2 only partly filled
        for p, b in enumerate(parts):
            arr.WritePart(p, b)
        arrs.append(arr)
        bufs.append(parts)
    strided = b"".join(bufs[a][p] for p in range(4) for a in range(3))
    inorder = b"".join(bufs[a][p] for a in range(3) for p in range(4))
    assert bytes(dfiles.BucketReader(arrs, dfiles.STRIDED).read()) == strided
    assert bytes(dfiles.BucketReader(arrs, dfiles.INORDER).read()) == inorder
    r = dfiles.BucketReader(arrs, dfiles.STRIDED)
    got, sizes = b"", []
    while True:
        refs = r.readRef(7)                                       # slices that straddle partitions and arrays
        if not refs:
            break
        sizes.append(sum(x.nbyte for x in refs))
        got += bytes(da.readPartRefs(refs))
    assert got == strided and all(s == 7 for s in sizes[:-1]) and sizes[-1] == len(strided) - 7 * (len(sizes) - 1)
    assert dfiles.BucketReader([], dfiles.STRIDED).readRef(10) == []
    with pytest.raises(ValueError):
        dfiles.BucketReader(arrs, 7)


@pytest.mark.parametrize("nworker,width,n", [(2, 8, 5000), (3, 8, 4099), (2, 16, 3000), (1, 4, 257)])
def test_sort_distrib_over_files_numpy_worker(da, dfiles, tmp_path, nworker, width, n):
    """SortDistribFromRaw (distrib.go:183-248) end to end with the numpy stand-in worker: the orchestration,
    the request format and the file exchange; nothing is left behind under the mount."""
    da.SetDistribMount(tmp_path)
    keys = np.random.default_rng(n).integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    keys[::7] = keys[0]
    out = dfiles.SortDistribFromRaw(keys.tobytes(), "t", nworker=nworker, width=width, worker=numpy_worker(da))
    assert np.array_equal(np.frombuffer(bytes(out), np.uint32), np.sort(keys))
    assert list(tmp_path.iterdir()) == []


def test_sort_distrib_worker_failure_is_reported(da, dfiles, tmp_path):
    da.SetDistribMount(tmp_path)
    with pytest.raises(RuntimeError, match="Worker failure on step 0, worker 1: boom"):
        dfiles.SortDistribFromRaw(np.arange(100, dtype=np.uint32).tobytes(), "f", nworker=2,
                                  worker=lambda req: {"success": not req["output"].endswith("worker1_output"), "err": "boom"},
                                  parallel=False)


def test_exact_boundaries_from_reference_quirks(da, oracle):
    """exactBoundaries() turns what gpuPartial returns (the reference's gpu_groups + back-fill, quirks
    included: oracle.gpu_partial) into the contract's boundaries (oracle.boundaries_exact) for every
    skew that triggers a quirk -- lowest groups empty, group 1 empty, single group, single key."""
    x = oracle.pcg_stream(30000)
    cases = {"uniform": x, "nogroup0": x | np.uint32(0x01010101), "nogroup1": x & ~np.uint32(0x01010101),
             "zeros": np.zeros(500, np.uint32), "ones": np.full(777, 0xFFFFFFFF, np.uint32),
             "few": (x % np.uint32(3)) * np.uint32(0x00010001), "sparse": x[:40], "high": x | np.uint32(0xF0F0F0F0),
             "one": np.array([0x00020000], np.uint32)}
    quirky = 0
    for name, k in cases.items():
        k = np.ascontiguousarray(k, dtype=np.uint32)
        for off, w in ((0, 8), (8, 8), (16, 16), (0, 16), (4, 2), (24, 8), (3, 12), (30, 2)):
            out, b = oracle.gpu_partial(k, off, w)
            want = oracle.boundaries_exact(k, off, w).tolist()
            quirky += b.tolist() != want
            assert da.exactBoundaries(out.tobytes(), b.tolist(), off, w) == want, (name, off, w)
    assert quirky > 10                                             # the inputs really exercise the deviations
    assert da.exactBoundaries(b"", [0, 0, 0, 0], 0, 2) == [0, 0, 0, 0]


def test_sort_distrib_over_files_survives_boundary_quirks(da, dfiles, oracle, tmp_path):
    """With 16-bit steps most groups are empty; partitions cut at the reference's quirky boundaries put
    keys into the wrong bucket and the regather loses stability (this input sorts wrong that way).  The
    worker's repair (exactBoundaries, as faas.f applies it) restores the contract."""
    da.SetDistribMount(tmp_path)
    keys = np.random.default_rng(3).integers(0, 1 << 32, 200003, dtype=np.uint64).astype(np.uint32)

    def worker(repair):
        def work(req):
            raw = da.readPartRefs(da.getPartRefs(req))
            out, b = oracle.gpu_partial(np.frombuffer(bytes(raw), np.uint32), req["offset"], req["width"])
            b = da.exactBoundaries(out.tobytes(), b.tolist(), req["offset"], req["width"]) if repair else b.tolist()
            da.writeOutput(req, out.tobytes(), b)
            return {"success": True, "err": ""}
        return work
    good = dfiles.SortDistribFromRaw(keys.tobytes(), "r", nworker=3, width=16, worker=worker(True), parallel=False)
    assert np.array_equal(np.frombuffer(bytes(good), np.uint32), np.sort(keys))
    bad = dfiles.SortDistribFromRaw(keys.tobytes(), "q", nworker=3, width=16, worker=worker(False), parallel=False)
    assert not np.array_equal(np.frombuffer(bytes(bad), np.uint32), np.sort(keys))     # what the quirk costs


def test_golden_arrays_written_by_the_reference(da, tmp_path):
    """tests/golden/arrays/* were written by the reference's own faasTest/pylibsort/data.py
    (tests/golden/make_golden_arrays.py, which also checked that the reference reads OUR arrays).
    Ours must read them back exactly and, for the same operations, write byte-identical files."""
    import shutil
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "arrays")
    want = json.load(open(os.path.join(gold, "expected.json")))
    shutil.copytree(gold, tmp_path / "arrays")
    for name, exp in want.items():
        arr = da.fileDistribArray.Open(tmp_path / "arrays" / name)
        assert arr.shape.caps == exp["caps"] and arr.shape.lens == exp["lens"], name
        assert [bytes(arr.ReadPart(p)).hex() for p in range(arr.shape.npart)] == exp["parts"], name
    # the same operations through our writer: identical bytes on disk
    exp = want["partial"]
    parts = [bytes.fromhex(h) for h in exp["parts"]]
    mine = da.fileDistribArray.Create(tmp_path / "mine", da.ArrayShape.fromCaps(exp["caps"]))
    mine.WritePart(0, parts[0][:4])
    mine.WritePart(3, parts[3])
    mine.WritePart(0, parts[0][4:])
    mine.WritePart(1, parts[1])
    mine.Close()
    da.SetDistribMount(tmp_path)
    da.writeOutput({"arrType": "file", "output": "b"}, bytes.fromhex(want["buckets"]["sorted"]), want["buckets"]["boundaries"])
    for ours, theirs in (("mine", "partial"), ("b", "buckets")):
        for f in ("meta.json", "data.dat"):
            assert open(tmp_path / ours / f, "rb").read() == open(os.path.join(gold, theirs, f), "rb").read(), (ours, f)
