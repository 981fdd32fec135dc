"""The reference's distributed sort over file distributed arrays, driven from Python.

Mirror of the Go driver's data path (benchmark/pkg/sort/helpers.go:22-140 ``BucketReader``;
distrib.go:90-179 ``SortDistribFromArr``; distrib.go:183-248 ``SortDistribFromRaw``) with the
FaaS sort worker (``faas.f``, the reference's faasTest/f.py) as the thing that sorts:

    for step in 0 .. 32/width:
        read the previous step's outputs bucket-major / worker-minor (STRIDED),
        cut that stream into nworker slices of partRefs,
        every worker partially sorts its slice by digit `step` on a GPU and writes one
        output array with a partition per bucket.

It is the like-for-like "same protocol, new backend" path: every exchange goes through files
exactly as with the reference (the NVLink paths -- distrib.py, libsortDistribSort -- are what
replaces it).  Workers run concurrently on threads; libsort's device pool hands each
concurrent gpuPartial call its own GPU.  Pure Python apart from the worker, so the
orchestration is tested on CPU with a numpy stand-in worker (tests/test_distrib_array.py).
"""
from concurrent.futures import ThreadPoolExecutor

from . import distrib_array as da

INORDER = 0   # array 0's partitions, then array 1's, ...
STRIDED = 1   # every array's partition 0, then every array's partition 1, ...  (bucket-major)


class BucketReader:
    """Iterates the partitions of several arrays as one byte stream (helpers.go:30-140)."""

    def __init__(self, arrays, order=STRIDED):
        if order not in (INORDER, STRIDED):
            raise ValueError("order must be INORDER or STRIDED")
        self.arrs = list(arrays)
        self.nArr = len(self.arrs)
        self.nPart = self.arrs[0].shape.npart if self.arrs else 0
        for a in self.arrs:
            if a.shape.npart != self.nPart:
                raise da.DistribArrayError("BucketReader needs the same number of partitions in every array")
        self.order = order
        self.arrX = self.partX = self.dataX = 0
        self.done = self.nArr == 0 or self.nPart == 0

    def _advance(self):
        self.dataX = 0
        if self.order == STRIDED:
            self.arrX += 1
            if self.arrX >= self.nArr:
                self.arrX, self.partX = 0, self.partX + 1
                self.done = self.partX >= self.nPart
        else:
            self.partX += 1
            if self.partX >= self.nPart:
                self.partX, self.arrX = 0, self.arrX + 1
                self.done = self.arrX >= self.nArr

    def readRef(self, sz):
        """partRefs covering the next sz bytes of the stream (fewer at its end; helpers.go:94-121)."""
        out, need = [], sz
        while not self.done and need > 0:
            have = self.arrs[self.arrX].shape.lens[self.partX] - self.dataX
            if have <= 0:
                self._advance()
                continue
            take = min(have, need)
            out.append(da.partRef(self.arrs[self.arrX], partID=self.partX, start=self.dataX, nbyte=take))
            self.dataX += take
            need -= take
            if self.dataX >= self.arrs[self.arrX].shape.lens[self.partX]:
                self._advance()
        return out

    def read(self, sz=-1):
        """The next sz bytes (all that is left for -1) as a bytearray (helpers.go:123-160)."""
        if sz < 0:
            sz = sum(sum(a.shape.lens) for a in self.arrs)
        return bytearray(da.readPartRefs(self.readRef(sz)))


def _as_request_refs(refs, names):
    return [{"arrayName": names[id(r.arr)], "partID": r.partID, "start": r.start, "nbyte": r.nbyte} for r in refs]


def SortDistribFromArr(arrName, sz, baseName, nworker=2, width=8, worker=None, parallel=True):
    """Distributed sort of the array `arrName` (under the current mount, sz valid bytes of uint32 keys).
    Returns the names of the nworker output arrays of the last step: their partitions, read
    bucket-major / worker-minor, are the sorted keys.  `worker(request) -> {"success", "err"}` is
    the FaaS sort worker (default: faas.f, i.e. a GPU partial sort).  Intermediate arrays (and the
    input, as in the reference) are destroyed."""
    if worker is None:
        from . import faas
        worker = faas.f
    if width < 1 or 32 % width:
        raise ValueError("width must divide 32")
    nstep = 32 // width
    nElem = sz // 4
    maxPerWorker = -(-nElem // nworker) * 4            # ceil: the last worker may get less
    names = [arrName]
    for step in range(nstep):
        inputs = [da.fileDistribArray.Open(da.FileDistribArrayMount / n) for n in names]
        lookup = {id(a): n for a, n in zip(inputs, names)}
        gen = BucketReader(inputs, STRIDED)
        reqs = []
        for w in range(nworker):
            refs = gen.readRef(maxPerWorker)
            reqs.append({"offset": step * width, "width": width, "arrType": "file", "input": _as_request_refs(refs, lookup),
                         "output": "%s_step%d_worker%d_output" % (baseName, step, w)})
        if not gen.done and gen.readRef(1):
            raise da.DistribArrayError("input generator still has data after the last worker")
        if parallel and nworker > 1:
            with ThreadPoolExecutor(nworker) as pool:
                resps = list(pool.map(worker, reqs))
        else:
            resps = [worker(r) for r in reqs]
        for w, resp in enumerate(resps):
            if not resp.get("success"):
                raise RuntimeError("Worker failure on step %d, worker %d: %s" % (step, w, resp.get("err")))
        da.closeOpenArrays()                            # the workers' cached handles of this step's inputs
        for a in inputs:
            a.Destroy()
        names = [r["output"] for r in reqs]
    return names


def SortDistribFromRaw(inRaw, baseName, nworker=2, width=8, worker=None, parallel=True):
    """Sort a native byte buffer of uint32 keys through file distributed arrays (distrib.go:183-248):
    returns the sorted bytes; every array created on the way is destroyed."""
    raw = memoryview(inRaw).cast("B")
    inName = baseName + "_input"
    arr = da.fileDistribArray.Create(da.FileDistribArrayMount / inName, da.ArrayShape.fromUniform(len(raw), 1))
    arr.WritePart(0, raw)
    arr.Close()
    outNames = SortDistribFromArr(inName, len(raw), baseName, nworker=nworker, width=width, worker=worker, parallel=parallel)
    outs = [da.fileDistribArray.Open(da.FileDistribArrayMount / n) for n in outNames]
    result = BucketReader(outs, STRIDED).read()
    for a in outs:
        a.Destroy()
    if len(result) != len(raw):
        raise da.DistribArrayError("sorted output has %d bytes, input had %d" % (len(result), len(raw)))
    return result
