"""File-backed distributed arrays: the exchange format either side of the sort path.

Same public names, on-disk format and errors as the reference's
faasTest/pylibsort/data.py:1-309 (and the Go twin benchmark/pkg/data/file.go:36-41,
146-202), so arrays written by either side open on the other:

    <mount>/<arrayName>/meta.json   {"Lens": [bytes valid per partition], "Caps": [bytes reserved per partition]}
    <mount>/<arrayName>/data.dat    partitions back to back at their CAPACITY offsets

A sort request (faas/inputs.go:13-33, faasTest/README.md:27-57) names its input as a
list of partition slices {"arrayName", "partID", "start", "nbyte"} and its output as an
array name; the sorter writes one output partition per radix bucket.

Pure Python, no GPU and no libsort needed (so it is importable on its own; the
``pylibsort`` package re-exports everything here).  I/O is positional
(os.pread / os.pwrite on one descriptor per array) and partition slices can be read
straight into a caller-provided buffer -- e.g. pinned memory that the GPU then DMAs.
"""
import json
import os
import threading
import pathlib
import shutil

import numpy as np

FileDistribArrayMount = pathlib.Path("/shared")      # reference data.py:16


class DistribArrayError(Exception):
    def __init__(self, cause):
        super().__init__(cause)
        self.cause = cause

    def __str__(self):
        return self.cause


def SetDistribMount(newRoot):
    """Directory under which request-named arrays live (default /shared; reference data.py:27-31)."""
    global FileDistribArrayMount
    FileDistribArrayMount = pathlib.Path(newRoot)


class ArrayShape:
    """Capacities and valid lengths (bytes) of an array's partitions; ``starts[i]`` is where
    partition i begins in data.dat, with one extra entry holding the total capacity."""

    def __init__(self, caps, lens):
        self.caps = [int(c) for c in caps]
        self.lens = [int(x) for x in lens]
        if len(self.caps) != len(self.lens):
            raise DistribArrayError("shape needs one length per capacity")
        self.npart = len(self.caps)
        self.starts = [0] * (self.npart + 1)
        for i, c in enumerate(self.caps):
            self.starts[i + 1] = self.starts[i] + c

    @classmethod
    def fromUniform(cls, cap, npart):
        return cls([cap] * npart, [0] * npart)

    @classmethod
    def fromCaps(cls, caps):
        return cls(list(caps), [0] * len(caps))


class DistribArray:
    """Interface of a distributed array (reference data.py:63-111)."""
    shape = None

    def Close(self):
        raise NotImplementedError

    def Destroy(self):
        raise NotImplementedError

    def ReadPart(self, partID, start=0, nbyte=-1):
        raise NotImplementedError

    def WritePart(self, partID, buf):
        raise NotImplementedError

    def ReadAll(self):
        raise NotImplementedError

    def WriteAll(self, buf):
        raise NotImplementedError


class fileDistribArray(DistribArray):
    """A distributed array stored as a directory (see the module docstring for the format)."""

    def __init__(self, rootPath):
        self.rootPath = pathlib.Path(rootPath)
        self.datPath = self.rootPath / "data.dat"
        self.metaPath = self.rootPath / "meta.json"
        self.closed = False
        self._fd = -1

    # -- life cycle ---------------------------------------------------------------
    @classmethod
    def Create(cls, rootPath, shape):
        arr = cls(rootPath)
        old = os.umask(0)              # world-accessible, like the reference (container uid mismatches)
        try:
            arr.rootPath.mkdir(0o777)  # an existing array is an error (FileExistsError), as in the reference
            arr._fd = os.open(arr.datPath, os.O_RDWR | os.O_CREAT, 0o666)
            os.close(os.open(arr.metaPath, os.O_WRONLY | os.O_CREAT, 0o666))
        finally:
            os.umask(old)
        arr.shape = ArrayShape(shape.caps, shape.lens)
        return arr

    @classmethod
    def Open(cls, rootPath):
        arr = cls(rootPath)
        if not arr.rootPath.exists():
            raise DistribArrayError("Array {} does not exist".format(rootPath))
        with open(arr.metaPath, "r") as f:
            meta = json.load(f)
        arr.shape = ArrayShape(meta["Caps"], meta["Lens"])
        arr._fd = os.open(arr.datPath, os.O_RDWR)
        return arr

    def Close(self):
        if self.closed:
            return                     # idempotent
        os.close(self._fd)
        self._fd = -1
        with open(self.metaPath, "w") as f:
            json.dump({"Lens": self.shape.lens, "Caps": self.shape.caps}, f)
        self.closed = True

    def Abandon(self):
        """Drop the handle WITHOUT writing the metadata: for a handle whose files were destroyed or
        replaced behind it (its lengths would overwrite the new array's meta.json)."""
        if not self.closed:
            os.close(self._fd)
            self._fd = -1
            self.closed = True

    def Destroy(self):
        if self._fd >= 0:
            os.close(self._fd)
            self._fd = -1
            self.closed = True
        shutil.rmtree(self.rootPath)

    # -- partitions -----------------------------------------------------------------
    def _pread_into(self, view, offset):
        done = 0
        while done < len(view):
            got = os.preadv(self._fd, [view[done:]], offset + done)
            if got == 0:               # hole at the end of a short file: reads as zeros
                view[done:] = bytes(len(view) - done)
                break
            done += got

    def ReadPart(self, partID, start=0, nbyte=-1, dest=None):
        """Bytes [start, start+nbyte) of a partition (nbyte -1: up to its valid length).  Returns a new
        bytearray, or fills the first nbyte bytes of ``dest`` (bytearray / memoryview / pinned buffer)."""
        have = self.shape.lens[partID]
        if nbyte == -1:
            nbyte = have - start
        if start > have or nbyte < 0 or start + nbyte > have:
            raise DistribArrayError("Read beyond end of partition {} (asked for {}+{}, limit {}".format(
                partID, start, nbyte, have))
        out = bytearray(nbyte) if dest is None else dest
        view = memoryview(out).cast("B")[:nbyte]
        self._pread_into(view, self.shape.starts[partID] + start)
        return out if dest is None else None

    def WritePart(self, partId, buf):
        """Append buf to a partition; it must fit the remaining capacity."""
        data = memoryview(buf).cast("B")
        room = self.shape.caps[partId] - self.shape.lens[partId]
        if len(data) > room:
            raise DistribArrayError("Wrote beyond end of partition (asked for {}b, limit {}b)".format(len(data), room))
        self._pwrite(data, self.shape.starts[partId] + self.shape.lens[partId])
        self.shape.lens[partId] += len(data)

    def _pwrite(self, data, offset):
        done = 0
        while done < len(data):
            done += os.pwrite(self._fd, data[done:], offset + done)

    # -- whole array ------------------------------------------------------------------
    def ReadAll(self, dest=None):
        """The whole data file, total-capacity bytes (partition i at shape.starts[i]; bytes past a
        partition's valid length are unspecified, past the end of a short file zero)."""
        total = self.shape.starts[self.shape.npart]
        out = bytearray(total) if dest is None else dest
        self._pread_into(memoryview(out).cast("B")[:total], 0)
        return memoryview(out) if dest is None else None

    def WriteAll(self, buf):
        """Overwrite the array; every partition is taken to be full afterwards."""
        data = memoryview(buf).cast("B")
        total = self.shape.starts[self.shape.npart]
        if len(data) != total:
            raise DistribArrayError("Buffer length {}b does not match array capacity {}b".format(len(data), total))
        self._pwrite(data, 0)
        self.shape.lens = list(self.shape.caps)


class partRef:
    """A slice of one partition of an array (reference data.py:221-233)."""

    def __init__(self, arr, partID=0, start=0, nbyte=-1):
        self.arr = arr
        self.partID = partID
        self.start = start
        self.nbyte = nbyte

    def read(self, dest=None):
        return self.arr.ReadPart(self.partID, start=self.start, nbyte=self.nbyte, dest=dest)


# arrays opened on behalf of requests, by name: one descriptor per array however many slices name it.
# Worker threads (faas.py) share the cache: every lookup and replacement happens under the lock.
openArrs = {}
_openArrsLock = threading.Lock()


def closeOpenArrays():
    """Close every array opened through getPartRefs (the reference leaves them open for the process lifetime)."""
    with _openArrsLock:
        for arr in openArrs.values():
            arr.Close()
        openArrs.clear()


def _still_the_same_file(arr):
    """A cached handle is valid only while its descriptor is the file now at that path: an array that
    was destroyed and re-created under the same name is another file (and possibly has new lengths)."""
    try:
        now, held = os.stat(arr.datPath), os.fstat(arr._fd)
    except OSError:
        return False
    # st_nlink == 0: the held file was unlinked (its inode number may already belong to the new file)
    return held.st_nlink > 0 and (now.st_ino, now.st_dev) == (held.st_ino, held.st_dev)


def _file_ref(entry):
    name = entry["arrayName"]
    with _openArrsLock:
        arr = openArrs.get(name)
        if (arr is None or arr.closed or arr.rootPath.parent != FileDistribArrayMount or not _still_the_same_file(arr)):
            if arr is not None:
                arr.Abandon()
            arr = fileDistribArray.Open(FileDistribArrayMount / name)
            openArrs[name] = arr
    nbyte = entry["nbyte"]
    if nbyte == -1:   # "the remainder of the partition (from start)", faasTest/README.md:52
        nbyte = arr.shape.lens[entry["partID"]] - entry["start"]
    return partRef(arr, partID=entry["partID"], start=entry["start"], nbyte=nbyte)


def getPartRefs(req):
    """partRefs for the 'input' list of a sort request (reference data.py:268-274)."""
    if req["arrType"] == "file":
        return [_file_ref(e) for e in req["input"]]
    raise ValueError("Invalid request type: " + str(req["arrType"]))


def readPartRefs(refs, dest=None):
    """Concatenation of the slices, as a writable memoryview (reference data.py:253-265).  With ``dest``
    (any writable buffer at least that long, e.g. pinned memory) the bytes land there instead."""
    total = sum(r.nbyte for r in refs)
    out = memoryview(bytearray(total) if dest is None else dest).cast("B")[:total]
    loc = 0
    for r in refs:
        r.read(dest=out[loc:loc + r.nbyte])
        loc += r.nbyte
    return out


def getOutputArray(req, shape):
    """A new array named by the request's 'output' field."""
    if req["arrType"] == "file":
        return fileDistribArray.Create(FileDistribArrayMount / req["output"], shape)
    raise ValueError("Invalid request type: " + str(req["arrType"]))


def exactBoundaries(buf, boundaries, offset, width):
    """Group boundaries as the reference's contract defines them (exclusive prefix sums of the group
    sizes; localTest/tests.cpp:44-68, faasTest/pylibsort/sort.py:67-91) from what gpuPartial returned.

    The reference's implementation -- which gpuPartial reproduces bit for bit -- deviates from that
    contract when the lowest groups are empty or group 1 is (libsort/sort.cu:14-27: thread 0 never
    writes; 381-387: the back-fill stops at group 2), e.g. it reports boundaries[1] == 0 although
    group 0 holds keys.  Consumers that cut the buffer into per-group partitions need the contract.
    O(number of groups): with g0 the group of the first sorted key, every boundary up to g0 is 0,
    boundaries above max(g0, 1) are already right, and boundaries[1] equals boundaries[2] when group 1
    is empty."""
    nElem = int(len(buf) / 4)
    out = [int(b) for b in boundaries]
    ngroups = len(out)
    if nElem == 0 or ngroups == 0:
        return [0] * ngroups
    first = int(np.frombuffer(memoryview(buf).cast("B")[:4], dtype=np.uint32)[0])
    g0 = (first >> offset) & ((1 << width) - 1)
    for g in range(min(g0 + 1, ngroups)):
        out[g] = 0
    if g0 == 0 and ngroups > 1 and out[1] == 0:       # group 1 is empty: it starts where group 2 starts
        out[1] = out[2] if ngroups > 2 else nElem
    return out


def writeOutput(req, rawBytes, boundaries):
    """Store a partially sorted buffer as the request's output array: one partition per bucket, bucket
    g spanning [4*boundaries[g], 4*boundaries[g+1]) (reference data.py:289-298)."""
    edges = np.asarray(boundaries, dtype=np.int64) * 4
    caps = np.diff(edges, append=len(rawBytes))
    out = getOutputArray(req, ArrayShape.fromCaps(caps.tolist()))
    out.WriteAll(rawBytes)
    out.Close()
