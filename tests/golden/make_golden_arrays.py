"""Golden file distributed arrays written by the REFERENCE's own Python code (faasTest/pylibsort/data.py),
plus a cross-check that the reference opens arrays written by our implementation.

Run in the build container (needs /root/reference; nothing here runs on the GPU box):

    LD_LIBRARY_PATH=$PWD/gpu-radix-sort_b200/libsort python tests/golden/make_golden_arrays.py

The unmodified reference package imports against our libsort.so (it only needs the library to load;
the array code is pure Python).  Output: tests/golden/arrays/<name>/{meta.json,data.dat} and
tests/golden/arrays/expected.json (what each partition must read back as, hex).
"""
import importlib
import json
import os
import pathlib
import shutil
import sys
import tempfile

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[2]
OUT = ROOT / "tests" / "golden" / "arrays"
sys.path.insert(0, "/root/reference/faasTest")
sys.path.insert(0, str(ROOT))
import pylibsort as ref  # noqa: E402  (the reference, unmodified)

ours = importlib.import_module("gpu-radix-sort_b200.distrib_array")


def main():
    if OUT.exists():
        shutil.rmtree(OUT)
    OUT.mkdir(parents=True)
    rng = np.random.default_rng(20261104)
    expected = {}

    # 1. partially filled partitions, appended in two goes, one empty partition
    arr = ref.fileDistribArray.Create(OUT / "partial", ref.ArrayShape.fromCaps([16, 8, 0, 24]))
    parts = [rng.integers(0, 256, n, dtype=np.uint8).tobytes() for n in (10, 3, 0, 24)]
    arr.WritePart(0, parts[0][:4])
    arr.WritePart(3, parts[3])
    arr.WritePart(0, parts[0][4:])
    arr.WritePart(1, parts[1])
    arr.Close()
    expected["partial"] = {"caps": [16, 8, 0, 24], "lens": [10, 3, 0, 24], "parts": [p.hex() for p in parts]}

    # 2. WriteAll on a uniform shape
    arr = ref.fileDistribArray.Create(OUT / "uniform", ref.ArrayShape.fromUniform(12, 3))
    whole = rng.integers(0, 256, 36, dtype=np.uint8).tobytes()
    arr.WriteAll(whole)
    arr.Close()
    expected["uniform"] = {"caps": [12, 12, 12], "lens": [12, 12, 12], "parts": [whole[i:i + 12].hex() for i in (0, 12, 24)]}

    # 3. writeOutput: one partition per bucket from (sorted bytes, boundaries), incl. empty buckets
    keys = np.sort(rng.integers(0, 4, 40, dtype=np.uint32) * np.uint32(3) % np.uint32(4))  # 2-bit groups, some empty
    keys = keys[keys != 2].astype(np.uint32)
    counts = np.bincount(keys & 3, minlength=4)
    bounds = (np.cumsum(counts) - counts).tolist()
    ref.SetDistribMount(OUT)
    ref.writeOutput({"arrType": "file", "output": "buckets"}, keys.tobytes(), bounds)
    expected["buckets"] = {"caps": (counts * 4).tolist(), "lens": (counts * 4).tolist(),
                           "parts": [keys[keys & 3 == g].tobytes().hex() for g in range(4)], "boundaries": bounds,
                           "sorted": keys.tobytes().hex()}

    json.dump(expected, open(OUT / "expected.json", "w"), indent=1)

    # cross-check: arrays written by OUR code, opened and read by the reference
    with tempfile.TemporaryDirectory() as t:
        t = pathlib.Path(t)
        mine = ours.fileDistribArray.Create(t / "mine", ours.ArrayShape.fromCaps([16, 8, 0, 24]))
        mine.WritePart(0, parts[0][:4])
        mine.WritePart(3, parts[3])
        mine.WritePart(0, parts[0][4:])
        mine.WritePart(1, parts[1])
        mine.Close()
        theirs = ref.fileDistribArray.Open(t / "mine")
        assert theirs.shape.caps == [16, 8, 0, 24] and theirs.shape.lens == [10, 3, 0, 24]
        assert [bytes(theirs.ReadPart(p)) for p in range(4)] == parts
        # byte-identical files for identical operations
        for f in ("meta.json", "data.dat"):
            assert open(t / "mine" / f, "rb").read() == open(OUT / "partial" / f, "rb").read(), f
        ours.SetDistribMount(t)
        ours.writeOutput({"arrType": "file", "output": "b"}, keys.tobytes(), bounds)
        for f in ("meta.json", "data.dat"):
            assert open(t / "b" / f, "rb").read() == open(OUT / "buckets" / f, "rb").read(), f
        ref.SetDistribMount(t)
        refs = ref.getPartRefs({"arrType": "file", "input": [{"arrayName": "b", "partID": 1, "start": 4, "nbyte": -1 + 5}]})
        assert bytes(ref.readPartRefs(refs)) == keys[keys & 3 == 1].tobytes()[4:8]
    print("golden arrays written to", OUT, "and cross-checked against the reference implementation")


if __name__ == "__main__":
    main()
