"""The sort worker of the reference's FaaS protocol (faasTest/f.py:47-72,156-190): a JSON request
names slices of file distributed arrays as input, a bit range, and an output array; the worker
partially sorts the concatenated slices on the GPU and writes one output partition per bucket.

    {"offset": 0, "width": 8, "arrType": "file",
     "input": [{"arrayName": "in0", "partID": 0, "start": 0, "nbyte": -1}, ...], "output": "out0"}

B200 side of it: the slices are read from the files straight into ONE pinned host buffer, which
gpuPartial then DMAs without a staging copy, and the sorted bytes are written back from that same buffer.
Run as ``python -m gpu-radix-sort_b200.faas`` it behaves like the reference's directInvoke(): request
on stdin, OL_SHARED_VOLUME names the array mount, JSON reply on stdout, exit status 0/1.
"""
import json
import os
import pathlib
import sys


def _pinned_buffer(nbyte):
    import torch
    t = torch.empty(max(nbyte, 1), dtype=torch.uint8).pin_memory()
    return t, memoryview(t.numpy())[:nbyte]


def f(event):
    """Handle one sort request; returns {"success": bool, "err": str} (reference f.py:47-72)."""
    if event["arrType"] != "file":
        return {"success": False, "err": "Function currently only supports file distributed arrays"}
    try:
        from . import pylibsort      # loads libsort.so and calls initLibSort() once; no GPU -> error reply
        refs = pylibsort.getPartRefs(event)
        keep, raw = _pinned_buffer(sum(r.nbyte for r in refs))
        pylibsort.readPartRefs(refs, dest=raw)
        boundaries = pylibsort.sortPartial(raw, event["offset"], event["width"])
        # gpuPartial is bit-exact with the reference, boundary quirks included; the
        # output partitions follow the contract (one partition per group, sized by the group), which is
        # what the consumers of these arrays -- BucketReader, checkPartial, CheckPartialArray -- rely on
        boundaries = pylibsort.exactBoundaries(raw, boundaries, event["offset"], event["width"])
    except Exception as e:  # noqa: BLE001 -- the protocol reports every failure in-band
        return {"success": False, "err": str(e)}
    pylibsort.writeOutput(event, raw, boundaries)
    del keep
    return {"success": True, "err": ""}


def selfTest(nElem=1 << 20, offset=0, width=8, narr=2, npart=2, mount=None):
    """The reference's selfTest (f.py:74-153) at a chosen size: narr input arrays of npart partitions,
    one request over all of them, output checked with checkPartial.  Returns the reply."""
    import tempfile
    from . import pylibsort
    nbyte = nElem * 4
    per_part = nbyte // (narr * npart)
    if per_part * narr * npart != nbyte or per_part % 4:
        raise ValueError("nElem must split evenly into narr*npart partitions of whole keys")
    data = pylibsort.generateInputs(nElem)
    with tempfile.TemporaryDirectory(dir=mount) as tdir:
        tdir = pathlib.Path(tdir)
        pylibsort.SetDistribMount(tdir)
        refs = []
        for a in range(narr):
            name = "faasSortTestIn%d" % a
            arr = pylibsort.fileDistribArray.Create(tdir / name, pylibsort.ArrayShape.fromUniform(per_part, npart))
            arr.WriteAll(data[a * npart * per_part:(a + 1) * npart * per_part])
            arr.Close()
            refs += [{"arrayName": name, "partID": p, "start": 0, "nbyte": -1} for p in range(npart)]
        req = {"offset": offset, "width": width, "arrType": "file", "input": refs, "output": "faasSortTestOut"}
        resp = f(req)
        if resp["success"]:
            out = pylibsort.fileDistribArray.Open(tdir / "faasSortTestOut")
            got, caps = out.ReadAll(), list(out.shape.caps)
            out.Destroy()
            pylibsort.checkPartial(data, got, caps, offset, width)
        pylibsort.closeOpenArrays()
    return resp


def directInvoke():
    """main(): request JSON on stdin, reply JSON on stdout, exit 0 on success (reference f.py:156-190)."""
    mount = os.environ.get("OL_SHARED_VOLUME", "")
    if mount == "":
        print(json.dumps({"success": False,
                          "err": "OL_SHARED_VOLUME not set, set it to the shared directory for distrib arrays"}))
        sys.exit(1)
    try:
        cmd = json.loads(sys.stdin.read())
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"success": False, "err": "Argument parsing error: " + str(e)}))
        sys.exit(1)
    from . import distrib_array
    distrib_array.SetDistribMount(pathlib.Path(mount))
    resp = f(cmd)
    print(json.dumps(resp))
    sys.exit(0 if resp["success"] else 1)


if __name__ == "__main__":
    directInvoke()
