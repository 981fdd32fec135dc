"""Runs in a fresh process with oracle/_ref/pyref on sys.path and LD_LIBRARY_PATH naming the
libsort.so under test: imports the reference's UNMODIFIED pylibsort package (faasTest/pylibsort,
staged by `make -C oracle ref_pylibsort`) and drives its sortFull / sortPartial / check* functions.
argv[1] = "script": run the reference's own testPyLibSort.py instead and report how it ended."""
import os
import runpy
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PYREF = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "pyref")
sys.path.insert(0, PYREF)


def loaded_libsort():
    return [l.split()[-1] for l in open("/proc/self/maps") if "libsort" in l and l.rstrip().endswith(".so")]


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "script":
        try:
            runpy.run_path(os.path.join(PYREF, "testPyLibSort.py"), run_name="__main__")
        except SystemExit as e:
            print("EXIT", e.code)
        except Exception as e:  # noqa: BLE001 -- the script's own defects surface here, identically for any library
            import traceback
            where = [f.name for f in traceback.extract_tb(e.__traceback__) if f.filename.endswith("testPyLibSort.py")]
            print("EXC", type(e).__name__, e, "in", where[-1] if where else "?")
        print("LIB", sorted(set(loaded_libsort())))
        return 0
    import pylibsort                                   # the reference's package; initLibSort() runs here
    assert os.path.realpath(pylibsort.__file__).startswith(os.path.realpath(PYREF)), pylibsort.__file__
    print("LIB", sorted(set(loaded_libsort())))
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import oracle
    rng = np.random.default_rng(7)
    for n in (1, 255, 1021, 4099, 100003, (1 << 20) + 3):
        k = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
        buf = bytearray(k.tobytes())
        pylibsort.sortFull(buf)
        got = pylibsort.bytesToInts(buf)
        assert np.array_equal(got, np.sort(k)), ("sortFull", n)
        if n <= 4099:
            pylibsort.checkSortFull(got, k)            # the reference's own checkers
            pylibsort.checkOrder(got)
        for off, w in ((0, 8), (4, 8), (16, 16), (0, 4)):
            buf = bytearray(k.tobytes())
            bounds = pylibsort.sortPartial(buf, off, w)
            want, wb = oracle.gpu_partial(k, off, w)
            assert np.array_equal(pylibsort.bytesToInts(buf), want), ("sortPartial", n, off, w)
            assert np.array_equal(np.array(bounds, dtype=np.uint32), wb), ("boundaries", n, off, w)
    print("WORKER PASS")
    return 0


if __name__ == "__main__":
    sys.exit(main())
