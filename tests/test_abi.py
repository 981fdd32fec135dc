"""CPU-side checks of the drop-in boundary: the library loads, exports every
symbol include/*.h declares, and the host-only entry points behave like the
reference's.  No GPU compute is called here."""
import ctypes
import os
import re
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")


def declared_symbols(header):
    txt = open(os.path.join(INC, header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return re.findall(r"LIBSORT_API\s+[\w\s\*]+?\b(\w+)\s*\(", txt)


def test_headers_compile_as_c99_and_cxx14(tmp_path):
    c = tmp_path / "t.c"
    c.write_text('#include "libsort_b200.h"\nint main(void){return 0;}\n')
    subprocess.check_call(["gcc", "-std=gnu99", "-Wall", "-Werror", "-fsyntax-only", "-I", INC, str(c)])
    cpp = tmp_path / "t.cpp"
    cpp.write_text('#include "libsort_b200.h"\nint main(){return 0;}\n')
    subprocess.check_call(["g++", "-std=c++14", "-Wall", "-Werror", "-fsyntax-only", "-I", INC, str(cpp)])


def test_library_exports_every_declared_symbol(libsort, pkg):
    legacy = declared_symbols("libsort.h")
    assert sorted(legacy) == sorted(pkg.binding.LEGACY_SYMBOLS)   # the reference's seven (libsort.h:14-32)
    for name in legacy + declared_symbols("libsort_b200.h"):
        assert hasattr(libsort, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", pkg.lib_path()], capture_output=True, text=True).stdout
    exported = set(l.split()[-1] for l in out.splitlines() if " T " in l)
    assert set(legacy) <= exported
    # nothing but the C ABI leaks (C++ symbols are hidden)
    assert not [s for s in exported if s.startswith("_Z")], "C++ symbols exported"


def test_populate_input_matches_reference_stream(libsort, oracle):
    """libsort/utils.cu:63-79: fixed seed, state continues across calls."""
    a = np.empty(1000, np.uint32)
    b = np.empty(500, np.uint32)
    libsort.populateInput(a.ctypes.data, a.size)
    libsort.populateInput(b.ctypes.data, b.size)
    # this is the only test in the CPU session that draws from the library stream
    assert np.array_equal(np.concatenate([a, b]), oracle.pcg_stream(1500))


def test_provided_cpu_is_a_sort(libsort, oracle):
    k = oracle.pcg_stream(10007)
    x = k.copy()
    assert libsort.providedCpu(x.ctypes.data, x.size) is True
    assert np.array_equal(x, oracle.cpu_sort(k))


def test_bool_returns_are_clean_ints(libsort, pkg):
    """ctypes callers that never set restype read a c_int (reference sort.py:101,118)."""
    raw = ctypes.CDLL(pkg.lib_path())
    x = np.arange(10, dtype=np.uint32)
    assert raw.providedCpu(ctypes.c_void_p(x.ctypes.data), ctypes.c_size_t(x.size)) == 1


def test_workspace_and_tile_queries(libsort):
    t = libsort.libsortTileKeys()
    assert t >= 1024 and t % 4 == 0
    assert libsort.libsortWorkspaceBytes(1 << 20) < libsort.libsortWorkspaceBytes(1 << 24)
    assert libsort.libsortErrorString(0) == b"success"
    assert libsort.libsortDeviceCount() >= 0


def test_pylibsort_mirror_names(pkg):
    """Same public names as reference faasTest/pylibsort/sort.py and data.py:313."""
    src = open(os.path.join(os.path.dirname(pkg.__file__), "pylibsort", "sort.py")).read()
    for name in ("sortFull", "sortPartial", "checkOrder", "checkSortFull", "checkPartial", "groupBits",
                 "bytesToInts", "sortException", "sortResultException"):
        assert re.search(r"^(def|class) %s\b" % name, src, flags=re.M), name
    data = open(os.path.join(os.path.dirname(pkg.__file__), "pylibsort", "data.py")).read()
    assert re.search(r"^def generateInputs\b", data, flags=re.M)
