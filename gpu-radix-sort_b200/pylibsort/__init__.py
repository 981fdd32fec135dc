"""Mirror of the reference's ``pylibsort`` package (faasTest/pylibsort/__init__.py:5-29).

Same module-level behaviour: the library is located and ``initLibSort()`` is
called exactly once when the package is imported (reference __init__.py:17-20),
and ``sort`` / ``data`` names are re-exported flat.  The library searched first
is this repo's own build; ``LD_LIBRARY_PATH`` is honoured like the reference
(so the unmodified reference package also works against our libsort.so).
"""
import types

from .. import binding


def __setup():
    s = types.SimpleNamespace()
    s.sortLib = binding.lib()          # raises when libsort.so is not built: no fallback
    binding.init_once()                # must be called exactly once per process
    return s


__state = __setup()

from .data import *   # noqa: E402,F401,F403
from .sort import *   # noqa: E402,F401,F403
