"""ai4e_b200 — Python side of the B200 classify engine (ctypes over libai4e_b200.so).

Importing this package never touches CUDA; the shared library is loaded on first use and
there is no CPU fallback (see `_lib.load`)."""
from . import _lib  # noqa: F401
from ._lib import Ai4eError, Ai4eLibraryError  # noqa: F401

__all__ = ["Ai4eError", "Ai4eLibraryError"]
