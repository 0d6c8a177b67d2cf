"""B200 compute ops: ctypes bindings of the in-tree sm_100a kernel library plus the PyTorch
reference path used on CPU (tests) -- see ``_lib.py`` for the loading / no-silent-fallback rule."""
from . import _lib  # noqa: F401
