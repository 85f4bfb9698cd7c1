"""Loader for the C++ reducer extension (``csrc/reducer/reducer.cpp``), built in-tree on demand."""
from __future__ import annotations

import importlib.util
import threading

_mod = None
_lock = threading.Lock()


def load():
    global _mod
    if _mod is not None:
        return _mod
    with _lock:
        if _mod is not None:
            return _mod
        import torch  # noqa: F401  (libtorch must be loaded before the extension)
        from .. import build

        path = build.reducer_path()
        if path is None:
            path = build.build_reducer()
        spec = importlib.util.spec_from_file_location("_reducer", str(path))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _mod = mod
        return mod
