"""CPU check of the drop-in boundary: libgsw.so loads here (no GPU needed) and exports every function that
include/gsw.h declares; the ctypes table binds each of them; compute calls fail loudly without a device."""
import ctypes as C
import os
import re

import pytest

from gsworld_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "gsw.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gsw_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    names = _declared()
    assert len(names) >= 18 and "gsw_render_frames" in names and "gsw_rasterize_forward" in names
    lib = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/gsw.h but not exported by libgsw.so"
        assert n in _lib.SYMBOLS, f"{n} not bound in gsworld_b200/_lib.py"
    assert sorted(_lib.SYMBOLS) == names
    assert _lib.load().gsw_version() == 100


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from gsworld_b200.engine import Engine
    with pytest.raises(RuntimeError):
        Engine("cpu")
    h = C.c_void_p()
    rc = _lib.load().gsw_create(0, C.byref(h))
    assert rc != 0 and _lib.load().gsw_last_error()
