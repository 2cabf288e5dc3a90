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


def _header():
    return open(os.path.join(ROOT, "include", "gsw.h")).read()


def _define(name):
    m = re.search(r"#define\s+%s\s+\(?(-?\d+)\)?" % name, _header())
    assert m, f"{name} not defined in include/gsw.h"
    return int(m.group(1))


def test_python_constants_and_struct_layout_follow_the_header():
    """The ctypes mirror is written by hand: constants, delivery flags and the field order of gsw_frames_desc must be
    the header's (a silent mismatch would shift every pointer of the descriptor)."""
    assert _lib.POSE_STRIDE == _define("GSW_POSE_STRIDE")
    assert _lib.GSW_E_CAPACITY == _define("GSW_E_CAPACITY")
    assert _lib.SEG_NONE == _define("GSW_SEG_NONE")
    src = re.sub(r"/\*.*?\*/", "", _header(), flags=re.S)
    body = re.search(r"typedef struct gsw_frames_desc\s*\{(.*?)\}\s*gsw_frames_desc;", src, flags=re.S)
    assert body, "gsw_frames_desc not found in include/gsw.h"
    fields = []
    for decl in body.group(1).split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):  # "int32_t a, b" declares several
            name = re.findall(r"([A-Za-z_][A-Za-z0-9_]*)\s*(?:\[\d+\])?\s*$", part.strip())
            assert name, decl
            fields.append(name[0])
    assert fields == [f[0] for f in _lib.FramesDesc._fields_]
    # delivery modes of Engine.render_frames(defer_delivery_join=...)
    eng = open(os.path.join(ROOT, "gsworld_b200", "engine.py")).read()
    m = re.search(r'\{"peers":\s*(\d+),\s*"after":\s*(\d+)\}', eng)
    assert m, "delivery flag table not found in engine.py"
    assert int(m.group(1)) == _define("GSW_DELIVER_NO_PEER_JOIN")
    assert int(m.group(2)) == _define("GSW_DELIVER_NO_JOIN") | _define("GSW_DELIVER_AFTER_BLEND")
