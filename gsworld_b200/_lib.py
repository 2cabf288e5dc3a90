"""ctypes binding of gsworld_b200/csrc/libgsw.so (the C ABI declared in include/gsw.h).

There is no CPU fallback: if the library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GSW_LIB_PATH: another build of the same sources (A/B runs of compile-time tuning constants)
LIB_PATH = os.environ.get("GSW_LIB_PATH") or os.path.join(_HERE, "csrc", "libgsw.so")
POSE_STRIDE = 20

GSW_E_CAPACITY = -3
SEG_NONE = -2  # GSW_SEG_NONE


class FramesDesc(C.Structure):
    _fields_ = [
        ("num_envs", C.c_int32), ("num_views", C.c_int32), ("num_frames", C.c_int32),
        ("height", C.c_int32), ("width", C.c_int32), ("antialiasing", C.c_int32),
        ("scale_modifier", C.c_float), ("near_plane", C.c_float), ("bg", C.c_float * 3),
        ("pose", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p), ("campos", C.c_void_p),
        ("tanfov", C.c_void_p), ("frame_env", C.c_void_p), ("frame_view", C.c_void_p),
        ("out_rgb_u8", C.c_void_p), ("out_color", C.c_void_p), ("out_invdepth", C.c_void_p),
        ("out_radii", C.c_void_p), ("host_rgb_u8", C.c_void_p), ("out_seg", C.c_void_p),
        ("peer_rgb_u8", C.POINTER(C.c_void_p)), ("num_peer_rgb_u8", C.c_int32), ("deliver_flags", C.c_int32),
    ]


# every symbol include/gsw.h declares: (name, restype, argtypes)
_vp, _i32, _i64, _f = C.c_void_p, C.c_int32, C.c_int64, C.c_float
SYMBOLS = {
    "gsw_version": (C.c_int, []),
    "gsw_last_error": (C.c_char_p, []),
    "gsw_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "gsw_destroy": (C.c_int, [_vp]),
    "gsw_scene_set": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int, _vp]),
    "gsw_scene_set_labels": (C.c_int, [_vp, _vp]),
    "gsw_render_frames": (C.c_int, [_vp, C.POINTER(FramesDesc), _vp]),
    "gsw_finish": (C.c_int, [_vp, C.POINTER(_i64), C.POINTER(_i64 * 4)]),
    "gsw_delivery_wait": (C.c_int, [_vp, _vp]),
    "gsw_delivery_gate": (C.c_int, [_vp, _vp]),
    "gsw_reserve_entries": (C.c_int, [_vp, _i64]),
    "gsw_set_depth_buckets": (C.c_int, [_vp, C.c_int]),
    "gsw_set_tile_culling": (C.c_int, [_vp, C.c_int]),
    "gsw_rasterize_forward": (C.c_int, [_vp, _i32, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _f, _vp,
                                        _vp, _vp, _vp, _vp, _f, _f, _i32, _i32, _f, _vp, _vp, _vp, _vp]),
    "gsw_rasterize_forward_static": (C.c_int, [_i32, _i32, _i32, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _f, _vp, _vp,
                                               _vp, _vp, _vp, _f, _f, _i32, _f, _vp, _vp, _vp, _vp, _vp]),
    "gsw_mark_visible": (C.c_int, [_vp, _i32, _vp, _vp, _f, _vp, _vp]),
    "gsw_rig_set": (C.c_int, [_vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gsw_compose_poses": (C.c_int, [_vp, C.c_int, _vp, _vp, C.c_int, _vp, _vp]),
    "gsw_compose_cameras": (C.c_int, [_vp, C.c_int, _vp, _vp, C.c_int, C.c_int, _vp, _f, _vp, _vp, _vp, _vp, _vp]),
    "gsw_kernel_launches": (C.c_longlong, []),
    "gsw_profile_enable": (C.c_int, [_vp, C.c_int]),
    "gsw_profile_read": (C.c_int, [_vp, C.POINTER(C.c_float), C.c_int]),
    "gsw_profile_counters": (C.c_int, [_vp, C.POINTER(_i64), C.c_int]),
    "gsw_debug_frame_attrs": (C.c_int, [_vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gsw_debug_frame_lists": (C.c_int, [_vp, _i32, _vp, _vp, _vp, _i64, C.POINTER(_i64), _vp]),
    "gsw_debug_bin_offsets": (C.c_int, [_vp, _vp, _i64, C.POINTER(_i64), _vp]),
}

_lib = None


class GswError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libgsw error {code}: {msg}")
        self.code = code


def load():
    """Load libgsw.so and bind every declared symbol.  Raises if the extension was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: the CUDA extension is not built (run `python -c 'import __graft_entry__ as g; "
                f"g.build()'` or `make -C gsworld_b200/csrc`). gsworld_b200 has no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int):
    if rc != 0:
        raise GswError(rc, load().gsw_last_error().decode("utf-8", "replace"))
