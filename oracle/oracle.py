"""ctypes front-end of the CPU oracle (TEST INFRASTRUCTURE — never imported by gsworld_b200/).

Loads oracle/libgsw_oracle.so (built by `make -C oracle`, see oracle/Makefile) and exposes numpy-level
wrappers.  See the header of gsw_oracle.cpp for what each stage follows in the reference and for the
parity status (rasterizer stages: PARITY UNPINNED — third-party source absent from /root/reference).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# GSW_ORACLE_LIB: an alternative build of gsw_oracle.cpp (tools/fma_audit.py compares a contracted build with this one)
_LIB_PATH = os.environ.get("GSW_ORACLE_LIB") or os.path.join(_HERE, "libgsw_oracle.so")
_lib = None

NEAR_GSWORLD = 0.05  # reference README.md:33 (hand patch of auxiliary.h)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "gsw_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_rasterize_forward.restype = C.c_longlong
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def _p(a, ty=C.c_void_p):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ty)


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def num_threads() -> int:
    return int(lib().orc_num_threads())


def transform_gaussians(xyz, log_scale, quat_raw, R, t, qR, kappa=None):
    """One (env, link/object) pair.  Returns dict(xyz, log_scale, quat, scale_act, quat_act)."""
    xyz, log_scale, quat_raw = _f32(xyz), _f32(log_scale), _f32(quat_raw)
    R, t, qR = _f32(R).reshape(9), _f32(t).reshape(3), _f32(qR).reshape(4)
    n = xyz.shape[0]
    out = dict(
        xyz=np.empty((n, 3), np.float32),
        log_scale=np.empty((n, 3), np.float32),
        quat=np.empty((n, 4), np.float32),
        scale_act=np.empty((n, 3), np.float32),
        quat_act=np.empty((n, 4), np.float32),
    )
    lib().orc_transform_gaussians(
        C.c_int(n), _p(xyz), _p(log_scale), _p(quat_raw), _p(R), _p(t), _p(qR),
        C.c_int(0 if kappa is None else 1), C.c_float(0.0 if kappa is None else float(kappa)),
        _p(out["xyz"]), _p(out["log_scale"]), _p(out["quat"]), _p(out["scale_act"]), _p(out["quat_act"]),
    )
    return out


def rasterize_forward(
    means3D, opacities, viewmatrix, projmatrix, campos, tanfovx, tanfovy, H, W, *,
    shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None,
    sh_degree=3, bg=(0.0, 0.0, 0.0), scale_modifier=1.0, antialiasing=False,
    near=NEAR_GSWORLD, debug=False, fragile_eps=1e-4, global_stable_sort=False,
):
    """One frame.  `scales`/`rotations`/`opacities` are the ACTIVATED tensors render() passes
    (gaussian_renderer/__init__.py:54-85).  Returns dict(color[3,H,W], radii[P], invdepth[1,H,W], D, ...)."""
    means3D = _f32(means3D).reshape(-1, 3)
    P = means3D.shape[0]
    opacities = _f32(opacities).reshape(-1)
    assert opacities.shape[0] == P
    if (shs is None) == (colors_precomp is None):
        raise Exception("Please provide excatly one of either SHs or precomputed colors!")
    if ((scales is None or rotations is None) and cov3D_precomp is None) or (
        (scales is not None or rotations is not None) and cov3D_precomp is not None
    ):
        raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
    shs = _f32(shs)
    M = 0
    if shs is not None:
        shs = shs.reshape(P, -1, 3)
        M = shs.shape[1]
    colors_precomp = _f32(colors_precomp)
    scales, rotations, cov3D_precomp = _f32(scales), _f32(rotations), _f32(cov3D_precomp)
    V = _f32(viewmatrix).reshape(16)
    PM = _f32(projmatrix).reshape(16)
    cam = _f32(campos).reshape(3)
    bg = _f32(bg).reshape(3)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    T = gx * gy

    out_color = np.empty((3, H, W), np.float32)
    out_radii = np.empty((P,), np.int32)
    out_invd = np.empty((1, H, W), np.float32)
    res = dict(color=out_color, radii=out_radii, invdepth=out_invd)
    dbg = {}
    cap = 0
    if debug:
        dbg = dict(
            tiles_touched=np.zeros((P,), np.int32), depth=np.zeros((P,), np.float32),
            xy=np.zeros((P, 2), np.float32), conic_opacity=np.zeros((P, 4), np.float32),
            rgb=np.zeros((P, 3), np.float32), ranges=np.zeros((T, 2), np.int32),
            n_contrib=np.zeros((H, W), np.int32), final_T=np.zeros((H, W), np.float32),
            fragile=np.zeros((H, W), np.uint8), best_id=np.zeros((H, W), np.int32),
            best_margin=np.zeros((H, W), np.float32),
        )
        cap = max(16, P * 16)
        dbg["keys"] = np.zeros((cap,), np.uint64)
        dbg["vals"] = np.zeros((cap,), np.uint32)
    pairs = C.c_longlong(0)

    def call():
        return lib().orc_rasterize_forward(
            C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(bg), C.c_int(W), C.c_int(H), _p(means3D),
            _p(shs), _p(colors_precomp), _p(opacities), _p(scales), C.c_float(scale_modifier), _p(rotations),
            _p(cov3D_precomp), _p(V), _p(PM), _p(cam), C.c_float(tanfovx), C.c_float(tanfovy),
            C.c_int(1 if antialiasing else 0), C.c_float(near), _p(out_color), _p(out_radii), _p(out_invd),
            _p(dbg.get("tiles_touched")), _p(dbg.get("depth")), _p(dbg.get("xy")), _p(dbg.get("conic_opacity")),
            _p(dbg.get("rgb")), _p(dbg.get("keys")), _p(dbg.get("vals")), C.c_longlong(cap), _p(dbg.get("ranges")),
            _p(dbg.get("n_contrib")), _p(dbg.get("final_T")), _p(dbg.get("fragile")), C.c_float(fragile_eps),
            C.byref(pairs), C.c_int(1 if global_stable_sort else 0), _p(dbg.get("best_id")),
            _p(dbg.get("best_margin")),
        )

    D = call()
    if D < 0:  # key capacity too small: retry with the exact size
        cap = int(dbg["tiles_touched"].astype(np.int64).sum()) + 16
        dbg["keys"] = np.zeros((cap,), np.uint64)
        dbg["vals"] = np.zeros((cap,), np.uint32)
        D = call()
    res["D"] = int(D)
    res["pairs"] = int(pairs.value)
    if debug:
        dbg["keys"] = dbg["keys"][:D]
        dbg["vals"] = dbg["vals"][:D]
        res.update(dbg)
    return res


def mark_visible(means3D, viewmatrix, near=NEAR_GSWORLD):
    means3D = _f32(means3D).reshape(-1, 3)
    out = np.zeros((means3D.shape[0],), np.uint8)
    lib().orc_mark_visible(C.c_int(means3D.shape[0]), _p(means3D), _p(_f32(viewmatrix).reshape(16)),
                           C.c_float(near), _p(out))
    return out.astype(bool)


def sh_to_rgb(means3D, campos, shs, deg=3):
    means3D = _f32(means3D).reshape(-1, 3)
    shs = _f32(shs).reshape(means3D.shape[0], -1, 3)
    out = np.empty((means3D.shape[0], 3), np.float32)
    lib().orc_sh_to_rgb(C.c_int(means3D.shape[0]), C.c_int(deg), C.c_int(shs.shape[1]), _p(means3D),
                        _p(_f32(campos).reshape(3)), _p(shs), _p(out))
    return out


def cov3d(scales, rotations, mod=1.0):
    scales = _f32(scales).reshape(-1, 3)
    rotations = _f32(rotations).reshape(-1, 4)
    out = np.empty((scales.shape[0], 6), np.float32)
    lib().orc_cov3d(C.c_int(scales.shape[0]), _p(scales), C.c_float(mod), _p(rotations), _p(out))
    return out


def to_uint8_hwc(chw):
    chw = _f32(chw)
    _, H, W = chw.shape
    out = np.empty((H, W, 3), np.uint8)
    lib().orc_to_uint8_hwc(C.c_int(H), C.c_int(W), _p(chw), _p(out))
    return out
