"""Restated reference-style GPU pipeline (COMPARISON POINT for bench.py; not product, not the reference's source).

`librefgpu.so` (baseline/ref_rasterizer/ref_rasterizer.cu) restates the published forward structure of the
absent diff-gaussian-rasterization submodule.  Two drivers, as BASELINE.md §3.2 asks:
  * as_gsworld_calls_it: the reference's torch statements on the device (oracle/torch_transform.RefPath with
    device="cuda": transform_gs_perlink, then per camera and env deepcopy + isin write-backs + getters'
    activations + one rasterizer call with its blocking readback + uint8 conversion);
  * rasterizer_only: pre-activated per-frame tensors, one rasterizer call per frame.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "ref_rasterizer", "librefgpu.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_LIB)
        _lib.refgpu_rasterize_forward.restype = C.c_longlong
    return _lib


def _p(t):
    return C.c_void_p(t.data_ptr())


def rasterize(means, shs, opac, scales, rots, view, proj, campos, tanx, tany, H, W, bg, near=0.05):
    P = means.shape[0]
    dev = means.device
    color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    invd = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    f = lambda t: t.contiguous().float()
    means, shs, opac, scales, rots = f(means), f(shs), f(opac), f(scales), f(rots)
    view, proj, campos, bg = f(view), f(proj), f(campos), f(bg)
    n = lib().refgpu_rasterize_forward(
        C.c_int(P), C.c_int(3), C.c_int(16), _p(bg), C.c_int(W), C.c_int(H), _p(means), _p(shs), _p(opac), _p(scales),
        C.c_float(1.0), _p(rots), _p(view), _p(proj), _p(campos), C.c_float(tanx), C.c_float(tany), C.c_float(near),
        _p(color), _p(radii), _p(invd), C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    if n < 0:
        raise RuntimeError("refgpu_rasterize_forward failed")
    return color, radii, invd, int(n)


def render_like_reference(cam, g, bg=None):
    """gaussian_renderer.render with the getters' activations (gaussian_model.py:102-130), rasterizer = librefgpu."""
    dev = g._xyz.device
    bg = torch.zeros(3, device=dev) if bg is None else bg
    scales = torch.exp(g._scaling)
    rots = torch.nn.functional.normalize(g._rotation)
    opac = torch.sigmoid(g._opacity)
    shs = torch.cat((g._features_dc, g._features_rest), dim=1)
    color, _, _, _ = rasterize(g._xyz, shs, opac.reshape(-1), scales, rots, cam["view"].to(dev), cam["proj"].to(dev),
                               cam["campos"].to(dev), cam["tanfovx"], cam["tanfovy"], cam["H"], cam["W"], bg)
    return color
