"""Drop-in for the `diff_gaussian_rasterization` Python package (forward only).

Keeps the names, field order, argument list, return tuple and exception text the reference relies on:
  GaussianRasterizationSettings  — 13 fields, pinned by the call site
                                   submodules/gaussian-splatting/gaussian_renderer/__init__.py:36-50
  GaussianRasterizer(raster_settings)(means3D, means2D, opacities, shs, colors_precomp, scales,
                                      rotations, cov3D_precomp) -> (color, radii, invdepth)
                                   — call site gaussian_renderer/__init__.py:52,102-110
The package itself is the absent submodule graphdeco-inria/diff-gaussian-rasterization@9c5c2028
(branch dr_aa) with GSWorld's near-plane patch 0.2 -> 0.05 (README.md:33).  The observation path is
forward-only (every loaded tensor has requires_grad=False, semantic_3dgs_wrapper.py:151-165), so no
backward is provided; `means2D` is accepted and ignored.  `SparseGaussianAdam` is deliberately absent,
which makes the wrapper's probe (gs_world_wrapper.py:22-26) choose separate_sh=False as with the
pinned rasterizer.
"""
from __future__ import annotations

from typing import NamedTuple

import torch
from torch import nn

from .engine import Engine

NEAR_PLANE = 0.05  # GSWorld README.md:33

_engines = {}


def _engine_for(device: torch.device) -> Engine:
    key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
    if key not in _engines:
        _engines[key] = Engine(torch.device("cuda", key[1]))
    return _engines[key]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings: GaussianRasterizationSettings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        rs = self.raster_settings
        with torch.no_grad():
            return _engine_for(positions.device).mark_visible(positions, rs.viewmatrix, NEAR_PLANE)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception("Please provide excatly one of either SHs or precomputed colors!")
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
            (scales is not None or rotations is not None) and cov3D_precomp is not None
        ):
            raise Exception("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!")
        if not means3D.is_cuda:
            raise RuntimeError("GaussianRasterizer: tensors must be CUDA tensors (there is no CPU rasterizer)")
        eng = _engine_for(means3D.device)
        with torch.no_grad():
            color, radii, invdepth = eng.rasterize_forward(
                means3D.detach().float(), opacities.detach().float().reshape(-1), rs.viewmatrix.float(),
                rs.projmatrix.float(), rs.campos.float(), rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width,
                rs.bg.float(),
                shs=None if shs is None else shs.detach().float(),
                colors_precomp=None if colors_precomp is None else colors_precomp.detach().float(),
                scales=None if scales is None else scales.detach().float(),
                rotations=None if rotations is None else rotations.detach().float(),
                cov3D_precomp=None if cov3D_precomp is None else cov3D_precomp.detach().float(),
                sh_degree=rs.sh_degree, scale_modifier=rs.scale_modifier, prefiltered=rs.prefiltered,
                antialiasing=rs.antialiasing, near=NEAR_PLANE)
            if rs.debug:
                torch.cuda.synchronize(means3D.device)
        return color, radii, invdepth
