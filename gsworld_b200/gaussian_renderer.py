"""render(): the reference's render wrapper, same signature and returned dict.

Mirrors submodules/gaussian-splatting/gaussian_renderer/__init__.py:18-128 for the observation path
(forward only): builds GaussianRasterizationSettings from the camera, applies the model getters'
activations, honours pipe.compute_cov3D_python / pipe.convert_SHs_python / override_color, calls
GaussianRasterizer and clamps.  `use_trained_exp` (training-time exposure) is not supported on this path.
"""
from __future__ import annotations

import math

import torch

from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

_C0 = 0.28209479177387814
_C1 = 0.4886025119029199
_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
       1.445305721320277, -0.5900435899266435]


def eval_sh(deg, sh, dirs):
    """SH -> value, degree <= 3 (utils/sh_utils.py:57-112); sh [...,C,K], dirs [...,3]."""
    assert 0 <= deg <= 3
    res = _C0 * sh[..., 0]
    if deg > 0:
        x, y, z = dirs[..., 0:1], dirs[..., 1:2], dirs[..., 2:3]
        res = res - _C1 * y * sh[..., 1] + _C1 * z * sh[..., 2] - _C1 * x * sh[..., 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = (res + _C2[0] * xy * sh[..., 4] + _C2[1] * yz * sh[..., 5] +
                   _C2[2] * (2.0 * zz - xx - yy) * sh[..., 6] + _C2[3] * xz * sh[..., 7] +
                   _C2[4] * (xx - yy) * sh[..., 8])
            if deg > 2:
                res = (res + _C3[0] * y * (3 * xx - yy) * sh[..., 9] + _C3[1] * xy * z * sh[..., 10] +
                       _C3[2] * y * (4 * zz - xx - yy) * sh[..., 11] +
                       _C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[..., 12] +
                       _C3[4] * x * (4 * zz - xx - yy) * sh[..., 13] + _C3[5] * z * (xx - yy) * sh[..., 14] +
                       _C3[6] * x * (xx - 3 * yy) * sh[..., 15])
    return res


def covariance_python(scaling, scaling_modifier, rotation):
    """strip_symmetric(L L^T), L = R(normalize(q)) diag(mod*s)  (gaussian_model.py:33-37, general_utils.py:64-110)."""
    q = rotation / torch.sqrt((rotation * rotation).sum(-1, keepdim=True))
    r, x, y, z = q.unbind(-1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)
    L = R * (scaling_modifier * scaling)[:, None, :]
    S = L @ L.transpose(1, 2)
    return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], -1)


def render(viewpoint_camera, pc, pipe, bg_color: torch.Tensor, scaling_modifier=1.0, separate_sh=False,
           override_color=None, use_trained_exp=False):
    if use_trained_exp:
        raise NotImplementedError("use_trained_exp is a training-side feature; the observation path passes False")
    if separate_sh:
        raise NotImplementedError("separate_sh needs the 3dgs_accel rasterizer; GSWorld pins dr_aa (separate_sh=False)")
    screenspace_points = torch.zeros_like(pc.get_xyz)
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=tanfovx, tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center, prefiltered=False,
        debug=pipe.debug, antialiasing=pipe.antialiasing)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    means3D, opacity = pc.get_xyz, pc.get_opacity
    scales = rotations = cov3D_precomp = None
    if pipe.compute_cov3D_python:
        cov3D_precomp = covariance_python(pc.get_scaling, scaling_modifier, pc._rotation)
    else:
        scales, rotations = pc.get_scaling, pc.get_rotation
    shs = colors_precomp = None
    if override_color is None:
        if pipe.convert_SHs_python:
            shs_view = pc.get_features.transpose(1, 2).view(-1, 3, (pc.max_sh_degree + 1) ** 2)
            dir_pp = pc.get_xyz - viewpoint_camera.camera_center.repeat(pc.get_features.shape[0], 1)
            dir_pp = dir_pp / dir_pp.norm(dim=1, keepdim=True)
            colors_precomp = torch.clamp_min(eval_sh(pc.active_sh_degree, shs_view, dir_pp) + 0.5, 0.0)
        else:
            shs = pc.get_features
    else:
        colors_precomp = override_color
    rendered_image, radii, depth_image = rasterizer(
        means3D=means3D, means2D=screenspace_points, shs=shs, colors_precomp=colors_precomp, opacities=opacity,
        scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp)
    rendered_image = rendered_image.clamp(0, 1)
    return {"render": rendered_image, "viewspace_points": screenspace_points,
            "visibility_filter": (radii > 0).nonzero(), "radii": radii, "depth": depth_image}


class PipelineParams:
    """Duck type of arguments.PipelineParams (gsworld/utils/gs_utils.py:72-78): all default False."""

    def __init__(self, convert_SHs_python=False, compute_cov3D_python=False, debug=False, antialiasing=False):
        self.convert_SHs_python = convert_SHs_python
        self.compute_cov3D_python = compute_cov3D_python
        self.debug = debug
        self.antialiasing = antialiasing
