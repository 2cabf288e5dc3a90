"""Generates tests/golden/*.npz by RUNNING THE REAL REFERENCE FUNCTIONS (container only).

Run from the repo root:  python tests/golden/make_golden.py
Needs /root/reference (read-only).  The GPU box never runs this; it only reads the committed .npz.

What is executed from the reference, loaded by file path (no package __init__, so none of the
ManiSkill / SAPIEN / open3d / plyfile imports of the surrounding packages run):
  gsworld/utils/gs_utils.py        transform_gaussians, inverse_sigmoid
  gsworld/utils/pcd_utils.py       extract_rigid_transform
  submodules/gaussian-splatting/utils/sh_utils.py        eval_sh
  submodules/gaussian-splatting/utils/general_utils.py   build_scaling_rotation, strip_symmetric
  submodules/gaussian-splatting/utils/graphics_utils.py  getWorld2View2, getProjectionMatrix
Stubs: `mani_skill.utils.geometry.rotation_conversions` (absent external dependency) is replaced by
the restatement in oracle/torch_transform.py, so the quaternion golden values depend on that
restatement (parity of those two helpers stays unpinned); `open3d`/`plyfile` are empty modules;
`device="cuda"` in general_utils is redirected to CPU.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
OUT = os.path.join(REPO, "tests", "golden")
sys.path.insert(0, REPO)

from oracle import torch_transform as ot  # noqa: E402


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _stub_modules():
    rc = types.ModuleType("mani_skill.utils.geometry.rotation_conversions")
    rc.matrix_to_quaternion = ot.matrix_to_quaternion
    rc.quaternion_multiply = ot.quaternion_multiply
    for name in ("mani_skill", "mani_skill.utils", "mani_skill.utils.geometry"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["mani_skill.utils.geometry.rotation_conversions"] = rc
    o3d = types.ModuleType("open3d")
    sys.modules.setdefault("open3d", o3d)
    ply = types.ModuleType("plyfile")
    ply.PlyData = ply.PlyElement = object
    sys.modules.setdefault("plyfile", ply)


class _CpuZeros:
    """general_utils hard-codes device='cuda'; run the same statements on CPU."""

    def __enter__(self):
        self._z = torch.zeros

        def zeros(*a, **k):
            k.pop("device", None)
            return self._z(*a, **k)

        torch.zeros = zeros

    def __exit__(self, *exc):
        torch.zeros = self._z


def main():
    _stub_modules()
    gs_utils = _load("ref_gs_utils", f"{REF}/gsworld/utils/gs_utils.py")
    pcd_utils = _load("ref_pcd_utils", f"{REF}/gsworld/utils/pcd_utils.py")
    sh_utils = _load("ref_sh_utils", f"{REF}/submodules/gaussian-splatting/utils/sh_utils.py")
    general = _load("ref_general_utils", f"{REF}/submodules/gaussian-splatting/utils/general_utils.py")
    graphics = _load("ref_graphics_utils", f"{REF}/submodules/gaussian-splatting/utils/graphics_utils.py")

    g = torch.Generator().manual_seed(20261106)
    rn = lambda *s: torch.randn(*s, generator=g)

    def rand_rot(B):
        q = rn(B, 4)
        q = q / q.norm(dim=-1, keepdim=True)
        w, x, y, z = q.unbind(-1)
        return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                            2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                            2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1).reshape(B, 3, 3)

    # ---- transform_gaussians: link case (no scale) and object case (scale), B in {1, 3} ----
    N, n = 64, 23
    model = types.SimpleNamespace(
        _xyz=rn(N, 3), _scaling=rn(N, 3) * 0.5 - 5.0, _rotation=rn(N, 4), _opacity=rn(N, 1, 1))
    idx = torch.randperm(N, generator=g)[:n].sort().values
    cases = {}
    for tag, B, with_scale in (("link_b1", 1, False), ("link_b3", 3, False), ("obj_b1", 1, True), ("obj_b3", 3, True)):
        R, t = rand_rot(B), rn(B, 3)
        scale = (torch.rand(B, generator=g) * 0.6 + 0.7) if with_scale else None
        xyz, sc, rot, op = gs_utils.transform_gaussians(model, idx, scale=scale, rot_mat=R, translation=t)
        cases[tag] = dict(R=R, t=t, scale=scale if with_scale else torch.zeros(0), xyz=xyz, scaling=sc,
                          rotation=rot, opacity=op)
    flat = {"model_xyz": model._xyz, "model_scaling": model._scaling, "model_rotation": model._rotation,
            "model_opacity": model._opacity, "idx": idx}
    for tag, d in cases.items():
        for k, v in d.items():
            flat[f"{tag}__{k}"] = v
    np.savez_compressed(os.path.join(OUT, "transform_gaussians.npz"), **{k: v.numpy() for k, v in flat.items()})

    # ---- extract_rigid_transform on similarity matrices (single and batched) ----
    B = 5
    M = torch.eye(4).repeat(B, 1, 1)
    s = torch.rand(B, generator=g) * 2 + 0.3
    M[:, :3, :3] = rand_rot(B) * s[:, None, None]
    M[:, :3, 3] = rn(B, 3)
    Mr, sc, R, t = pcd_utils.extract_rigid_transform(M)
    M1r, s1, R1, t1 = pcd_utils.extract_rigid_transform(M[0])
    np.savez_compressed(os.path.join(OUT, "extract_rigid.npz"), M=M.numpy(), M_rigid=Mr.numpy(), scales=sc.numpy(),
                        R=R.numpy(), t=t.numpy(), single_M_rigid=M1r.numpy(), single_scale=s1.numpy(),
                        single_R=R1.numpy(), single_t=t1.numpy())

    # ---- SH colour (render()'s convert_SHs_python branch, gaussian_renderer/__init__.py:76-80) and cov3D ----
    P = 96
    xyz = rn(P, 3)
    campos = rn(3)
    feats = torch.cat([rn(P, 1, 3), rn(P, 15, 3) * 0.1], 1)  # (P,16,3) = get_features layout
    out = {"xyz": xyz.numpy(), "campos": campos.numpy(), "features": feats.numpy()}
    for deg in range(4):
        shs_view = feats.transpose(1, 2).view(-1, 3, 16)
        d = xyz - campos.repeat(P, 1)
        d = d / d.norm(dim=1, keepdim=True)
        rgb = torch.clamp_min(sh_utils.eval_sh(deg, shs_view, d) + 0.5, 0.0)
        out[f"rgb_deg{deg}"] = rgb.numpy()
    scales = torch.exp(rn(P, 3) * 0.5 - 4.0)
    rots = rn(P, 4)  # un-normalised: build_rotation normalises internally (general_utils.py:79-81)
    with _CpuZeros():
        L = general.build_scaling_rotation(1.0 * scales, rots)
        cov = general.strip_symmetric(L @ L.transpose(1, 2))
        L2 = general.build_scaling_rotation(1.7 * scales, rots)
        cov2 = general.strip_symmetric(L2 @ L2.transpose(1, 2))
    out.update(scales=scales.numpy(), rots=rots.numpy(), cov3d_mod1=cov.numpy(), cov3d_mod1p7=cov2.numpy())
    np.savez_compressed(os.path.join(OUT, "sh_cov.npz"), **out)

    # ---- camera matrices ----
    R = rand_rot(1)[0].numpy().astype(np.float64)
    T = rn(3).numpy().astype(np.float64)
    w2v = graphics.getWorld2View2(R, T)
    w2v_ts = graphics.getWorld2View2(R, T, np.array([0.1, -0.2, 0.3]), 1.5)
    fx, fy = 0.9, 0.7
    proj = graphics.getProjectionMatrix(znear=0.01, zfar=100.0, fovX=fx, fovY=fy)
    np.savez_compressed(os.path.join(OUT, "camera.npz"), R=R, T=T, w2v=w2v, w2v_ts=w2v_ts, fovx=fx, fovy=fy,
                        proj=proj.numpy())
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
