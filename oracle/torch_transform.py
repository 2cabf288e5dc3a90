"""CPU ORACLE (TEST INFRASTRUCTURE) — the reference's torch-level path, restated.

Follows, statement by statement, what GSWorldWrapper does around the rasterizer:
  * transform_gs_perlink        gsworld/mani_skill/utils/wrappers/gs_world_wrapper.py:110-162
  * transform_gaussians         gsworld/utils/gs_utils.py:283-385 (+ :169-170, :225-231)
  * extract_rigid_transform     gsworld/utils/pcd_utils.py:224-252
  * cam_maniskill2gs            gs_world_wrapper.py:277-325 (+ scene/cameras.py:80-89,
                                utils/graphics_utils.py:38-71)
  * _render_gsworld             gs_world_wrapper.py:232-275 (deepcopy + isin write-back + render + uint8)
  * render                      submodules/gaussian-splatting/gaussian_renderer/__init__.py:18-128
with the rasterizer call replaced by oracle.rasterize_forward (the CUDA rasterizer has no CPU path).
Pinned by tests/golden/transform_*.npz, generated from the real reference functions
(tests/golden/make_golden.py).  The two ManiSkill quaternion helpers are external
(mani_skill==3.0.0b15, not in the tree): restated from the published pytorch3d algorithm.

This is also the "reference torch transform path on host cores" that bench.py times as cpu_baseline.
Never imported by gsworld_b200/.
"""
from __future__ import annotations

import copy
import math

import numpy as np
import torch

from . import oracle as orc


# ---- external helpers (pytorch3d algorithm) ---------------------------------------------
def quaternion_multiply(a, b):
    aw, ax, ay, az = a.unbind(-1)
    bw, bx, by, bz = b.unbind(-1)
    o = torch.stack(
        (aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
         aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw), -1)
    return torch.where(o[..., :1] < 0, -o, o)


def matrix_to_quaternion(m):
    b = m.shape[:-2]
    m = m.reshape(-1, 3, 3)
    m00, m01, m02 = m[:, 0, 0], m[:, 0, 1], m[:, 0, 2]
    m10, m11, m12 = m[:, 1, 0], m[:, 1, 1], m[:, 1, 2]
    m20, m21, m22 = m[:, 2, 0], m[:, 2, 1], m[:, 2, 2]
    tr = torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22, 1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22], -1)
    q_abs = torch.where(tr > 0, torch.sqrt(tr.clamp_min(0)), torch.zeros_like(tr))
    rows = [
        torch.stack([q_abs[:, 0] ** 2, m21 - m12, m02 - m20, m10 - m01], -1),
        torch.stack([m21 - m12, q_abs[:, 1] ** 2, m10 + m01, m02 + m20], -1),
        torch.stack([m02 - m20, m10 + m01, q_abs[:, 2] ** 2, m12 + m21], -1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[:, 3] ** 2], -1),
    ]
    cand = torch.stack(rows, 1) / (2.0 * torch.maximum(q_abs, torch.tensor(0.1, dtype=q_abs.dtype, device=q_abs.device))[..., None])
    best = q_abs.argmax(-1)
    return cand[torch.arange(m.shape[0], device=m.device), best].reshape(b + (4,))


def pose7_to_matrix(pose):
    p, q = pose[:, :3], pose[:, 3:7]
    q = q / q.norm(dim=-1, keepdim=True)
    w, x, y, z = q.unbind(-1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1).reshape(-1, 3, 3)
    M = torch.eye(4, dtype=pose.dtype, device=pose.device).repeat(pose.shape[0], 1, 1)
    M[:, :3, :3] = R
    M[:, :3, 3] = p
    return M


# ---- gs_utils.py / pcd_utils.py ---------------------------------------------------------
def inverse_sigmoid(x):
    return torch.log(x / (1 - x))


def extract_rigid_transform(M):
    single = M.ndim == 2
    if single:
        M = M.unsqueeze(0)
    A, t = M[:, :3, :3], M[:, :3, 3]
    U, S, Vh = torch.linalg.svd(A)
    scales = S.mean(dim=1)
    R = U @ Vh
    Mr = torch.eye(4, dtype=M.dtype, device=M.device).repeat(M.shape[0], 1, 1)
    Mr[:, :3, :3] = R
    Mr[:, :3, 3] = t
    return (Mr[0], scales[0], R[0], t[0]) if single else (Mr, scales, R, t)


def _rot_quat(converted, r):
    norm = r.norm(dim=-1, keepdim=True)
    return quaternion_multiply(converted, r / norm) * norm


def transform_gaussians(g, selected_indices, scale=None, rot_mat=None, translation=None):
    """scale -> rotate -> translate on the selected rows; (B,·) batch semantics of the reference."""
    xyz = g._xyz[selected_indices]
    scaling = g._scaling[selected_indices]
    rotation = g._rotation[selected_indices]
    opac = g._opacity[selected_indices]
    if scale is not None:
        if scale.dim() == 0:
            xyz = xyz * scale
            scaling = inverse_sigmoid(torch.exp(scaling) * scale)
        else:
            B, N = scale.size(0), xyz.size(0)
            xyz = xyz.unsqueeze(0).expand(B, N, 3) * scale[:, None, None]
            scaling = inverse_sigmoid(torch.exp(scaling.unsqueeze(0).expand(B, N, 3)) * scale[:, None, None])
    if rot_mat is not None:
        quat_r = matrix_to_quaternion(rot_mat)
        if rot_mat.size(0) == 1:
            if xyz.dim() == 2:
                xyz = torch.bmm(rot_mat.expand(xyz.size(0), 3, 3), xyz.unsqueeze(-1)).squeeze(-1)
            else:
                B, N = xyz.shape[:2]
                xyz = torch.matmul(rot_mat.expand(B, N, 3, 3), xyz.unsqueeze(-1)).squeeze(-1)
        elif rot_mat.size(0) == xyz.size(0) and xyz.dim() == 2:
            xyz = torch.bmm(rot_mat, xyz.unsqueeze(-1)).squeeze(-1)
        else:
            B, N = rot_mat.size(0), xyz.size(-2)
            pts = xyz if xyz.dim() == 3 else xyz.unsqueeze(0).expand(B, N, 3)
            xyz = torch.matmul(rot_mat.unsqueeze(1).expand(B, N, 3, 3), pts.unsqueeze(-1)).squeeze(-1)
        if rotation.numel() > 0:
            if quat_r.size(0) == rotation.size(0) and xyz.dim() == 2:
                rotation = _rot_quat(quat_r, rotation)
            else:
                B, N = quat_r.size(0), rotation.size(0)
                rotation = _rot_quat(quat_r[:, None, :].expand(B, N, 4).reshape(B * N, 4),
                                     rotation[None].expand(B, N, 4).reshape(B * N, 4)).view(B, N, 4)
    if translation is not None:
        if translation.dim() == 1:
            xyz = xyz + translation
        elif xyz.dim() == 2:
            xyz = xyz.unsqueeze(0).expand(translation.size(0), xyz.size(0), 3) + translation[:, None, :]
        else:
            xyz = xyz + translation[:, None, :]
    return xyz, scaling, rotation, opac


# ---- gs_world_wrapper.py ----------------------------------------------------------------
class RefPath:
    """The reference wrapper's per-step logic over a duck-typed env (CPU torch)."""

    def __init__(self, env, scene, sim2gs_arm, link_labels, scan_link_mats, obj_sim2gs, obj_labels,
                 obj_offset, obj_scale, xarm_offset=None, device="cpu", rasterizer=None):
        # device="cpu": the oracle.  device="cuda" + a GPU `rasterizer(cam, model) -> color (3,H,W) tensor`:
        # the same statements on the device, used only by bench.py's restated-reference GPU comparison point.
        self.env = env
        self.num_envs = env.num_envs
        self.scene = scene
        self.device = torch.device(device)
        self.rasterizer = rasterizer
        self.S = torch.as_tensor(sim2gs_arm, dtype=torch.float32, device=self.device)
        self.gs_semantics = dict(link_labels)
        self.gs_link_pose_mats = [torch.as_tensor(m, dtype=torch.float32, device=self.device).reshape(1, 4, 4)
                                  for m in scan_link_mats]
        self.obj_sim2gs, self.obj_labels = obj_sim2gs, obj_labels
        self.obj_offset, self.obj_scale = obj_offset, obj_scale
        self.xarm_offset = xarm_offset
        self.rigid_sim2real, self.scale_sim2real, _, _ = extract_rigid_transform(self.S)
        self.gs_movable_pts = {}

    def link_transforms(self):
        """(B,L,4,4) S @ M_link @ inv(M_scan) @ inv(S)   (gs_world_wrapper.py:115-120)."""
        out = []
        for idx, link in enumerate(self.env.agent.robot.get_links()):
            m = link.pose.to_transformation_matrix().to(self.device, torch.float32).clone()
            if self.xarm_offset is not None:
                for i in range(3):
                    m[:, i, 3] += self.xarm_offset[i]
            out.append(self.S @ m @ torch.linalg.inv(self.gs_link_pose_mats[idx]) @ torch.linalg.inv(self.S))
        return torch.stack(out, 1)

    def object_transforms(self):
        """dict name -> (rigid (B,4,4), kappa (B,))   (gs_world_wrapper.py:135-156)."""
        out = {}
        actors = self.env.get_state_dict().get("actors", {})
        for k in actors:
            if k not in self.obj_sim2gs:
                continue
            mat = pose7_to_matrix(actors[k][:, :7].to(self.device, torch.float32))
            if k in self.obj_offset:
                for i in range(3):
                    mat[:, i, 3] += self.obj_offset[k][i]
            full = self.S @ mat @ torch.linalg.inv(torch.as_tensor(self.obj_sim2gs[k], dtype=torch.float32,
                                                                   device=self.device))
            rigid, scale, _, _ = extract_rigid_transform(full)
            out[k] = (rigid, scale * self.obj_scale[k])
        return out

    def transform_gs_perlink(self):
        g = self.scene
        sem = g._semantics.long().squeeze(-1)
        lt = self.link_transforms()
        for idx, link in enumerate(self.env.agent.robot.get_links()):
            target = torch.as_tensor(self.gs_semantics[link.name], device=sem.device).long().reshape(-1)
            mask = torch.isin(sem, target)
            self.gs_movable_pts[link.name] = transform_gaussians(
                g, torch.where(mask)[0], None, lt[:, idx, :3, :3], lt[:, idx, :3, 3])
        for k, (rigid, kappa) in self.object_transforms().items():
            idxs = torch.where(g._semantics == self.obj_labels[k])[0]
            self.gs_movable_pts[k] = transform_gaussians(g, idxs, kappa, rigid[:, :3, :3], rigid[:, :3, 3])
            self.gs_semantics[k] = self.obj_labels[k]
        return self.gs_movable_pts

    def cameras(self, strict=True):
        """cam name -> list over envs of dict(view, proj, campos, tanfovx, tanfovy, H, W).
        strict=True reproduces the reference's use of env 0 for every env (gs_world_wrapper.py:280-281)."""
        params = self.env.get_sensor_params()
        out = {}
        for cam, p in params.items():
            H, W = self.env.sensor_hw[cam]
            per_env = []
            for e in range(self.num_envs):
                src = 0 if strict else e
                ext = p["extrinsic_cv"][src].to(self.device, torch.float32)
                K = p["intrinsic_cv"][src].to(self.device, torch.float32)
                w2c = torch.vstack([ext, torch.tensor([0, 0, 0, 1], dtype=ext.dtype, device=ext.device)])
                c2w = torch.linalg.inv(w2c)
                fovx = 2 * torch.arctan(W / (2 * K[0, 0]))
                fovy = 2 * torch.arctan(H / (2 * K[1, 1]))
                c2w[:3, 3] = c2w[:3, 3] * self.scale_sim2real
                real_w2c = torch.linalg.inv(self.rigid_sim2real @ c2w)
                R = real_w2c[:3, :3].T.cpu().numpy()
                T = real_w2c[:3, 3].cpu().numpy()
                per_env.append(camera_tensors(R, T, float(fovx), float(fovy), H, W))
            out[cam] = per_env
        return out

    def render_gsworld(self, strict=True, return_aux=False, debug=False):
        """{cam: uint8 (B,H,W,3)} exactly as _render_gsworld builds it (gs_world_wrapper.py:232-275)."""
        cams = self.cameras(strict)
        outs, aux = {}, {}
        base_sem = self.scene._semantics.long().squeeze(-1)
        for cam, per_env in cams.items():
            frames, auxes = [], []
            for i in range(self.num_envs):
                g = copy.deepcopy(self.scene)
                for key, pts in self.gs_movable_pts.items():
                    mask = torch.isin(base_sem, torch.as_tensor(self.gs_semantics[key],
                                                                device=base_sem.device).long().reshape(-1))
                    if pts[0].shape[0] == self.num_envs:
                        g._xyz[mask] = pts[0][i]
                    if pts[1].shape[0] == self.num_envs:
                        g._scaling[mask] = pts[1][i]
                    if pts[2].shape[0] == self.num_envs:
                        g._rotation[mask] = pts[2][i]
                    if pts[3].shape[0] == self.num_envs:
                        g._opacity[mask] = pts[3][i]
                if self.rasterizer is not None:  # device path: render() + the wrapper's uint8 conversion in torch
                    rendering = self.rasterizer(per_env[i], g).clamp(0, 1)
                    gs_render = rendering.permute(1, 2, 0).unsqueeze(dim=0)
                    frames.append((gs_render * 255).clamp(0, 255).to(torch.uint8))
                    continue
                r = render_ref(per_env[i], g, debug=debug)
                frames.append(orc.to_uint8_hwc(np.clip(r["color"], 0.0, 1.0))[None])
                auxes.append(r)
            outs[cam] = torch.vstack(frames) if self.rasterizer is not None else np.concatenate(frames, 0)
            aux[cam] = auxes
        return (outs, aux) if return_aux else outs


def camera_tensors(R, T, fovx, fovy, H, W, znear=0.01, zfar=100.0):
    """scene/cameras.py:80-89 + utils/graphics_utils.py:38-71 -> transposed matrices."""
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = R.transpose()
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    C2W = np.linalg.inv(Rt)
    Rt = np.float32(np.linalg.inv(C2W))  # trans=0, scale=1 round trip of getWorld2View2
    wv = torch.tensor(Rt).transpose(0, 1)
    th, tw = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = th * znear, tw * znear
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    proj = P.transpose(0, 1)
    full = wv.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0)
    center = wv.inverse()[3, :3]
    return dict(view=wv.contiguous(), proj=full.contiguous(), campos=center.contiguous(),
                tanfovx=math.tan(fovx * 0.5), tanfovy=math.tan(fovy * 0.5), H=H, W=W, FoVx=fovx, FoVy=fovy)


def render_ref(cam, g, bg=(0.0, 0.0, 0.0), antialiasing=False, debug=False):
    """gaussian_renderer.render with getters' activations, rasterizer = CPU oracle."""
    scales = torch.exp(g._scaling)
    rots = torch.nn.functional.normalize(g._rotation)
    opac = torch.sigmoid(g._opacity)
    shs = torch.cat((g._features_dc, g._features_rest), dim=1)
    return orc.rasterize_forward(
        g._xyz.numpy(), opac.reshape(-1).numpy(), cam["view"].numpy(), cam["proj"].numpy(), cam["campos"].numpy(),
        cam["tanfovx"], cam["tanfovy"], cam["H"], cam["W"], shs=shs.numpy(), scales=scales.numpy(),
        rotations=rots.numpy(), sh_degree=3, bg=bg, antialiasing=antialiasing, debug=debug)
