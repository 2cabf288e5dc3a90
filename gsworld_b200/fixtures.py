"""Seeded stand-ins for what the reference gets from ManiSkill/SAPIEN and from HuggingFace assets.

The reference has no tests or fixtures for the observation path; the real scans are not
in its tree (README.md:6).  This module provides
  * GaussianScene      – duck-type of the merged Semantic3DGSWrapper after load
                         (semantic_3dgs_wrapper.py:100-167; tensor layouts of SURVEY.md Appendix A),
  * make_scene(...)    – the synthetic scene of SURVEY.md §8(d) / BASELINE.md §3.3,
  * FakeEnv            – the slice of the ManiSkill BaseEnv API GSWorldWrapper touches
                         (gs_world_wrapper.py:71,115-116,135-139,238,288-289).
Nothing here is on the timed path.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from types import SimpleNamespace
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from . import rig_tables as rt

# --------------------------------------------------------------------------------------
# kinematic tables: (link name, parent index, origin xyz, origin rpy, joint type, axis, qpos index)
# Approximate Franka FR3 + UMI gripper and UFactory xArm6 + gripper chains, in URDF link order.
# --------------------------------------------------------------------------------------
_H = math.pi / 2
_CHAINS = {
    "fr3_umi": [
        ("base", -1, (0, 0, 0), (0, 0, 0), "fixed", None, -1),
        ("fr3_link0", 0, (0, 0, 0), (0, 0, 0), "fixed", None, -1),
        ("fr3_link1", 1, (0, 0, 0.333), (0, 0, 0), "rev", (0, 0, 1), 0),
        ("fr3_link2", 2, (0, 0, 0), (-_H, 0, 0), "rev", (0, 0, 1), 1),
        ("fr3_link3", 3, (0, -0.316, 0), (_H, 0, 0), "rev", (0, 0, 1), 2),
        ("fr3_link4", 4, (0.0825, 0, 0), (_H, 0, 0), "rev", (0, 0, 1), 3),
        ("fr3_link5", 5, (-0.0825, 0.384, 0), (-_H, 0, 0), "rev", (0, 0, 1), 4),
        ("fr3_link6", 6, (0, 0, 0), (_H, 0, 0), "rev", (0, 0, 1), 5),
        ("fr3_link7", 7, (0.088, 0, 0), (_H, 0, 0), "rev", (0, 0, 1), 6),
        ("fr3_link8", 8, (0, 0, 0.107), (0, 0, 0), "fixed", None, -1),
        ("fr3_hand", 9, (0, 0, 0), (0, 0, -math.pi / 4), "fixed", None, -1),
        ("fr3_hand_tcp", 10, (0, 0, 0.1034), (0, 0, 0), "fixed", None, -1),
        ("fr3_leftfinger", 10, (0, 0, 0.0584), (0, 0, 0), "pri", (0, 1, 0), 7),
        ("fr3_rightfinger", 10, (0, 0, 0.0584), (0, 0, 0), "pri", (0, -1, 0), 8),
    ],
    "xarm6_uf_gripper": [
        ("world", -1, (0, 0, 0), (0, 0, 0), "fixed", None, -1),
        ("link_base", 0, (0, 0, 0), (0, 0, 0), "fixed", None, -1),
        ("link1", 1, (0, 0, 0.267), (0, 0, 0), "rev", (0, 0, 1), 0),
        ("link2", 2, (0, 0, 0), (-_H, 0, 0), "rev", (0, 0, 1), 1),
        ("link3", 3, (0.0535, -0.2845, 0), (0, 0, 0), "rev", (0, 0, 1), 2),
        ("link4", 4, (0.0775, 0.3425, 0), (-_H, 0, 0), "rev", (0, 0, 1), 3),
        ("link5", 5, (0, 0, 0), (_H, 0, 0), "rev", (0, 0, 1), 4),
        ("link6", 6, (0.076, 0.097, 0), (-_H, 0, 0), "rev", (0, 0, 1), 5),
        ("xarm_gripper_base_link", 7, (0, 0, 0), (0, 0, 0), "fixed", None, -1),
        ("left_outer_knuckle", 8, (0, 0.035, 0.059098), (0, 0, 0), "rev", (1, 0, 0), 6),
        ("left_finger", 9, (0, 0.035465, 0.042039), (0, 0, 0), "rev", (-1, 0, 0), 7),
        ("left_inner_knuckle", 8, (0, 0.02, 0.074098), (0, 0, 0), "rev", (1, 0, 0), 8),
        ("right_outer_knuckle", 8, (0, -0.035, 0.059098), (0, 0, 0), "rev", (-1, 0, 0), 9),
        ("right_finger", 12, (0, -0.035465, 0.042039), (0, 0, 0), "rev", (1, 0, 0), 10),
        ("right_inner_knuckle", 8, (0, -0.02, 0.074098), (0, 0, 0), "rev", (-1, 0, 0), 11),
        ("xarm_hand_tcp", 8, (0, 0, 0.172), (0, 0, 0), "fixed", None, -1),
    ],
}
_EE_LINK = {"fr3_umi": "fr3_hand_tcp", "xarm6_uf_gripper": "xarm_gripper_base_link"}
_FAMILY = {"fr3_umi": "fr3", "xarm6_uf_gripper": "xarm"}
_WRIST_MOUNT = {"fr3_umi": rt.WRIST2EEF, "xarm6_uf_gripper": rt.XARM_WRIST2BASE}
_THIRD_MOUNT = {"fr3_umi": rt.RIGHT2BASE, "xarm6_uf_gripper": rt.XARM_RIGHT2BASE}


def link_names(robot: str) -> List[str]:
    return [c[0] for c in _CHAINS[robot]]


def _rpy_to_mat(r, p, y):
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def _axis_angle(axis, ang):
    """Rodrigues, batched over `ang` (B,) -> (B,3,3)."""
    a = np.asarray(axis, np.float64)
    a = a / np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    s, c = np.sin(ang)[:, None, None], np.cos(ang)[:, None, None]
    return np.eye(3)[None] + s * K[None] + (1 - c) * (K @ K)[None]


def forward_kinematics(robot: str, qpos: np.ndarray) -> np.ndarray:
    """qpos (B,Q) -> link poses (B,L,4,4) float32, URDF link order, robot base at the sim origin."""
    qpos = np.atleast_2d(np.asarray(qpos, np.float64))
    B = qpos.shape[0]
    chain = _CHAINS[robot]
    out = np.zeros((B, len(chain), 4, 4))
    for i, (_n, parent, xyz, rpy, jt, axis, qi) in enumerate(chain):
        T0 = np.eye(4)
        T0[:3, :3] = _rpy_to_mat(*rpy)
        T0[:3, 3] = xyz
        Tj = np.tile(np.eye(4), (B, 1, 1))
        if jt == "rev":
            Tj[:, :3, :3] = _axis_angle(axis, qpos[:, qi])
        elif jt == "pri":
            Tj[:, :3, 3] = np.asarray(axis)[None] * qpos[:, qi][:, None]
        local = T0[None] @ Tj
        out[:, i] = local if parent < 0 else out[:, parent] @ local
    return out.astype(np.float32)


# --------------------------------------------------------------------------------------
# scene
# --------------------------------------------------------------------------------------
class GaussianScene:
    """Loaded-model duck type (raw, un-activated parameters; none require grad)."""

    def __init__(self, xyz, features_dc, features_rest, opacity, scaling, rotation, semantics):
        self._xyz = xyz                      # (N,3)
        self._features_dc = features_dc      # (N,1,3)
        self._features_rest = features_rest  # (N,15,3)
        self._opacity = opacity              # (N,1,1) logit   (semantic_3dgs_wrapper.py:154)
        self._scaling = scaling              # (N,3) log
        self._rotation = rotation            # (N,4) wxyz, un-normalised
        self._semantics = semantics          # (N,1) float, integer valued, -1 background
        self.max_sh_degree = 3
        self.active_sh_degree = 3

    # getters with the activations of scene/gaussian_model.py:39-47,102-130
    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_scaling(self):
        return torch.exp(self._scaling)

    @property
    def get_rotation(self):
        return torch.nn.functional.normalize(self._rotation)

    @property
    def get_opacity(self):
        return torch.sigmoid(self._opacity)

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)

    @property
    def get_features_dc(self):
        return self._features_dc

    @property
    def get_features_rest(self):
        return self._features_rest

    def to(self, device):
        return GaussianScene(*[t.to(device) for t in (self._xyz, self._features_dc, self._features_rest,
                                                      self._opacity, self._scaling, self._rotation,
                                                      self._semantics)])

    @property
    def num_points(self):
        return int(self._xyz.shape[0])


@dataclass
class SceneSpec:
    robot: str = "fr3_umi"
    n_total: int = 200_000
    per_link: int = 500
    objects: Tuple[str, ...] = ("006_mustard_bottle", "snack_box")
    per_object: int = 7000
    seed: int = 1234


def _apply_sim(S: np.ndarray, pts: np.ndarray) -> np.ndarray:
    return pts @ S[:3, :3].T.astype(np.float64) + S[:3, 3].astype(np.float64)


def make_scene(spec: SceneSpec, device="cpu") -> GaussianScene:
    """Synthetic merged scene (background + labelled robot links + labelled objects), GS frame."""
    rng = np.random.default_rng(spec.seed)
    fam = _FAMILY[spec.robot]
    S = rt.SIM2GS_ROBOT[fam].astype(np.float64)
    s_arm = float(np.cbrt(abs(np.linalg.det(S[:3, :3]))))
    chain = _CHAINS[spec.robot]
    labels_of = rt.LINK_LABELS[fam]
    scan = forward_kinematics(spec.robot, rt.ROBOT_SCAN_QPOS[spec.robot][None])[0].astype(np.float64)

    xyz_parts, ls_parts, lab_parts = [], [], []
    # robot links: capsule 6cm x 6cm x 25cm in the link's scan-pose frame
    for li, c in enumerate(chain):
        lab = labels_of[c[0]]
        labs = lab if isinstance(lab, list) else [lab]
        n = spec.per_link
        local = np.stack([rng.uniform(-0.03, 0.03, n), rng.uniform(-0.03, 0.03, n), rng.uniform(-0.125, 0.125, n)], 1)
        sim = local @ scan[li, :3, :3].T + scan[li, :3, 3]
        xyz_parts.append(_apply_sim(S, sim))
        ls_parts.append(rng.normal(math.log(0.004), 0.5, (n, 3)) + math.log(s_arm))
        lab_parts.append(np.asarray(labs)[rng.integers(0, len(labs), n)].astype(np.float64))
    # objects: 10 cm box in the object's own frame, stored in the object's GS frame
    for name in spec.objects:
        So = rt.SIM2GS_OBJECT[name].astype(np.float64)
        s_obj = float(np.cbrt(abs(np.linalg.det(So[:3, :3]))))
        n = spec.per_object
        local = rng.uniform(-0.05, 0.05, (n, 3))
        xyz_parts.append(_apply_sim(So, local))
        ls_parts.append(rng.normal(math.log(0.004), 0.5, (n, 3)) + math.log(s_obj))
        lab_parts.append(np.full(n, float(rt.OBJECT_LABELS[name])))
    n_mov = sum(p.shape[0] for p in xyz_parts)
    n_bg = spec.n_total - n_mov
    if n_bg < 0:
        raise ValueError(f"n_total={spec.n_total} smaller than movable set {n_mov}")
    n_slab = int(round(0.7 * n_bg))
    n_shell = n_bg - n_slab
    slab = np.stack([rng.uniform(-0.2, 1.2, n_slab), rng.uniform(-0.7, 0.7, n_slab), rng.uniform(-0.05, 0.6, n_slab)], 1)
    d = rng.normal(size=(n_shell, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True) + 1e-12
    shell = d * rng.uniform(1.5, 4.0, (n_shell, 1))
    bg_xyz = _apply_sim(S, np.concatenate([slab, shell], 0))
    bg_ls = np.concatenate([rng.normal(math.log(0.004), 0.5, (n_slab, 3)),
                            rng.normal(math.log(0.02), 0.6, (n_shell, 3))], 0) + math.log(s_arm)
    xyz = np.concatenate([bg_xyz] + xyz_parts, 0)
    ls = np.concatenate([bg_ls] + ls_parts, 0)
    lab = np.concatenate([np.full(n_bg, -1.0)] + lab_parts, 0)
    # interleave so movable Gaussians are scattered through the arrays like a real labelled scan
    perm = rng.permutation(spec.n_total)
    xyz, ls, lab = xyz[perm], ls[perm], lab[perm]
    N = spec.n_total
    quat = rng.normal(size=(N, 4))
    opac = rng.normal(1.0, 2.0, (N, 1, 1))
    dc = rng.normal(0.0, 1.0, (N, 1, 3))
    rest = rng.normal(0.0, 0.1, (N, 15, 3))
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(device)
    return GaussianScene(f(xyz), f(dc), f(rest), f(opac), f(ls), f(quat), f(lab[:, None]))


# --------------------------------------------------------------------------------------
# fake env
# --------------------------------------------------------------------------------------
class _Pose:
    def __init__(self, mat):
        self._m = mat

    def to_transformation_matrix(self):
        return self._m


class _Link:
    def __init__(self, name, owner, idx):
        self.name, self._o, self._i = name, owner, idx

    @property
    def pose(self):
        return _Pose(self._o._link_mats[:, self._i].clone())


class _Robot:
    def __init__(self, name, owner):
        self.name = name
        self._links = [_Link(n, owner, i) for i, n in enumerate(link_names(name))]

    def get_links(self):
        return self._links


class FakeEnv:
    """Minimal ManiSkill-like env: seeded joint/object motion, two sensors, no physics, no renderer.

    `sensor_hw[cam] = (H, W)` replaces get_sensor_images() (SAPIEN Vulkan render; unavailable on this box).
    """

    def __init__(self, robot="fr3_umi", num_envs=1, cameras=("right_cam", "wrist_cam"), hw=(128, 128),
                 objects=("006_mustard_bottle", "snack_box"), seed=0, device="cpu"):
        self.num_envs = num_envs
        self.device = torch.device(device)
        self.robot_uid = robot
        self.agent = SimpleNamespace(uid=robot, robot=_Robot(robot, self))
        self.cameras = tuple(cameras)
        self.static_cameras = tuple(c for c in self.cameras if c != "wrist_cam")  # world-fixed mounts
        self.sensor_hw = {c: tuple(hw) for c in self.cameras}
        self.objects = tuple(objects)
        self._rng = np.random.default_rng(seed)
        self._q0 = np.tile(rt.ROBOT_TASK_INIT_QPOS[robot][None].astype(np.float64), (num_envs, 1))
        self._q = self._q0.copy()
        self._obj = {}
        self.reset(seed=seed)

    # -- gym-ish surface --
    @property
    def unwrapped(self):
        return self

    @property
    def base_env(self):
        return self

    def reset(self, seed=None, options=None):
        if seed is not None:
            self._rng = np.random.default_rng(seed)
        nq = self._q0.shape[1]
        self._q = self._q0 + self._rng.uniform(-0.3, 0.3, (self.num_envs, nq)) * (np.arange(nq) < 7)[None]
        for k, name in enumerate(self.objects):
            p = np.stack([self._rng.uniform(0.4, 0.7, self.num_envs), self._rng.uniform(-0.2, 0.2, self.num_envs),
                          np.full(self.num_envs, 0.03 + 0.01 * k)], 1)
            yaw = self._rng.uniform(-math.pi, math.pi, self.num_envs)
            q = np.stack([np.cos(yaw / 2), np.zeros_like(yaw), np.zeros_like(yaw), np.sin(yaw / 2)], 1)
            self._obj[name] = np.concatenate([p, q, np.zeros((self.num_envs, 6))], 1)
        self._refresh()
        return {"sensor_data": {c: {} for c in self.cameras}}, {}

    def step(self, action=None):
        nq = self._q.shape[1]
        self._q = self._q + self._rng.uniform(-0.02, 0.02, self._q.shape) * (np.arange(nq) < 7)[None]
        for name in self.objects:
            self._obj[name][:, :2] += self._rng.uniform(-0.002, 0.002, (self.num_envs, 2))
        self._refresh()
        obs = {"sensor_data": {c: {} for c in self.cameras}}
        z = torch.zeros(self.num_envs, device=self.device)
        return obs, z, z.bool(), z.bool(), {}

    def _refresh(self):
        self._link_mats = torch.from_numpy(forward_kinematics(self.robot_uid, self._q)).to(self.device)

    # -- what render_current_step asks of a ManiSkill BaseEnv (gs_world_wrapper.py:200-230) --
    def get_info(self):
        return {"success": torch.zeros(self.num_envs, dtype=torch.bool, device=self.device)}

    def get_obs(self, info=None):
        return {"sensor_data": {c: {} for c in self.cameras}}

    def get_reward(self, obs=None, action=None, info=None):
        return torch.zeros(self.num_envs, device=self.device)

    # -- the calls GSWorldWrapper makes --
    def get_state_dict(self):
        return {"actors": {k: torch.from_numpy(v.astype(np.float32)).to(self.device) for k, v in self._obj.items()}}

    def get_sensor_params(self):
        out = {}
        names = link_names(self.robot_uid)
        ee = names.index(_EE_LINK[self.robot_uid])
        for cam in self.cameras:
            H, W = self.sensor_hw[cam]
            K = torch.tensor(rt.RS_D435I_RGB_K).clone()
            K[0, :] *= W / rt.RS_D435I_RES[0]
            K[1, :] *= H / rt.RS_D435I_RES[1]
            if cam == "wrist_cam":
                c2w = self._link_mats[:, ee].cpu() @ torch.tensor(_WRIST_MOUNT[self.robot_uid])
            else:
                c2w = torch.tensor(_THIRD_MOUNT[self.robot_uid])[None].repeat(self.num_envs, 1, 1)
            w2c = torch.linalg.inv(c2w)
            out[cam] = {
                "extrinsic_cv": w2c[:, :3, :].contiguous().to(self.device),
                "intrinsic_cv": K[None].repeat(self.num_envs, 1, 1).to(self.device),
                "cam2world_gl": c2w.to(self.device),
            }
        return out
