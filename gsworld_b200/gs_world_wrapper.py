"""GSWorldWrapper — the reference env wrapper's observation path on the batched B200 engine.

Same public surface as gsworld/mani_skill/utils/wrappers/gs_world_wrapper.py:
  GSWorldWrapper(env, robot_pipe, scene_gs_cfg_name, device, ...)          (:34-107)
  .transform_gs_perlink(splats, gs_init_qpos=None, target_mat=None, eef_pos=None)   (:110-162)
  ._render_gsworld(renderer='gaussian-splatting') -> {cam: uint8 (B,H,W,3)}         (:232-275)
  .cam_maniskill2gs(config_maniskill, device, cam_names) -> {cam: Camera}           (:277-325)
  .step / .reset / .render_current_step                                             (:176-230)
What changes underneath: instead of a deepcopy of the model, L+K isin scans and B*C serial render()
calls with a host sync each, transform_gs_perlink only composes the (B, L+K) pose table
(S @ M_link @ inv(M_scan) @ inv(S); objects through extract_rigid_transform) and _render_gsworld makes ONE
batched libgsw launch set that transforms the movable Gaussians and rasterizes every (env, camera) frame.
Physics (env.step) stays with the wrapped env.

Scene loading (GaussianModelMerger + PLY, gs_world_wrapper.py:75-79) and the scan-pose probe through a
second ManiSkill env (:82-103) are outside the hot path: pass `merged_model` and `scan_link_mats`.

strict_reference=True reproduces reference quirk Q1: every env is rendered through env
0's camera (gs_world_wrapper.py:280-281).  False uses each env's own camera.
"""
from __future__ import annotations

import math
from collections.abc import Mapping
from typing import Dict, List, Optional, Union

import numpy as np
import torch

from . import _lib
from . import rig_tables as rt
from .cameras import Camera
from .engine import Engine
from .rigid import extract_rigid_transform, pose7_to_matrix
from .rotation_conversions import matrix_to_quaternion, quaternion_multiply

try:  # the reference subclasses gym.Wrapper; gymnasium is optional here
    import gymnasium as _gym

    _Base = _gym.Wrapper
except Exception:  # pragma: no cover - gymnasium absent in this image
    _gym = None

    class _Base:  # minimal stand-in with the same delegation behaviour
        def __init__(self, env):
            self.env = env

        def __getattr__(self, name):
            if name.startswith("_"):
                raise AttributeError(name)
            return getattr(self.env, name)

        @property
        def unwrapped(self):
            return getattr(self.env, "unwrapped", self.env)


class _HostStage:
    """Pinned staging for per-step HOST inputs (link / object / camera poses of a CPU simulator).

    The step fills views of a pinned arena and pushes it with ONE asynchronous H2D copy, instead of one
    synchronous pageable copy per tensor.  Two arenas alternate, each guarded by an event recorded after its
    copy, so the host may prepare step k+1 while step k's copy is still in flight.  Device-resident inputs
    (GPU simulation) never come here."""

    def __init__(self, n_floats: int, device):
        self.h = [torch.empty(n_floats, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.d = torch.empty(n_floats, dtype=torch.float32, device=device)
        self.ev = [torch.cuda.Event(), torch.cuda.Event()]
        self.i = 0

    def begin(self) -> torch.Tensor:
        self.i ^= 1
        self.ev[self.i].synchronize()  # the copy that last read this arena has completed
        return self.h[self.i]

    def push(self) -> torch.Tensor:
        self.d.copy_(self.h[self.i], non_blocking=True)
        self.ev[self.i].record()
        return self.d


class _MovablePts(Mapping):
    """The reference's `gs_movable_pts` dict (gs_world_wrapper.py:131,160): name -> (xyz (B,n,3), scaling (n,3) for
    links / (B,n,3) for objects, rotation (B,n,4), opacity (n,1,1)), raw (un-activated) parameters in the GS frame.

    The batched engine never materialises these tensors (the transform is fused into the rasterizer's preprocess
    kernel), so this legacy view computes an entry from the current pose table with a few torch statements the first
    time somebody reads it after a step (gs_utils.py:283-385 order: scale, rotate, translate; Q4 inverse_sigmoid
    scaling; Q12 quaternion re-scaling).  Always batched (B, n, .), also for B == 1 (the reference's B == 1 shapes
    are a quirk of its bmm branches).  Not on the hot path."""

    def __init__(self, wrap):
        self._w = wrap
        self._cache = {}

    def invalidate(self):
        self._cache.clear()

    def _names(self):
        w = self._w
        return [n for n in w._link_names if n in w.gs_semantics] + list(w._obj_names)

    def __iter__(self):
        return iter(self._names())

    def __len__(self):
        return len(self._names())

    def __getitem__(self, name):
        if name in self._cache:
            return self._cache[name]
        w = self._w
        if name in w._obj_names:
            s = len(w._link_names) + w._obj_names.index(name)
        elif name in w._link_names and name in w.gs_semantics:
            s = w._link_names.index(name)
        else:
            raise KeyError(name)
        g = w.initial_merger_robot
        idx = torch.where(w._slot == s)[0]
        P = w._pose[:, s]
        xyz, ls, q, op = (a[idx].to(P.device) for a in (g._xyz, g._scaling, g._rotation, g._opacity))
        R, t, qR, kappa = P[:, 0:9].reshape(-1, 3, 3), P[:, 9:12], P[:, 12:16], P[:, 16]
        B = P.shape[0]
        if bool((P[:, 17] != 0).any()):
            xb = xyz[None] * kappa[:, None, None]
            u = torch.exp(ls)[None] * kappa[:, None, None]
            scaling = torch.log(u / (1 - u))
        else:
            xb = xyz[None].expand(B, -1, 3)
            scaling = ls
        xb = torch.matmul(R[:, None], xb[..., None]).squeeze(-1) + t[:, None, :]
        norm = q.norm(dim=-1, keepdim=True)
        rot = quaternion_multiply(qR[:, None, :].expand(B, q.shape[0], 4), (q / norm)[None].expand(B, -1, 4)) * norm
        self._cache[name] = (xb, scaling, rot, op)
        return self._cache[name]


class GSWorldWrapper(_Base):
    def __init__(self, env, robot_pipe, scene_gs_cfg_name: str = "franka_merger",
                 device: Union[str, torch.device] = "cuda", log_state: bool = False,
                 state_log_path: str = "./exp_log", cam_randomization: bool = False, *, merged_model=None,
                 scan_link_mats=None, strict_reference: bool = True, engine: Optional[Engine] = None,
                 fused_host: bool = True, cfg_dir: Optional[str] = None, asset_dir: Optional[str] = None,
                 inv_depth_obs: bool = False, segmentation_obs: bool = False, tile_culling: bool = True):
        super().__init__(env)
        self.num_envs = env.num_envs
        self.robot_pipe = robot_pipe
        self.device = torch.device(device)
        self.scene_gs_cfg_name = scene_gs_cfg_name
        self.log_state = log_state
        self.cam_randomization = cam_randomization
        self.strict_reference = strict_reference
        # True: wait for each launch and auto-grow the tile-entry workspace on overflow (safe default).
        # False: fully asynchronous steps; call self._engine.finish() yourself to validate.
        self.sync_each_step = True
        # True: pose table and camera matrices come from one libgsw launch each (no host sync).
        # False: the reference's own torch / numpy statements (slow; kept for validation).
        self.fused_host = fused_host
        fam = rt.robot_family(scene_gs_cfg_name)
        self.gs_semantics = dict(rt.LINK_LABELS[fam])
        self.sim2gs_arm_trans = torch.tensor(rt.SIM2GS_ROBOT[fam], dtype=torch.float32, device=self.device)
        self.obj_gs_semantics = rt.OBJECT_LABELS
        self.rigid_sim2real, self.scale_sim2real, _, _ = extract_rigid_transform(self.sim2gs_arm_trans)
        # True: also expose the rasterizer's inverse-depth image as obs['sensor_data'][cam]['inv_depth']
        # (float32 (B,H,W,1)); the reference computes it (render()'s "depth") and drops it.
        self.inv_depth_obs = inv_depth_obs
        # True: obs['sensor_data'][cam]['segmentation'] = int32 (B,H,W,1), the scene label (link / object label,
        # -1 background) of the Gaussian with the largest blending weight at each pixel, -2 where nothing was
        # drawn.  Stands in for SAPIEN's segmentation image (Vulkan; unavailable here).  SURVEY.md §8(f) row 3.
        self.segmentation_obs = segmentation_obs
        # Optional caller-owned destination of the uint8 observations, shape (n_cams, B, H, W, 3): a device tensor,
        # or a PINNED HOST tensor, in which case the blend kernel's epilogue writes the pixels straight into
        # host memory over PCIe while later tiles are still being composited (no separate D2H copy).
        self.obs_out = None
        # Optional PINNED HOST tensor (n_cams, B, H, W, 3) uint8: the observations are also delivered there, copied
        # group by group while later frames are still being blended (gsw_frames_desc.host_rgb_u8).
        self.obs_host = None
        # Optional list of device tensors (n_cams, B, H, W, 3) uint8 on this or on peer GPUs (mapped into this
        # process, e.g. dist.PeerObservationExchange.destinations()): the observations are also pushed there group
        # by group on the copy stream while later frames are still being blended (gsw_frames_desc.peer_rgb_u8).
        self.obs_peers = None
        # With obs_host / obs_peers: True = do not make the render stream wait for the delivery (GSW_DELIVER_NO_JOIN);
        # "peers" = wait for the host copy but not for the peer pushes (GSW_DELIVER_NO_PEER_JOIN).  The caller orders
        # consumers with engine.delivery_wait() and must give a persistent obs_out.
        self.defer_delivery_join = False
        if merged_model is None and cfg_dir is not None:
            # reference path: GaussianModelMerger().load_models_from_config(CFG_DIR/<name>.json) + merge_models()
            # (gs_world_wrapper.py:75-79)
            import os
            from .gaussian_merger import GaussianModelMerger
            merger = GaussianModelMerger(device=self.device, asset_dir=asset_dir)
            merger.load_models_from_config(os.path.join(cfg_dir, f"{scene_gs_cfg_name}.json"))
            merged_model = merger.merge_models()
        if merged_model is None or scan_link_mats is None:
            raise NotImplementedError(
                "pass merged_model= (or cfg_dir= / asset_dir= for the PLY merger) and scan_link_mats= (link poses at "
                "the scan qpos; the reference probes them through a second ManiSkill env, gs_world_wrapper.py:82-103)")
        self.initial_merger_robot = merged_model
        self.gs_link_pose_mats = [torch.as_tensor(m, dtype=torch.float32, device=self.device).reshape(1, 4, 4)
                                  for m in scan_link_mats]
        self.gs_movable_pts = _MovablePts(self)  # lazy legacy view, see _MovablePts
        self._engine = engine if engine is not None else Engine(self.device)
        # exact tile culling (include/gsw.h): observations are bit-identical with it on or off; off keeps the
        # reference's tile lists for the parity taps
        self._engine.set_tile_culling(bool(tile_culling))
        self._setup_rig()

    # ------------------------------------------------------------------ set-up (once)
    @property
    def base_env(self):
        return self.env.unwrapped

    def _setup_rig(self):
        """Label -> pose-table slot, constant matrix products, scene upload with activations done once."""
        be, dev = self.base_env, self.device
        links = be.agent.robot.get_links()
        self._link_names = [l.name for l in links]
        actors = be.get_state_dict().get("actors", {}) if hasattr(be, "get_state_dict") else {}
        self._obj_names = [k for k in actors if k in rt.SIM2GS_OBJECT]
        L, K = len(links), len(self._obj_names)
        self.num_slots = L + K
        g = self.initial_merger_robot
        sem = g._semantics.long().squeeze(-1)
        slot = torch.full_like(sem, -1, dtype=torch.int32)
        self._slot = slot  # filled below; kept for the lazy gs_movable_pts view
        for i, name in enumerate(self._link_names):  # later keys overwrite earlier ones, as the write-back loop does
            if name not in self.gs_semantics:
                continue
            target = torch.as_tensor(self.gs_semantics[name], device=sem.device).long().reshape(-1)
            slot[torch.isin(sem, target)] = i
        for k, name in enumerate(self._obj_names):
            slot[sem == int(self.obj_gs_semantics[name])] = L + k
        # constants of gs_world_wrapper.py:120 and :146
        S = self.sim2gs_arm_trans
        self._S_inv = torch.linalg.inv(S)
        self._scan_inv = torch.cat([torch.linalg.inv(m) for m in self.gs_link_pose_mats], 0)  # (L,4,4)
        self._is_xarm = "xarm" in str(be.agent.uid)
        self._xarm_off = torch.tensor(rt.OBJECT_OFFSET["xarm_arm"], dtype=torch.float32, device=dev)
        self._obj_S_inv = [torch.linalg.inv(torch.tensor(rt.SIM2GS_OBJECT[n], dtype=torch.float32, device=dev))
                           for n in self._obj_names]
        self._obj_off = [torch.tensor(rt.OBJECT_OFFSET.get(n, [0.0, 0.0, 0.0]), dtype=torch.float32, device=dev)
                         for n in self._obj_names]
        self._obj_scale = [float(rt.OBJECT_SCALE[n]) for n in self._obj_names]
        # activations exactly as the model getters (gaussian_model.py:102-130), once instead of per frame
        f = lambda t: t.to(dev, torch.float32).contiguous()
        self._engine.set_scene(
            f(g._xyz), f(torch.exp(g._scaling)), f(torch.nn.functional.normalize(g._rotation)),
            f(torch.sigmoid(g._opacity).reshape(-1)), f(torch.cat((g._features_dc, g._features_rest), dim=1)),
            f(g._scaling), f(g._rotation), slot.to(dev).contiguous(), self.num_slots)
        if self.segmentation_obs:
            self._engine.set_labels(sem.to(dev, torch.int32).contiguous())
        self._pose = torch.zeros((self.num_envs, self.num_slots, _lib.POSE_STRIDE), dtype=torch.float32, device=dev)
        self._frame_maps = {}
        self._engine.set_rig(
            S, self._S_inv, self._scan_inv, self._xarm_off if self._is_xarm else None,
            torch.stack(self._obj_S_inv) if K else None, torch.stack(self._obj_off) if K else None,
            torch.tensor(self._obj_scale) if K else None)
        self._rigid_host = self.rigid_sim2real.detach().to("cpu", torch.float32).contiguous()
        self._scale_host = float(self.scale_sim2real)
        self._cam_bufs = {}
        self._stages = {}  # pinned staging arenas for host-resident per-step inputs

    # ------------------------------------------------------------------ per step
    def transform_gs_perlink(self, splats=None, gs_init_qpos=None, target_mat=None, eef_pos=None):
        """Compose the per-(env, link/object) rigid transforms for this step.

        Same matrix algebra as the reference (gs_world_wrapper.py:115-120, :139-156); the Gaussian-level
        part (gs_utils.transform_gaussians) is fused into the rasterizer's preprocess kernel."""
        be, dev = self.base_env, self.device
        self.gs_movable_pts.invalidate()
        links = be.agent.robot.get_links()
        mats = [l.pose.to_transformation_matrix() for l in links]
        if self.fused_host and mats[0].device.type == "cpu" and torch.device(dev).type == "cuda":
            # host simulator: one pinned arena, one asynchronous H2D copy for all link and object poses
            B, L, K = self.num_envs, len(mats), len(self._obj_names)
            st = self._stages.get("pose")
            if st is None:
                st = self._stages["pose"] = _HostStage(B * L * 16 + B * K * 7, dev)
            h = st.begin()
            hM = h[:B * L * 16].view(B, L, 4, 4)
            if all(m.dtype == torch.float32 for m in mats):
                torch.stack(mats, 1, out=hM)
            else:
                for j, m in enumerate(mats):
                    hM[:, j].copy_(m)
            if K:
                actors = be.get_state_dict()["actors"]
                hO = h[B * L * 16:].view(B, K, 7)
                for k, n in enumerate(self._obj_names):
                    hO[:, k].copy_(actors[n][:, :7])
                    self.gs_semantics[n] = self.obj_gs_semantics[n]
            d = st.push()
            return self._engine.compose_poses(d[:B * L * 16].view(B, L, 4, 4),
                                              d[B * L * 16:].view(B, K, 7) if K else None, self._pose)
        M = torch.stack(mats, 1)  # (B,L,4,4)
        M = M.to(dev, torch.float32, non_blocking=True).contiguous()
        if self.fused_host:
            obj = None
            if self._obj_names:
                actors = be.get_state_dict()["actors"]
                obj = torch.stack([actors[n][:, :7] for n in self._obj_names], 1)
                obj = obj.to(dev, torch.float32, non_blocking=True).contiguous()
                for name in self._obj_names:
                    self.gs_semantics[name] = self.obj_gs_semantics[name]
            return self._engine.compose_poses(M, obj, self._pose)
        if self._is_xarm:
            M = M.clone()
            M[:, :, :3, 3] += self._xarm_off
        A = self.sim2gs_arm_trans @ M @ self._scan_inv[None] @ self._S_inv
        L = len(links)
        P = self._pose
        P.zero_()
        R = A[..., :3, :3]
        P[:, :L, 0:9] = R.reshape(self.num_envs, L, 9)
        P[:, :L, 9:12] = A[..., :3, 3]
        P[:, :L, 12:16] = matrix_to_quaternion(R)
        if self._obj_names:
            actors = be.get_state_dict()["actors"]
            for k, name in enumerate(self._obj_names):
                mat = pose7_to_matrix(actors[name][:, :7].to(dev, torch.float32))
                mat[:, :3, 3] += self._obj_off[k]
                full = self.sim2gs_arm_trans @ mat @ self._obj_S_inv[k]
                rigid, scale, _, _ = extract_rigid_transform(full)
                Ro = rigid[:, :3, :3]
                P[:, L + k, 0:9] = Ro.reshape(self.num_envs, 9)
                P[:, L + k, 9:12] = rigid[:, :3, 3]
                P[:, L + k, 12:16] = matrix_to_quaternion(Ro)
                P[:, L + k, 16] = scale * self._obj_scale[k]
                P[:, L + k, 17] = 1.0
                self.gs_semantics[name] = self.obj_gs_semantics[name]
        return P

    def _sensor_hw(self, cam_name):
        be = self.base_env
        hw = getattr(be, "sensor_hw", None)
        if hw is not None:
            return hw[cam_name]
        img = be.get_sensor_images()[cam_name]["rgb"]  # reference path (gs_world_wrapper.py:288-289)
        return int(img.shape[1]), int(img.shape[2])

    def cam_maniskill2gs(self, config_maniskill: dict, device, cam_names: list) -> Dict[str, Camera]:
        """ManiSkill sensor params -> 3DGS cameras, env 0 only (reference behaviour, :277-325)."""
        return {c: self._camera_for(config_maniskill[c], c, 0, device) for c in cam_names}

    def _camera_for(self, params, cam_name, env_idx, device) -> Camera:
        ext = params["extrinsic_cv"][env_idx].to(device)
        K = params["intrinsic_cv"][env_idx].to(device)
        w2c = torch.vstack([ext, torch.tensor([0, 0, 0, 1], dtype=ext.dtype, device=ext.device)])
        c2w = torch.linalg.inv(w2c)
        img_h, img_w = self._sensor_hw(cam_name)
        fovx = 2 * torch.arctan(img_w / (2 * K[0, 0]))
        fovy = 2 * torch.arctan(img_h / (2 * K[1, 1]))
        c2w[:3, 3] = c2w[:3, 3] * self.scale_sim2real
        real_w2c = torch.linalg.inv(self.rigid_sim2real.to(device) @ c2w)
        R = real_w2c[:3, :3].T
        T = real_w2c[:3, 3]
        return Camera(resolution=(img_w, img_h), colmap_id=0, R=R.cpu().numpy(), T=T.cpu().numpy(),
                      FoVx=float(fovx), FoVy=float(fovy), depth_params=None, image=None, invdepthmap=None,
                      image_name=cam_name, uid=0, data_device=device)

    def _views(self, cam_names):
        """Stack the view tensors the engine consumes and the frame -> view map.

        fused_host (default): gsw_compose_cameras, one launch per camera, no host sync.  Otherwise the reference's
        own torch / numpy statements (bit-identical camera matrices; validation path): strict_reference -> one
        view per camera from env 0's parameters, else each env's own camera, where a camera whose parameters are
        identical for every env (fixed third-person mount) collapses to a single view."""
        params = self.base_env.get_sensor_params()
        B = self.num_envs
        if self.fused_host:
            return self._views_fused(params, cam_names)
        if self.strict_reference:
            cams = [self._camera_for(params[c], c, 0, self.device) for c in cam_names]
            self.gs_cam = dict(zip(cam_names, cams))
            view = torch.stack([c.world_view_transform for c in cams])
            proj = torch.stack([c.full_proj_transform for c in cams])
            campos = torch.stack([c.camera_center for c in cams])
            tanfov = torch.tensor([[math.tan(c.FoVx * 0.5), math.tan(c.FoVy * 0.5)] for c in cams],
                                  dtype=torch.float32, device=self.device)
            fv = torch.arange(len(cam_names), dtype=torch.int32, device=self.device).repeat_interleave(B)
            return view.contiguous(), proj.contiguous(), campos.contiguous(), tanfov, fv
        # each env's own camera through the same reference statements (validation path: one small torch / numpy
        # round trip per camera); a camera whose parameters are identical for every env collapses to one view
        cams, fvs = [], []
        for c in cam_names:
            ext, K = params[c]["extrinsic_cv"], params[c]["intrinsic_cv"]
            shared = bool(((ext == ext[:1]).all() & (K == K[:1]).all()).item())
            base = len(cams)
            cams += [self._camera_for(params[c], c, e, self.device) for e in range(1 if shared else B)]
            fvs.append(torch.full((B,), base, dtype=torch.int32, device=self.device) if shared
                       else base + torch.arange(B, dtype=torch.int32, device=self.device))
        view = torch.stack([c.world_view_transform for c in cams])
        proj = torch.stack([c.full_proj_transform for c in cams])
        campos = torch.stack([c.camera_center for c in cams])
        tanfov = torch.tensor([[math.tan(c.FoVx * 0.5), math.tan(c.FoVy * 0.5)] for c in cams],
                              dtype=torch.float32, device=self.device)
        return view.contiguous(), proj.contiguous(), campos.contiguous(), tanfov, torch.cat(fvs).contiguous()

    def _views_fused(self, params, cam_names):
        """Camera matrices through gsw_compose_cameras: no host sync.  strict -> env 0's camera for every env;
        otherwise one view per (camera, env) unless the env exposes `static_cameras` naming world-fixed ones."""
        B, dev = self.num_envs, self.device
        static = set(getattr(self.base_env, "static_cameras", ()))
        key = (tuple(cam_names), self.strict_reference)
        if key not in self._cam_bufs:
            counts = [1 if (self.strict_reference or c in static) else B for c in cam_names]
            V = sum(counts)
            # view numbering: world-fixed cameras first.  The engine keeps the static lists of a leading run of
            # views whose camera has not changed since the previous step (static-list cache), so the views that
            # never move go in front of the ones that move every step.  Frames keep the order camera * B + env.
            bases, base = [0] * len(cam_names), 0
            for ci in sorted(range(len(cam_names)), key=lambda i: cam_names[i] not in static):
                bases[ci] = base
                base += counts[ci]
            fv = [torch.full((B,), b, dtype=torch.int32) if n == 1 else b + torch.arange(B, dtype=torch.int32)
                  for b, n in zip(bases, counts)]
            z = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
            self._cam_bufs[key] = dict(counts=counts, bases=bases, view=z(V, 16), proj=z(V, 16), campos=z(V, 3),
                                       tanfov=z(V, 2), fv=torch.cat(fv).to(dev))
        cb = self._cam_bufs[key]
        staged = None
        first = params[cam_names[0]]["extrinsic_cv"]
        if first.device.type == "cpu" and torch.device(dev).type == "cuda":
            # host simulator: all cameras' extrinsics and intrinsics through one pinned arena / one H2D copy
            ne = first.shape[-2] * first.shape[-1]
            tot = sum(cb["counts"]) * (ne + 9)
            st = self._stages.get(("cam",) + key)
            if st is None:
                st = self._stages[("cam",) + key] = _HostStage(tot, dev)
            h = st.begin()
            o, spans = 0, []
            for c, n in zip(cam_names, cb["counts"]):
                h[o:o + n * ne].view(n, *first.shape[-2:]).copy_(params[c]["extrinsic_cv"][:n])
                h[o + n * ne:o + n * (ne + 9)].view(n, 3, 3).copy_(params[c]["intrinsic_cv"][:n])
                spans.append((o, o + n * ne, o + n * (ne + 9)))
                o += n * (ne + 9)
            d = st.push()
            staged = [(d[a:b].view(-1, *first.shape[-2:]), d[b:e].view(-1, 3, 3)) for a, b, e in spans]
        for ci, (c, n) in enumerate(zip(cam_names, cb["counts"])):
            base = cb["bases"][ci]
            H, W = self._sensor_hw(c)
            if staged is not None:
                ext, K = staged[ci]
            else:
                ext = params[c]["extrinsic_cv"][:n].to(dev, torch.float32, non_blocking=True).contiguous()
                K = params[c]["intrinsic_cv"][:n].to(dev, torch.float32, non_blocking=True).contiguous()
            sl = slice(base, base + n)
            self._engine.compose_cameras(ext, K, H, W, self._rigid_host, self._scale_host, cb["view"][sl],
                                         cb["proj"][sl], cb["campos"][sl], cb["tanfov"][sl])
        return cb["view"], cb["proj"], cb["campos"], cb["tanfov"], cb["fv"]

    def _maps(self, n_cams):
        if n_cams not in self._frame_maps:  # frame = cam * B + env
            self._frame_maps[n_cams] = torch.arange(self.num_envs, dtype=torch.int32).repeat(n_cams).to(self.device)
        return self._frame_maps[n_cams]

    def _render_gsworld(self, renderer="gaussian-splatting"):
        if renderer != "gaussian-splatting":
            return ValueError(f"{renderer} is not supported. Official renderer also supports gsplat.")
        cam_names = list(self.base_env.get_sensor_params().keys())
        sizes = {self._sensor_hw(c) for c in cam_names}
        groups = [[c for c in cam_names if self._sensor_hw(c) == hw] for hw in sorted(sizes)]
        gs_renders = {}
        for names in groups:  # one launch set per distinct resolution (the reference sensors share one)
            H, W = self._sensor_hw(names[0])
            view, proj, campos, tanfov, fv = self._views(names)
            fe = self._maps(len(names))
            dst = None
            if self.obs_out is not None and len(groups) == 1:
                dst = self.obs_out.view(len(names) * self.num_envs, H, W, 3)
            out = self._engine.render_frames(
                self._pose, view, proj, campos, tanfov, fe, fv, H, W, num_envs=self.num_envs, bg=(0.0, 0.0, 0.0),
                antialiasing=bool(self.robot_pipe.antialiasing), want_u8=True, want_invdepth=self.inv_depth_obs,
                want_seg=self.segmentation_obs, out_rgb_u8=dst, finish=self.sync_each_step,
                host_rgb_u8=None if (self.obs_host is None or len(groups) != 1) else
                self.obs_host.view(len(names) * self.num_envs, H, W, 3),
                peer_rgb_u8=None if (not self.obs_peers or len(groups) != 1) else
                [t.view(len(names) * self.num_envs, H, W, 3) for t in self.obs_peers],
                defer_delivery_join=(self.defer_delivery_join if dst is not None else False))
            rgb = out["rgb_u8"].view(len(names), self.num_envs, H, W, 3)
            for i, c in enumerate(names):
                gs_renders[c] = rgb[i]
            if self.inv_depth_obs:
                inv = out["invdepth"].view(len(names), self.num_envs, H, W, 1)
                self.last_inv_depth = {c: inv[i] for i, c in enumerate(names)}
            if self.segmentation_obs:
                seg = out["seg"].view(len(names), self.num_envs, H, W, 1)
                self.last_segmentation = {c: seg[i] for i, c in enumerate(names)}
        return {c: gs_renders[c] for c in cam_names}

    # ------------------------------------------------------------------ gym surface
    def step(self, action):
        obs, reward, terminated, truncated, info = self.env.step(action)
        self.transform_gs_perlink(self.initial_merger_robot)
        self._write_obs(obs)
        return obs, reward, terminated, truncated, info

    def reset(self, *, seed=None, options=None):
        obs, info = self.env.reset(seed=seed, options=options)
        self.transform_gs_perlink(self.initial_merger_robot)
        self._write_obs(obs)
        return obs, info

    def _write_obs(self, obs):
        for cam_name, gs_render in self._render_gsworld().items():
            obs["sensor_data"][cam_name]["rgb"] = gs_render
            if self.inv_depth_obs:
                obs["sensor_data"][cam_name]["inv_depth"] = self.last_inv_depth[cam_name]
            if self.segmentation_obs:
                obs["sensor_data"][cam_name]["segmentation"] = self.last_segmentation[cam_name]

    def render(self):
        return self.env.render()

    def render_current_step(self):
        be = self.base_env
        info = be.get_info()
        obs = be.get_obs(info)
        reward = be.get_reward(obs=obs, action=None, info=info)
        if "success" in info:
            terminated = torch.logical_or(info["success"], info["fail"]) if "fail" in info else info["success"].clone()
        else:
            terminated = info["fail"].clone() if "fail" in info else torch.zeros(self.num_envs, dtype=bool,
                                                                                 device=self.device)
        self._write_obs(obs)
        return obs, reward, terminated, torch.zeros(self.num_envs, dtype=bool, device=self.device), info
