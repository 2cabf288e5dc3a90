"""Python handle on a libgsw context: scene upload + batched frame rendering on the caller's stream.

torch is used for device memory and streams only; all arithmetic on the hot path is in libgsw.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _chk(t: torch.Tensor, dtype, name, device):
    if t.device != device:
        raise ValueError(f"{name}: expected device {device}, got {t.device}")
    if t.dtype != dtype:
        raise ValueError(f"{name}: expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name}: must be contiguous")
    return t


class Engine:
    NEAR_GSWORLD = 0.05  # GSWorld's patched near plane (reference README.md:33)

    def __init__(self, device="cuda:0"):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("gsworld_b200.Engine needs a CUDA device; there is no CPU path")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.lib = _lib.load()
        h = C.c_void_p()
        _lib.check(self.lib.gsw_create(self.device.index, C.byref(h)))
        self._h = h
        self._keep = {}
        self.N = 0
        self.num_slots = 0
        self.last_stats = None

    def close(self):
        if getattr(self, "_h", None):
            self.lib.gsw_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # -- scene --
    def set_scene(self, xyz, scale_act, quat_act, opacity_act, sh, log_scale_raw=None, quat_raw=None, slot=None,
                  num_slots=0):
        """All tensors on self.device, fp32 contiguous; layouts as include/gsw.h (gsw_scene_set)."""
        d, f32 = self.device, torch.float32
        N = xyz.shape[0]
        xyz = _chk(xyz.reshape(N, 3), f32, "xyz", d)
        scale_act = _chk(scale_act.reshape(N, 3), f32, "scale_act", d)
        quat_act = _chk(quat_act.reshape(N, 4), f32, "quat_act", d)
        opacity_act = _chk(opacity_act.reshape(N), f32, "opacity_act", d)
        sh = _chk(sh.reshape(N, 16, 3), f32, "sh", d)
        if slot is not None:
            slot = _chk(slot.reshape(N), torch.int32, "slot", d)
            log_scale_raw = _chk(log_scale_raw.reshape(N, 3), f32, "log_scale_raw", d)
            quat_raw = _chk(quat_raw.reshape(N, 4), f32, "quat_raw", d)
        self._keep = dict(xyz=xyz, scale_act=scale_act, quat_act=quat_act, opacity_act=opacity_act, sh=sh,
                          log_scale_raw=log_scale_raw, quat_raw=quat_raw, slot=slot)
        _lib.check(self.lib.gsw_scene_set(self._h, N, _ptr(xyz), _ptr(scale_act), _ptr(quat_act), _ptr(opacity_act),
                                          _ptr(sh), _ptr(log_scale_raw), _ptr(quat_raw), _ptr(slot), int(num_slots),
                                          self._stream()))
        self.N, self.num_slots = N, int(num_slots)

    def set_labels(self, labels):
        """Per-Gaussian int32 labels (original order) for the segmentation output; kept alive here, used in place."""
        labels = _chk(labels.reshape(self.N), torch.int32, "labels", self.device)
        self._keep["labels"] = labels
        _lib.check(self.lib.gsw_scene_set_labels(self._h, _ptr(labels)))

    def set_depth_buckets(self, nb: int):
        _lib.check(self.lib.gsw_set_depth_buckets(self._h, int(nb)))

    def set_tile_culling(self, on: bool):
        """Exact tile culling (include/gsw.h): drop tile entries that provably stay below alpha 1/255 on the
        whole tile.  Images and radii are unchanged bit for bit; the tile lists become sub-lists."""
        _lib.check(self.lib.gsw_set_tile_culling(self._h, 1 if on else 0))

    def reserve_entries(self, n: int):
        _lib.check(self.lib.gsw_reserve_entries(self._h, int(n)))

    # -- per-step pose / camera algebra (one launch each) --
    def set_rig(self, S, S_inv, scan_inv, link_offset, obj_S_inv, obj_offset, obj_scale):
        """Host fp32 tensors: S, S_inv (4,4); scan_inv (L,4,4); link_offset (3,) or None;
        obj_S_inv (K,4,4); obj_offset (K,3); obj_scale (K,)."""
        h = lambda t: None if t is None else t.detach().to("cpu", torch.float32).contiguous()
        S, S_inv, scan_inv, link_offset = h(S), h(S_inv), h(scan_inv), h(link_offset)
        obj_S_inv, obj_offset, obj_scale = h(obj_S_inv), h(obj_offset), h(obj_scale)
        L = 0 if scan_inv is None else scan_inv.shape[0]
        K = 0 if obj_S_inv is None else obj_S_inv.shape[0]
        self._rig_host = (S, S_inv, scan_inv, link_offset, obj_S_inv, obj_offset, obj_scale)
        p = lambda t: None if t is None or t.numel() == 0 else C.c_void_p(t.data_ptr())
        _lib.check(self.lib.gsw_rig_set(self._h, L, K, p(S), p(S_inv), p(scan_inv), p(link_offset), p(obj_S_inv),
                                        p(obj_offset), p(obj_scale)))
        self.rig_L, self.rig_K = L, K

    def compose_poses(self, link_mats, obj_pose, out):
        """link_mats (B,L,4,4) fp32 device; obj_pose (B,K,>=7) fp32 device or None; out (B,L+K,20)."""
        B = out.shape[0]
        stride = 7 if obj_pose is None else int(obj_pose.shape[-1])
        _lib.check(self.lib.gsw_compose_poses(self._h, B, _ptr(link_mats), _ptr(obj_pose), stride, _ptr(out),
                                              self._stream()))
        return out

    def compose_cameras(self, ext, K, H, W, rigid_sim2real_host, scale_sim2real, view, proj, campos, tanfov):
        n = ext.shape[0]
        _lib.check(self.lib.gsw_compose_cameras(self._h, n, _ptr(ext), _ptr(K), int(H), int(W),
                                                C.c_void_p(rigid_sim2real_host.data_ptr()), float(scale_sim2real),
                                                _ptr(view), _ptr(proj), _ptr(campos), _ptr(tanfov), self._stream()))

    # -- frames --
    def render_frames(self, pose, viewmatrix, projmatrix, campos, tanfov, frame_env, frame_view, H, W, *,
                      num_envs, bg=(0.0, 0.0, 0.0), antialiasing=False, scale_modifier=1.0, near=NEAR_GSWORLD,
                      out_rgb_u8=None, out_color=None, out_invdepth=None, out_radii=None, want_u8=True,
                      want_color=False, want_invdepth=False, want_radii=False, finish=True, host_rgb_u8=None,
                      peer_rgb_u8=None, defer_delivery_join=False,
                      want_seg=False):
        """Enqueue one batched transform+rasterize launch set.  Returns dict of output tensors.

        pose [B,S,20] fp32; viewmatrix/projmatrix [V,4,4]; campos [V,3]; tanfov [V,2];
        frame_env/frame_view [F] int32.  With finish=True (default) waits for the launch, and if the
        tile-entry workspace overflowed grows it and re-runs (so results are always valid on return)."""
        d, f32 = self.device, torch.float32
        V, F = viewmatrix.shape[0], frame_env.shape[0]
        viewmatrix = _chk(viewmatrix.reshape(V, 16), f32, "viewmatrix", d)
        projmatrix = _chk(projmatrix.reshape(V, 16), f32, "projmatrix", d)
        campos = _chk(campos.reshape(V, 3), f32, "campos", d)
        tanfov = _chk(tanfov.reshape(V, 2), f32, "tanfov", d)
        frame_env = _chk(frame_env.reshape(F), torch.int32, "frame_env", d)
        frame_view = _chk(frame_view.reshape(F), torch.int32, "frame_view", d)
        if pose is not None:
            pose = _chk(pose.reshape(num_envs, self.num_slots, _lib.POSE_STRIDE), f32, "pose", d)
        if want_u8 and out_rgb_u8 is None:
            out_rgb_u8 = torch.empty((F, H, W, 3), dtype=torch.uint8, device=d)
        elif out_rgb_u8 is not None:
            # a device tensor, or pinned host memory (unified addressing: the kernel writes it over PCIe)
            ok_dev = out_rgb_u8.device == d or (out_rgb_u8.device.type == "cpu" and out_rgb_u8.is_pinned())
            if not ok_dev or out_rgb_u8.dtype != torch.uint8 or not out_rgb_u8.is_contiguous() \
                    or out_rgb_u8.numel() != F * H * W * 3:
                raise ValueError("out_rgb_u8 must be a contiguous uint8 (F,H,W,3) tensor on the device or in pinned "
                                 "host memory")
        if want_color and out_color is None:
            out_color = torch.empty((F, 3, H, W), dtype=f32, device=d)
        if want_invdepth and out_invdepth is None:
            out_invdepth = torch.empty((F, 1, H, W), dtype=f32, device=d)
        if want_radii and out_radii is None:
            out_radii = torch.empty((F, self.N), dtype=torch.int32, device=d)
        desc = _lib.FramesDesc()
        desc.num_envs, desc.num_views, desc.num_frames = int(num_envs), int(V), int(F)
        desc.height, desc.width, desc.antialiasing = int(H), int(W), int(bool(antialiasing))
        desc.scale_modifier, desc.near_plane = float(scale_modifier), float(near)
        desc.bg[0], desc.bg[1], desc.bg[2] = float(bg[0]), float(bg[1]), float(bg[2])
        desc.pose = None if pose is None else pose.data_ptr()
        desc.viewmatrix, desc.projmatrix = viewmatrix.data_ptr(), projmatrix.data_ptr()
        desc.campos, desc.tanfov = campos.data_ptr(), tanfov.data_ptr()
        desc.frame_env, desc.frame_view = frame_env.data_ptr(), frame_view.data_ptr()
        desc.out_rgb_u8 = None if out_rgb_u8 is None else out_rgb_u8.data_ptr()
        desc.out_color = None if out_color is None else out_color.data_ptr()
        desc.out_invdepth = None if out_invdepth is None else out_invdepth.data_ptr()
        desc.out_radii = None if out_radii is None else out_radii.data_ptr()
        out_seg = torch.empty((F, H, W), dtype=torch.int32, device=d) if want_seg else None
        desc.out_seg = None if out_seg is None else out_seg.data_ptr()
        if host_rgb_u8 is not None:
            if not (host_rgb_u8.device.type == "cpu" and host_rgb_u8.is_pinned() and host_rgb_u8.is_contiguous()
                    and host_rgb_u8.dtype == torch.uint8 and host_rgb_u8.numel() == F * H * W * 3):
                raise ValueError("host_rgb_u8 must be a contiguous pinned uint8 (F,H,W,3) host tensor")
            desc.host_rgb_u8 = host_rgb_u8.data_ptr()
        else:
            desc.host_rgb_u8 = None
        if peer_rgb_u8:
            # device buffers (F,H,W,3) uint8 on this or on peer GPUs (mapped into this process): every finished frame
            # group is pushed to each of them on the copy stream (gsw_frames_desc.peer_rgb_u8)
            for t in peer_rgb_u8:
                if not (t.is_cuda and t.is_contiguous() and t.dtype == torch.uint8 and t.numel() == F * H * W * 3):
                    raise ValueError("peer_rgb_u8 entries must be contiguous uint8 (F,H,W,3) CUDA tensors")
            arr = (C.c_void_p * len(peer_rgb_u8))(*[t.data_ptr() for t in peer_rgb_u8])
            desc.peer_rgb_u8 = C.cast(arr, C.POINTER(C.c_void_p))
            desc.num_peer_rgb_u8 = len(peer_rgb_u8)
            self._peer_keep = (arr, list(peer_rgb_u8))
        else:
            desc.peer_rgb_u8 = None
            desc.num_peer_rgb_u8 = 0
        # GSW_DELIVER_NO_JOIN (1) / GSW_DELIVER_NO_PEER_JOIN (2, defer_delivery_join="peers") /
        # GSW_DELIVER_NO_JOIN | GSW_DELIVER_AFTER_BLEND (5, defer_delivery_join="after": asynchronous producers)
        desc.deliver_flags = {"peers": 2, "after": 5}.get(defer_delivery_join, 1 if defer_delivery_join else 0)
        self._last_inputs = (pose, viewmatrix, projmatrix, campos, tanfov, frame_env, frame_view)
        self._dbg_N = None
        for _attempt in range(4):
            _lib.check(self.lib.gsw_render_frames(self._h, C.byref(desc), self._stream()))
            if not finish:
                break
            if self.finish(raise_on_overflow=False):
                break
        else:
            raise _lib.GswError(_lib.GSW_E_CAPACITY, "tile-entry workspace kept overflowing")
        return dict(rgb_u8=out_rgb_u8, color=out_color, invdepth=out_invdepth, radii=out_radii, seg=out_seg)

    def delivery_wait(self, stream=None):
        """Make `stream` (a torch stream; default: the current one) wait for the last delivery to host / peers."""
        st = self._stream() if stream is None else C.c_void_p(stream.cuda_stream)
        _lib.check(self.lib.gsw_delivery_wait(self._h, st))

    def delivery_gate(self, stream=None):
        """The next launch's peer pushes start only after the work enqueued so far on `stream`
        (a torch stream; default: the current one) has completed."""
        st = self._stream() if stream is None else C.c_void_p(stream.cuda_stream)
        _lib.check(self.lib.gsw_delivery_gate(self._h, st))

    def finish(self, raise_on_overflow=True) -> bool:
        """Wait for the last launch.  Returns True if it was valid; on overflow grows the workspace
        (caller must re-launch) and returns False, or raises."""
        need = C.c_int64(0)
        stats = (C.c_int64 * 4)()
        rc = self.lib.gsw_finish(self._h, C.byref(need), C.byref(stats))
        self.last_stats = dict(D=stats[0], staged=stats[1], cap=stats[2], launches=stats[3])
        if rc == _lib.GSW_E_CAPACITY:
            self.reserve_entries(int(need.value * 1.5) + 1024)
            if raise_on_overflow:
                _lib.check(rc)
            return False
        _lib.check(rc)
        return True

    # -- single frame, upstream argument list --
    def rasterize_forward(self, means3D, opacities, viewmatrix, projmatrix, campos, tanfovx, tanfovy, H, W, bg, *,
                          shs=None, colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None,
                          sh_degree=3, scale_modifier=1.0, prefiltered=False, antialiasing=False,
                          near=NEAR_GSWORLD):
        d, f32 = self.device, torch.float32
        P = means3D.shape[0]
        c = lambda t, n: None if t is None else _chk(t.contiguous(), f32, n, d)
        means3D, opacities = c(means3D, "means3D"), c(opacities, "opacities")
        shs, colors_precomp = c(shs, "shs"), c(colors_precomp, "colors_precomp")
        scales, rotations, cov3D_precomp = c(scales, "scales"), c(rotations, "rotations"), c(cov3D_precomp, "cov3D")
        viewmatrix, projmatrix = c(viewmatrix, "viewmatrix"), c(projmatrix, "projmatrix")
        campos, bg = c(campos, "campos"), c(bg, "bg")
        M = 0 if shs is None else int(shs.reshape(P, -1, 3).shape[1]) if P > 0 else 16
        self._dbg_N = P
        color = torch.zeros((3, H, W), dtype=f32, device=d)
        radii = torch.zeros((P,), dtype=torch.int32, device=d)
        invd = torch.zeros((1, H, W), dtype=f32, device=d)
        for _attempt in range(4):
            _lib.check(self.lib.gsw_rasterize_forward(
                self._h, P, int(sh_degree), M, _ptr(bg), int(W), int(H), _ptr(means3D), _ptr(shs),
                _ptr(colors_precomp), _ptr(opacities), _ptr(scales), float(scale_modifier), _ptr(rotations),
                _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos), float(tanfovx),
                float(tanfovy), int(bool(prefiltered)), int(bool(antialiasing)), float(near), _ptr(color),
                _ptr(radii), _ptr(invd), self._stream()))
            if P == 0 or self.finish(raise_on_overflow=False):
                break
        else:
            raise _lib.GswError(_lib.GSW_E_CAPACITY, "tile-entry workspace kept overflowing")
        return color, radii, invd

    def mark_visible(self, positions, viewmatrix, near=NEAR_GSWORLD):
        P = positions.shape[0]
        out = torch.zeros((P,), dtype=torch.uint8, device=self.device)
        _lib.check(self.lib.gsw_mark_visible(self._h, P, _ptr(positions.contiguous()), _ptr(viewmatrix.contiguous()),
                                             float(near), _ptr(out), self._stream()))
        return out.bool()

    # -- measurement hooks --
    STAGES = ("preprocess_static", "preprocess_dynamic", "scan", "scatter", "sort", "gather", "blend", "radii")

    def profile(self, on: bool):
        _lib.check(self.lib.gsw_profile_enable(self._h, int(bool(on))))

    def stage_ms(self):
        buf = (C.c_float * 8)()
        _lib.check(self.lib.gsw_profile_read(self._h, buf, 8))
        return {k: float(buf[i]) for i, k in enumerate(self.STAGES)}

    def counters(self):
        """Work counters of the last launch: blend pair evaluations, entries staged."""
        buf = (C.c_int64 * 6)()
        _lib.check(self.lib.gsw_profile_counters(self._h, buf, 6))
        return {"blend_pairs": int(buf[0]), "blend_staged": int(buf[1]), "static_chunks_live": int(buf[2]),
                "visible_static": int(buf[3]), "visible_dynamic": int(buf[4]), "static_chunk_items": int(buf[5])}

    def kernel_launches(self) -> int:
        return int(self.lib.gsw_kernel_launches())

    # -- parity taps --
    def debug_frame_attrs(self, frame: int):
        N, d = self._last_N(), self.device
        o = dict(radii=torch.zeros(N, dtype=torch.int32, device=d),
                 tiles_touched=torch.zeros(N, dtype=torch.int32, device=d),
                 depth=torch.zeros(N, dtype=torch.float32, device=d), xy=torch.zeros((N, 2), device=d),
                 conic_opacity=torch.zeros((N, 4), device=d), rgb=torch.zeros((N, 3), device=d))
        _lib.check(self.lib.gsw_debug_frame_attrs(self._h, int(frame), _ptr(o["radii"]), _ptr(o["tiles_touched"]),
                                                  _ptr(o["depth"]), _ptr(o["xy"]), _ptr(o["conic_opacity"]),
                                                  _ptr(o["rgb"]), self._stream()))
        return o

    def _last_N(self):
        return getattr(self, "_dbg_N", None) or self.N

    def debug_bin_offsets(self):
        """Exclusive (list, tile) bin offsets of the last launch as a numpy uint32 array [nbins + 1]."""
        import numpy as np

        n = C.c_int64(0)
        buf = np.zeros((1,), np.uint32)
        rc = self.lib.gsw_debug_bin_offsets(self._h, buf.ctypes.data, 0, C.byref(n), self._stream())
        buf = np.zeros((int(n.value) + 1,), np.uint32)
        _lib.check(self.lib.gsw_debug_bin_offsets(self._h, buf.ctypes.data, int(n.value) + 1, C.byref(n),
                                                  self._stream()))
        return buf

    def debug_frame_lists(self, frame: int, T: int):
        d = self.device
        cap = 1 << 16
        while True:
            ranges = torch.zeros((T, 2), dtype=torch.int32, device=d)
            keys = torch.zeros((cap,), dtype=torch.int64, device=d)
            vals = torch.zeros((cap,), dtype=torch.int32, device=d)
            cnt = C.c_int64(0)
            rc = self.lib.gsw_debug_frame_lists(self._h, int(frame), _ptr(ranges), _ptr(keys), _ptr(vals), cap,
                                                C.byref(cnt), self._stream())
            if rc == _lib.GSW_E_CAPACITY:
                cap = int(cnt.value) + 16
                continue
            _lib.check(rc)
            n = int(cnt.value)
            return dict(ranges=ranges, keys=keys[:n], vals=vals[:n], D=n)
