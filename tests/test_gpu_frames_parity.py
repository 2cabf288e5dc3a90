"""GPU parity of the batched entry point (gsw_render_frames) through the GSWorldWrapper mirror: fused
per-link / per-object transform + rasterization of every (env, camera) frame vs the CPU oracle run frame
by frame on the same pose table and cameras."""
import math

import numpy as np
import pytest
import torch

from oracle import oracle as orc
from tests.helpers import assert_image_close, oracle_frame_inputs

pytestmark = pytest.mark.gpu


def _make(robot, cfg_name, B, hw, objects, strict, n_total=24_000, seed=11, nb=0):
    from gsworld_b200 import fixtures as fx
    from gsworld_b200.gaussian_renderer import PipelineParams
    from gsworld_b200.gs_world_wrapper import GSWorldWrapper
    dev = torch.device("cuda:0")
    scene = fx.make_scene(fx.SceneSpec(robot=robot, n_total=n_total, per_link=250, objects=objects, per_object=1200,
                                       seed=seed))
    env = fx.FakeEnv(robot, num_envs=B, hw=hw, objects=objects, seed=seed, device=dev)
    scan = fx.forward_kinematics(robot, fx.rt.ROBOT_SCAN_QPOS[robot][None])[0]
    wrap = GSWorldWrapper(env, PipelineParams(), cfg_name, device=dev, merged_model=scene.to(dev),
                          scan_link_mats=list(scan), strict_reference=strict, tile_culling=False)  # reference lists
    if nb:
        wrap._engine.set_depth_buckets(nb)
    return scene, env, wrap


def _check_frames(scene, env, wrap, H, W, exact_objects=True):
    eng = wrap._engine
    torch.cuda.synchronize()
    pose = wrap._pose.cpu().numpy()
    slot = eng._keep["slot"].cpu().numpy()
    names = list(env.cameras)
    params = env.get_sensor_params()
    T = ((H + 15) // 16) * ((W + 15) // 16)
    # second launch with float outputs for the tolerance checks
    view, proj, campos, tanfov, fv = wrap._views(names)
    fe = wrap._maps(len(names))
    out = eng.render_frames(wrap._pose, view, proj, campos, tanfov, fe, fv, H, W, num_envs=env.num_envs,
                            want_u8=True, want_color=True, want_invdepth=True, want_radii=True)
    torch.cuda.synchronize()
    n_obj_mismatch = 0
    for ci, name in enumerate(names):
        for e in range(env.num_envs):
            f = ci * env.num_envs + e
            v = int(fv[f])
            fin = oracle_frame_inputs(scene, slot, pose[e], len(wrap._link_names))
            want = orc.rasterize_forward(fin["xyz"], fin["opacities"], view[v].cpu().numpy(), proj[v].cpu().numpy(),
                                         campos[v].cpu().numpy(), float(tanfov[v, 0]), float(tanfov[v, 1]), H, W,
                                         shs=fin["shs"], scales=fin["scales"], rotations=fin["rotations"], debug=True)
            radii = out["radii"][f].cpu().numpy()
            attrs = {k: t.cpu().numpy() for k, t in eng.debug_frame_attrs(f).items()}
            lists = eng.debug_frame_lists(f, T)
            is_obj = slot >= len(wrap._link_names)
            # link + static Gaussians: exact.  Object Gaussians go through expf(logf(.)) (reference quirk Q4);
            # CUDA libdevice and glibc may differ by an ulp there, which can flip a ceil() in rare cases.
            np.testing.assert_array_equal(radii[~is_obj], want["radii"][~is_obj])
            mism = int((radii[is_obj] != want["radii"][is_obj]).sum())
            n_obj_mismatch += mism
            if mism == 0:
                np.testing.assert_array_equal(attrs["tiles_touched"], want["tiles_touched"])
                assert lists["D"] == want["D"]
                np.testing.assert_array_equal(lists["keys"].cpu().numpy().view(np.uint64), want["keys"])
                np.testing.assert_array_equal(lists["vals"].cpu().numpy().view(np.uint32), want["vals"])
                np.testing.assert_array_equal(lists["ranges"].cpu().numpy(), want["ranges"])
                vis = (want["radii"] > 0) & ~is_obj
                bits = lambda a: np.ascontiguousarray(a, np.float32).view(np.uint32)
                np.testing.assert_array_equal(bits(attrs["depth"][vis]), bits(want["depth"][vis]))
                np.testing.assert_array_equal(bits(attrs["xy"][vis]), bits(want["xy"][vis]))
                np.testing.assert_array_equal(bits(attrs["rgb"][vis]), bits(want["rgb"][vis]))
            frag = want["fragile"].astype(bool)
            assert_image_close(out["color"][f].cpu().numpy(), want["color"], frag, what=f"color f{f}")
            assert_image_close(out["invdepth"][f, 0].cpu().numpy(), want["invdepth"][0], frag, what=f"invdepth f{f}")
            ref8 = orc.to_uint8_hwc(np.clip(want["color"], 0, 1)).astype(np.int32)
            got8 = out["rgb_u8"][f].cpu().numpy().astype(np.int32)
            assert (np.abs(got8 - ref8) <= 1)[~frag].all()
            assert (got8 == ref8).mean() > 0.98
    assert n_obj_mismatch <= 2, f"{n_obj_mismatch} object-Gaussian radii differ from the oracle"


@pytest.mark.parametrize("strict", [True, False], ids=["strict_env0_camera", "per_env_camera"])
def test_fr3_two_cameras(strict):
    scene, env, wrap = _make("fr3_umi", "fr3_pnp_box", 3, (64, 48), ("006_mustard_bottle", "snack_box"), strict)
    obs, _ = wrap.reset(seed=5)
    _check_frames(scene, env, wrap, 64, 48)
    obs, *_ = wrap.step(None)
    assert obs["sensor_data"]["right_cam"]["rgb"].shape == (3, 64, 48, 3)
    assert obs["sensor_data"]["right_cam"]["rgb"].dtype == torch.uint8
    _check_frames(scene, env, wrap, 64, 48)


def test_xarm_offset_and_bread_scale():
    # xarm: +0.05 z on every link (Q8); bread_slice: object_scale 0.95 and a 13x object frame (Q4 path)
    scene, env, wrap = _make("xarm6_uf_gripper", "xarm6_align", 2, (48, 80), ("005_tomato_soup_can", "bread_slice"),
                             True, seed=13)
    wrap.reset(seed=2)
    _check_frames(scene, env, wrap, 48, 80)


def test_depth_buckets_do_not_change_results():
    scene, env, wrap = _make("fr3_umi", "fr3_pnp_box", 2, (64, 64), ("006_mustard_bottle",), True, nb=8)
    wrap.reset(seed=9)
    _check_frames(scene, env, wrap, 64, 64)


def test_entry_workspace_overflow_is_detected_and_recovered():
    scene, env, wrap = _make("fr3_umi", "fr3_pnp_box", 2, (64, 64), ("snack_box",), True)
    eng = wrap._engine
    # force a tiny capacity by constructing a fresh engine and reserving too little is not possible
    # (capacity only grows), so check the protocol instead: stats are reported and D + merged <= cap.
    wrap.reset(seed=1)
    st = eng.last_stats
    assert st["D"] > 0 and st["D"] <= st["cap"] and st["staged"] > 0


def test_wrapper_from_ply_config_and_inverse_depth_obs(tmp_path):
    """Scene through the PLY + semantics loader (reference config format) renders byte-identically to the same
    scene passed in memory; the optional inverse-depth observation has the reference's 'depth' values."""
    import json
    import os
    from gsworld_b200 import fixtures as fx
    from gsworld_b200.gaussian_merger import save_ply
    from gsworld_b200.gaussian_renderer import PipelineParams
    from gsworld_b200.gs_world_wrapper import GSWorldWrapper
    dev = torch.device("cuda:0")
    objects = ("006_mustard_bottle",)
    scene = fx.make_scene(fx.SceneSpec(robot="fr3_umi", n_total=9000, per_link=200, objects=objects, per_object=800,
                                       seed=5))
    os.makedirs(tmp_path / "scene")
    save_ply(scene, str(tmp_path / "scene" / "all.ply"), with_semantics=False)
    np.save(tmp_path / "scene" / "all_semantics_gs.npy", scene._semantics.reshape(-1).numpy().astype(np.int64))
    (tmp_path / "fr3_pnp_box.json").write_text(json.dumps({"models": [
        {"data_path": "./scene/all.ply", "semantic_labels": "./scene/all_semantics_gs.npy", "transformation": []}]}))
    env = fx.FakeEnv("fr3_umi", num_envs=2, hw=(48, 48), objects=objects, seed=3, device=dev)
    scan = fx.forward_kinematics("fr3_umi", fx.rt.ROBOT_SCAN_QPOS["fr3_umi"][None])[0]
    w_mem = GSWorldWrapper(env, PipelineParams(), "fr3_pnp_box", device=dev, merged_model=scene.to(dev),
                           scan_link_mats=list(scan))
    w_ply = GSWorldWrapper(env, PipelineParams(), "fr3_pnp_box", device=dev, cfg_dir=str(tmp_path),
                           asset_dir=str(tmp_path), scan_link_mats=list(scan), inv_depth_obs=True)
    o1, _ = w_mem.reset(seed=3)
    o2, _ = w_ply.reset(seed=3)
    for cam in env.cameras:
        assert torch.equal(o1["sensor_data"][cam]["rgb"], o2["sensor_data"][cam]["rgb"])
        d = o2["sensor_data"][cam]["inv_depth"]
        assert d.shape == (2, 48, 48, 1) and d.dtype == torch.float32 and float(d.min()) >= 0.0 and float(d.max()) > 0.0


@pytest.mark.parametrize("B,wait_value", [(3, True), (3, False), (11, True), (5, "after")])
def test_host_delivery_matches_device_output(B, wait_value, monkeypatch):
    """gsw_frames_desc.host_rgb_u8: delivery overlapped with the blend must hand the host exactly the frames the
    plain path leaves on the device.  wait_value=True: one blend launch + in-kernel group flags + a stream memory
    wait on the copy stream (frame counts below and above the 8 groups); False: the fallback, grouped launches
    with events; "after": delivery after a plain blend launch, render stream not joined.  Several steps in a row, so
    the epoch / counters are reused."""
    if not wait_value:
        monkeypatch.setenv("GSW_NO_WAIT_VALUE", "1")  # read by gsw_create
    scene, env, wrap = _make("fr3_umi", "fr3_pnp_box", B, (64, 48), ("006_mustard_bottle",), True)
    wrap.reset(seed=5)
    pinned = torch.zeros((len(env.cameras), B, 64, 48, 3), dtype=torch.uint8).pin_memory()
    for _ in range(3):
        wrap.obs_host = None
        obs = wrap.step(None)[0]
        torch.cuda.synchronize()
        want = {c: obs["sensor_data"][c]["rgb"].clone() for c in env.cameras}
        pinned.zero_()
        wrap.obs_host = pinned
        if wait_value == "after":  # GSW_DELIVER_AFTER_BLEND | GSW_DELIVER_NO_JOIN: one copy after a plain blend launch
            wrap.defer_delivery_join = "after"  # (deferred joins need a caller-owned output buffer)
            wrap.obs_out = torch.zeros((len(env.cameras), B, 64, 48, 3), dtype=torch.uint8, device="cuda:0")
        out = wrap._render_gsworld()
        if wait_value == "after":
            wrap._engine.delivery_wait()  # the render stream did not wait for the copy
            wrap.defer_delivery_join = False
            wrap.obs_out = None
        torch.cuda.current_stream().synchronize()
        for i, c in enumerate(env.cameras):
            assert torch.equal(out[c].cpu(), want[c].cpu())
            assert torch.equal(pinned[i], want[c].cpu())


def test_cfg4_like_pour_mustard_objects_128():
    # BASELINE config 4 shape at test size: fr3_umi, 128x128, 2 cameras, mustard bottle + bread slice
    # (object_scale 0.95, 13x object frame: the kappa path of quirk Q4), per-env cameras
    scene, env, wrap = _make("fr3_umi", "fr3_pour_mustard", 4, (128, 128), ("006_mustard_bottle", "bread_slice"),
                             False, n_total=40_000, seed=17)
    wrap.reset(seed=4)
    _check_frames(scene, env, wrap, 128, 128)


@pytest.mark.parametrize("strict", [True, False], ids=["shared_views", "per_env_views"])
def test_segmentation_observation_matches_oracle_argmax(strict):
    """out_seg: label of the Gaussian with the largest blending weight.  Checked against the oracle's best_id on
    every pixel that is neither threshold-fragile nor a near tie between the two largest weights."""
    from gsworld_b200 import fixtures as fx
    from gsworld_b200.gaussian_renderer import PipelineParams
    from gsworld_b200.gs_world_wrapper import GSWorldWrapper
    dev = torch.device("cuda:0")
    objects = ("006_mustard_bottle", "snack_box")
    scene = fx.make_scene(fx.SceneSpec(robot="fr3_umi", n_total=24_000, per_link=250, objects=objects, per_object=1200,
                                       seed=23))
    env = fx.FakeEnv("fr3_umi", num_envs=3, hw=(64, 64), objects=objects, seed=23, device=dev)
    scan = fx.forward_kinematics("fr3_umi", fx.rt.ROBOT_SCAN_QPOS["fr3_umi"][None])[0]
    wrap = GSWorldWrapper(env, PipelineParams(), "fr3_pnp_box", device=dev, merged_model=scene.to(dev),
                          scan_link_mats=list(scan), strict_reference=strict, segmentation_obs=True)
    obs, _ = wrap.reset(seed=7)
    torch.cuda.synchronize()
    labels = scene._semantics.reshape(-1).numpy().astype(np.int32)
    pose = wrap._pose.cpu().numpy()
    slot = wrap._engine._keep["slot"].cpu().numpy()
    names = list(env.cameras)
    view, proj, campos, tanfov, fv = wrap._views(names)
    checked = 0
    for ci, name in enumerate(names):
        seg = obs["sensor_data"][name]["segmentation"]
        assert seg.shape == (3, 64, 64, 1) and seg.dtype == torch.int32
        for e in range(3):
            v = int(fv[ci * 3 + e])
            fin = oracle_frame_inputs(scene, slot, pose[e], 14)
            want = orc.rasterize_forward(fin["xyz"], fin["opacities"], view[v].cpu().numpy(), proj[v].cpu().numpy(),
                                         campos[v].cpu().numpy(), float(tanfov[v, 0]), float(tanfov[v, 1]), 64, 64,
                                         shs=fin["shs"], scales=fin["scales"], rotations=fin["rotations"], debug=True)
            exp = np.where(want["best_id"] >= 0, labels[np.maximum(want["best_id"], 0)], -2)
            ok = (~want["fragile"].astype(bool)) & (want["best_margin"] > 1e-3)
            got = seg[e, :, :, 0].cpu().numpy()
            assert (got[ok] == exp[ok]).all()
            checked += int(ok.sum())
    assert checked > 0.9 * 2 * 3 * 64 * 64


@pytest.mark.gpu
@pytest.mark.parametrize("strict", [True, False])
def test_host_simulator_inputs_match_device_inputs(strict):
    """A CPU simulator hands the wrapper host tensors (link / object / camera poses): they go through the pinned
    staging arenas and one asynchronous copy each.  Same seed on a device-resident env must give the same pose
    table, the same cameras and byte-identical observations, over several steps (both arenas get reused)."""
    from gsworld_b200 import fixtures as fx
    from gsworld_b200.gaussian_renderer import PipelineParams
    from gsworld_b200.gs_world_wrapper import GSWorldWrapper
    dev = torch.device("cuda:0")
    objects = ("006_mustard_bottle", "snack_box")
    scene = fx.make_scene(fx.SceneSpec(robot="fr3_umi", n_total=20_000, per_link=200, objects=objects, per_object=900,
                                       seed=31)).to(dev)
    scan = fx.forward_kinematics("fr3_umi", fx.rt.ROBOT_SCAN_QPOS["fr3_umi"][None])[0]
    from types import SimpleNamespace

    class HostView:
        """The same FakeEnv numbers, handed out as CPU tensors (what a CPU simulator would give)."""

        def __init__(self, env):
            self._e = env
            links = env.agent.robot.get_links()
            mk = lambda l: SimpleNamespace(name=l.name, pose=SimpleNamespace(
                to_transformation_matrix=lambda: l.pose.to_transformation_matrix().cpu()))
            self.agent = SimpleNamespace(uid=env.agent.uid, robot=SimpleNamespace(
                name=env.agent.robot.name, get_links=lambda: [mk(l) for l in links]))

        def __getattr__(self, name):
            return getattr(self._e, name)

        @property
        def unwrapped(self):
            return self

        def get_state_dict(self):
            return {"actors": {k: v.cpu() for k, v in self._e.get_state_dict()["actors"].items()}}

        def get_sensor_params(self):
            return {c: {k: v.cpu() for k, v in p.items()} for c, p in self._e.get_sensor_params().items()}

    wraps = []
    for host in (True, False):
        env = fx.FakeEnv("fr3_umi", num_envs=3, hw=(48, 64), objects=objects, seed=31, device=dev)
        wraps.append(GSWorldWrapper(HostView(env) if host else env, PipelineParams(), "fr3_pnp_box", device=dev,
                                    merged_model=scene, scan_link_mats=list(scan), strict_reference=strict))
    obs = [w.reset(seed=5)[0] for w in wraps]
    for step in range(4):
        if step:
            obs = [w.step(None)[0] for w in wraps]
        torch.cuda.synchronize()
        assert "pose" in wraps[0]._stages and "pose" not in wraps[1]._stages
        assert torch.equal(wraps[0]._pose, wraps[1]._pose)
        for name in wraps[0].base_env.cameras:
            a, b = obs[0]["sensor_data"][name]["rgb"], obs[1]["sensor_data"][name]["rgb"]
            assert a.device.type == "cuda" and torch.equal(a, b)
