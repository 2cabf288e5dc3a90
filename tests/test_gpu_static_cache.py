"""Static-list cache across steps (a world-fixed camera's static lists are built
once, not once per step).  Views whose camera is bitwise the one of the previous launch keep their static records,
bins, sorted keys, gathered splats and static-only tiles; everything else is rebuilt.

Each step is rendered by two engines over the same scene and env: one with the cache (its history is the steps
before), one with GSW_NO_STATIC_CACHE=1 set around its launches (everything rebuilt every step).  The uint8 frames,
float colour bits and radii must be identical, and the work counters must show that the cache did skip work.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

H = W = 128
B = 3


@pytest.fixture()
def rig():
    from gsworld_b200 import fixtures as fx
    from gsworld_b200.gaussian_renderer import PipelineParams
    from gsworld_b200.gs_world_wrapper import GSWorldWrapper
    dev = torch.device("cuda:0")
    objects = ("006_mustard_bottle", "snack_box")
    scene = fx.make_scene(fx.SceneSpec(robot="fr3_umi", n_total=150_000, per_link=1500, objects=objects,
                                       per_object=2500, seed=77))
    env = fx.FakeEnv("fr3_umi", num_envs=B, hw=(H, W), objects=objects, seed=5, device=dev)
    scan = fx.forward_kinematics("fr3_umi", fx.rt.ROBOT_SCAN_QPOS["fr3_umi"][None])[0]
    scene_dev = scene.to(dev)
    mk = lambda strict: GSWorldWrapper(env, PipelineParams(), "fr3_pnp_box", device=dev, merged_model=scene_dev,
                                       scan_link_mats=list(scan), strict_reference=strict, tile_culling=True)
    saved = fx._THIRD_MOUNT["fr3_umi"]
    yield fx, env, mk
    fx._THIRD_MOUNT["fr3_umi"] = saved


def _render(wrap, cache):
    names = list(wrap.base_env.cameras)
    wrap.transform_gs_perlink(None)
    view, proj, campos, tanfov, fv = wrap._views(names)
    fe = wrap._maps(len(names))
    if not cache:
        os.environ["GSW_NO_STATIC_CACHE"] = "1"
    try:
        out = wrap._engine.render_frames(wrap._pose, view, proj, campos, tanfov, fe, fv, H, W,
                                         num_envs=wrap.num_envs, want_u8=True, want_color=True, want_radii=True,
                                         want_invdepth=True)
        torch.cuda.synchronize()
    finally:
        os.environ.pop("GSW_NO_STATIC_CACHE", None)
    keep = {k: out[k].clone() for k in ("rgb_u8", "color", "radii", "invdepth")}
    return keep, wrap._engine.counters(), wrap._engine.last_stats["D"], int(view.shape[0])


def _same(a, b):
    assert torch.equal(a["rgb_u8"], b["rgb_u8"])
    assert torch.equal(a["color"].view(torch.int32), b["color"].view(torch.int32))
    assert torch.equal(a["invdepth"].view(torch.int32), b["invdepth"].view(torch.int32))
    assert torch.equal(a["radii"], b["radii"])


def _move_third_person(fx, dx):
    m = np.array(fx._THIRD_MOUNT["fr3_umi"], dtype=np.float32, copy=True)
    m[0, 3] += dx
    fx._THIRD_MOUNT["fr3_umi"] = m  # a new array: the module's constant is put back by the fixture


@pytest.mark.parametrize("strict", [False, True])
def test_kept_static_lists_change_no_byte(rig, strict):
    fx, env, mk = rig
    cached, fresh = mk(strict), mk(strict)
    # step 0: nothing to keep yet
    a, ca, Da, V = _render(cached, True)
    b, cb, Db, _ = _render(fresh, False)
    assert V == (2 if strict else 1 + B)
    _same(a, b)
    assert Da == Db and ca["static_chunks_live"] == cb["static_chunks_live"] > 0
    # step 1: joints move, so the wrist cameras move; the third-person camera (view 0) does not
    env.step()
    a, ca, Da, _ = _render(cached, True)
    b, cb, Db, _ = _render(fresh, False)
    _same(a, b)
    assert Da == Db
    assert 0 < ca["static_chunks_live"] < cb["static_chunks_live"]  # view 0 was not walked again
    # step 2: nothing moves at all: no static chunk is processed, the frames are those of a full rebuild
    a2, ca2, Da2, _ = _render(cached, True)
    b2, cb2, _, _ = _render(fresh, False)
    _same(a2, b2)
    _same(a2, a)
    assert ca2["static_chunks_live"] == 0 and ca2["visible_static"] == 0 and Da2 == Da
    assert cb2["static_chunks_live"] == cb["static_chunks_live"]
    # step 3: the third-person camera itself moves: view 0 changes, so no prefix of views is left to keep
    _move_third_person(fx, 0.05)
    env.step()
    a3, ca3, Da3, _ = _render(cached, True)
    b3, cb3, Db3, _ = _render(fresh, False)
    _same(a3, b3)
    assert Da3 == Db3 and ca3["static_chunks_live"] == cb3["static_chunks_live"] > 0
    assert not torch.equal(a3["rgb_u8"], a2["rgb_u8"])
    # step 4: still again
    a4, ca4, _, _ = _render(cached, True)
    _same(a4, a3)
    assert ca4["static_chunks_live"] == 0


def test_scene_or_setting_change_drops_the_kept_lists(rig):
    fx, env, mk = rig
    cached, fresh = mk(True), mk(True)
    _render(cached, True)
    a, ca, _, _ = _render(cached, True)
    assert ca["static_chunks_live"] == 0
    # exact tile culling off: other lists, so everything is rebuilt once, then kept again
    cached._engine.set_tile_culling(False)
    fresh._engine.set_tile_culling(False)
    a, ca, Da, _ = _render(cached, True)
    b, cb, Db, _ = _render(fresh, False)
    _same(a, b)
    assert Da == Db and ca["static_chunks_live"] == cb["static_chunks_live"] > 0
    a, ca, _, _ = _render(cached, True)
    _same(a, b)
    assert ca["static_chunks_live"] == 0


def test_world_fixed_cameras_are_numbered_first(rig):
    """Sensor order (wrist, third-person): the wrapper still gives the world-fixed camera view 0, so its lists are kept."""
    fx, env, mk = rig
    env.cameras = ("wrist_cam", "right_cam")
    env.sensor_hw = {c: (H, W) for c in env.cameras}
    for strict in (True, False):
        cached, fresh = mk(strict), mk(strict)
        _render(cached, True)
        env.step()
        a, ca, Da, _ = _render(cached, True)
        b, cb, Db, _ = _render(fresh, False)
        _same(a, b)
        assert Da == Db and 0 < ca["static_chunks_live"] < cb["static_chunks_live"]


def test_static_chunk_size_changes_no_byte(rig):
    """The static layer's chunk (items per work-list entry) is picked per launch from the number of views; every
    choice must give the same bytes (GSW_IPB_STATIC forces one; read at every launch)."""
    fx, env, mk = rig
    for strict in (True, False):
        ref, _, Dref, _ = _render(mk(strict), False)
        for ipb in ("256", "512", "1024", "2048"):
            os.environ["GSW_IPB_STATIC"] = ipb
            try:
                w = mk(strict)
                a, _, Da, _ = _render(w, True)
                env_state = env._q.copy()
                b, cb, Db, _ = _render(w, True)  # second launch: kept lists, same chunk size
            finally:
                os.environ.pop("GSW_IPB_STATIC", None)
            assert (env._q == env_state).all()
            _same(a, ref)
            _same(b, ref)
            assert Da == Dref == Db and cb["static_chunks_live"] == 0
