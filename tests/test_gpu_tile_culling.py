"""Exact tile culling (gsw_set_tile_culling, include/gsw.h).

With culling on, a (Gaussian, tile) entry that provably stays below alpha 1/255 on every pixel of the tile is
never created.  The reference's blend loop skips such an entry without touching any state, so:

* colour, inverse depth, uint8 observations, radii and the per-Gaussian tile-count tap are BIT-IDENTICAL to the
  run without culling (which the other parity tests pin to the oracle);
* every tile list is an order-preserving sub-list of the reference's list (= of the oracle's);
* every omitted entry is dead by the ORACLE's own numbers: recomputed here in float64 from the oracle's
  pixel position / conic / opacity taps, its alpha stays below 1/255 (or its power above 0) on every pixel
  of the tile that lies inside the image.
"""
import numpy as np
import pytest
import torch

from oracle import oracle as orc
from tests.helpers import look_at_camera, random_gaussians

pytestmark = pytest.mark.gpu


@pytest.fixture()
def eng():
    from gsworld_b200.engine import Engine
    e = Engine("cuda:0")
    yield e
    e.close()


def _gpu(eng, g, cam, H, W, **kw):
    dev = eng.device
    color, radii, invd = eng.rasterize_forward(
        g["xyz"].to(dev), g["opacities"].to(dev), cam["view"].to(dev), cam["proj"].to(dev), cam["campos"].to(dev),
        cam["tanfovx"], cam["tanfovy"], H, W, torch.zeros(3, device=dev), shs=g["shs"].to(dev),
        scales=g["scales"].to(dev), rotations=g["rotations"].to(dev), **kw)
    T = ((H + 15) // 16) * ((W + 15) // 16)
    attrs = {k: v.cpu().numpy() for k, v in eng.debug_frame_attrs(0).items()}
    lists = {k: (v.cpu().numpy() if torch.is_tensor(v) else v) for k, v in eng.debug_frame_lists(0, T).items()}
    return color.cpu().numpy(), radii.cpu().numpy(), invd.cpu().numpy(), attrs, lists


def _max_alpha_on_tile(xy, conic_o, tile, gx, H, W):
    """float64 restatement of the reference's per-pixel alpha for one Gaussian over one tile's in-image pixels."""
    tx, ty = tile % gx, tile // gx
    xs = np.arange(tx * 16, min(tx * 16 + 16, W), dtype=np.float64)
    ys = np.arange(ty * 16, min(ty * 16 + 16, H), dtype=np.float64)
    dx = float(xy[0]) - xs[None, :]
    dy = float(xy[1]) - ys[:, None]
    A, B, C, o = (float(v) for v in conic_o)
    power = -0.5 * (A * dx * dx + C * dy * dy) - B * dx * dy
    alpha = np.minimum(0.99, o * np.exp(np.minimum(power, 0.0)))
    alpha[power > 0.0] = 0.0
    return float(alpha.max()) if alpha.size else 0.0


CASES = [
    dict(n=6000, H=64, W=64, seed=1),
    dict(n=20000, H=128, W=96, seed=2),                       # ragged image
    dict(n=4000, H=96, W=96, seed=3, log_scale_std=1.5),       # very elongated splats: most of the square is dead
    dict(n=2000, H=64, W=64, seed=13, log_scale_mean=-0.5),    # huge splats: rectangles of > 64 tiles never occur at
                                                               # 4x4 tiles, lists are long
    dict(n=1500, H=256, W=256, seed=14, log_scale_mean=-0.7),  # rectangles of > 64 tiles: tested again by the scatter
    dict(n=6000, H=64, W=64, seed=4, antialiasing=True),
    dict(n=6000, H=64, W=64, seed=7, scale_modifier=1.7),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items()))
def test_culling_changes_no_pixel_and_drops_only_dead_entries(eng, case):
    case = dict(case)
    n, H, W, seed = case.pop("n"), case.pop("H"), case.pop("W"), case.pop("seed")
    g = random_gaussians(n, seed=seed, log_scale_mean=case.pop("log_scale_mean", np.log(0.05)),
                         log_scale_std=case.pop("log_scale_std", 0.6))
    g["opacities"] = g["opacities"] * torch.rand(n, generator=torch.Generator().manual_seed(seed))  # many faint ones
    cam = look_at_camera(H, W, eye=(0.3, -0.2, -3.0))
    eng.set_tile_culling(False)
    c0, r0, d0, a0, l0 = _gpu(eng, g, cam, H, W, **case)
    eng.set_tile_culling(True)
    c1, r1, d1, a1, l1 = _gpu(eng, g, cam, H, W, **case)
    # images, radii, taps: bit-identical
    assert np.array_equal(c0.view(np.uint32), c1.view(np.uint32))
    assert np.array_equal(d0.view(np.uint32), d1.view(np.uint32))
    assert np.array_equal(r0, r1)
    for k in a0:
        assert np.array_equal(a0[k], a1[k]), k
    # lists: order-preserving sub-lists, tile by tile
    assert l1["D"] < l0["D"], "nothing was culled: the case does not exercise the feature"
    k0, v0 = l0["keys"].view(np.uint64), l0["vals"].view(np.uint32)
    k1, v1 = l1["keys"].view(np.uint64), l1["vals"].view(np.uint32)
    pair0 = (k0 >> np.uint64(32)) << np.uint64(32) | v0.astype(np.uint64)  # (tile, Gaussian) of every full-list entry
    pair1 = (k1 >> np.uint64(32)) << np.uint64(32) | v1.astype(np.uint64)
    assert len(np.unique(pair0)) == len(pair0)
    keep = np.isin(pair0, pair1)
    assert int(keep.sum()) == len(pair1)
    np.testing.assert_array_equal(k0[keep], k1)  # same keys, same order
    np.testing.assert_array_equal(v0[keep], v1)
    T = l0["ranges"].shape[0]
    cnt1 = np.bincount((k1 >> np.uint64(32)).astype(np.int64), minlength=T)
    np.testing.assert_array_equal(l1["ranges"][:, 1] - l1["ranges"][:, 0], cnt1)
    # every dropped entry is dead by the oracle's numbers
    want = orc.rasterize_forward(g["xyz"].numpy(), g["opacities"].numpy(), cam["view"].numpy(), cam["proj"].numpy(),
                                 cam["campos"].numpy(), cam["tanfovx"], cam["tanfovy"], H, W, shs=g["shs"].numpy(),
                                 scales=g["scales"].numpy(), rotations=g["rotations"].numpy(), debug=True, **case)
    assert want["D"] == l0["D"]
    gx = (W + 15) // 16
    dropped = np.nonzero(~keep)[0]
    rng = np.random.default_rng(seed)
    if len(dropped) > 4000:  # bounded run time: a random sample of the dropped entries
        dropped = rng.choice(dropped, 4000, replace=False)
    worst = 0.0
    for e in dropped:
        tile, idx = int(k0[e] >> np.uint64(32)), int(v0[e])
        worst = max(worst, _max_alpha_on_tile(want["xy"][idx], want["conic_opacity"][idx], tile, gx, H, W))
    assert worst < 1.0 / 255.0, f"a dropped entry reaches alpha {worst}"


@pytest.mark.parametrize("strict", [True, False])
def test_wrapper_observations_identical_with_culling(strict):
    from gsworld_b200 import fixtures as fx
    from gsworld_b200.gaussian_renderer import PipelineParams
    from gsworld_b200.gs_world_wrapper import GSWorldWrapper
    dev = torch.device("cuda:0")
    objects = ("006_mustard_bottle", "snack_box")
    scene = fx.make_scene(fx.SceneSpec(robot="fr3_umi", n_total=60_000, per_link=600, objects=objects, per_object=2000,
                                       seed=41)).to(dev)
    scan = fx.forward_kinematics("fr3_umi", fx.rt.ROBOT_SCAN_QPOS["fr3_umi"][None])[0]
    wraps = []
    for cull in (False, True):
        env = fx.FakeEnv("fr3_umi", num_envs=4, hw=(128, 128), objects=objects, seed=41, device=dev)
        w = GSWorldWrapper(env, PipelineParams(), "fr3_pnp_box", device=dev, merged_model=scene,
                           scan_link_mats=list(scan), strict_reference=strict, inv_depth_obs=True,
                           segmentation_obs=True, tile_culling=cull)
        wraps.append(w)
    obs = [w.reset(seed=3)[0] for w in wraps]
    for step in range(3):
        if step:
            obs = [w.step(None)[0] for w in wraps]
        torch.cuda.synchronize()
        assert wraps[1]._engine.last_stats["D"] < wraps[0]._engine.last_stats["D"]
        for cam in wraps[0].base_env.cameras:
            for key in ("rgb", "inv_depth", "segmentation"):
                assert torch.equal(obs[0]["sensor_data"][cam][key], obs[1]["sensor_data"][cam][key]), (cam, key)
