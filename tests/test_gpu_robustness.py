"""One Engine reused across changing shapes, scenes and degenerate layer splits: grow-only workspace, re-upload,
no movable Gaussians, only movable Gaussians.  Radii (exact) and colour (1e-3) against the oracle each time."""
import numpy as np
import pytest
import torch

from oracle import oracle as orc
from tests.helpers import assert_image_close, look_at_camera, random_gaussians

pytestmark = pytest.mark.gpu


def _identity_pose(B, S, dev):
    p = torch.zeros((B, S, 20), device=dev)
    p[..., 0] = p[..., 4] = p[..., 8] = 1.0   # R = I
    p[..., 12] = 1.0                           # qR = (1,0,0,0)
    return p


def _render_and_check(eng, g, slot, S, B, H, W, eye, seed):
    dev = eng.device
    cam = look_at_camera(H, W, eye=eye)
    d = lambda t: t.to(dev).contiguous()
    eng.set_scene(d(g["xyz"]), d(g["scales"]), d(g["rotations"]), d(g["opacities"]), d(g["shs"]), d(g["log_scale"]),
                  d(g["quat_raw"]), None if slot is None else d(slot), S)
    view = d(cam["view"])[None]
    proj = d(cam["proj"])[None]
    campos = d(cam["campos"])[None]
    tanfov = torch.tensor([[cam["tanfovx"], cam["tanfovy"]]], device=dev)
    fe = torch.arange(B, dtype=torch.int32, device=dev)
    fv = torch.zeros(B, dtype=torch.int32, device=dev)
    out = eng.render_frames(_identity_pose(B, max(S, 1), dev) if S else None, view, proj, campos, tanfov, fe, fv, H, W,
                            num_envs=B, want_u8=True, want_color=True, want_radii=True)
    torch.cuda.synchronize()
    # identity poses: movable Gaussians keep their place; their quaternion goes through normalise(q/|q|*|q|)
    q = g["quat_raw"]
    rot = g["rotations"].clone()
    if slot is not None:
        m = slot >= 0
        nrm = torch.sqrt((q * q).sum(-1, keepdim=True))
        q2 = (q / nrm)
        q2 = torch.where(q2[:, :1] < 0, -q2, q2) * nrm
        rot[m] = (q2 / torch.sqrt((q2 * q2).sum(-1, keepdim=True)).clamp_min(1e-12))[m]
    want = orc.rasterize_forward(g["xyz"].numpy(), g["opacities"].numpy(), cam["view"].numpy(), cam["proj"].numpy(),
                                 cam["campos"].numpy(), cam["tanfovx"], cam["tanfovy"], H, W, shs=g["shs"].numpy(),
                                 scales=g["scales"].numpy(), rotations=rot.numpy(), debug=True)
    for f in range(B):
        mism = (out["radii"][f].cpu().numpy() != want["radii"]).sum()
        assert mism <= (0 if slot is None else 2), f"{mism} radii differ (frame {f})"
        assert_image_close(out["color"][f].cpu().numpy(), want["color"], want["fragile"].astype(bool))


def test_engine_reuse_across_shapes_and_scenes():
    from gsworld_b200.engine import Engine
    eng = Engine("cuda:0")
    gen = torch.Generator().manual_seed(3)
    for n, H, W, B, mode in [(3000, 64, 64, 2, "mixed"), (12000, 128, 96, 3, "mixed"), (500, 32, 48, 1, "static"),
                             (4000, 80, 64, 2, "movable"), (20000, 64, 64, 4, "mixed"), (300, 256, 256, 2, "static")]:
        g = random_gaussians(n, seed=n)
        if mode == "static":
            slot, S = None, 0
        elif mode == "movable":
            slot, S = torch.randint(0, 3, (n,), generator=gen, dtype=torch.int32), 3
        else:
            slot = torch.randint(-1, 4, (n,), generator=gen, dtype=torch.int32)
            slot[torch.rand(n, generator=gen) < 0.7] = -1
            S = 4
        _render_and_check(eng, g, slot, S, B, H, W, (0.2, -0.1, -3.0), n)
    eng.close()


def test_invalid_arguments_are_rejected_with_messages():
    from gsworld_b200 import _lib
    from gsworld_b200.engine import Engine
    eng = Engine("cuda:0")
    dev = eng.device
    z = torch.zeros(16, device=dev)
    i32 = torch.zeros(4, dtype=torch.int32, device=dev)
    with pytest.raises(_lib.GswError, match="gsw_scene_set first"):
        eng.render_frames(None, z.view(1, 16), z.view(1, 16), z[:3].view(1, 3), z[:2].view(1, 2), i32[:1], i32[:1], 32,
                          32, num_envs=1)
    with pytest.raises(ValueError):
        eng.set_scene(z[:12].view(4, 3).cpu(), z[:12].view(4, 3), z.view(4, 4), z[:4], torch.zeros(4, 16, 3, device=dev))
    with pytest.raises(_lib.GswError, match="power of two"):
        eng.set_depth_buckets(3)
    eng.close()


def test_very_large_image_uses_global_atomic_binning_path():
    """20608 tiles: the per-CTA shared-memory tile histogram no longer fits, so counting/scatter fall back to
    global atomics; results must not change."""
    from gsworld_b200.engine import Engine
    eng = Engine("cuda:0")
    dev = eng.device
    H, W = 2048, 2576
    g = random_gaussians(3000, seed=77, log_scale_mean=np.log(0.03))
    cam = look_at_camera(H, W, eye=(0.1, 0.2, -3.0))
    want = orc.rasterize_forward(g["xyz"].numpy(), g["opacities"].numpy(), cam["view"].numpy(), cam["proj"].numpy(),
                                 cam["campos"].numpy(), cam["tanfovx"], cam["tanfovy"], H, W, shs=g["shs"].numpy(),
                                 scales=g["scales"].numpy(), rotations=g["rotations"].numpy(), debug=True)
    color, radii, invd = eng.rasterize_forward(
        g["xyz"].to(dev), g["opacities"].to(dev), cam["view"].to(dev), cam["proj"].to(dev), cam["campos"].to(dev),
        cam["tanfovx"], cam["tanfovy"], H, W, torch.zeros(3, device=dev), shs=g["shs"].to(dev),
        scales=g["scales"].to(dev), rotations=g["rotations"].to(dev))
    torch.cuda.synchronize()
    np.testing.assert_array_equal(radii.cpu().numpy(), want["radii"])
    T = ((H + 15) // 16) * ((W + 15) // 16)
    lists = eng.debug_frame_lists(0, T)
    assert lists["D"] == want["D"]
    np.testing.assert_array_equal(lists["vals"].cpu().numpy().view(np.uint32), want["vals"])
    assert_image_close(color.cpu().numpy(), want["color"], want["fragile"].astype(bool))
    # exact tile culling on the same global-atomic path: same pixels, fewer entries
    eng.set_tile_culling(True)
    color2, radii2, invd2 = eng.rasterize_forward(
        g["xyz"].to(dev), g["opacities"].to(dev), cam["view"].to(dev), cam["proj"].to(dev), cam["campos"].to(dev),
        cam["tanfovx"], cam["tanfovy"], H, W, torch.zeros(3, device=dev), shs=g["shs"].to(dev),
        scales=g["scales"].to(dev), rotations=g["rotations"].to(dev))
    torch.cuda.synchronize()
    assert torch.equal(color.view(torch.int32), color2.view(torch.int32)) and torch.equal(radii, radii2)
    assert eng.debug_frame_lists(0, T)["D"] < want["D"]
    eng.close()
