"""GPU parity of the single-frame C-ABI entry point (gsw_rasterize_forward) against the CPU oracle.

Integer / bit quantities (radii, tiles touched, depth bits, pixel positions, sorted keys, value order,
tile ranges) must be exact.  Colour and inverse depth: max-abs 1e-3 (north_star tolerance) on every pixel
the oracle does not flag as sitting on a discontinuous threshold (alpha == 1/255, T == 1e-4), PSNR >= 50 dB.
"""
import numpy as np
import pytest
import torch

from oracle import oracle as orc
from tests.helpers import assert_image_close, look_at_camera, psnr, random_gaussians

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from gsworld_b200.engine import Engine
    e = Engine("cuda:0")
    yield e
    e.close()


def _run_both(eng, g, cam, H, W, *, bg=(0.0, 0.0, 0.0), antialiasing=False, sh_degree=3, scale_modifier=1.0,
              near=0.05, use_cov=False, use_colors=False):
    dev = eng.device
    kw_o, kw_g = {}, {}
    cov = colors = None
    if use_cov:
        cov = torch.tensor(orc.cov3d(g["scales"].numpy(), g["rotations"].numpy(), scale_modifier))
        kw_o["cov3D_precomp"], kw_g["cov3D_precomp"] = cov.numpy(), cov.to(dev)
    else:
        kw_o.update(scales=g["scales"].numpy(), rotations=g["rotations"].numpy())
        kw_g.update(scales=g["scales"].to(dev), rotations=g["rotations"].to(dev))
    if use_colors:
        colors = torch.rand(g["xyz"].shape[0], 3, generator=torch.Generator().manual_seed(5))
        kw_o["colors_precomp"], kw_g["colors_precomp"] = colors.numpy(), colors.to(dev)
    else:
        kw_o["shs"], kw_g["shs"] = g["shs"].numpy(), g["shs"].to(dev)
    want = orc.rasterize_forward(g["xyz"].numpy(), g["opacities"].numpy(), cam["view"].numpy(), cam["proj"].numpy(),
                                 cam["campos"].numpy(), cam["tanfovx"], cam["tanfovy"], H, W, sh_degree=sh_degree,
                                 bg=bg, scale_modifier=scale_modifier, antialiasing=antialiasing, near=near,
                                 debug=True, **kw_o)
    color, radii, invd = eng.rasterize_forward(
        g["xyz"].to(dev), g["opacities"].to(dev), cam["view"].to(dev), cam["proj"].to(dev), cam["campos"].to(dev),
        cam["tanfovx"], cam["tanfovy"], H, W, torch.tensor(bg, dtype=torch.float32, device=dev), sh_degree=sh_degree,
        scale_modifier=scale_modifier, antialiasing=antialiasing, near=near, **kw_g)
    T = ((H + 15) // 16) * ((W + 15) // 16)
    attrs = {k: v.cpu().numpy() for k, v in eng.debug_frame_attrs(0).items()}
    lists = {k: (v.cpu().numpy() if torch.is_tensor(v) else v) for k, v in eng.debug_frame_lists(0, T).items()}
    return want, color.cpu().numpy(), radii.cpu().numpy(), invd.cpu().numpy(), attrs, lists


def _assert_exact(want, radii, attrs, lists):
    np.testing.assert_array_equal(radii, want["radii"])
    np.testing.assert_array_equal(attrs["radii"], want["radii"])
    np.testing.assert_array_equal(attrs["tiles_touched"], want["tiles_touched"])
    vis = want["radii"] > 0
    bits = lambda a: np.ascontiguousarray(a, np.float32).view(np.uint32)
    np.testing.assert_array_equal(bits(attrs["depth"][vis]), bits(want["depth"][vis]))
    np.testing.assert_array_equal(bits(attrs["xy"][vis]), bits(want["xy"][vis]))
    np.testing.assert_array_equal(bits(attrs["conic_opacity"][vis]), bits(want["conic_opacity"][vis]))
    np.testing.assert_array_equal(bits(attrs["rgb"][vis]), bits(want["rgb"][vis]))
    assert lists["D"] == want["D"]
    np.testing.assert_array_equal(lists["keys"].view(np.uint64), want["keys"])
    np.testing.assert_array_equal(lists["vals"].view(np.uint32), want["vals"])
    np.testing.assert_array_equal(lists["ranges"], want["ranges"])


CASES = [
    dict(n=4000, H=64, W=64, seed=1),
    dict(n=20000, H=128, W=96, seed=2),            # ragged: W not a multiple of 16
    dict(n=3000, H=50, W=70, seed=3, bg=(0.2, 0.5, 0.9)),
    dict(n=6000, H=64, W=64, seed=4, antialiasing=True),
    dict(n=6000, H=64, W=64, seed=5, use_cov=True),
    dict(n=6000, H=64, W=64, seed=6, use_colors=True),
    dict(n=6000, H=64, W=64, seed=7, scale_modifier=1.7),
    dict(n=6000, H=64, W=64, seed=8, near=0.2),
    dict(n=5000, H=64, W=64, seed=9, sh_degree=0),
    dict(n=5000, H=64, W=64, seed=10, sh_degree=1),
    dict(n=5000, H=64, W=64, seed=11, sh_degree=2),
    dict(n=40000, H=256, W=256, seed=12, log_scale_mean=-4.5),
    dict(n=2000, H=64, W=64, seed=13, log_scale_mean=-0.5),  # huge splats: every tile, long lists
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items()))
def test_forward_matches_oracle(eng, case):
    case = dict(case)
    n, H, W, seed = case.pop("n"), case.pop("H"), case.pop("W"), case.pop("seed")
    lsm = case.pop("log_scale_mean", np.log(0.05))
    g = random_gaussians(n, seed=seed, log_scale_mean=lsm)
    cam = look_at_camera(H, W, eye=(0.3, -0.2, -3.0))
    want, color, radii, invd, attrs, lists = _run_both(eng, g, cam, H, W, **case)
    _assert_exact(want, radii, attrs, lists)
    frag = want["fragile"].astype(bool)
    assert_image_close(color, want["color"], frag, what="color")
    assert_image_close(invd[0], want["invdepth"][0], frag, what="invdepth")
    ok = ~frag
    assert psnr(color[:, ok], want["color"][:, ok]) >= 50.0


def test_camera_inside_cloud_near_plane_cull(eng):
    g = random_gaussians(8000, seed=21, spread=2.0)
    cam = look_at_camera(64, 64, eye=(0.0, 0.0, 0.0), target=(0.0, 0.0, 1.0))
    want, color, radii, invd, attrs, lists = _run_both(eng, g, cam, 64, 64)
    assert (want["radii"] == 0).sum() > 1000  # many behind the camera / inside the near plane
    _assert_exact(want, radii, attrs, lists)


def test_empty_and_all_culled(eng):
    dev = eng.device
    cam = look_at_camera(32, 48)
    z = lambda *s: torch.zeros(*s, device=dev)
    color, radii, invd = eng.rasterize_forward(z(0, 3), z(0), cam["view"].to(dev), cam["proj"].to(dev),
                                               cam["campos"].to(dev), cam["tanfovx"], cam["tanfovy"], 32, 48,
                                               torch.tensor([0.3, 0.3, 0.3], device=dev), shs=z(0, 16, 3),
                                               scales=z(0, 3), rotations=z(0, 4))
    assert color.shape == (3, 32, 48) and float(color.abs().max()) == 0.0  # P == 0: untouched zeros, like upstream
    assert radii.numel() == 0
    g = random_gaussians(500, seed=3)
    g["xyz"][:, 2] -= 50.0  # everything behind the camera
    want, color, radii, invd, attrs, lists = _run_both(eng, g, cam, 32, 48, bg=(0.3, 0.6, 0.1))
    assert want["D"] == 0 and lists["D"] == 0 and int(radii.max()) == 0
    np.testing.assert_allclose(color, want["color"], atol=0)  # pure background


def test_argument_validation_matches_upstream_messages(eng):
    from gsworld_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    dev = eng.device
    cam = look_at_camera(32, 32)
    g = random_gaussians(10, seed=1, device=dev)
    rs = GaussianRasterizationSettings(32, 32, cam["tanfovx"], cam["tanfovy"], torch.zeros(3, device=dev), 1.0,
                                       cam["view"].to(dev), cam["proj"].to(dev), 3, cam["campos"].to(dev), False,
                                       False, False)
    r = GaussianRasterizer(rs)
    with pytest.raises(Exception, match="excatly one of either SHs"):
        r(g["xyz"], None, g["opacities"], shs=None, colors_precomp=None, scales=g["scales"], rotations=g["rotations"])
    with pytest.raises(Exception, match="exactly one of either scale/rotation"):
        r(g["xyz"], None, g["opacities"], shs=g["shs"], scales=g["scales"], rotations=None)
    color, radii, invd = r(g["xyz"], None, g["opacities"], shs=g["shs"], scales=g["scales"], rotations=g["rotations"])
    assert color.shape == (3, 32, 32) and radii.shape == (10,) and invd.shape == (1, 32, 32)
    vis = r.markVisible(g["xyz"])
    np.testing.assert_array_equal(vis.cpu().numpy(), orc.mark_visible(g["xyz"].cpu().numpy(), cam["view"].numpy()))
