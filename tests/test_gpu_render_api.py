"""gaussian_renderer.render() — the reference's render wrapper (gaussian_renderer/__init__.py:18-128) on the
B200 rasterizer: returned dict, the three `pipe` routes and override_color.

The reference offers two in-tree cross-checks through the SAME rasterizer: with
pipe.compute_cov3D_python the 3-D covariance is computed in torch and passed as cov3D_precomp, with
pipe.convert_SHs_python the SH colour is evaluated in torch and passed as colors_precomp.  Torch's matmul / SH
polynomial round differently from the kernel's glm-ordered products, so those routes agree within the colour
tolerance (1e-3) and on all but a handful of radii; fed with the CPU oracle's own no-FMA cov3D / SH colour
(same expression order as the kernel) the precomputed routes must reproduce the fused route BIT FOR BIT.
"""
import numpy as np
import pytest
import torch

from tests.helpers import look_at_camera, random_gaussians

pytestmark = pytest.mark.gpu


class _Cam:
    def __init__(self, c, dev):
        self.FoVx, self.FoVy = c["FoVx"], c["FoVy"]
        self.image_height, self.image_width = c["H"], c["W"]
        self.world_view_transform = c["view"].to(dev)
        self.full_proj_transform = c["proj"].to(dev)
        self.camera_center = c["campos"].to(dev)


def _scene(n, seed, dev):
    from gsworld_b200.fixtures import GaussianScene
    g = random_gaussians(n, seed=seed)
    sh = g["shs"]
    op = torch.logit(g["opacities"].clamp(1e-4, 1 - 1e-4)).reshape(n, 1, 1)
    return GaussianScene(g["xyz"], sh[:, :1].contiguous(), sh[:, 1:].contiguous(), op, g["log_scale"], g["quat_raw"],
                         torch.full((n, 1), -1.0)).to(dev)


@pytest.mark.parametrize("antialiasing", [False, True])
def test_render_pipe_routes_agree(antialiasing):
    from gsworld_b200.gaussian_renderer import PipelineParams, render
    dev = torch.device("cuda:0")
    H, W = 96, 128
    pc = _scene(8000, 3, dev)
    cam = _Cam(look_at_camera(H, W, eye=(0.2, -0.1, -3.0)), dev)
    bg = torch.tensor([0.2, 0.1, 0.0], device=dev)
    base = render(cam, pc, PipelineParams(antialiasing=antialiasing), bg)
    assert set(base) == {"render", "viewspace_points", "visibility_filter", "radii", "depth"}
    assert base["render"].shape == (3, H, W) and base["depth"].shape == (1, H, W) and base["radii"].shape == (8000,)
    assert base["radii"].dtype == torch.int32 and float(base["render"].min()) >= 0 and float(base["render"].max()) <= 1
    assert torch.equal(base["visibility_filter"], (base["radii"] > 0).nonzero())
    assert base["viewspace_points"].shape == (8000, 3)
    vis = int((base["radii"] > 0).sum())
    assert vis > 1000
    for kw in (dict(compute_cov3D_python=True), dict(convert_SHs_python=True),
               dict(compute_cov3D_python=True, convert_SHs_python=True)):
        other = render(cam, pc, PipelineParams(antialiasing=antialiasing, **kw), bg)
        flips = int((other["radii"] != base["radii"]).sum())
        assert flips <= max(2, vis // 2000), f"{kw}: {flips} radii differ"
        if flips == 0:
            d = (other["render"] - base["render"]).abs().max().item()
            assert d <= 1e-3, f"{kw}: colour differs by {d}"
            assert (other["depth"] - base["depth"]).abs().max().item() <= 1e-3
    red = torch.zeros(8000, 3, device=dev)
    red[:, 0] = 1.0
    over = render(cam, pc, PipelineParams(antialiasing=antialiasing), bg, override_color=red)
    assert torch.equal(over["radii"], base["radii"])
    covered = base["depth"][0] > 0.05
    assert float(over["render"][1][covered].max()) <= float(bg[1]) + 1e-6  # only the background has green
    with pytest.raises(NotImplementedError):
        render(cam, pc, PipelineParams(), bg, separate_sh=True)
    sm = render(cam, pc, PipelineParams(antialiasing=antialiasing), bg, scaling_modifier=1.5)
    assert int((sm["radii"] > base["radii"]).sum()) > vis // 2


def test_precomputed_routes_with_oracle_arithmetic_are_bit_identical():
    """cov3D_precomp / colors_precomp computed with the oracle's (= the kernel's) expression order."""
    from gsworld_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from oracle import oracle as orc
    dev = torch.device("cuda:0")
    H, W = 80, 112
    g = random_gaussians(6000, seed=9)
    c = look_at_camera(H, W, eye=(-0.3, 0.2, -2.6))
    bg = torch.zeros(3, device=dev)
    rs = GaussianRasterizationSettings(H, W, c["tanfovx"], c["tanfovy"], bg, 1.0, c["view"].to(dev), c["proj"].to(dev),
                                       3, c["campos"].to(dev), False, False, False)
    ras = GaussianRasterizer(rs)
    t = {k: v.to(dev) for k, v in g.items()}
    zeros = torch.zeros_like(t["xyz"])
    fused = ras(t["xyz"], zeros, t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    cov = torch.from_numpy(orc.cov3d(g["scales"].numpy(), g["rotations"].numpy(), 1.0)).to(dev)
    rgb = torch.from_numpy(orc.sh_to_rgb(g["xyz"].numpy(), c["campos"].numpy(), g["shs"].numpy(), 3)).to(dev)
    for kw in (dict(shs=t["shs"], cov3D_precomp=cov), dict(colors_precomp=rgb, scales=t["scales"], rotations=t["rotations"]),
               dict(colors_precomp=rgb, cov3D_precomp=cov)):
        other = ras(t["xyz"], zeros, t["opacities"], **kw)
        assert torch.equal(other[1], fused[1]), f"{sorted(kw)}: radii"
        assert torch.equal(other[0].view(torch.int32), fused[0].view(torch.int32)), f"{sorted(kw)}: colour bits"
        assert torch.equal(other[2].view(torch.int32), fused[2].view(torch.int32)), f"{sorted(kw)}: depth bits"
    vis = ras.markVisible(t["xyz"])
    assert vis.dtype == torch.bool and torch.equal(vis.cpu(), torch.from_numpy(orc.mark_visible(g["xyz"].numpy(), c["view"].numpy())))
