"""gsw_rasterize_forward_static — the entry point for callers that may not allocate at render time
(real-time loops, plugins on a fixed GPU memory budget, hosts that own the allocator) — must give the
same frame as gsw_rasterize_forward.  Both are called through the C ABI on the same inputs, twice in a
row with different scenes, so that anything the shared fixed scratch array keeps from the first call
would show up in the second.  Tolerances are those of the other parity tests: radii exact, colour and
inverse depth max-abs 1e-3 (the two paths run the same kernels, so they normally agree bit for bit).
Collected last on purpose (file name), and it must also pass when run alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from tests.helpers import look_at_camera, random_gaussians

pytestmark = pytest.mark.gpu


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _static_forward(lib, g, cam, H, W, bg, dev, stream):
    P = g["xyz"].shape[0]
    color = torch.full((3, H, W), -7.0, device=dev)       # poisoned outputs: every pixel must be written
    radii = torch.full((P,), -7, dtype=torch.int32, device=dev)
    invd = torch.full((1, H, W), -7.0, device=dev)
    status = torch.full((4,), -7, dtype=torch.int64, device=dev)
    t = {k: v.to(dev).contiguous() for k, v in g.items()}
    view, proj, campos = cam["view"].to(dev).contiguous(), cam["proj"].to(dev).contiguous(), cam["campos"].to(dev)
    rc = lib.gsw_rasterize_forward_static(
        P, 3, 16, _ptr(bg), W, H, _ptr(t["xyz"]), _ptr(t["shs"]), None, _ptr(t["opacities"]), _ptr(t["scales"]), 1.0,
        _ptr(t["rotations"]), None, _ptr(view), _ptr(proj), _ptr(campos), cam["tanfovx"], cam["tanfovy"], 0, 0.05,
        _ptr(color), _ptr(radii), _ptr(invd), _ptr(status), stream)
    assert rc == 0, lib.gsw_last_error()
    torch.cuda.synchronize()
    return color, radii, invd, status


def test_static_scratch_forward_matches_allocating_forward_twice():
    from gsworld_b200 import _lib
    from gsworld_b200.engine import Engine
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    lib = _lib.load()
    eng = Engine(dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    # two different frames back to back: the second is smaller and sparser than the first, so stale bins,
    # keys or records from call 1 would change call 2's image if anything were left uncleared
    for n, H, W, seed, eye in ((20000, 128, 96, 31, (0.3, -0.2, -3.0)), (3000, 64, 80, 32, (-0.5, 0.4, -2.5))):
        g = random_gaussians(n, seed=seed)
        cam = look_at_camera(H, W, eye=eye)
        want_c, want_r, want_d = eng.rasterize_forward(
            g["xyz"].to(dev), g["opacities"].to(dev), cam["view"].to(dev), cam["proj"].to(dev), cam["campos"].to(dev),
            cam["tanfovx"], cam["tanfovy"], H, W, bg, shs=g["shs"].to(dev), scales=g["scales"].to(dev),
            rotations=g["rotations"].to(dev))
        torch.cuda.synchronize()
        got_c, got_r, got_d, status = _static_forward(lib, g, cam, H, W, bg, dev, stream)
        assert int(status[2]) == 0 and int(status[0]) == eng.last_stats["D"]
        np.testing.assert_array_equal(got_r.cpu().numpy(), want_r.cpu().numpy())
        np.testing.assert_allclose(got_c.cpu().numpy(), want_c.cpu().numpy(), atol=1e-3, rtol=0)
        np.testing.assert_allclose(got_d.cpu().numpy(), want_d.cpu().numpy(), atol=1e-3, rtol=0)
    eng.close()


def test_static_scratch_forward_rejects_sizes_beyond_its_fixed_array():
    from gsworld_b200 import _lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    lib = _lib.load()
    z = torch.zeros(16, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    args = lambda P, W, H: (P, 3, 16, _ptr(z), W, H, _ptr(z), _ptr(z), None, _ptr(z), _ptr(z), 1.0, _ptr(z), None,
                            _ptr(z), _ptr(z), _ptr(z), 1.0, 1.0, 0, 0.05, None, None, None, None, stream)
    assert lib.gsw_rasterize_forward_static(*args(32769, 64, 64)) == -1     # too many Gaussians
    assert b"exceeds the fixed" in lib.gsw_last_error()
    assert lib.gsw_rasterize_forward_static(*args(100, 1024, 512)) == -1    # 2048 tiles > 1024
    assert b"exceeds the fixed" in lib.gsw_last_error()


def test_static_scratch_forward_matches_the_cpu_oracle():
    """Direct check against the CPU oracle (not only against the allocating entry point)."""
    from gsworld_b200 import _lib
    from oracle import oracle as orc
    from tests.helpers import assert_image_close
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    lib = _lib.load()
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    bg = torch.tensor([0.0, 0.1, 0.2], device=dev)
    H, W = 96, 112
    g = random_gaussians(6000, seed=41)
    cam = look_at_camera(H, W, eye=(0.2, 0.1, -2.8))
    got_c, got_r, got_d, status = _static_forward(lib, g, cam, H, W, bg, dev, stream)
    want = orc.rasterize_forward(g["xyz"].numpy(), g["opacities"].numpy(), cam["view"].numpy(), cam["proj"].numpy(),
                                 cam["campos"].numpy(), cam["tanfovx"], cam["tanfovy"], H, W, shs=g["shs"].numpy(),
                                 scales=g["scales"].numpy(), rotations=g["rotations"].numpy(), bg=(0.0, 0.1, 0.2),
                                 debug=True)
    np.testing.assert_array_equal(got_r.cpu().numpy(), want["radii"])
    assert int(status[0]) == want["D"] and int(status[2]) == 0
    frag = want["fragile"].astype(bool)
    assert_image_close(got_c.cpu().numpy(), want["color"], frag, what="color")
    assert_image_close(got_d[0].cpu().numpy(), want["invdepth"][0], frag, what="invdepth")


def test_static_scratch_forward_allocates_nothing_and_handles_empty_input():
    """Free device memory is the same before and after a call (outputs are allocated by the test beforehand),
    and P == 0 leaves zeros like the allocating entry point."""
    from gsworld_b200 import _lib
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    lib = _lib.load()
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    bg = torch.tensor([0.3, 0.3, 0.3], device=dev)
    H, W = 64, 64
    g = random_gaussians(4000, seed=43)
    cam = look_at_camera(H, W)
    _static_forward(lib, g, cam, H, W, bg, dev, stream)  # first call on the device: looks the array's address up
    t = {k: v.to(dev).contiguous() for k, v in g.items()}
    view, proj, campos = cam["view"].to(dev).contiguous(), cam["proj"].to(dev).contiguous(), cam["campos"].to(dev)
    color = torch.empty((3, H, W), device=dev)
    radii = torch.empty((4000,), dtype=torch.int32, device=dev)
    invd = torch.empty((1, H, W), device=dev)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    for _ in range(3):
        rc = lib.gsw_rasterize_forward_static(
            4000, 3, 16, _ptr(bg), W, H, _ptr(t["xyz"]), _ptr(t["shs"]), None, _ptr(t["opacities"]), _ptr(t["scales"]),
            1.0, _ptr(t["rotations"]), None, _ptr(view), _ptr(proj), _ptr(campos), cam["tanfovx"], cam["tanfovy"], 0,
            0.05, _ptr(color), _ptr(radii), _ptr(invd), None, stream)
        assert rc == 0, lib.gsw_last_error()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, f"device memory changed across the calls: {free0} -> {free1}"
    # P == 0
    color.fill_(-7.0)
    invd.fill_(-7.0)
    status = torch.full((4,), -7, dtype=torch.int64, device=dev)
    rc = lib.gsw_rasterize_forward_static(0, 3, 16, _ptr(bg), W, H, None, None, None, None, None, 1.0, None, None,
                                          _ptr(view), _ptr(proj), _ptr(campos), cam["tanfovx"], cam["tanfovy"], 0, 0.05,
                                          _ptr(color), None, _ptr(invd), _ptr(status), stream)
    assert rc == 0, lib.gsw_last_error()
    torch.cuda.synchronize()
    assert float(color.abs().max()) == 0.0 and float(invd.abs().max()) == 0.0 and int(status.abs().max()) == 0


def test_static_scratch_forward_reports_entry_overflow_and_recovers():
    """The entry count depends on the data: a frame that needs more (Gaussian, tile) entries than the fixed array
    holds comes back as plain background with out_status[2] set and out_status[0] = the count it needed; the next
    call with a frame that fits is unaffected."""
    from gsworld_b200 import _lib
    from gsworld_b200.engine import Engine
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    lib = _lib.load()
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    bg = torch.tensor([0.25, 0.5, 0.75], device=dev)
    H = W = 512  # 1024 tiles, the most the entry point accepts
    g = random_gaussians(1500, seed=47, spread=0.3, log_scale_mean=0.0, log_scale_std=0.1)  # metre-sized splats
    cam = look_at_camera(H, W, eye=(0.0, 0.0, -3.0))
    color, radii, invd, status = _static_forward(lib, g, cam, H, W, bg, dev, stream)
    assert int(status[2]) != 0 and int(status[0]) > 262144
    want_bg = bg.view(3, 1, 1).expand(3, H, W)
    assert torch.equal(color, want_bg.contiguous()) and float(invd.abs().max()) == 0.0
    assert int((radii > 0).sum()) > 0  # radii are still the projected ones
    # a frame that fits, right after the overflow
    g2 = random_gaussians(5000, seed=48)
    cam2 = look_at_camera(96, 96)
    eng = Engine(dev)
    want_c, want_r, want_d = eng.rasterize_forward(
        g2["xyz"].to(dev), g2["opacities"].to(dev), cam2["view"].to(dev), cam2["proj"].to(dev), cam2["campos"].to(dev),
        cam2["tanfovx"], cam2["tanfovy"], 96, 96, bg, shs=g2["shs"].to(dev), scales=g2["scales"].to(dev),
        rotations=g2["rotations"].to(dev))
    got_c, got_r, got_d, st2 = _static_forward(lib, g2, cam2, 96, 96, bg, dev, stream)
    assert int(st2[2]) == 0
    np.testing.assert_array_equal(got_r.cpu().numpy(), want_r.cpu().numpy())
    np.testing.assert_allclose(got_c.cpu().numpy(), want_c.cpu().numpy(), atol=1e-3, rtol=0)
    eng.close()
