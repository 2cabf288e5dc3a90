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
