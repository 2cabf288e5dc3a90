"""Pins the CPU oracle against golden vectors produced by the reference's own functions
(tests/golden/make_golden.py documents which reference code generated each file)."""
import os

import numpy as np
import torch

from oracle import oracle as orc
from oracle import torch_transform as ot


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name)))


def _model(g):
    import types
    return types.SimpleNamespace(_xyz=torch.tensor(g["model_xyz"]), _scaling=torch.tensor(g["model_scaling"]),
                                 _rotation=torch.tensor(g["model_rotation"]), _opacity=torch.tensor(g["model_opacity"]))


def test_torch_restatement_matches_reference_transform_gaussians(golden_dir):
    g = _load(golden_dir, "transform_gaussians.npz")
    m, idx = _model(g), torch.tensor(g["idx"])
    for tag in ("link_b1", "link_b3", "obj_b1", "obj_b3"):
        R, t = torch.tensor(g[f"{tag}__R"]), torch.tensor(g[f"{tag}__t"])
        sc = torch.tensor(g[f"{tag}__scale"]) if tag.startswith("obj") else None
        xyz, scaling, rot, op = ot.transform_gaussians(m, idx, sc, R, t)
        # same torch ops in the same order on the same machine: exact
        assert torch.equal(xyz, torch.tensor(g[f"{tag}__xyz"])), tag
        assert torch.equal(scaling, torch.tensor(g[f"{tag}__scaling"])), tag
        assert torch.equal(rot, torch.tensor(g[f"{tag}__rotation"])), tag
        assert torch.equal(op, torch.tensor(g[f"{tag}__opacity"])), tag


def test_c_oracle_transform_matches_reference(golden_dir):
    """Scalar C restatement vs the reference's batched torch result.  fp32 summation order / libm
    differ from torch's vectorised kernels, so the bar is a few ulp (tolerance written here)."""
    g = _load(golden_dir, "transform_gaussians.npz")
    idx = g["idx"]
    xyz, ls, q = g["model_xyz"][idx], g["model_scaling"][idx], g["model_rotation"][idx]
    for tag in ("link_b1", "link_b3", "obj_b1", "obj_b3"):
        R, t = g[f"{tag}__R"], g[f"{tag}__t"]
        qR = ot.matrix_to_quaternion(torch.tensor(R)).numpy()
        for b in range(R.shape[0]):
            kappa = float(g[f"{tag}__scale"][b]) if tag.startswith("obj") else None
            o = orc.transform_gaussians(xyz, ls, q, R[b], t[b], qR[b], kappa)
            np.testing.assert_allclose(o["xyz"], g[f"{tag}__xyz"][b], rtol=2e-6, atol=2e-6)
            np.testing.assert_allclose(o["quat"], g[f"{tag}__rotation"][b], rtol=2e-6, atol=2e-6)
            want_ls = g[f"{tag}__scaling"][b] if tag.startswith("obj") else g[f"{tag}__scaling"]
            np.testing.assert_allclose(o["log_scale"], want_ls, rtol=2e-6, atol=2e-6)
            np.testing.assert_allclose(o["scale_act"], np.exp(want_ls), rtol=3e-6)
            want_q = torch.nn.functional.normalize(torch.tensor(g[f"{tag}__rotation"][b])).numpy()
            np.testing.assert_allclose(o["quat_act"], want_q, rtol=2e-6, atol=2e-6)


def test_extract_rigid_transform(golden_dir):
    g = _load(golden_dir, "extract_rigid.npz")
    Mr, sc, R, t = ot.extract_rigid_transform(torch.tensor(g["M"]))
    assert torch.equal(Mr, torch.tensor(g["M_rigid"]))
    assert torch.equal(sc, torch.tensor(g["scales"]))
    assert torch.equal(R, torch.tensor(g["R"])) and torch.equal(t, torch.tensor(g["t"]))
    M1, s1, R1, t1 = ot.extract_rigid_transform(torch.tensor(g["M"][0]))
    assert torch.equal(M1, torch.tensor(g["single_M_rigid"])) and torch.equal(s1, torch.tensor(g["single_scale"]))


def test_sh_colour_matches_reference_eval_sh(golden_dir):
    g = _load(golden_dir, "sh_cov.npz")
    for deg in range(4):
        rgb = orc.sh_to_rgb(g["xyz"], g["campos"], g["features"], deg)
        np.testing.assert_allclose(rgb, g[f"rgb_deg{deg}"], rtol=1e-5, atol=2e-6)


def test_cov3d_matches_reference_build_scaling_rotation(golden_dir):
    g = _load(golden_dir, "sh_cov.npz")
    # the rasterizer receives get_rotation = normalize(q); build_rotation normalises internally
    qn = torch.nn.functional.normalize(torch.tensor(g["rots"])).numpy()
    for mod, key in ((1.0, "cov3d_mod1"), (1.7, "cov3d_mod1p7")):
        cov = orc.cov3d(g["scales"], qn, mod)
        np.testing.assert_allclose(cov, g[key], rtol=2e-5, atol=1e-9)


def test_camera_tensors_match_reference_graphics_utils(golden_dir):
    g = _load(golden_dir, "camera.npz")
    cam = ot.camera_tensors(g["R"], g["T"], float(g["fovx"]), float(g["fovy"]), 48, 64)
    np.testing.assert_array_equal(cam["view"].numpy(), g["w2v"].T)
    P = g["proj"]
    full = torch.tensor(g["w2v"]).T.unsqueeze(0).bmm(torch.tensor(P).T.unsqueeze(0)).squeeze(0)
    np.testing.assert_array_equal(cam["proj"].numpy(), full.numpy())
