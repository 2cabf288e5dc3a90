"""gsw_compose_poses / gsw_compose_cameras (one launch each) against the reference's own torch / numpy
statements as mirrored in GSWorldWrapper(fused_host=False) and in oracle/torch_transform.py.
Float-tolerance stage: the reference goes through LAPACK inverses and an SVD; bar = 2e-5 absolute."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _wrap(robot, cfg, objects, fused, strict, B=5, seed=4):
    from gsworld_b200 import fixtures as fx
    from gsworld_b200.gaussian_renderer import PipelineParams
    from gsworld_b200.gs_world_wrapper import GSWorldWrapper
    dev = torch.device("cuda:0")
    scene = fx.make_scene(fx.SceneSpec(robot=robot, n_total=6000, per_link=100, objects=objects, per_object=300,
                                       seed=seed))
    env = fx.FakeEnv(robot, num_envs=B, hw=(48, 64), objects=objects, seed=seed, device=dev)
    scan = fx.forward_kinematics(robot, fx.rt.ROBOT_SCAN_QPOS[robot][None])[0]
    w = GSWorldWrapper(env, PipelineParams(), cfg, device=dev, merged_model=scene.to(dev), scan_link_mats=list(scan),
                       strict_reference=strict, fused_host=fused)
    return scene, env, w, scan


@pytest.mark.parametrize("robot,cfg,objects", [
    ("fr3_umi", "fr3_pnp_box", ("006_mustard_bottle", "snack_box")),
    ("xarm6_uf_gripper", "xarm6_align", ("005_tomato_soup_can", "dtc_green_can", "bread_slice")),
])
def test_pose_table_matches_reference_algebra(robot, cfg, objects):
    from gsworld_b200 import rig_tables as rt
    from oracle import torch_transform as ot
    scene, env, w_fused, scan = _wrap(robot, cfg, objects, True, True)
    _, _, w_torch, _ = _wrap(robot, cfg, objects, False, True)
    w_torch.env = w_fused.env  # same env state
    pf = w_fused.transform_gs_perlink(None).clone().cpu().numpy()
    pt = w_torch.transform_gs_perlink(None).clone().cpu().numpy()
    np.testing.assert_allclose(pf[..., :12], pt[..., :12], atol=2e-5, rtol=0)   # R, t
    np.testing.assert_allclose(pf[..., 16:18], pt[..., 16:18], atol=2e-5, rtol=1e-5)  # kappa, flag
    # quaternions: same rotation (the largest-component branch can differ only at exact ties)
    dots = np.abs((pf[..., 12:16] * pt[..., 12:16]).sum(-1))
    assert dots.min() > 1 - 1e-5
    # and against the CPU oracle's statement-by-statement restatement of the reference
    fam = rt.robot_family(cfg)
    ref = ot.RefPath(env, scene, rt.SIM2GS_ROBOT[fam], rt.LINK_LABELS[fam], list(scan), rt.SIM2GS_OBJECT,
                     rt.OBJECT_LABELS, rt.OBJECT_OFFSET, rt.OBJECT_SCALE,
                     xarm_offset=rt.OBJECT_OFFSET["xarm_arm"] if fam == "xarm" else None)
    lt = ref.link_transforms().numpy()
    L = lt.shape[1]
    np.testing.assert_allclose(pf[:, :L, :9].reshape(-1, L, 3, 3), lt[:, :, :3, :3], atol=2e-5)
    np.testing.assert_allclose(pf[:, :L, 9:12], lt[:, :, :3, 3], atol=2e-5)
    for k, (name, (rigid, kappa)) in enumerate(ref.object_transforms().items()):
        np.testing.assert_allclose(pf[:, L + k, :9].reshape(-1, 3, 3), rigid[:, :3, :3].numpy(), atol=2e-5)
        np.testing.assert_allclose(pf[:, L + k, 9:12], rigid[:, :3, 3].numpy(), atol=2e-5)
        np.testing.assert_allclose(pf[:, L + k, 16], kappa.numpy(), rtol=1e-5)


@pytest.mark.parametrize("strict", [True, False])
def test_camera_matrices_match_reference_path(strict):
    objects = ("006_mustard_bottle",)
    scene, env, w_fused, scan = _wrap("fr3_umi", "fr3_pnp_box", objects, True, strict)
    _, _, w_ref, _ = _wrap("fr3_umi", "fr3_pnp_box", objects, False, strict)
    w_ref.env = w_fused.env
    names = list(env.cameras)
    vf = [t.cpu().numpy() for t in w_fused._views(names)]
    vr = [t.cpu().numpy() for t in w_ref._views(names)]
    np.testing.assert_array_equal(vf[4], vr[4])  # frame -> view map
    for a, b, what in zip(vf[:4], vr[:4], ("view", "proj", "campos", "tanfov")):
        np.testing.assert_allclose(a.reshape(b.shape), b, atol=2e-5, rtol=2e-6, err_msg=what)
    if strict:  # and the oracle's numpy restatement of cam_maniskill2gs + Camera
        from gsworld_b200 import rig_tables as rt
        from oracle import torch_transform as ot
        ref = ot.RefPath(env, scene, rt.SIM2GS_ROBOT["fr3"], rt.LINK_LABELS["fr3"], list(scan), rt.SIM2GS_OBJECT,
                         rt.OBJECT_LABELS, rt.OBJECT_OFFSET, rt.OBJECT_SCALE)
        cams = ref.cameras(strict=True)
        for i, c in enumerate(names):
            np.testing.assert_allclose(vf[0][i].reshape(4, 4), cams[c][0]["view"].numpy(), atol=2e-5)
            np.testing.assert_allclose(vf[1][i].reshape(4, 4), cams[c][0]["proj"].numpy(), atol=2e-5, rtol=2e-6)
            np.testing.assert_allclose(vf[2][i], cams[c][0]["campos"].numpy(), atol=2e-5)
            np.testing.assert_allclose(vf[3][i], [cams[c][0]["tanfovx"], cams[c][0]["tanfovy"]], rtol=2e-6)
