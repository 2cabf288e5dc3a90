#!/usr/bin/env python
"""BASELINE config 5: rasterizer scaling sweep on one GPU (Gaussians x resolution x envs), strict cameras.
Prints one JSON line per point: frames/s of this framework (inputs resident) and, where it finishes quickly,
of the restated reference-style rasterizer (baseline/, one call per frame).  Not a bench value contract."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from gsworld_b200 import fixtures as fx, rig_tables as rt
from gsworld_b200.gaussian_renderer import PipelineParams
from gsworld_b200.gs_world_wrapper import GSWorldWrapper

dev = torch.device("cuda:0")
points = [(250_000, 256, 64), (1_000_000, 128, 64), (1_000_000, 256, 1), (1_000_000, 256, 16), (1_000_000, 256, 64),
          (1_000_000, 256, 256), (1_000_000, 512, 64), (4_000_000, 256, 64), (4_000_000, 512, 16)]
if len(sys.argv) > 1:
    points = points[: int(sys.argv[1])]
scenes = {}
for N, res, B in points:
    if N not in scenes:
        scenes.clear()
        torch.cuda.empty_cache()
        scenes[N] = fx.make_scene(fx.SceneSpec(robot=bench.ROBOT, n_total=N, per_link=4000, objects=bench.OBJECTS,
                                               per_object=7000, seed=bench.SEED)).to(dev)
    scene = scenes[N]
    base = fx.FakeEnv(bench.ROBOT, num_envs=B, cameras=bench.CAMERAS, hw=(res, res), objects=bench.OBJECTS, seed=bench.SEED)
    scan = fx.forward_kinematics(bench.ROBOT, rt.ROBOT_SCAN_QPOS[bench.ROBOT][None])[0]
    env = bench.ReplayEnv(base, 4, dev, True)
    wrap = GSWorldWrapper(env, PipelineParams(), "fr3_pnp_box", device=dev, merged_model=scene, scan_link_mats=list(scan))

    def step():
        env.advance()
        wrap.transform_gs_perlink(None)
        return wrap._render_gsworld()

    for _ in range(6):
        step()
    wrap.sync_each_step = False
    torch.cuda.synchronize()
    steps = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    assert wrap._engine.finish(raise_on_overflow=False)
    ms = e0.elapsed_time(e1) / steps
    F = B * len(bench.CAMERAS)
    line = {"N": N, "res": res, "envs": B, "frames_per_step": F, "ms_per_step": round(ms, 3),
            "frames_per_s": round(F / ms * 1e3, 1), "D": wrap._engine.last_stats["D"]}
    try:  # restated reference-style rasterizer, one call per frame, rasterizer only
        from baseline import ref_gpu
        from oracle import torch_transform as ot
        ref = ot.RefPath(base, scene, rt.SIM2GS_ROBOT["fr3"], rt.LINK_LABELS["fr3"], list(scan), rt.SIM2GS_OBJECT,
                         rt.OBJECT_LABELS, rt.OBJECT_OFFSET, rt.OBJECT_SCALE, device=dev)
        cams = ref.cameras(True)
        sc, ro = torch.exp(scene._scaling), torch.nn.functional.normalize(scene._rotation)
        op = torch.sigmoid(scene._opacity).reshape(-1)
        shs = torch.cat((scene._features_dc, scene._features_rest), 1).contiguous()
        bg = torch.zeros(3, device=dev)
        per = [{k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in cams[c][0].items()} for c in bench.CAMERAS]
        n = 0
        for c in per:
            ref_gpu.rasterize(scene._xyz, shs, op, sc, ro, c["view"], c["proj"], c["campos"], c["tanfovx"], c["tanfovy"], res, res, bg)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(4):
            for c in per:
                ref_gpu.rasterize(scene._xyz, shs, op, sc, ro, c["view"], c["proj"], c["campos"], c["tanfovx"], c["tanfovy"], res, res, bg)
                n += 1
        torch.cuda.synchronize()
        line["restated_reference_rasterizer_only_frames_per_s"] = round(n / (time.perf_counter() - t0), 1)
    except Exception as e:
        line["restated_reference_rasterizer_only_frames_per_s"] = repr(e)[:80]
    print(json.dumps(line), flush=True)
    del wrap, env
    torch.cuda.empty_cache()
