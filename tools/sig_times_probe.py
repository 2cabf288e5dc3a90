import os, sys
sys.path.insert(0, "/root/repo" if os.path.exists("/root/repo/bench.py") else os.getcwd())
import torch, bench
from gsworld_b200 import fixtures as fx, rig_tables as rt
from gsworld_b200.gaussian_renderer import PipelineParams
from gsworld_b200.gs_world_wrapper import GSWorldWrapper
dev = torch.device("cuda:0")
scene = bench.make_scene().to(dev)
base = fx.FakeEnv(bench.ROBOT, num_envs=64, cameras=bench.CAMERAS, hw=(256, 256), objects=bench.OBJECTS, seed=1)
scan = fx.forward_kinematics(bench.ROBOT, rt.ROBOT_SCAN_QPOS[bench.ROBOT][None])[0]
env = bench.ReplayEnv(base, 8, dev, False)
wrap = GSWorldWrapper(env, PipelineParams(), "fr3_pnp_box", device=dev, merged_model=scene, scan_link_mats=list(scan))
pinned = torch.empty((2, 64, 256, 256, 3), dtype=torch.uint8).pin_memory()
wrap.obs_host = pinned
for i in range(6):
    env.advance(); wrap.transform_gs_perlink(None); wrap._render_gsworld(); torch.cuda.synchronize()
