"""e2e leg (host inputs, pinned host observations, sync every step) for one delivery grouping (GSW_SIG_GROUPS)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
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
def step():
    env.advance(); wrap.transform_gs_perlink(None); return wrap._render_gsworld()
for _ in range(10): step()
best = 1e9
for rep in range(3):
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(50):
        step(); torch.cuda.current_stream().synchronize()
    best = min(best, 1e3 * (time.perf_counter() - t) / 50)
print(f"GSW_SIG_GROUPS={os.environ.get('GSW_SIG_GROUPS')}: {best:.3f} ms/step")
