"""e2e leg only (host inputs, pinned host observations, sync every step), for experiments on the delivery path."""
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
def step():
    env.advance(); wrap.transform_gs_perlink(None); return wrap._render_gsworld()
for mode in ("none", "library", "torch"):
    wrap.obs_host = pinned if mode == "library" else None
    for _ in range(10): step()
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(50):
        out = step()
        if mode == "torch":
            for i, c in enumerate(bench.CAMERAS): pinned[i].copy_(out[c], non_blocking=True)
        torch.cuda.current_stream().synchronize()
    print(f"{mode}: {1e3 * (time.perf_counter() - t) / 50:.3f} ms/step")

# where the host time of a synchronous e2e step goes: time from step start to the entry of the C call that
# launches the heavy kernels, time inside that call (all launches), and the wait for completion
eng = wrap._engine
orig = eng.lib.gsw_render_frames
marks = {}
def timed_call(*a):
    marks["enter"] = time.perf_counter()
    r = orig(*a)
    marks["exit"] = time.perf_counter()
    return r
eng.lib.gsw_render_frames = timed_call
wrap.obs_host = pinned
pre, inside, wait, tr = [], [], [], []
for _ in range(40):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    env.advance(); wrap.transform_gs_perlink(None); t1 = time.perf_counter()
    wrap._render_gsworld(); torch.cuda.current_stream().synchronize(); t2 = time.perf_counter()
    tr.append(t1 - t0); pre.append(marks["enter"] - t0); inside.append(marks["exit"] - marks["enter"]); wait.append(t2 - marks["exit"])
med = lambda v: 1e3 * sorted(v)[len(v) // 2]
print(f"e2e step: transform_gs_perlink {med(tr):.3f} ms; start -> gsw_render_frames entry {med(pre):.3f} ms; inside the call "
      f"{med(inside):.3f} ms; wait after it {med(wait):.3f} ms")
