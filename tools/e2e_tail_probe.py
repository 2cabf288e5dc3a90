"""Where the e2e tail comes from: the synchronous step with (a) no delivery, (b) delivery to a DEVICE buffer (same
signalling and joins, copies at HBM speed), (c) delivery to pinned host memory."""
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
devbuf = torch.empty((2, 64, 256, 256, 3), dtype=torch.uint8, device=dev)
out_buf = torch.empty((2, 64, 256, 256, 3), dtype=torch.uint8, device=dev)
def step():
    env.advance(); wrap.transform_gs_perlink(None); return wrap._render_gsworld()
for mode in ("none", "device", "host", "host_after"):
    wrap.obs_out = out_buf
    wrap.obs_host = pinned if mode == "host" else None
    wrap.obs_peers = [devbuf] if mode == "device" else None
    for _ in range(10): step()
    best = 1e9
    for rep in range(3):
        torch.cuda.synchronize(); t = time.perf_counter()
        for _ in range(50):
            step()
            if mode == "host_after": pinned.copy_(out_buf, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        best = min(best, 1e3 * (time.perf_counter() - t) / 50)
    print(f"{mode}: {best:.3f} ms/step")
