import sys, time, cProfile, pstats, io
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
from gsworld_b200 import fixtures as fx, rig_tables as rt
from gsworld_b200.gaussian_renderer import PipelineParams
from gsworld_b200.gs_world_wrapper import GSWorldWrapper
dev = torch.device("cuda:0")
scene = bench.make_scene().to(dev)
base = fx.FakeEnv(bench.ROBOT, num_envs=64, cameras=bench.CAMERAS, hw=(256,256), objects=bench.OBJECTS, seed=1)
scan = fx.forward_kinematics(bench.ROBOT, rt.ROBOT_SCAN_QPOS[bench.ROBOT][None])[0]
RESIDENT = 'host' not in sys.argv  # `python tools/prof_host.py host` = pinned host inputs (the e2e leg)
env = bench.ReplayEnv(base, 8, dev, RESIDENT)
wrap = GSWorldWrapper(env, PipelineParams(), "fr3_pnp_box", device=dev, merged_model=scene, scan_link_mats=list(scan))
def step():
    env.advance(); wrap.transform_gs_perlink(None); return wrap._render_gsworld()
for _ in range(5): step()
wrap.sync_each_step = False
torch.cuda.synchronize()
t=time.perf_counter()
for _ in range(100): step()
t1=time.perf_counter(); torch.cuda.synchronize(); t2=time.perf_counter()
print(f"host issue {1e3*(t1-t)/100:.3f} ms/step, total {1e3*(t2-t)/100:.3f} ms/step")
hs = []
for _ in range(30):  # queue empty at every step: pure host cost of issuing one step
    torch.cuda.synchronize(); a = time.perf_counter(); step(); hs.append(time.perf_counter() - a)
hs.sort(); print(f"host cost of issuing one step with an empty queue: median {1e3*hs[len(hs)//2]:.3f} ms, min {1e3*hs[0]:.3f} ms")
ta, tb, tc = [], [], []
for _ in range(30):  # where the host time of a synchronous step goes
    torch.cuda.synchronize(); a = time.perf_counter(); env.advance(); wrap.transform_gs_perlink(None)
    b = time.perf_counter(); wrap._render_gsworld(); c = time.perf_counter(); torch.cuda.synchronize(); d = time.perf_counter()
    ta.append(b - a); tb.append(c - b); tc.append(d - c)
med = lambda v: 1e3 * sorted(v)[len(v) // 2]
print(f"sync step: transform_gs_perlink {med(ta):.3f} ms, _render_gsworld issue {med(tb):.3f} ms, wait {med(tc):.3f} ms")
pr = cProfile.Profile(); pr.enable()
for _ in range(100): step()
pr.disable(); torch.cuda.synchronize()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(22); print(s.getvalue()[:4500])
