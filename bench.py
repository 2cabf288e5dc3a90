#!/usr/bin/env python
"""bench.py — rendered frames/s of the GSWorld observation path (transform + 3DGS rasterization).

Contract: `python bench.py --gpus N --steps K --warmup W [--impl reference]` prints ONE JSON line on rank 0.
Workload = BASELINE.json configs[2]: PnpBoxFr3Env-v1-shaped step, fr3_umi, 64 envs x 2 cameras
(wrist + third-person) x 256x256, synthetic 1M-Gaussian scene (SURVEY.md §8d generator), per GPU.
A "step" = compose the (env, link/object) pose table + render every (env, camera) frame to uint8.
Multi-GPU: envs are sharded by rank (cfg3: 64 per rank, weak scaling, plus a `strong` leg with 64 envs in total;
other --config values shard their stated env count), scene replicated, and every step delivers the uint8
observations to the policy rank (the only exchange on the path): copy-engine pushes into peer memory overlapped
with the blend (gsworld_b200/dist.py), verified once against an NCCL all-gather; NCCL all-gather as the fallback.

--impl reference: the reference's own path restated on the host cores (oracle/: torch transform with
deepcopy + isin write-back per frame, CPU rasterizer) on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np
import torch

# BASELINE.json configs: cfg3 is the one `metric` is quoted on and the default.  cfg1/2/4 are the other
# named configurations (same step, other robot / sizes); cfg5 is one point of the rasterizer sweep, set with
# --n / --res / --envs (tools/sweep.py runs the whole grid).  `envs` is the TOTAL env count BASELINE.json states.
CONFIGS = {
    "cfg1": dict(robot="xarm6_uf_gripper", scene_cfg="xarm6_align", family="xarm", envs=1, cams=("right_cam",),
                 hw=128, n=200_000, per_link=500, objects=("005_tomato_soup_can", "dtc_green_can"), seed=1235),
    "cfg2": dict(robot="xarm6_uf_gripper", scene_cfg="xarm6_align", family="xarm", envs=16, cams=("right_cam",),
                 hw=256, n=500_000, per_link=2000, objects=("005_tomato_soup_can", "dtc_green_can"), seed=1236),
    "cfg3": dict(robot="fr3_umi", scene_cfg="fr3_pnp_box", family="fr3", envs=64, cams=("right_cam", "wrist_cam"),
                 hw=256, n=1_000_000, per_link=4000, objects=("006_mustard_bottle", "snack_box"), seed=1237),
    "cfg4": dict(robot="fr3_umi", scene_cfg="fr3_pour_mustard", family="fr3", envs=256, cams=("right_cam", "wrist_cam"),
                 hw=128, n=1_000_000, per_link=4000, objects=("006_mustard_bottle", "bread_slice"), seed=1238),
    "cfg5": dict(robot="fr3_umi", scene_cfg="fr3_pnp_box", family="fr3", envs=64, cams=("right_cam", "wrist_cam"),
                 hw=256, n=1_000_000, per_link=4000, objects=("006_mustard_bottle", "snack_box"), seed=1237),
}
H = W = 256
ENVS_PER_GPU = 64
N_GAUSS = 1_000_000
ROBOT = "fr3_umi"
SCENE_CFG = "fr3_pnp_box"
FAMILY = "fr3"
PER_LINK = 4000
OBJECTS = ("006_mustard_bottle", "snack_box")
CAMERAS = ("right_cam", "wrist_cam")
SEED = 1234 + 3
CONFIG = "cfg3"
SHARD = "weak"  # cfg3 (the driver's line): 64 envs on EVERY GPU; the other configs shard their stated env count
METRIC = "rendered frames/sec (envs x cams, 256^2, 1M Gaussians)"
WORKLOAD = "cfg3: fr3_umi, 64 envs/GPU x 2 cams (wrist + third-person), 256x256, synthetic 1M-Gaussian scene"


def select_config(name, world, n=None, res=None, envs=None):
    """Point the module-level workload description at one of BASELINE.json's configs."""
    global H, W, ENVS_PER_GPU, N_GAUSS, ROBOT, SCENE_CFG, FAMILY, PER_LINK, OBJECTS, CAMERAS, SEED, CONFIG, SHARD, WORKLOAD
    global METRIC
    c = dict(CONFIGS[name])
    if name == "cfg5":
        c["n"], c["hw"], c["envs"] = n or c["n"], res or c["hw"], envs or c["envs"]
    CONFIG = name
    H = W = c["hw"]
    N_GAUSS, ROBOT, SCENE_CFG, FAMILY, PER_LINK = c["n"], c["robot"], c["scene_cfg"], c["family"], c["per_link"]
    OBJECTS, CAMERAS, SEED = c["objects"], c["cams"], c["seed"]
    if name == "cfg3":
        SHARD, ENVS_PER_GPU = "weak", c["envs"]
    else:
        from gsworld_b200.dist import shard_envs
        SHARD = "strong"
        ENVS_PER_GPU = max(hi - lo for lo, hi in (shard_envs(c["envs"], r, world) for r in range(world)))
        if c["envs"] % world:
            raise SystemExit(f"{name}: {c['envs']} envs do not split evenly over {world} GPUs")
    if name != "cfg3":
        METRIC = f"rendered frames/sec (envs x cams, {H}^2, {N_GAUSS // 1000}k Gaussians)"
    per = f"{c['envs']} envs/GPU" if SHARD == "weak" else f"{c['envs']} envs in total ({ENVS_PER_GPU}/GPU)"
    WORKLOAD = (f"{name}: {ROBOT}, {per} x {len(CAMERAS)} cam{'s' if len(CAMERAS) > 1 else ''}, {H}x{W}, "
                f"synthetic {N_GAUSS // 1000}k-Gaussian scene")


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def make_scene():
    from gsworld_b200 import fixtures as fx
    return fx.make_scene(fx.SceneSpec(robot=ROBOT, n_total=N_GAUSS, per_link=PER_LINK, objects=OBJECTS, per_object=7000,
                                      seed=SEED))


class ReplayEnv:
    """Pre-generated env states (seeded FakeEnv roll-out) replayed step by step, so the timed region holds
    no kinematics/physics.  resident=True hands out device tensors; False hands out pinned host tensors
    (the wrapper's .to(device) is then the per-step H2D copy)."""

    def __init__(self, base, n_states, device, resident, cam_base=None):
        from types import SimpleNamespace
        self.num_envs = base.num_envs
        self.sensor_hw = base.sensor_hw
        self.cameras = base.cameras
        self.static_cameras = base.static_cameras
        self._names = [l.name for l in base.agent.robot.get_links()]
        self._states = []
        for _ in range(n_states):
            base.step(None)
            mats = torch.stack([l.pose.to_transformation_matrix() for l in base.agent.robot.get_links()], 1).float()
            actors = {k: v.clone().float() for k, v in base.get_state_dict()["actors"].items()}
            src = base
            if cam_base is not None:  # cameras of the GLOBAL env 0 (rank 0's first env), rolled in lockstep
                cam_base.step(None)
                src = cam_base
            params = {c: {k: v.clone().float() for k, v in p.items()} for c, p in src.get_sensor_params().items()}
            st = dict(mats=mats, actors=actors, params=params)
            self._states.append(self._place(st, device, resident))
        self._i = 0
        env = self

        class _P:
            def __init__(s, j):
                s.j = j

            def to_transformation_matrix(s):
                return env._states[env._i]["mats"][:, s.j]

        self.agent = SimpleNamespace(uid=base.agent.uid, robot=SimpleNamespace(
            name=base.agent.robot.name,
            get_links=lambda: [SimpleNamespace(name=n, pose=_P(j)) for j, n in enumerate(self._names)]))
        self.h2d_bytes = 0 if resident else sum(t.numel() * t.element_size() for t in self._tensors(self._states[0]))

    @staticmethod
    def _tensors(st):
        yield st["mats"]
        yield from st["actors"].values()
        for p in st["params"].values():
            yield p["extrinsic_cv"]
            yield p["intrinsic_cv"]

    @staticmethod
    def _place(st, device, resident):
        f = (lambda t: t.to(device)) if resident else (lambda t: t.cpu().pin_memory())
        return dict(mats=f(st["mats"]), actors={k: f(v) for k, v in st["actors"].items()},
                    params={c: {k: f(v) for k, v in p.items()} for c, p in st["params"].items()})

    @property
    def unwrapped(self):
        return self

    def advance(self):
        self._i = (self._i + 1) % len(self._states)

    def get_state_dict(self):
        return {"actors": self._states[self._i]["actors"]}

    def get_sensor_params(self):
        return self._states[self._i]["params"]


class ClockSampler(threading.Thread):
    """SM clock / throttle reasons sampled (NVML, every 20 ms, rank 0 only) while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    @staticmethod
    def _physical_index(index):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[index])
            except Exception:
                pass
        return index

    def run(self):
        nv = self.nv
        while not self._stop_evt.is_set():
            try:
                if nv is not None:
                    sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(
                        nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                    self.rows.append((float(sm), int(r)))
                else:
                    out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader,nounits",
                                          "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                    a, b = [float(x) for x in out.strip().split(",")]
                    self.max_sm = b
                    self.rows.append((a, 0))
            except Exception:
                pass
            self._stop_evt.wait(0.02)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=3)
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        reasons = set()
        for _, r in self.rows:
            for bit, n in names.items():
                if r & bit:
                    reasons.add(n)
        sm = [a for a, _ in self.rows]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(getattr(self, "max_sm", 0)) or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_of(stage):
    """DRAM read+write bytes per step for a stage from the committed ncu --set full capture (newest round), or None.
    Only meaningful for the configuration the capture was taken on (cfg3, strict cameras)."""
    if CONFIG != "cfg3":
        return None
    for name in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                return json.load(f)["dram_bytes_per_step"].get(stage)
        except Exception:
            continue
    return None


def pair_roofline(pairs, kernel_ms, sm_mhz, ref_pairs, traffic):
    """the blend against FP32 FMA + MUFU throughput.  `pairs` = (splat, pixel) pair evaluations the
    kernel itself counted for one step (gsw_profile_counters), ~22 fp32 flop + 1 ex2 each."""
    if not pairs or kernel_ms <= 0:
        return None
    clk = (sm_mhz or 1965.0) * 1e6
    peak_fma = 148 * 128 * 2 * clk / 1e12          # TFLOP/s fp32 (non-tensor)
    peak_mufu = 148 * 16 * clk / 1e12              # T ex2/s
    per_s = pairs / (kernel_ms * 1e-3)
    ach_fma, ach_mufu = per_s * 22 / 1e12, per_s / 1e12
    frac_fma, frac_mufu = ach_fma / peak_fma, ach_mufu / peak_mufu
    out = {"bound": "sm_fp32_mufu", "kernel": "blend", "achieved": ach_fma, "peak": peak_fma, "unit": "TFLOP/s",
           "frac": max(frac_fma, frac_mufu), "frac_fma": frac_fma, "frac_mufu": frac_mufu,
           "pairs_per_step": int(pairs), "achieved_pairs_per_s": per_s, "peak_fma": peak_fma,
           "peak_mufu": peak_mufu, "achieved_mufu": ach_mufu, "flop_per_pair": 22, "kernel_ms": kernel_ms,
           "peak_source": "148 SMs x 128 fp32 lanes x 2 x SM clock; MUFU 16 / clk / SM (SURVEY.md section 8d)",
           "traffic": traffic,
           "note": "pairs_per_step is the kernel's own count of per-pixel alpha evaluations (after exact tile culling, "
                   "the per-warp footprint cull and early termination)"}
    if ref_pairs:
        out["reference_pairs_per_step"] = int(ref_pairs["pairs_per_step"])
        out["reference_pairs_source"] = ref_pairs["source"]
        out["work_avoided_by_culling"] = 1.0 - pairs / ref_pairs["pairs_per_step"]
        out["reference_pairs_per_s"] = ref_pairs["pairs_per_step"] / (kernel_ms * 1e-3)
    return out


def reference_pairs_sample(scene, frames_per_cam):
    """Pair evaluations of the REFERENCE algorithm for one step: the oracle's sum over pixels of n_contrib (how far
    each pixel walks its tile's list before it terminates), measured on env 0's frame of every camera and scaled to
    the step's frames (the oracle needs seconds per frame at this size)."""
    from gsworld_b200 import fixtures as fx
    from gsworld_b200 import rig_tables as rt
    from oracle import torch_transform as ot
    env = fx.FakeEnv(ROBOT, num_envs=1, cameras=CAMERAS, hw=(H, W), objects=OBJECTS, seed=SEED)
    scan = fx.forward_kinematics(ROBOT, rt.ROBOT_SCAN_QPOS[ROBOT][None])[0]
    ref = ot.RefPath(env, scene, rt.SIM2GS_ROBOT[FAMILY], rt.LINK_LABELS[FAMILY], list(scan), rt.SIM2GS_OBJECT,
                     rt.OBJECT_LABELS, rt.OBJECT_OFFSET, rt.OBJECT_SCALE,
                     xarm_offset=rt.OBJECT_OFFSET["xarm_arm"] if "xarm" in ROBOT else None)
    env.step(None)
    ref.transform_gs_perlink()
    _, aux = ref.render_gsworld(strict=True, return_aux=True)
    per_cam = {c: int(aux[c][0]["pairs"]) for c in CAMERAS}
    return {"pairs_per_step": sum(per_cam.values()) * frames_per_cam, "per_frame": per_cam,
            "source": f"oracle (sum over pixels of n_contrib) on env 0's {len(CAMERAS)} frame(s) x {frames_per_cam}"}


def issue_roofline(stage_ms, sm_mhz):
    """The blend against the bound it actually has: warp instructions issued per second vs the SMs' issue
    capacity (148 SMs x 4 schedulers x 1 warp instruction per clock).  Instruction count per step from the
    committed ncu capture (same command, same workload), kernel time measured live."""
    try:
        inst = 0.0
        src = "r02_ncu_full_strict.txt" if os.path.exists(os.path.join(ROOT, "profiles", "r02_ncu_full_strict.txt")) \
            else "r01_ncu_full_strict.txt"
        if CONFIG != "cfg3":
            return None
        with open(os.path.join(ROOT, "profiles", src)) as f:
            hdr = [h.split("[")[0].strip() for h in f.readline().split("|")]
            for ln in f:
                c = [x.strip() for x in ln.split("|")]
                if c and c[0].startswith("blend_kernel"):
                    inst += float(c[hdr.index("warp_inst")])
        if inst <= 0 or stage_ms <= 0:
            return None
        peak = 148 * 4 * (sm_mhz or 1965.0) * 1e6 / 1e9
        ach = inst / (stage_ms * 1e-3) / 1e9
        return {"bound": "sm_issue", "kernel": "blend", "achieved": ach, "peak": peak, "unit": "G warp-inst/s",
                "frac": ach / peak, "warp_inst_per_step": inst,
                "source": f"instruction count: profiles/{src}; time: live stage events"}
    except Exception:
        return None


def hbm_roofline(stages, alg, peak, compulsory=None):
    """The largest HBM-facing stage (the dominant kernel, blend, is bound by fp32 / MUFU issue instead).  `achieved` uses
    the bytes that MUST cross HBM (inputs once per step + what the kernel writes; re-reads of the 4.8 MB movable set by
    every frame are L2 hits, not HBM traffic); the per-frame upper bound is kept beside it."""
    cands = {k: stages[k] for k in ("gather", "scatter", "preprocess_static", "preprocess_dynamic") if stages.get(k, 0) > 0}
    if not cands:
        return None
    k = max(cands, key=cands.get)
    b = (compulsory or {}).get(k, alg[k])
    ach = b / (stages[k] * 1e-3) / 1e9
    out = {"bound": "hbm", "kernel": k, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
           "algorithmic_bytes": b, "traffic": traffic_of(k), "kernel_ms": stages[k],
           "upper_bound_bytes_if_every_frame_reread_its_inputs": alg[k]}
    if out["traffic"]:
        out["traffic_ratio"] = out["traffic"] / max(1, b)
    return out


def cpu_reference_sample(scene, n_envs, steps, warmup, strict=True):
    """The reference path on host cores: torch transform + per-frame deepcopy/isin + CPU rasterizer."""
    from gsworld_b200 import fixtures as fx
    from gsworld_b200 import rig_tables as rt
    from oracle import oracle as orc
    from oracle import torch_transform as ot
    torch.set_num_threads(os.cpu_count() or 1)
    env = fx.FakeEnv(ROBOT, num_envs=n_envs, cameras=CAMERAS, hw=(H, W), objects=OBJECTS, seed=SEED)
    scan = fx.forward_kinematics(ROBOT, rt.ROBOT_SCAN_QPOS[ROBOT][None])[0]
    ref = ot.RefPath(env, scene, rt.SIM2GS_ROBOT[FAMILY], rt.LINK_LABELS[FAMILY], list(scan), rt.SIM2GS_OBJECT,
                     rt.OBJECT_LABELS, rt.OBJECT_OFFSET, rt.OBJECT_SCALE,
                     xarm_offset=rt.OBJECT_OFFSET["xarm_arm"] if "xarm" in ROBOT else None)
    times = []
    for i in range(warmup + steps):
        env.step(None)
        t0 = time.perf_counter()
        ref.transform_gs_perlink()
        ref.render_gsworld(strict=strict)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    frames = n_envs * len(CAMERAS)
    return frames * len(times) / sum(times), float(np.mean(times) * 1e3), orc.num_threads()


def gpu_reference_restated(scene_cpu, dev, n_envs=4, steps=3):
    """Reference-STYLE pipeline on the same GPU (baseline/: restated, per-frame launches, CUB sort, a blocking
    readback per frame).  The real rasterizer's source is absent from the reference tree, so this is a
    labelled restatement, reported beside the headline, never used by the product."""
    from baseline import ref_gpu
    from gsworld_b200 import fixtures as fx
    from gsworld_b200 import rig_tables as rt
    from oracle import torch_transform as ot
    env = fx.FakeEnv(ROBOT, num_envs=n_envs, cameras=CAMERAS, hw=(H, W), objects=OBJECTS, seed=SEED, device=dev)
    scan = fx.forward_kinematics(ROBOT, rt.ROBOT_SCAN_QPOS[ROBOT][None])[0]
    scene = scene_cpu.to(dev)
    ref = ot.RefPath(env, scene, rt.SIM2GS_ROBOT[FAMILY], rt.LINK_LABELS[FAMILY], list(scan), rt.SIM2GS_OBJECT,
                     rt.OBJECT_LABELS, rt.OBJECT_OFFSET, rt.OBJECT_SCALE, device=dev,
                     xarm_offset=rt.OBJECT_OFFSET["xarm_arm"] if "xarm" in ROBOT else None,
                     rasterizer=ref_gpu.render_like_reference)
    frames = n_envs * len(CAMERAS)

    def full_step():
        ref.transform_gs_perlink()
        return ref.render_gsworld(strict=True)

    full_step()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(steps):
        env.step(None)
        full_step()
    torch.cuda.synchronize(dev)
    fps_full = frames * steps / (time.perf_counter() - t0)
    # rasterizer only: pre-activated tensors of the un-posed scene, one call per frame
    cams = ref.cameras(True)
    sc, ro = torch.exp(scene._scaling), torch.nn.functional.normalize(scene._rotation)
    op = torch.sigmoid(scene._opacity).reshape(-1)
    shs = torch.cat((scene._features_dc, scene._features_rest), 1).contiguous()
    bg = torch.zeros(3, device=dev)
    per = [cams[c][0] for c in CAMERAS]
    per = [{k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in c.items()} for c in per]

    def ras_only():
        for c in per:
            for _ in range(n_envs):
                ref_gpu.rasterize(scene._xyz, shs, op, sc, ro, c["view"], c["proj"], c["campos"], c["tanfovx"],
                                  c["tanfovy"], H, W, bg)

    ras_only()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(steps):
        ras_only()
    torch.cuda.synchronize(dev)
    fps_ras = frames * steps / (time.perf_counter() - t0)
    # integration level 1: OUR GaussianRasterizer shim called the same way, one frame per
    # call through gsw_rasterize_forward (no layers, no culling hierarchy, a host wait per call for the overflow check)
    from gsworld_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    zeros = torch.zeros_like(scene._xyz)

    def ours_single():
        for c in per:
            rs = GaussianRasterizationSettings(H, W, c["tanfovx"], c["tanfovy"], bg, 1.0, c["view"], c["proj"], 3,
                                               c["campos"], False, False, False)
            ras = GaussianRasterizer(rs)
            for _ in range(n_envs):
                ras(scene._xyz, zeros, op, shs=shs, scales=sc, rotations=ro)

    ours_single()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(steps):
        ours_single()
    torch.cuda.synchronize(dev)
    fps_ours_single = frames * steps / (time.perf_counter() - t0)
    return {"as_gsworld_calls_it": fps_full, "rasterizer_only": fps_ras, "unit": "frames/s",
            "ours_single_frame_api": fps_ours_single,
            "sample": f"{n_envs} envs x {len(CAMERAS)} cams x {steps} steps, same scene/cameras, 1 GPU",
            "what": "RESTATED reference-style CUDA pipeline (baseline/ref_rasterizer), not the reference's binary"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    scene = make_scene()
    n_envs = 1
    fps, ms, cores = cpu_reference_sample(scene, n_envs, args.steps, args.warmup)
    sample = f"sample: {n_envs} env x {len(CAMERAS)} cams per step (of {ENVS_PER_GPU} envs), same scene and cameras"
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{WORKLOAD} -- {sample}", "cameras": "strict (env-0 camera, reference behaviour)",
                   "what": "CPU port of the reference path (oracle/): torch transform with deepcopy + isin write-back "
                           "per frame, CPU rasterizer; per-frame cost does not depend on the env count"},
        "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch.distributed as dist

    from gsworld_b200 import fixtures as fx
    from gsworld_b200 import rig_tables as rt
    from gsworld_b200.gaussian_renderer import PipelineParams
    from gsworld_b200.gs_world_wrapper import GSWorldWrapper

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        from gsworld_b200.dist import bind_to_gpu_numa_node
        numa = bind_to_gpu_numa_node(local)  # before any pinned allocation: host buffers local to the GPU's socket
        if os.environ.get("GSW_BENCH_DEBUG"):
            log(f"[rank {rank}] {numa}")
        dist.init_process_group("nccl", device_id=dev)
    scene_cpu = make_scene()
    scene = scene_cpu.to(dev)

    def run_legs(B, lean):
        """All timed legs for B envs on this GPU.  lean: the resident-input strict leg only (strong-scaling point)."""
        F = B * len(CAMERAS)
        base = fx.FakeEnv(ROBOT, num_envs=B, cameras=CAMERAS, hw=(H, W), objects=OBJECTS, seed=SEED + 17 * rank)
        scan = fx.forward_kinematics(ROBOT, rt.ROBOT_SCAN_QPOS[ROBOT][None])[0]
        n_states = 8

        def build(resident, strict):
            # strict mode = the reference's behaviour: EVERY env of the batch is rendered through env 0's camera
            # (gs_world_wrapper.py:280-281).  With the batch sharded over ranks that is the GLOBAL env 0, so every
            # rank takes its camera parameters from a replica of rank 0's env roll-out.
            cam_base = None
            if strict and world > 1 and rank != 0:
                cam_base = fx.FakeEnv(ROBOT, num_envs=B, cameras=CAMERAS, hw=(H, W), objects=OBJECTS, seed=SEED)
                for _ in range(build.calls * n_states):
                    cam_base.step(None)
            env = ReplayEnv(base, n_states, dev, resident, cam_base=cam_base)
            build.calls += 1
            wrap = GSWorldWrapper(env, PipelineParams(), SCENE_CFG, device=dev, merged_model=scene,
                                  scan_link_mats=list(scan), strict_reference=strict)
            for _ in range(n_states):  # size the grow-only workspace over every replay state (not timed)
                env.advance()
                wrap.transform_gs_perlink(None)
                wrap._render_gsworld()
            return env, wrap

        # Multi-GPU: every step ends with ONE all-gather of the uint8 observations.  It is issued asynchronously
        # into one of two buffers, so the exchange of step k overlaps the rendering of step k+1; a buffer is
        # only reused after the collective that read/wrote it has completed.
        nbuf = 2
        # "peer" (default where torch symmetric memory works): libgsw's peer stream pushes every finished frame group
        # into the POLICY rank's buffer (rank 0) with the copy engines, overlapped with the
        # blend; arrival is a symmetric-memory signal per producer, buffer reuse a credit signal back (dist.py).
        # "peer_all": the same into every rank's buffer + one device-side barrier per step.  "all": NCCL all-gather;
        # "root": NCCL gather to rank 0; "none": no exchange
        GATHER = os.environ.get("GSW_BENCH_GATHER", "peer")
        NO_DATA = os.environ.get("GSW_BENCH_XCHG_NO_DATA") == "1"
        build.calls = 0
        xchg = None
        if world > 1 and GATHER in ("peer", "peer_all"):
            from gsworld_b200.dist import PeerObservationExchange
            ok = 0
            try:
                if PeerObservationExchange.available():
                    xchg = PeerObservationExchange((len(CAMERAS), B, H, W, 3), dev, nbuf=nbuf,
                                                   roots=[0] if GATHER == "peer" else None)
                    ok = 1
            except Exception as e:  # noqa: BLE001
                log(f"[rank {rank}] symmetric memory unavailable ({e!r}); falling back to the NCCL all-gather")
            flag = torch.tensor([ok], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if int(flag.item()) == 0:
                xchg, GATHER = None, "all"
        gathered = [torch.empty((world, len(CAMERAS), B, H, W, 3), dtype=torch.uint8, device=dev) for _ in range(nbuf)] \
            if world > 1 and xchg is None and GATHER in ("all", "root") else None
        pending = [None] * nbuf
        counter = [0]
        consumer = torch.cuda.Stream(device=dev) if xchg is not None and xchg.is_root else None
        verified = [None]

        def consume(k):
            """The policy side: wait for step k's batch on its own stream, then hand the buffer back."""
            if consumer is not None and k >= 0:
                with torch.cuda.stream(consumer):
                    xchg.wait_ready(k)
                    xchg.release(k)

        def step(env, wrap):
            env.advance()
            wrap.transform_gs_perlink(None)
            if xchg is not None:
                k = counter[0]
                counter[0] += 1
                xchg.begin(k, wrap._engine)
                wrap.obs_out = xchg.local_out(k)  # composite straight into this rank's slot of its own buffer
                # (diagnostic: GSW_BENCH_XCHG_NO_DATA=1 keeps the signals and credits but pushes no frames)
                wrap.obs_peers = [] if NO_DATA else xchg.destinations(k)
                # asynchronous legs: the render stream does not wait for the delivery; the host-synchronous e2e leg
                # waits for its own host copy (it reads the frames right after) but not for the pushes to the policy rank
                # (asynchronous legs deliver after the blend: the pushes run beside the next step's kernels, and the
                # blend carries no completion signals; GSW_BENCH_DELIVER=overlap restores the in-blend delivery)
                wrap.defer_delivery_join = "peers" if wrap.sync_each_step else (
                    True if os.environ.get("GSW_BENCH_DELIVER") == "overlap" else "after")
                out = wrap._render_gsworld()
                consume(k - 1)  # one step behind, while step k is still being composited
                xchg.publish(k, wrap._engine)  # after xchg.wait_ready(k), xchg.gathered(k) holds every rank's frames
                return out
            wrap.obs_out = wrap.obs_peers = None
            wrap.defer_delivery_join = False
            out = wrap._render_gsworld()
            if world > 1 and GATHER in ("all", "root"):
                k = counter[0] % nbuf
                counter[0] += 1
                if pending[k] is not None:
                    pending[k].wait()
                mine = gathered[k][rank]
                for i, c in enumerate(CAMERAS):
                    mine[i].copy_(out[c])
                if GATHER == "root":  # only the policy rank (0) needs the whole batch
                    pending[k] = dist.gather(mine, list(gathered[k].unbind(0)) if rank == 0 else None, dst=0, async_op=True)
                else:
                    pending[k] = dist.all_gather_into_tensor(gathered[k].view(-1), mine.view(-1), async_op=True)
            return out

        def verify_exchange(env, wrap):
            """Once, before anything is timed: the batch the peer exchange delivers to the policy rank is byte-identical
            to an NCCL all-gather of the same frames (the 2-GPU pytest cannot run on a 1-GPU test box)."""
            if xchg is None or NO_DATA:
                return
            wrap.sync_each_step = False
            ok = True
            for _ in range(3):  # both buffers, one of them twice
                out = step(env, wrap)
                k = counter[0] - 1
                mine = torch.stack([out[c] for c in CAMERAS]).contiguous()
                want = torch.empty((world,) + tuple(mine.shape), dtype=torch.uint8, device=dev)
                dist.all_gather_into_tensor(want.view(-1), mine.view(-1))
                if xchg.is_root:
                    xchg.wait_ready(k)
                    torch.cuda.synchronize(dev)
                    ok = ok and bool(torch.equal(xchg.gathered(k), want))
                    xchg.release(k)
                torch.cuda.synchronize(dev)
            flag = torch.tensor([1 if ok else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            verified[0] = bool(int(flag.item()))
            if not verified[0]:
                raise RuntimeError("peer observation exchange differs from the NCCL all-gather")

        def drain():
            if xchg is not None:
                consume(counter[0] - 1)
                xchg.stream.synchronize()
                if consumer is not None:
                    consumer.synchronize()
            for k in range(nbuf):
                if pending[k] is not None:
                    pending[k].wait()
                    pending[k] = None

        def barrier():
            if world > 1:
                drain()
                dist.barrier()
            torch.cuda.synchronize(dev)

        timed_launches = [0]
        rank_ms = [0.0]

        def timed(env, wrap, steps, warmup, d2h=None):
            for _attempt in range(4):
                ms = timed_once(env, wrap, steps, warmup, d2h)
                if ms is not None:
                    return ms
                warmup = max(warmup, n_states)  # the workspace was grown: prime every replay state again, re-measure
            raise RuntimeError("tile-entry workspace kept overflowing inside the timed region")

        def timed_once(env, wrap, steps, warmup, d2h=None):
            wrap.sync_each_step = True  # warm-up validates (and if needed grows) the tile-entry workspace
            for _ in range(warmup):
                out = step(env, wrap)
                if d2h is not None and not isinstance(d2h, str):
                    for i, c in enumerate(CAMERAS):
                        d2h[i].copy_(out[c], non_blocking=True)
            wrap.sync_each_step = d2h is not None  # resident-input runs are fully asynchronous
            if not wrap.sync_each_step and steps > 0:
                for _ in range(2):  # untimed: the asynchronous path (peer exchange, deferred delivery) in its own mode
                    step(env, wrap)
            barrier()
            prof = steps > 0 and os.environ.get("GSW_PROFILE_RANGE") == "1"  # ncu --profile-from-start off
            if prof:
                torch.cuda.profiler.start()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            launches_before = wrap._engine.kernel_launches()
            e0.record()
            h0 = time.perf_counter()
            for _ in range(steps):
                out = step(env, wrap)
                if d2h is not None:
                    if not isinstance(d2h, str):
                        for i, c in enumerate(CAMERAS):
                            d2h[i].copy_(out[c], non_blocking=True)
                    torch.cuda.current_stream().synchronize()  # the caller reads the observation every step
            host_ms = (time.perf_counter() - h0) * 1e3
            timed_launches[0] = wrap._engine.kernel_launches() - launches_before  # kernels of the timed region only
            if world > 1:
                drain()  # the last exchanges belong to the timed steps
            e1.record()
            barrier()
            if prof:
                torch.cuda.profiler.stop()
            ok = wrap._engine.finish(raise_on_overflow=False)
            if world > 1:  # every rank must take the same decision
                flag = torch.tensor([0 if ok else 1], device=dev)
                dist.all_reduce(flag, op=dist.ReduceOp.MAX)
                ok = int(flag.item()) == 0
            if not ok:
                return None  # a launch in the timed region overflowed the tile-entry workspace: number invalid
            ms = e0.elapsed_time(e1)
            if os.environ.get("GSW_BENCH_DEBUG") and steps > 0:
                log(f"[rank {rank}] {ms / steps:.3f} ms/step on this rank, host loop {host_ms / steps:.3f} ms/step, "
                    f"D={wrap._engine.last_stats['D']}")
            rank_ms[0] = ms  # this rank's own device time (the line reports the max over ranks)
            if world > 1:
                t = torch.tensor([ms], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            return ms

        results = {}
        sampler = None
        modes = (("strict", not args.quick_per_env),) if args.quick else (("strict", True), ("per_env", False))
        if lean:
            modes = (("strict", True),)
        for mode, strict in modes:
            env, wrap = build(True, strict)
            eng = wrap._engine
            if verified[0] is None and xchg is not None:
                verify_exchange(env, wrap)
            if os.environ.get("GSW_NB"):
                eng.set_depth_buckets(int(os.environ["GSW_NB"]))
            if os.environ.get("GSW_TILE_CULL"):
                eng.set_tile_culling(os.environ["GSW_TILE_CULL"] == "1")
            if mode == "strict":
                sampler = ClockSampler(local)  # every rank samples its own GPU; rank 0 reports them all
                timed(env, wrap, 0, args.warmup)
                sampler.start()
                ms = timed(env, wrap, args.steps, 0)
                clocks = sampler.stop()
                if world > 1:
                    every = [None] * world
                    dist.all_gather_object(every, {"sm_mhz": clocks["sm_mhz"], "reasons": clocks["reasons"],
                                                   "ms_per_step": rank_ms[0] / max(args.steps, 1)})
                    clocks["per_rank"] = every
                    sm_all = [c["sm_mhz"] for c in every if c["sm_mhz"]]
                    if sm_all:
                        clocks["sm_mhz_min_over_ranks"] = min(sm_all)
                    clocks["reasons"] = sorted(set(sum((c["reasons"] for c in every), [])))
                launches = timed_launches[0]
            else:
                ms = timed(env, wrap, args.steps, args.warmup)
            # stage breakdown + traffic counts (outside the timed region)
            eng.profile(True)
            acc = {}
            for _ in range(3):
                step(env, wrap)
                for k, v in eng.stage_ms().items():
                    acc[k] = acc.get(k, 0.0) + v / 3
            eng.profile(False)
            cnts = eng.counters()  # of the last (profiled) step
            pairs = cnts["blend_pairs"]
            if os.environ.get("GSW_BENCH_DEBUG") and rank == 0:  # bin-size histogram of the last step (tuning aid)
                import numpy as np
                off = eng.debug_bin_offsets().astype(np.int64)
                n = np.diff(off)
                Vn = int(wrap._views(list(CAMERAS))[0].shape[0])
                T = (len(n)) // (Vn + F)
                for name, part in (("static", n[:Vn * T]), ("dynamic", n[Vn * T:])):
                    edges = [0, 1, 65, 129, 257, 513, 1025, 2049, 4097, 16385, 1 << 30]
                    h, _ = np.histogram(part, bins=edges)
                    w, _ = np.histogram(part, bins=edges, weights=part)
                    log(f"[bins {mode} {name}] total {int(part.sum())} in {len(part)} bins; per size class "
                        f"(upper edge: bins/keys): " + ", ".join(f"<{e}: {int(a)}/{int(b)}" for e, a, b in
                                                                zip(edges[1:], h, w)))
            st = dict(eng.last_stats)
            Vn = int(wrap._views(list(CAMERAS))[0].shape[0])
            off = eng.debug_bin_offsets()
            fv = wrap._views(list(CAMERAS))[4].cpu().numpy()
            Tn = (len(off) - 1) // (Vn + F)
            shared = [v for v in range(Vn) if int((fv == v).sum()) >= 2]
            d_shared = int(sum(int(off[(v + 1) * Tn]) - int(off[v * Tn]) for v in shared))
            results[mode] = dict(ms=ms, fps=world * F * args.steps / (ms * 1e-3), stages=acc, stats=st, V=Vn,
                                 D_shared_static=d_shared, pairs=pairs, chunks_live=cnts["static_chunks_live"],
                                 vis_s=cnts["visible_static"], vis_d=cnts["visible_dynamic"],
                                 chunk_items=cnts["static_chunk_items"],
                                 chunks_total=Vn * (-(-int((scene_cpu._semantics.reshape(-1) < 0).sum()) //
                                                      max(cnts["static_chunk_items"], 1))))
            if mode == "strict" and not args.quick and not lean and not os.environ.get("GSW_TILE_CULL"):
                # the same step with the reference's full tile lists (exact tile culling off); same pixels
                eng.set_tile_culling(False)
                ms_full = timed(env, wrap, args.steps, args.warmup)
                results[mode]["reference_tile_lists"] = {
                    "value": world * F * args.steps / (ms_full * 1e-3), "ms_per_step": ms_full / args.steps,
                    "D": int(eng.last_stats["D"])}
                eng.set_tile_culling(True)
                # ... and with the static-list cache off: the world-fixed camera's static lists rebuilt every step
                os.environ["GSW_NO_STATIC_CACHE"] = "1"
                try:
                    ms_nc = timed(env, wrap, args.steps, args.warmup)
                finally:
                    os.environ.pop("GSW_NO_STATIC_CACHE", None)
                results[mode]["static_cache_off"] = {"value": world * F * args.steps / (ms_nc * 1e-3),
                                                     "ms_per_step": ms_nc / args.steps}
            if mode == "strict":
                results[mode].update(clocks=clocks, launches=launches)
                if lean:
                    continue
                if args.quick:
                    results["e2e"] = dict(ms=float("nan"), fps=float("nan"), h2d=0, d2h=0)
                    results["per_env"] = results["strict"]
                    continue
                # e2e: host (pinned) inputs every step, uint8 observations read back to pinned host memory
                env_h, wrap_h = build(False, True)
                pinned = torch.empty((len(CAMERAS), B, H, W, 3), dtype=torch.uint8).pin_memory()
                # the uint8 observations land in pinned host memory: the library sends the frames group by group on its
                # copy stream as soon as the blend kernel has signalled that group complete (obs_host)
                wrap_h.obs_host = pinned
                ms_e = timed(env_h, wrap_h, args.steps, args.warmup, d2h="library")
                wrap_h.obs_host = None
                results["e2e"] = dict(ms=ms_e, fps=world * F * args.steps / (ms_e * 1e-3), h2d=env_h.h2d_bytes,
                                      d2h=pinned.numel())
                del env_h, wrap_h
            del env, wrap
            torch.cuda.empty_cache()

        return results, xchg is not None, verified[0]

    results, used_peer, exchange_verified = run_legs(ENVS_PER_GPU, False)
    strong = None
    if world > 1 and CONFIG == "cfg3" and not args.quick and ENVS_PER_GPU % world == 0:
        # BASELINE.json states cfg3 as 64 envs over 1/2/4/8 GPUs: the same step with the 64 envs SHARDED (strong scaling)
        rs, _, _ = run_legs(ENVS_PER_GPU // world, True)
        strong = {"envs_total": ENVS_PER_GPU, "envs_per_gpu": ENVS_PER_GPU // world, "value": rs["strict"]["fps"],
                  "ms_per_step": rs["strict"]["ms"] / args.steps, "scaling": "strong",
                  "stages_ms": {k: round(v, 4) for k, v in rs["strict"]["stages"].items()}}
    B = ENVS_PER_GPU
    F = B * len(CAMERAS)
    if rank == 0:
        r = results["strict"]
        peak, peak_src = peaks()
        Nm = int((scene_cpu._semantics.reshape(-1) >= 0).sum())
        D, R = r["stats"]["D"], r["stats"]["staged"]
        stages = r["stages"]
        dom = max(stages, key=stages.get)
        # algorithmic bytes per launch of each stage: N static reads once per VIEW,
        # movable Gaussians once per FRAME, D tile entries built, R entries staged by the blend.
        V = r["V"]
        alg = {
            "preprocess_static": V * N_GAUSS * (12 + 12 + 16 + 4 + 4 + 4),
            "preprocess_dynamic": F * Nm * (12 + 12 + 12 + 16 + 4 + 4 + 4 + 4),
            "scatter": (V * N_GAUSS + F * Nm) * 4 + D * 8,
            "sort": D * (8 + 8),
            # only the static lists of views shared by several frames are materialised as splat streams
            "gather": r["D_shared_static"] * (8 + 4 + 48 + 48),
            # every staged entry: key + position lookup + 48 B of its record (fetched by the blend itself for all
            # dynamic lists and unshared static lists; a 48 B splat for shared static lists), + the uint8 frames
            "blend": R * (8 + 4 + 48) + F * H * W * 3,
        }
        # what must cross HBM: every input once per step (the per-frame re-reads of the movable set hit L2), plus the
        # 64 B record + 16 B tile mask + 4 B radius of every visible item and the 192 B SH row it reads
        Ns_ = N_GAUSS - Nm
        compulsory = {
            "preprocess_dynamic": Nm * (68 + 192) + r.get("vis_d", 0) * (64 + 4),
            "preprocess_static": Ns_ * 48 + r.get("vis_s", 0) * (192 + 64 + 4),
        }
        ach = alg.get(dom, 0) / (stages[dom] * 1e-3) / 1e9 if stages[dom] > 0 else 0.0
        cpu = None
        ref_pairs = None
        if world == 1 and not args.no_cpu_baseline and not args.quick:
            fps_c, ms_c, cores = cpu_reference_sample(scene_cpu, 1, 3, 1)
            cpu = {"value": fps_c, "unit": "frames/s", "cores": cores, "kind": "port",
                   "sample": f"sample: 1 env x {len(CAMERAS)} cams per step (of {B} envs), 3 steps, same "
                             f"{N_GAUSS // 1000}k-Gaussian scene, oracle/ port"}
            try:
                ref_pairs = reference_pairs_sample(scene_cpu, B)
            except Exception as e:  # noqa: BLE001
                log(f"reference pair count unavailable: {e!r}")
        sm_mhz = (r["clocks"] or {}).get("sm_mhz")
        hbm_dom = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                   "traffic": traffic_of(dom), "peak_source": peak_src, "kernel_ms": stages[dom],
                   "algorithmic_bytes": alg.get(dom, 0)}
        if hbm_dom["traffic"]:
            hbm_dom["traffic_ratio"] = hbm_dom["traffic"] / max(1, alg.get(dom, 0))
        if dom == "blend":
            # the dominant kernel is bound by fp32 + MUFU issue, not by HBM: report it against that roofline (SURVEY.md
            # section 8d) and keep the byte view beside it
            roof = pair_roofline(r["pairs"], stages["blend"], sm_mhz, ref_pairs, traffic_of("blend")) or hbm_dom
            if roof is not hbm_dom:
                roof["hbm_view"] = {k: hbm_dom[k] for k in ("achieved", "peak", "unit", "frac", "algorithmic_bytes")}
                if roof["traffic"]:
                    roof["traffic_ratio"] = roof["traffic"] / max(1, alg["blend"])
        else:
            roof = hbm_dom
        gpu_ref = None
        if world == 1 and not args.quick and not args.no_gpu_reference:
            try:
                gpu_ref = gpu_reference_restated(scene_cpu, dev)
            except Exception as e:  # the comparison point must never break the bench line
                gpu_ref = {"unavailable": repr(e)[:200]}
        line = {
            "metric": METRIC, "value": r["fps"], "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms"] / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "name": CONFIG, "envs_per_gpu": B, "frames_per_step_per_gpu": F, "views": V,
                       "cameras": "strict (env-0 camera for every env = reference behaviour, gs_world_wrapper.py:280-281)",
                       "tile_lists": "exact tile culling on (wrapper default): entries that stay below alpha 1/255 on a "
                                     "whole tile are never binned; pixels and radii bit-identical to the "
                                     "reference's full lists (reference_tile_lists = the same step with it off)",
                       "l2": "inputs larger than L2 (236 MB scene + GB-scale workspace per step), no explicit flush",
                       "parity_mode": "tile_culling_on (reference_tile_lists = the contract-conformant lists, same pixels)",
                       "parallelism": f"env-sharded x{world}, scene replicated, uint8 observations delivered to the "
                                      "policy rank by "
                                      + ("copy-engine pushes into peer memory (NVLink), overlapped with the blend"
                                         if used_peer else "an NCCL collective")},
            "e2e": {"value": results["e2e"]["fps"], "unit": "frames/s", "h2d_bytes_per_step": results["e2e"]["h2d"],
                    "d2h_bytes_per_step": results["e2e"]["d2h"], "ms_per_step": results["e2e"]["ms"] / args.steps},
            "gpu_launches": r["launches"],
            "clocks": r["clocks"],
            "roofline": roof,
            "roofline_hbm_kernel": hbm_roofline(stages, alg, peak, compulsory),
            "roofline_issue": issue_roofline(stages.get("blend", 0.0), (r["clocks"] or {}).get("sm_mhz")),
            "stages_ms": {k: round(v, 4) for k, v in stages.items()},
            "stages_note": "measured with the stages serialised on one stream; the timed step runs the static-layer "
                           "kernels on a side stream next to the dynamic-layer ones, so the sum exceeds ms_per_step",
            "stage_gbs": {k: round(alg[k] / (stages[k] * 1e-3) / 1e9, 1) for k in alg if stages.get(k, 0) > 0},
            "counts": {"N": N_GAUSS, "movable": Nm, "D": D, "D_shared_static": r["D_shared_static"],
                       "blend_staged": R, "visible_static": r.get("vis_s"), "visible_dynamic": r.get("vis_d")},
            "reference_tile_lists": r.get("reference_tile_lists"),
            "static_list_cache": {"on": os.environ.get("GSW_NO_STATIC_CACHE") is None,
                                  "what": "static lists / static-only tiles of a view are kept from the previous step "
                                          "while its camera is bitwise unchanged (the world-fixed third-person camera; "
                                          "the wrist camera and every movable Gaussian change each step); same bytes out",
                                  "off": r.get("static_cache_off")},
            "per_env_cameras": {"value": results["per_env"]["fps"], "ms_per_step": results["per_env"]["ms"] / args.steps,
                                "views": results["per_env"]["V"],
                                "stages_ms": {k: round(v, 4) for k, v in results["per_env"]["stages"].items()},
                                "counts": dict(results["per_env"]["stats"], static_chunks_live=results["per_env"].get("chunks_live"),
                                               static_chunks_total=results["per_env"].get("chunks_total"),
                                               static_chunk_items=results["per_env"].get("chunk_items"))},
            "cpu_baseline": cpu,
            "gpu_reference_restated": gpu_ref,
        }
        line["scaling"] = SHARD
        if world > 1:
            line["exchange_verified"] = exchange_verified
            line["strong"] = strong
        pe = results["per_env"]
        if pe is not r and pe.get("pairs"):
            line["per_env_cameras"]["roofline"] = pair_roofline(pe["pairs"], pe["stages"]["blend"], sm_mhz, None, None)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-reference", action="store_true")
    ap.add_argument("--quick-per-env", action="store_true", help="with --quick: profile the per-env-camera mode")
    ap.add_argument("--quick", action="store_true",
                    help="profiling aid: strict mode only, no e2e / per-env / cpu legs (line is NOT a bench value)")
    ap.add_argument("--config", default="cfg3", choices=sorted(CONFIGS),
                    help="BASELINE.json configuration (default cfg3, the one the metric is quoted on)")
    ap.add_argument("--n", type=int, default=None, help="cfg5: Gaussians")
    ap.add_argument("--res", type=int, default=None, help="cfg5: image side")
    ap.add_argument("--envs", type=int, default=None, help="cfg5: total envs")
    args = ap.parse_args()
    select_config(args.config, int(os.environ.get("WORLD_SIZE", "1")), args.n, args.res, args.envs)
    if args.steps is None:
        args.steps = 10 if args.impl == "reference" else 20
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
