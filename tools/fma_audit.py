#!/usr/bin/env python
"""Exposure of the integer parity quantities (radii, tile counts) to fp32 contraction, on the cfg3 scene.  CPU only.

The reference rasterizer is built with nvcc's default FMA contraction; its source is absent from /root/reference, so
the exact contraction pattern cannot be reproduced.  Our kernel and the oracle round every fp32 operation once
(-fmad=false / -ffp-contract=off).  This script puts numbers on what that choice can change:

  1. the oracle built twice from the same source, -ffp-contract=off (the pinned build) and -ffp-contract=fast -mfma
     (gcc fuses every a*b+c it finds): how many Gaussians change radius / tile count between the two;
  2. from the pinned build's taps, in float64: how many visible Gaussians have 3*sqrt(lambda_max) within k ulp(fp32)
     of an integer (the ceil() that gives the radius) or a tile-rectangle bound within k ulp of an integer (the
     (int) casts of getRect), k = 1, 4, 16.  Those are the ones ANY re-association could flip.

Usage: python tools/fma_audit.py [--n 1000000] [--out profiles/r02_fma_audit.json]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def taps(n_total, hw):
    """Both cfg3 cameras of env 0 (strict mode) through the reference-path restatement + the loaded oracle build."""
    import numpy as np
    from gsworld_b200 import fixtures as fx
    from gsworld_b200 import rig_tables as rt
    from oracle import torch_transform as ot
    objects = ("006_mustard_bottle", "snack_box")
    scene = fx.make_scene(fx.SceneSpec(robot="fr3_umi", n_total=n_total, per_link=4000 if n_total >= 200_000 else 250,
                                       objects=objects, per_object=7000 if n_total >= 200_000 else 1200, seed=1237))
    env = fx.FakeEnv("fr3_umi", num_envs=1, hw=(hw, hw), objects=objects, seed=1237)
    scan = fx.forward_kinematics("fr3_umi", rt.ROBOT_SCAN_QPOS["fr3_umi"][None])[0]
    ref = ot.RefPath(env, scene, rt.SIM2GS_ROBOT["fr3"], rt.LINK_LABELS["fr3"], list(scan), rt.SIM2GS_OBJECT,
                     rt.OBJECT_LABELS, rt.OBJECT_OFFSET, rt.OBJECT_SCALE)
    ref.transform_gs_perlink()
    _, aux = ref.render_gsworld(strict=True, return_aux=True, debug=True)
    out = {}
    for cam, frames in aux.items():
        r = frames[0]
        out[cam] = {k: np.asarray(r[k]) for k in ("radii", "tiles_touched", "xy", "conic_opacity")}
    return out


def near_boundary(t, hw):
    import numpy as np
    vis = t["radii"] > 0
    A, B, C = (t["conic_opacity"][vis, k].astype(np.float64) for k in range(3))
    det = A * C - B * B  # conic = inverse of the dilated 2x2 covariance
    a, c, b = C / det, A / det, -B / det
    mid = 0.5 * (a + c)
    lam = mid + np.sqrt(np.maximum(0.1, mid * mid - (a * c - b * b)))
    x = 3.0 * np.sqrt(lam)
    ulp = np.spacing(x.astype(np.float32)).astype(np.float64)
    d_ceil = np.minimum(x - np.floor(x), np.ceil(x) - x)
    r = t["radii"][vis].astype(np.float64)
    px, py = t["xy"][vis, 0].astype(np.float64), t["xy"][vis, 1].astype(np.float64)
    d_rect = np.full_like(x, np.inf)
    u_rect = np.ones_like(x)
    for v in ((px - r) / 16.0, (py - r) / 16.0, (px + r + 15.0) / 16.0, (py + r + 15.0) / 16.0):
        inside = (v > 0) & (v < hw / 16)  # casts outside the grid are clamped: a flip there changes nothing
        d = np.where(inside, np.abs(v - np.round(v)), np.inf)
        better = d < d_rect
        d_rect = np.where(better, d, d_rect)
        u_rect = np.where(better, np.spacing(np.abs(v).astype(np.float32)).astype(np.float64), u_rect)
    res = {"visible": int(vis.sum())}
    for k in (1, 4, 16):
        res[f"ceil_within_{k}ulp"] = int((d_ceil < k * ulp).sum())
        res[f"rect_cast_within_{k}ulp"] = int((d_rect < k * u_rect).sum())
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--hw", type=int, default=256)
    ap.add_argument("--out", default=None)
    ap.add_argument("--child", default=None)
    args = ap.parse_args()
    if args.child:  # run under GSW_ORACLE_LIB: dump the taps
        import numpy as np
        t = taps(args.n, args.hw)
        np.savez(args.child, **{f"{cam}/{k}": v for cam, d in t.items() for k, v in d.items()})
        return
    import numpy as np
    tmp = tempfile.mkdtemp(prefix="fma_audit_")
    fused = os.path.join(tmp, "libgsw_oracle_fma.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=fast", "-mfma", "-fno-fast-math", "-fopenmp", "-fPIC",
                           "-shared", "-o", fused, os.path.join(ROOT, "oracle", "gsw_oracle.cpp")])
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])
    dumps = {}
    for tag, libp in (("pinned", None), ("fma", fused)):
        env = dict(os.environ)
        if libp:
            env["GSW_ORACLE_LIB"] = libp
        else:
            env.pop("GSW_ORACLE_LIB", None)
        path = os.path.join(tmp, f"{tag}.npz")
        subprocess.check_call([sys.executable, os.path.abspath(__file__), "--n", str(args.n), "--hw", str(args.hw),
                               "--child", path], env=env)
        dumps[tag] = np.load(path)
    report = {"scene": f"cfg3 generator, N={args.n}, {args.hw}x{args.hw}, env 0, both cameras", "cameras": {}}
    for cam in sorted({k.split("/")[0] for k in dumps["pinned"].files}):
        p = {k: dumps["pinned"][f"{cam}/{k}"] for k in ("radii", "tiles_touched", "xy", "conic_opacity")}
        f = {k: dumps["fma"][f"{cam}/{k}"] for k in ("radii", "tiles_touched")}
        row = near_boundary(p, args.hw)
        row["radii_changed_by_gcc_fma_contraction"] = int((p["radii"] != f["radii"]).sum())
        row["tile_counts_changed_by_gcc_fma_contraction"] = int((p["tiles_touched"] != f["tiles_touched"]).sum())
        row["visibility_changed_by_gcc_fma_contraction"] = int(((p["radii"] > 0) != (f["radii"] > 0)).sum())
        report["cameras"][cam] = row
    txt = json.dumps(report, indent=1)
    print(txt)
    if args.out:
        with open(args.out, "w") as fh:
            fh.write(txt + "\n")


if __name__ == "__main__":
    main()
