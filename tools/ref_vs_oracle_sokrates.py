#!/usr/bin/env python
"""The reference build (oracle/_ref, unmodified sources) and the C restatement (oracle/psg_oracle.c) side by side on the
reference's shipped demo data, one outer iteration, block by block:

    python tools/ref_vs_oracle_sokrates.py [--frames 6] [--out profiles/r1_sokrates_ref_vs_oracle.json]

Real 1139 x 1709 RGB-D frames, 128^3 grid, the demo's settings (config_skorates.json).  Reports, after fusion and after every
block, whether the two hold bit-identical state (volume, light, poses) and bit-identical right-hand sides / solutions.  Runs
where /root/reference exists (the build container); the reference's membership test is a linear scan, so keep --frames small
(time grows with S^2 F).  Normals come from the restatement of NormalEstimator (bit-equal to OpenCV, tests/test_oracle_normals.py)
because the reference build here has no OpenCV.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=6)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r1_sokrates_ref_vs_oracle.json"))
    a = ap.parse_args()
    import cv2

    import run_sokrates as rs
    from oracle import pyoracle, pyref
    data = rs.STAGED if os.path.isdir(rs.STAGED) else rs.REF
    K = np.loadtxt(os.path.join(data, "intrinsics.txt"), dtype=np.float32)[:3, :3].copy()
    poses = rs.load_poses(os.path.join(data, "pose.txt"))
    unit, vs = np.float32(1.0 / 1000), np.float32(0.004)
    T = np.float32(5.0) * vs
    F, G = a.frames, 128
    d16 = [cv2.imread(os.path.join(data, f"depth{i:06d}.png"), cv2.IMREAD_ANYDEPTH | cv2.IMREAD_ANYCOLOR) for i in range(1, F + 1)]
    c8 = [cv2.imread(os.path.join(data, f"color{i:06d}.png")) for i in range(1, F + 1)]
    H, W = d16[0].shape
    centroid = pyoracle.compute_centroid(K, (d16[0].astype(np.float32) * unit).astype(np.float32), poses[0])
    on = pyoracle.OrNormalEstimator(W, H, K, 11)
    rv = pyref.RefVolume([G] * 3, vs, centroid, T, 0.5, 3.5)
    ro = pyref.RefOptimizer(rv, "SH1", vs, K, "/tmp/", reg_n=10.0)
    ov = pyoracle.OrVolume([G] * 3, vs, centroid, T, 0.5, 3.5, 64)
    oo = pyoracle.OrOptimizer(ov, "SH1", vs, K, reg_n=10.0)
    colors = []
    for f in range(F):
        depth = pyoracle.convert_depth_u16(d16[f], unit)
        colors.append(pyoracle.convert_color_u8(c8[f]))
        nx, ny, nz = on.compute(depth)
        R, t = pyref.se3_from_matrix(poses[f])
        rv.integrate(colors[-1], depth, nx, ny, nz, on.n_sq_inv, K, poses[f], f > 0)
        ov.integrate(colors[-1], depth, nx, ny, nz, on.n_sq_inv, K, R, t, f > 0)
    eq = lambda x, y: bool(np.array_equal(np.ascontiguousarray(x).view(np.uint32), np.ascontiguousarray(y).view(np.uint32)))
    out = {"data": "data/sokrates-mvs", "frames": F, "grid": G, "image": [H, W], "steps": []}
    out["fused_volume_bit_equal"] = eq(rv.download(), ov.download())
    fi = np.arange(F, dtype=np.int32)
    kp = np.stack(poses[:F]); imgs = np.stack(colors)
    ro.set_keyframes(fi, kp, imgs); ro.init(); ro.init_albedo()
    oo.set_keyframes(fi, kp, imgs); oo.init(); oo.init_albedo()
    out["surface_voxels"] = int(ro.S)
    out["surface_list_equal"] = bool(np.array_equal(ro.surface(), oo.surface()))
    er, eo = ro.energy(), oo.energy()
    out["energy0"] = {"reference_float": float(er[0]), "restatement_float": float(eo[0]), "equal": bool(er[0] == eo[0] and er[1] == eo[1] and er[2] == eo[2]),
                      "active_pairs": int(er[2])}
    for blk in ("albedo", "light", "dist", "pose"):
        t0 = time.perf_counter()
        r1 = ro.block(blk, apply=True, damping=1.0, normal_reg=True)
        t_ref = time.perf_counter() - t0
        t0 = time.perf_counter()
        r2 = oo.block(blk, apply=True, damping=1.0, normal_reg=True)
        t_or = time.perf_counter() - t0
        er, eo = ro.energy(), oo.energy()
        out["steps"].append({"block": blk, "n": int(len(r1["b"])), "rhs_bit_equal": eq(r1["b"], r2["b"]), "solution_bit_equal": eq(r1["delta"], r2["delta"]),
                             "cg_iterations": [int(r1["iters"]), int(r2["iters"])], "cg_info": [int(r1["info"]), int(r2["info"])],
                             "volume_bit_equal": eq(rv.download(), ov.download()), "light_bit_equal": eq(ro.light(), oo.light()),
                             "poses_bit_equal": eq(ro.poses(), oo.poses()), "energy_after": [float(er[0]), float(eo[0])], "energy_equal": bool(er[0] == eo[0] and er[1] == eo[1]),
                             "seconds_reference": round(t_ref, 1), "seconds_restatement": round(t_or, 1)})
        print(json.dumps(out["steps"][-1], default=float), flush=True)
    with open(a.out, "w") as f:
        json.dump(out, f, indent=1, default=float)
    print(json.dumps({k: v for k, v in out.items() if k != "steps"}, default=float))


if __name__ == "__main__":
    main()
