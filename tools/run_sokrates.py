#!/usr/bin/env python
"""BASELINE configs[0]: the reference's shipped demo (data/sokrates-mvs, config/config_skorates.json) through this library.

    python tools/run_sokrates.py [--data DIR] [--frames N] [--max-it K] [--check] [--save DIR]

Follows main_ps.cpp for `datatype multiview` with a pose file: intrinsics.txt, pose.txt (tx ty tz qx qy qz qw), frames
color%06d.png / depth%06d.png (16-bit, 1/1000 m), 128^3 grid placed at the centroid of the first depth map, voxel size 0.004,
T = 5 voxels, zmin 0.5, zmax 3.5, every frame a keyframe (sharpness threshold 0), SH1, Cauchy lambda 0.2, reg norm 10,
damping 1, convergence 5e-3, at most 100 iterations, all four blocks on.  The frames enter as the loader read them (16-bit
depth, 8-bit colour): conversions, normal estimation, fusion and the optimizer all run on the device.

--check additionally runs the CPU restatement (oracle/) on the same inputs and compares the fused volume, the surface list
and the first energies.  The data set is not part of this repository: --data defaults to oracle/_ref/sokrates-mvs (staged by
`tools/run_sokrates.py --stage`, git-ignored) and falls back to /root/reference/data/sokrates-mvs.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
STAGED = os.path.join(ROOT, "oracle", "_ref", "sokrates-mvs")
REF = "/root/reference/data/sokrates-mvs"


def quat_to_matrix(q):
    """Eigen::Quaternionf::toRotationMatrix (float arithmetic), q = (x, y, z, w)."""
    x, y, z, w = (np.float32(v) for v in q)
    two = np.float32(2)
    tx, ty, tz = two * x, two * y, two * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    one = np.float32(1)
    return np.array([[one - (tyy + tzz), txy - twz, txz + twy],
                     [txy + twz, one - (txx + tzz), tyz - twx],
                     [txz - twy, tyz + twx, one - (txx + tyy)]], np.float32)


def load_poses(path):
    poses = []
    for line in open(path):
        v = line.split()
        if len(v) < 8:
            continue
        P = np.eye(4, dtype=np.float32)
        P[:3, 3] = np.array(v[1:4], np.float32)
        P[:3, :3] = quat_to_matrix(v[4:8])
        poses.append(P)
    return poses


def se3_parts(P):
    """What main_ps.cpp's SE3(poses[i]) hands to update(): the rotation goes matrix -> unit quaternion -> matrix.  Uses the
    reference build's Sophus when it is present (oracle/_ref), else the matrix as read."""
    try:
        from oracle import pyref
        if pyref.available():
            return pyref.se3_from_matrix(P)
    except Exception:
        pass
    return P[:3, :3].copy(), P[:3, 3].copy()


def compute_centroid(K, depth, P):
    """main_ps.cpp:346-375: one serial float sum per component, in pixel order."""
    fx_inv, fy_inv = np.float32(1) / K[0, 0], np.float32(1) / K[1, 1]
    h, w = depth.shape
    ys, xs = np.nonzero(depth > 0)
    z = depth[ys, xs]
    x0 = (xs.astype(np.float32) - K[0, 2]) * fx_inv
    y0 = (ys.astype(np.float32) - K[1, 2]) * fy_inv
    pts = np.stack([x0 * z, y0 * z, z], 1).astype(np.float32)
    R, t = P[:3, :3], P[:3, 3]
    # R * p + t per pixel in float (a0 + (a1 + a2)), then the serial sum: cumsum in float32 is sequential
    w3 = np.stack([R[a, 0] * pts[:, 0] + (R[a, 1] * pts[:, 1] + R[a, 2] * pts[:, 2]) + t[a] for a in range(3)], 1).astype(np.float32)
    c = np.cumsum(w3, axis=0, dtype=np.float32)[-1]
    return (c / np.float32(len(z))).astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--data", default=None)
    ap.add_argument("--frames", type=int, default=34)
    ap.add_argument("--max-it", type=int, default=100)
    ap.add_argument("--check", action="store_true", help="compare fusion / surface / first energies with the CPU restatement")
    ap.add_argument("--save", default=None)
    ap.add_argument("--stage", action="store_true", help="copy the first --frames frames of the data set to oracle/_ref/sokrates-mvs and exit")
    a = ap.parse_args()

    if a.stage:
        os.makedirs(STAGED, exist_ok=True)
        for name in ["intrinsics.txt", "pose.txt"] + [f"{k}{i:06d}.png" for i in range(1, a.frames + 1) for k in ("color", "depth")]:
            shutil.copy(os.path.join(REF, name), os.path.join(STAGED, name))
        print("staged", a.frames, "frames in", STAGED)
        return

    import cv2

    data = a.data or (STAGED if os.path.isdir(STAGED) else REF)
    K = np.loadtxt(os.path.join(data, "intrinsics.txt"), dtype=np.float32)[:3, :3].copy()
    poses = load_poses(os.path.join(data, "pose.txt"))
    unit = np.float32(1.0 / 1000)
    vs = np.float32(0.004)
    T = np.float32(5.0) * vs
    G = 128
    d16s, c8s = [], []
    t0 = time.perf_counter()
    for i in range(1, a.frames + 1):
        d = cv2.imread(os.path.join(data, f"depth{i:06d}.png"), cv2.IMREAD_ANYDEPTH | cv2.IMREAD_ANYCOLOR)
        c = cv2.imread(os.path.join(data, f"color{i:06d}.png"))
        if d is None or c is None:
            break
        d16s.append(np.ascontiguousarray(d, np.uint16)); c8s.append(np.ascontiguousarray(c, np.uint8))
    F = len(d16s)
    H, W = d16s[0].shape
    t_load = time.perf_counter() - t0
    depth0 = (d16s[0].astype(np.float32) * unit).astype(np.float32)

    import torch

    import psgradientsdf_b200 as psg

    centroid = psg.compute_centroid(K, depth0, poses[0])            # psg_compute_centroid: host code of the library
    assert np.array_equal(centroid, compute_centroid(K, depth0, poses[0])), "library centroid differs from the numpy restatement"

    t0 = time.perf_counter()
    nest = psg.NormalEstimator(W, H, K, 11)
    vol = psg.VolumetricGradSdf([G, G, G], vs, centroid, T, 64); vol.set_zmin(0.5); vol.set_zmax(3.5)
    opt = psg.Optimizer(vol, vs, K, model="SH1", loss="cauchy", lam=0.2, reg_n=10.0)
    torch.cuda.synchronize()
    t_setup = time.perf_counter() - t0
    t0 = time.perf_counter()
    for f in range(F):
        if f:
            vol.increase_counter()
        R, t = se3_parts(poses[f])
        vol.update_raw(c8s[f], d16s[f], unit, K, R, t, nest)
    torch.cuda.synchronize()
    t_fuse = time.perf_counter() - t0
    fi = np.arange(F, dtype=np.int32)
    key_poses = np.stack(poses[:F])
    t0 = time.perf_counter()
    opt.set_keyframes_u8(fi, key_poses, np.stack(c8s))
    opt.init()
    torch.cuda.synchronize()
    t_init = time.perf_counter() - t0
    S = opt.S
    out = {"data": data, "frames": F, "image": [H, W], "centroid": centroid.tolist(), "surface_voxels": S,
           "load_s": round(t_load, 2), "setup_s": round(t_setup, 3), "fuse_ms_per_frame": round(1e3 * t_fuse / F, 3),
           "handoff_s": round(t_init, 3)}

    if a.check:
        from oracle import pyoracle
        on = pyoracle.OrNormalEstimator(W, H, K, 11)
        ov = pyoracle.OrVolume([G, G, G], vs, centroid, T, 0.5, 3.5, 64)
        oo = pyoracle.OrOptimizer(ov, "SH1", vs, K, reg_n=10.0)
        t0 = time.perf_counter()
        colors = []
        for f in range(F):
            depth = pyoracle.convert_depth_u16(d16s[f], unit)
            color = pyoracle.convert_color_u8(c8s[f])
            colors.append(color)
            nx, ny, nz = on.compute(depth)
            R, t = se3_parts(poses[f])
            ov.integrate(color, depth, nx, ny, nz, on.n_sq_inv, K, R, t, f > 0)
        t_cpu_fuse = time.perf_counter() - t0
        a_o, a_g = ov.download(), vol.download()
        out["check_fused_volume_bit_equal"] = bool(np.array_equal(a_o.view(np.uint32), a_g.view(np.uint32)))
        out["check_fused_voxels"] = int(np.count_nonzero(a_o[:, 4]))
        oo.set_keyframes(fi, key_poses, np.stack(colors)); oo.init()
        out["check_surface_equal"] = bool(np.array_equal(oo.surface(), opt.surface()))
        out["check_keyframe_visibility_equal"] = bool(np.array_equal(oo.vis_keyframes(), opt.vis_keyframes()))
        oo.init_albedo(); opt.init_albedo()
        eo = oo.energy(); eg = opt.energy()
        out["check_energy0"] = {"oracle_f64": eo[1], "gpu": eg[0], "active_pairs_equal": bool(eo[2] == eg[1]),
                                "rel_diff": abs(eo[1] - eg[0]) / abs(eo[1])}
        out["cpu_front_end_and_fusion_ms_per_frame"] = round(1e3 * t_cpu_fuse / F, 1)
        # restore the state alternatingOptimize expects (it runs initAlbedo itself)

    save = None
    if a.save:
        os.makedirs(a.save, exist_ok=True)
        save = a.save.rstrip("/") + "/"
    t0 = time.perf_counter()
    res = opt.alternatingOptimize(True, True, True, True, max_it=a.max_it, conv_threshold=5e-3, damping=1.0, upsample=False,
                                  save_path=save)
    torch.cuda.synchronize()
    out["loop_s"] = round(time.perf_counter() - t0, 3)
    out["result"] = {k: v for k, v in res.items() if k != "iteration_ms"}
    out["iteration_ms_first"] = [round(x, 3) for x in res["iteration_ms"][:6]]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
