#!/usr/bin/env python
"""The reference's own optimizer (oracle/_ref, unmodified sources) on its shipped demo, for BASELINE configs[0]:

    python tools/ref_sokrates_cpu.py [--max-it 1] [--out profiles/r1_sokrates_reference_log.txt]

Runs where /root/reference exists (the build container).  Normals come from the CPU restatement of NormalEstimator (bit-equal
to OpenCV, tests/test_oracle_normals.py) because the reference build here has no OpenCV.  Slow by construction: the
reference's surface-membership test is a linear scan, one outer iteration takes many minutes.
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-it", type=int, default=1)
    ap.add_argument("--frames", type=int, default=34)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r1_sokrates_reference_log.txt"))
    a = ap.parse_args()
    import cv2

    import run_sokrates as rs
    from oracle import pyoracle, pyref
    data = rs.REF
    K = np.loadtxt(os.path.join(data, "intrinsics.txt"), dtype=np.float32)[:3, :3].copy()
    poses = rs.load_poses(os.path.join(data, "pose.txt"))
    unit, vs = np.float32(1.0 / 1000), np.float32(0.004)
    T = np.float32(5.0) * vs
    F = a.frames
    d16 = [cv2.imread(os.path.join(data, f"depth{i:06d}.png"), cv2.IMREAD_ANYDEPTH | cv2.IMREAD_ANYCOLOR) for i in range(1, F + 1)]
    c8 = [cv2.imread(os.path.join(data, f"color{i:06d}.png")) for i in range(1, F + 1)]
    H, W = d16[0].shape
    depth0 = (d16[0].astype(np.float32) * unit).astype(np.float32)
    centroid = rs.compute_centroid(K, depth0, poses[0])
    on = pyoracle.OrNormalEstimator(W, H, K, 11)
    tmp = "/tmp/sokrates_ref/"
    os.makedirs(tmp, exist_ok=True)
    rv = pyref.RefVolume([128] * 3, vs, centroid, T, 0.5, 3.5)
    ro = pyref.RefOptimizer(rv, "SH1", vs, K, tmp, reg_n=10.0, max_it=a.max_it, conv=5e-3)
    colors = []
    for f in range(F):
        depth = pyoracle.convert_depth_u16(d16[f], unit)
        color = pyoracle.convert_color_u8(c8[f])
        colors.append(color)
        nx, ny, nz = on.compute(depth)
        rv.integrate(color, depth, nx, ny, nz, on.n_sq_inv, K, poses[f], f > 0)
    ro.set_keyframes(np.arange(F, dtype=np.int32), np.stack(poses[:F]), np.stack(colors))
    ro.init()
    t0 = time.perf_counter()
    conv, sec, out = ro.alternating(True, True, True, True)
    wall = time.perf_counter() - t0
    doc = open(tmp + "optimizer_doc.txt").read() if os.path.exists(tmp + "optimizer_doc.txt") else ""
    with open(a.out, "w") as f:
        f.write(f"# reference alternatingOptimize on data/sokrates-mvs, {F} keyframes, max_it={a.max_it}: S={ro.S}, {wall:.1f} s wall on 1 core\n")
        f.write(doc)
    print("S", ro.S, "wall", wall, "converged", conv)


if __name__ == "__main__":
    main()
