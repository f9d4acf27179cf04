#!/usr/bin/env python
"""Writes the binary input of examples/voxelps_host.cpp from the synthetic scene (layout in that file's header)."""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def write(path, sc, G, vs, T, model="SH1", max_it=3, conv=1e-9, damping=1.0, reg_n=10.0, reg_l=0.0, se3=None):
    F, H, W = sc.depth.shape
    f32 = np.float32
    with open(path, "wb") as f:
        np.array([F, H, W, G, {"SH1": 0, "SH2": 1, "LED": 2}[model], max_it], np.int32).tofile(f)
        np.array([vs, T, 0.5, 3.5, conv, damping, reg_n, reg_l], f32).tofile(f)
        np.ascontiguousarray(sc.K, f32).tofile(f)
        np.ascontiguousarray(sc.n_sq_inv, f32).tofile(f)
        for i in range(F):
            R, t = se3(sc.poses_gt[i]) if se3 else (sc.poses_gt[i][:3, :3], sc.poses_gt[i][:3, 3])
            for a in (R, t, sc.poses_init[i], sc.color[i], sc.depth[i], sc.nx[i], sc.ny[i], sc.nz[i]):
                np.ascontiguousarray(a, f32).tofile(f)


if __name__ == "__main__":
    from psgradientsdf_b200 import synth
    G = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    sc = synth.make_scene(F=6, H=120, W=160, model="SH1")
    vs, T = synth.grid_for_scene(G)
    write(sys.argv[1], sc, G, vs, T)
    print("wrote", sys.argv[1])
