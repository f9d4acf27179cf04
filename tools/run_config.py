#!/usr/bin/env python
"""Run one BASELINE.json configuration end to end through psg_opt_alternating_optimize and print one JSON summary.

    python tools/run_config.py --grid 256 --model LED --frames 30 --max-it 8
    python tools/run_config.py --grid 256 --model SH2 --upsample --max-it 8          # 256^3 -> 512^3 at iteration 5
    python tools/run_config.py --grid 1024 --frames 64 --max-it 3                    # scaling-sweep shape on one GPU

Checks the size-independent properties the parity tests use at small sizes (finite energies, surface list sorted and
inside the band, visibility words consistent with the list) and reports per-iteration times and device memory.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=128)
    ap.add_argument("--frames", type=int, default=30)
    ap.add_argument("--model", default="SH1", choices=["SH1", "SH2", "LED"])
    ap.add_argument("--max-it", type=int, default=8)
    ap.add_argument("--upsample", action="store_true")
    ap.add_argument("--damping", type=float, default=None)
    ap.add_argument("--save", default=None)
    a = ap.parse_args()

    import torch

    import psgradientsdf_b200 as psg
    from psgradientsdf_b200 import synth

    sc = synth.make_scene(F=a.frames, H=480, W=640, model=a.model)
    vs, T = synth.grid_for_scene(a.grid)
    G = a.grid
    led = a.model == "LED"
    # config_basket_LED.json layout: reg norm 0.1, reg laplacian 5, damping 3; SH configs: reg norm 10, damping 1
    reg_n, reg_l = (0.1, 5.0) if led else (10.0, 0.0)
    damping = a.damping if a.damping is not None else (3.0 if led else 1.0)
    t0 = time.perf_counter()
    gv = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, max(64, a.frames))
    gv.set_zmin(0.5); gv.set_zmax(3.5)
    go = psg.Optimizer(gv, vs, sc.K, model=a.model, reg_n=reg_n, reg_l=reg_l)
    for f in range(a.frames):
        if f > 0:
            gv.increase_counter()
        gv.update(sc.color[f], sc.depth[f], sc.K, sc.poses_gt[f][:3, :3].copy(), sc.poses_gt[f][:3, 3].copy(),
                  sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
    torch.cuda.synchronize()
    t_fuse = time.perf_counter() - t0
    fi = np.arange(a.frames, dtype=np.int32)
    t0 = time.perf_counter()
    go.set_keyframes(fi, sc.poses_init, sc.color)
    go.init()
    torch.cuda.synchronize()
    t_init = time.perf_counter() - t0
    S0 = go.S
    lin = go.surface()
    assert np.all(np.diff(lin) > 0), "surface list is not strictly ascending"
    if a.save:
        os.makedirs(a.save, exist_ok=True)
    t0 = time.perf_counter()
    res = go.alternatingOptimize(True, True, True, True, max_it=a.max_it, conv_threshold=0.0, damping=damping,
                                 upsample=a.upsample, save_path=(a.save.rstrip("/") + "/") if a.save else None)
    torch.cuda.synchronize()
    t_loop = time.perf_counter() - t0
    S1 = go.S
    E, A = go.energy()
    free, total = torch.cuda.mem_get_info()
    out = {"grid": G, "model": a.model, "frames": a.frames, "upsample": a.upsample, "surface_before": S0, "surface_after": S1,
           "active_pairs_after": A, "energy_after": E, "fuse_s": round(t_fuse, 3), "init_s": round(t_init, 3),
           "loop_s": round(t_loop, 3), "device_mem_used_gb": round((total - free) / 2**30, 2), "result": res}
    assert np.isfinite(E), "energy is not finite"
    print(json.dumps(out))


if __name__ == "__main__":
    main()
