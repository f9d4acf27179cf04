"""Diagnostic (GPU): per-block relative energy gap between the CUDA path and the C restatement, in its reference mode
(serial float sums) and in its accum64 mode, for a list of configurations.  Prints one JSON line per configuration.

    python tools/parity_gap.py small            # 40^3, 6 frames 160x120, all three models, two iterations
    python tools/parity_gap.py bench            # 128^3 SH1 x 30 keyframes 640x480 (the bench workload), one iteration
    python tools/parity_gap.py led256           # 256^3 LED x 30 (configs[2])
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import psgradientsdf_b200 as psg   # noqa: E402
from oracle import pyoracle        # noqa: E402
from psgradientsdf_b200 import synth   # noqa: E402


def order(model):
    return ["light", "albedo", "dist", "pose"] if model == "LED" else ["albedo", "light", "dist", "pose"]


def run(model, G, F, H, W, iters, accum64, split, reg_n=10.0, reg_l=None, damping=1.0, strict=False):
    reg_l = (5.0 if model == "LED" else 0.0) if reg_l is None else reg_l
    sc = synth.make_scene(F=F, H=H, W=W, model=model)
    vs, T = synth.grid_for_scene(G)
    t0 = time.time()
    ov = pyoracle.OrVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5, 64)
    oo = pyoracle.OrOptimizer(ov, model, vs, sc.K, reg_n=reg_n, reg_l=reg_l)
    gv = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64)
    gv.set_zmin(0.5); gv.set_zmax(3.5)
    go = psg.Optimizer(gv, vs, sc.K, model=model, reg_n=reg_n, reg_l=reg_l)
    for f in range(F):
        R, t = sc.poses_gt[f][:3, :3].copy(), sc.poses_gt[f][:3, 3].copy()
        ov.integrate(sc.color[f], sc.depth[f], sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv, sc.K, R, t, f > 0)
        if f > 0:
            gv.increase_counter()
        gv.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
    t_fuse = time.time() - t0
    fi = np.arange(F, dtype=np.int32)
    for o in (oo, go):
        o.set_keyframes(fi, sc.poses_init, sc.color); o.init()
    same_surface = bool(np.array_equal(oo.surface(), go.surface()))
    if model == "LED":
        go.set_light(oo.light())
    if accum64:
        oo.set_accum64(True)
    if strict and hasattr(go, "set_strict"):
        go.set_strict(True)
    go.set_sweep_split(1)
    oo.init_albedo(); go.init_albedo()
    go.set_sweep_split(split)
    out = {"model": model, "G": G, "F": F, "S": int(go.S), "accum64": accum64, "split": split, "strict": strict,
           "same_surface": same_surface, "fuse_s": round(t_fuse, 1), "gaps": []}
    eo, eg = oo.energy(), go.energy()
    out["A"] = int(eg[1]); out["gap0"] = abs(eo[1] - eg[0]) / abs(eo[1])
    kw = dict(damping=damping, normal_reg=reg_n != 0, laplacian_reg=reg_l != 0)
    t_or = 0.0
    for it in range(iters):
        for blk in order(model):
            t0 = time.time(); r1 = oo.block(blk, apply=True, **kw); t_or += time.time() - t0
            r2 = go.step(blk, **kw)
            eo, eg = oo.energy(), go.energy()
            out["gaps"].append({"it": it, "blk": blk, "gap": abs(eo[1] - eg[0]) / abs(eo[1]), "A_equal": eo[2] == eg[1],
                                "iters": [int(r1["iters"]), int(r2["iters"])], "ok": [int(r1["info"] == 0), int(r2["cg_success"])]})
    out["oracle_block_s"] = round(t_or, 1)
    a, b = ov.download()[oo.surface()], gv.download()[oo.surface()]
    out["d_rms_vox"] = float(np.sqrt(np.mean((a[:, 0].astype(np.float64) - b[:, 0]) ** 2))) / vs
    out["a_rms"] = float(np.sqrt(np.mean((a[:, 5:8].astype(np.float64) - b[:, 5:8]) ** 2)))
    out["max_gap"] = max(g["gap"] for g in out["gaps"])
    print(json.dumps(out), flush=True)
    return out


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "small"
    strict = "strict" in sys.argv
    if what == "small":
        for model in ("SH1", "LED", "SH2"):
            for accum64, split in ((False, 0), (True, 0), (True, 1)):
                run(model, 40, 6, 120, 160, 2, accum64, split, strict=strict)
    elif what == "bench":
        for accum64, split in ((False, 0), (True, 0), (True, 1)):
            run("SH1", 128, 30, 480, 640, 1, accum64, split, strict=strict)
    elif what == "led256":
        for accum64, split in ((False, 0), (True, 0)):
            run("LED", 256, 30, 480, 640, 1, accum64, split, reg_n=0.1, reg_l=5.0, damping=3.0, strict=strict)
