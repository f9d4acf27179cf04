"""Diagnostic (not a test): runs the CUDA path next to the C oracle (and the reference harness when present) on a small
synthetic scene and prints per-stage differences.  Usage: python tools/gpu_diag.py [MODEL] [G] [F]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import psgradientsdf_b200 as psg
from oracle import pyoracle, pyref
from psgradientsdf_b200 import synth

model = sys.argv[1] if len(sys.argv) > 1 else "SH1"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 32
F = int(sys.argv[3]) if len(sys.argv) > 3 else 6
H, W = (120, 160) if G <= 48 else (240, 320)
sc = synth.make_scene(F=F, H=H, W=W, model=model)
vs, T = synth.grid_for_scene(G)
reg_l = 5.0 if model == "LED" else 0.0

ov = pyoracle.OrVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5, 64)
oo = pyoracle.OrOptimizer(ov, model, vs, sc.K, reg_n=10.0, reg_l=reg_l)
gv = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64)
gv.set_zmin(0.5); gv.set_zmax(3.5)
go = psg.Optimizer(gv, vs, sc.K, model=model, reg_n=10.0, reg_l=reg_l)


def se3(pose):
    if pyref.available():
        return pyref.se3_from_matrix(pose)
    return pose[:3, :3].copy(), pose[:3, 3].copy()


t0 = time.time()
for f in range(F):
    R, t = se3(sc.poses_gt[f])
    ov.integrate(sc.color[f], sc.depth[f], sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv, sc.K, R, t, f > 0)
    if f > 0:
        gv.increase_counter()
    gv.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)


def cmp(name, x, y):
    x = np.asarray(x); y = np.asarray(y)
    if x.dtype == np.float32 and y.dtype == np.float32:
        eq = np.array_equal(x.view(np.uint32), y.view(np.uint32))
    else:
        eq = np.array_equal(x, y)
    d = np.abs(x.astype(np.float64) - y.astype(np.float64))
    mx = d.max() if d.size else 0.0
    sc_ = max(np.abs(x.astype(np.float64)).max() if d.size else 0.0, 1e-30)
    print(f"  {name}: bit-equal={eq} maxabs={mx:.3e} rel={mx / sc_:.3e} nneq={(x != y).sum()}")


a, b = ov.download(), gv.download()
cmp("fused volume", a, b)
cmp("vis words", ov.vis()[0], gv.vis())
fi = np.arange(F, dtype=np.int32)
oo.set_keyframes(fi, sc.poses_init, sc.color); oo.init()
go.set_keyframes(fi, sc.poses_init, sc.color); go.init()
print("S", oo.S, go.S)
cmp("surface", oo.surface(), go.surface())
cmp("vis keyframes", oo.vis_keyframes(), go.vis_keyframes())
cmp("light0", oo.light(), go.light())
oo.init_albedo(); go.init_albedo()
cmp("albedo init", ov.download(), gv.download())
po, pg = oo.probe_pairs(), go.probe_pairs()
for nm, x, y in zip(["m", "n", "ok", "I", "rend"], po, pg):
    cmp(nm, x, y)
ro, rg = oo.residual(), go.residual()
cmp("residual", ro[0], rg[0]); cmp("weight", ro[1], rg[1])
eo, eg = oo.energy(), go.energy()
print("E oracle(float, f64, A)", eo, " gpu", eg, " rel", abs(eo[1] - eg[0]) / eo[1])
print("reg energies oracle", oo.normal_energy(), oo.laplacian_energy(), oo.albedo_reg_energy(), " gpu", go.reg_energies())
order = ["light", "albedo", "dist", "pose"] if model == "LED" else ["albedo", "light", "dist", "pose"]
for it in range(2):
    for blk in order:
        kw = dict(damping=1.0, normal_reg=True, laplacian_reg=(model == "LED"))
        r1 = oo.block(blk, apply=True, **kw)
        r2 = go.step(blk, **kw)
        print(blk, "n", r1["b"].size, r2["n"], "info", r1["info"], r2["cg_success"], "iters", r1["iters"], r2["iters"], "err", r1["err"], r2["err"], "applied", r2["applied"], "accepted", r2["n_accepted"])
        if blk in ("light", "pose"):
            Hg, bg, dg = go.last_system(blk)
            Ho = r1["H"].toarray()
            bs = Hg.shape[1]
            Hgd = np.zeros_like(Ho)
            for k in range(Hg.shape[0]):
                Hgd[k * bs:(k + 1) * bs, k * bs:(k + 1) * bs] = Hg[k]
            cmp("H", Ho.astype(np.float64), Hgd); cmp("b", r1["b"].astype(np.float64), bg); cmp("delta", r1["delta"].astype(np.float64), dg)
        else:
            bg, dg = go.last_vectors(blk)
            cmp("b", r1["b"], bg); cmp("delta", r1["delta"], dg)
        cmp("volume", ov.download(), gv.download())
        cmp("light", oo.light(), go.light()); cmp("poses", oo.poses(), go.poses())
        eo, eg = oo.energy(), go.energy()
        print("  E oracle", eo[1], "gpu", eg[0], "rel", abs(eo[1] - eg[0]) / eo[1], "A", eo[2], eg[1])
print("kernel launches", psg.kernel_launches(), "time", time.time() - t0)
