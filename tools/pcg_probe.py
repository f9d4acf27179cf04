"""Diagnostic: iteration count and device time of the distance solve for both PCG kernels at bench size."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import psgradientsdf_b200 as psg
from psgradientsdf_b200 import synth
import bench

class A: pass
a = A(); a.grid = int(sys.argv[1]) if len(sys.argv) > 1 else 128; a.frames = int(sys.argv[2]) if len(sys.argv) > 2 else 30
a.model = "SH1"; a.height = 480; a.width = 640
sc, vs, T = bench.build_scene(a)
G, F = a.grid, a.frames
gv = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64); gv.set_zmin(0.5); gv.set_zmax(3.5)
go = psg.Optimizer(gv, vs, sc.K, model="SH1", reg_n=10.0)
for f in range(F):
    R, t = bench.se3_parts(sc.poses_gt[f])
    if f: gv.increase_counter()
    gv.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
go.set_keyframes(np.arange(F, dtype=np.int32), sc.poses_init, sc.color); go.init(); go.init_albedo()
print("S", go.S, "E", go.energy())
for mode in (0, 2, 0, 2):
    go.set_pcg_mode(mode)
    psg.profile_reset(); psg.profile_enable(True)
    info = go.step("dist", damping=1.0, normal_reg=True)
    tab = psg.profile_table(); psg.profile_enable(False)
    print("mode", mode, "iters", info["iters"], "err", info["err"], {k: round(v[1] * 1e3, 1) for k, v in tab.items()})
for split in (1, 2, 4, 8):
    go.set_sweep_split(split)
    psg.profile_reset(); psg.profile_enable(True)
    for _ in range(5): go.energy()
    tab = psg.profile_table(); psg.profile_enable(False)
    print("split", split, {k: round(v[1] * 1e3 / v[0], 1) for k, v in tab.items()})
