"""Measurement (one GPU): the sharded path with the ranks of a job as threads of this process against the single-GPU path on the
same workload -- per-kernel device times from the library's launch timer.  On one device the ranks share the SMs, so the
sweeps cannot speed up; what this isolates is the cost of the cross-rank barriers inside the distributed distance solve.

    python tools/mg_probe.py GRID WORLD [pcg_mode]
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import psgradientsdf_b200 as psg   # noqa: E402
from psgradientsdf_b200 import synth   # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 256
world = int(sys.argv[2]) if len(sys.argv) > 2 else 2
mode = int(sys.argv[3]) if len(sys.argv) > 3 else 0
F = 30
sc = synth.make_scene(F=F, H=480, W=640, model="SH1")
vs, T = synth.grid_for_scene(G)


def build():
    gv = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64)
    gv.set_zmin(0.5); gv.set_zmax(3.5)
    go = psg.Optimizer(gv, vs, sc.K, model="SH1", reg_n=10.0)
    for f in range(F):
        R, t = sc.poses_gt[f][:3, :3].copy(), sc.poses_gt[f][:3, 3].copy()
        if f:
            gv.increase_counter()
        gv.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
    go.set_keyframes(np.arange(F, dtype=np.int32), sc.poses_init, sc.color); go.init()
    return gv, go


def table():
    return {k: round(v[1] / v[0], 4) for k, v in psg.profile_table().items()}


gv1, go1 = build()
go1.init_albedo()
for _ in range(3):
    go1.iteration(damping=1.0, normal_reg=True)
psg.profile_reset(); psg.profile_enable(True)
for _ in range(5):
    info1 = go1.iteration(damping=1.0, normal_reg=True)
single = table()
psg.profile_enable(False)
out = {"grid": G, "S": go1.S, "single": single, "pcg_iters": info1["dist"]["iters"], "single_kernel": go1.last_pcg_kernel()}
go1.set_pcg_mode(8)      # the dataflow kernel on one GPU
for _ in range(3):
    go1.iteration(damping=1.0, normal_reg=True)
psg.profile_reset(); psg.profile_enable(True)
for _ in range(5):
    info8 = go1.iteration(damping=1.0, normal_reg=True)
out["single_dataflow"] = {"k_dist_pcg": table().get("k_dist_pcg"), "pcg_iters": info8["dist"]["iters"], "kernel": go1.last_pcg_kernel()}
psg.profile_enable(False)
go1.set_pcg_mode(0)
pairs = [build() for _ in range(world)]
opts = [p[1] for p in pairs]
psg.comm_init_local(opts)
psg.run_ranks(lambda r: (opts[r].set_shard_policy(1), opts[r].set_pcg_mode(mode)), world)
psg.run_ranks(lambda r: opts[r].init_albedo(), world)
for _ in range(3):
    psg.run_ranks(lambda r: opts[r].iteration(damping=1.0, normal_reg=True), world)
psg.profile_reset(); psg.profile_enable(True)
for _ in range(5):
    infos = psg.run_ranks(lambda r: opts[r].iteration(damping=1.0, normal_reg=True), world)
out["local_ranks"] = {"world": world, "mode": mode, "kernels": table(), "pcg_iters": infos[0]["dist"]["iters"],
                      "rows": [o.row_range() for o in opts]}
psg.profile_enable(False)
print(json.dumps(out))
psg.run_ranks(lambda r: opts[r].comm_destroy(), world)
