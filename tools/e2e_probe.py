"""Measurement: where the time of the end-to-end hand-off goes (host wall clock per call, device synchronised after each),
float and 8-bit keyframes, with and without the overlapped upload."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import psgradientsdf_b200 as psg   # noqa: E402
from psgradientsdf_b200 import synth   # noqa: E402

G, F = 128, 30
sc = synth.make_scene(F=F, H=480, W=640, model="SH1")
vs, T = synth.grid_for_scene(G)
gv = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64); gv.set_zmin(0.5); gv.set_zmax(3.5)
go = psg.Optimizer(gv, vs, sc.K, model="SH1", reg_n=10.0)
for f in range(F):
    R, t = sc.poses_gt[f][:3, :3].copy(), sc.poses_gt[f][:3, 3].copy()
    if f:
        gv.increase_counter()
    gv.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
fi = np.arange(F, dtype=np.int32)
img = torch.from_numpy(np.ascontiguousarray(sc.color)).pin_memory().numpy()
img8 = torch.from_numpy(np.ascontiguousarray(np.round(sc.color * 255.0).astype(np.uint8))).pin_memory().numpy()
poses = torch.from_numpy(np.ascontiguousarray(sc.poses_init)).pin_memory().numpy()
out = {}
for overlap in (True, False):
    go.set_upload_overlap(overlap)
    for name, call in (("f32", lambda: go.set_keyframes(fi, poses, img)), ("u8", lambda: go.set_keyframes_u8(fi, poses, img8))):
        rows = []
        for rep in range(12):
            torch.cuda.synchronize()
            t0 = time.perf_counter(); call(); t1 = time.perf_counter()
            go.init(); t2 = time.perf_counter()
            go.init_albedo(); t3 = time.perf_counter()
            go.iteration(damping=1.0, normal_reg=True); t4 = time.perf_counter()
            go.poses(); go.light(); torch.cuda.synchronize(); t5 = time.perf_counter()
            rows.append([round(1e3 * x, 3) for x in (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t5 - t0)])
        out[f"{name} overlap={overlap}"] = {"set_keyframes/init/init_albedo/iteration/readback/total ms, reps 2..": rows[2:]}
print(json.dumps(out))
