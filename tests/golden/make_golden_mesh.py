#!/usr/bin/env python
"""Golden vectors for surface extraction, produced by the reference build (oracle/_ref):

    python tests/golden/make_golden_mesh.py      # writes tests/golden/mesh.npz

A 20^3 SH1 scene is fused and handed to the reference optimizer (init + initAlbedo, so the voxels carry colours); the volume is
cropped the way Optimizer::extract_mesh crops it (OptimizerAux.cpp:278-342) and the reference's own MarchingCubes
(third/mesh/MarchingCubes.cpp, unmodified, through ref_marching_cubes) triangulates it.  The reference's extract_mesh() file for
the same state is written beside it and checked against the arrays (6 significant digits), so the crop done here is the crop the
reference does.  Run in the build container only (needs /root/reference at oracle build time).
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import pyref                                     # noqa: E402
from psgradientsdf_b200 import synth                         # noqa: E402
from tests.common import handoff, make_ref, scene            # noqa: E402

G = 20
sc = scene("SH1")
vs, _ = synth.grid_for_scene(G)
rv, _ = make_ref("SH1", G, sc)
tmp = tempfile.mkdtemp() + "/"
ro = pyref.RefOptimizer(rv, "SH1", vs, sc.K, tmp, reg_n=10.0)
handoff(ro, sc); ro.init_albedo()
aos = rv.download()
vol = aos.reshape(G, G, G, 8)                                # [k, j, i]
near = ~(np.abs(vol[..., 0]).astype(np.float64) > np.sqrt(3.0) * np.float64(np.float32(vs)))
kk, jj, ii = np.nonzero(near)
lo = np.array([ii.min(), jj.min(), kk.min()], np.int32); hi = np.array([ii.max(), jj.max(), kk.max()], np.int32)
crop = vol[lo[2]:hi[2] + 1, lo[1]:hi[1] + 1, lo[0]:hi[0] + 1]
dims = (hi - lo + 1).astype(np.int32)
f32 = np.float32
rgb = [(f32(255) * crop[..., 5 + c]).astype(np.int32).astype(np.uint8) for c in range(3)]
size = (f32(vs) * dims.astype(f32)).astype(f32)
org = (-f32(vs) * lo.astype(f32)).astype(f32)
v, c, f = pyref.marching_cubes(dims, size, org, -crop[..., 0], crop[..., 4], rgb[0], rgb[1], rgb[2])
assert ro.extract_mesh("g")
lines = open(tmp + "g_mesh.ply").read().split("\n")
assert lines[2] == f"element vertex {len(v)}", (lines[2], len(v))
body = lines[12:12 + len(v)]
for i in (0, len(v) // 2, len(v) - 1):
    assert body[i] == "%g %g %g %d %d %d" % (*v[i], *c[i]), (body[i], v[i], c[i])
np.savez_compressed(os.path.join(HERE, "mesh.npz"), aos=aos, dims=np.array([G, G, G], np.int32), vs=f32(vs),
                    box=np.concatenate([lo, hi]), vertices=v, colors=c)
print("triangles", len(f), "box", lo, hi, os.path.getsize(os.path.join(HERE, "mesh.npz")), "bytes")
