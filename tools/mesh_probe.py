#!/usr/bin/env python
"""Fuses the small test scene into a G^3 volume and extracts the mesh a few times: a short program to put under ncu
(`ncu --set full -k regex:k_mesh -c 8 python tools/mesh_probe.py 512`) or to time the extraction alone."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402,F401

import psgradientsdf_b200 as psg  # noqa: E402
from tests.common import make_gpu, scene  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 512
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
gv, go = make_gpu("SH1", G, scene("SH1"))
for _ in range(reps):
    t0 = time.perf_counter()
    v, c = gv.mesh()
    dt = time.perf_counter() - t0
print({"grid": G, "triangles": len(v) // 3, "box": [a.tolist() for a in gv.mesh_box()], "ms": round(1e3 * dt, 3)})
