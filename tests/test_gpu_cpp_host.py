"""GPU: a plain C++ host (examples/voxelps_host.cpp, built next to the library) drives the path through the C ABI in the order
of the reference's main_ps.cpp; its files must equal, byte for byte, what the same calls made through the ctypes binding write."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from tests.common import handoff, make_gpu, scene, se3

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "psgradientsdf_b200", "voxelps_host")


def test_cpp_host_writes_the_same_files(tmp_path):
    if not os.path.exists(EXE):
        pytest.skip("psgradientsdf_b200/voxelps_host not built (python -m psgradientsdf_b200.build)")
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import write_host_input
    from psgradientsdf_b200 import synth
    model, G = "SH1", 32
    sc = scene(model)
    vs, T = synth.grid_for_scene(G)
    cdir = tmp_path / "cpp"; cdir.mkdir(); pdir = tmp_path / "py"; pdir.mkdir()
    write_host_input.write(str(tmp_path / "in.bin"), sc, G, vs, T, model=model, max_it=3, conv=1e-9, se3=se3)
    r = subprocess.run([EXE, str(tmp_path / "in.bin"), str(cdir) + "/"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert out["iterations"] == 3 and out["kernel_launches"] > 20 and out["init_mesh_triangles"] > 1000

    gv, go = make_gpu(model, G, sc)
    assert gv.extract_mesh(str(pdir / "init_mesh.ply")) and gv.extract_pc(str(pdir / "init_pointcloud.ply"))
    handoff(go, sc)
    res = go.alternatingOptimize(True, True, True, True, max_it=3, conv_threshold=1e-9, damping=1.0, save_path=str(pdir) + "/")
    assert res["iterations"] == 3 and go.S == out["surface_voxels"]
    names = sorted(os.listdir(pdir))
    assert names == sorted(os.listdir(cdir)) and "after_iter_3_mesh.ply" in names and "optimizer_doc.txt" in names, names
    for n in names:
        assert (pdir / n).read_bytes() == (cdir / n).read_bytes(), n
