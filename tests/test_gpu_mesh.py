"""GPU: surface extraction on the device through the C ABI, against the CPU restatement
(oracle/psg_oracle_mesh.c, itself pinned to the reference build by tests/test_oracle_mesh.py) and the committed golden mesh."""
import os

import numpy as np
import pytest

from tests.common import MODELS, assert_bit_equal, handoff, make_gpu, scene

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_golden_mesh_on_device():
    """tests/golden/mesh.npz (made by the reference build): crop box, vertices and colours bit for bit."""
    import psgradientsdf_b200 as psg
    g = np.load(os.path.join(ROOT, "tests", "golden", "mesh.npz"))
    vs = np.float32(g["vs"])
    vol = psg.VolumetricGradSdf(g["dims"], vs, [0, 0, 0], np.float32(5) * vs, 64)
    vol.upload(g["aos"])
    lo, hi = vol.mesh_box()
    assert_bit_equal(np.concatenate([lo, hi]), g["box"], "crop box")
    v, c = vol.mesh()
    assert_bit_equal(v, g["vertices"], "vertices"); assert_bit_equal(c, g["colors"], "colours")


@pytest.mark.parametrize("model", MODELS)
def test_mesh_matches_oracle(model):
    from oracle import pyoracle
    from psgradientsdf_b200 import synth
    G = 48
    sc = scene(model)
    vs, _ = synth.grid_for_scene(G)
    gv, go = make_gpu(model, G, sc)
    handoff(go, sc); go.init_albedo()
    aos = gv.download()
    lo_o, hi_o = pyoracle.mesh_bbox(aos, [G, G, G], vs)
    lo, hi = gv.mesh_box()
    assert np.array_equal(lo, lo_o) and np.array_equal(hi, hi_o)
    for white in (False, True):
        vo, co = pyoracle.extract_mesh(aos, [G, G, G], vs, white=white)
        v, c = gv.mesh(white=white)
        assert len(vo) > 3000
        assert_bit_equal(v, vo, "vertices"); assert_bit_equal(c, co, "colours")
    # one distance step moves the surface: the mesh follows the optimizer's state
    go.step("dist", damping=1.0, normal_reg=True)
    aos2 = gv.download()
    vo, co = pyoracle.extract_mesh(aos2, [G, G, G], vs)
    v, c = gv.mesh()
    assert not np.array_equal(aos2[:, 0], aos[:, 0])
    assert_bit_equal(v, vo, "vertices after a distance step"); assert_bit_equal(c, co, "colours after a distance step")


def test_mesh_edge_cases():
    """Exact zeros, values inside the 1e-7 snap band, equal neighbours, unseen voxels (weight 0), a non-cubic grid whose crop box
    touches the volume faces, and the empty cases."""
    import psgradientsdf_b200 as psg
    from oracle import pyoracle
    rng = np.random.default_rng(7)
    dims = [21, 13, 17]
    n = int(np.prod(dims))
    vs = np.float32(0.01)
    aos = np.zeros((n, 8), np.float32)
    d = rng.normal(0, 0.008, n).astype(np.float32)
    d[rng.random(n) < 0.05] = 0.0
    d[rng.random(n) < 0.05] = np.float32(5e-8)
    d[rng.random(n) < 0.05] = np.float32(-5e-8)
    d[rng.random(n) < 0.10] = np.float32(0.004)
    d[rng.random(n) < 0.10] = np.float32(-0.004)
    aos[:, 0] = d
    aos[:, 1:4] = rng.normal(0, 1, (n, 3))
    aos[:, 4] = (rng.random(n) > 0.03) * rng.integers(1, 9, n)
    aos[:, 5:8] = rng.random((n, 3))
    vol = psg.VolumetricGradSdf(dims, vs, [0, 0, 0], np.float32(5) * vs, 64)
    vol.upload(aos)
    lo_o, hi_o = pyoracle.mesh_bbox(aos, dims, vs)
    lo, hi = vol.mesh_box()
    assert np.array_equal(lo, lo_o) and np.array_equal(hi, hi_o)
    vo, co = pyoracle.extract_mesh(aos, dims, vs)
    v, c = vol.mesh()
    assert len(vo) > 3000
    assert_bit_equal(v, vo, "vertices"); assert_bit_equal(c, co, "colours")
    # nothing near the surface: empty crop box, no triangles, no file
    far = aos.copy(); far[:, 0] = 1.0
    vol.upload(far)
    assert vol.mesh_box() is None and len(vol.mesh()[0]) == 0
    assert vol.extract_mesh("/tmp/psg_should_not_exist.ply") is False and not os.path.exists("/tmp/psg_should_not_exist.ply")
    # a surface, but nobody saw it (all weights 0): a crop box and no triangle
    unseen = aos.copy(); unseen[:, 4] = 0.0
    vol.upload(unseen)
    assert vol.mesh_box() is not None and len(vol.mesh()[0]) == 0
    # a box thinner than 3 voxels has no cells
    thin = far.copy(); thin.reshape(dims[2], dims[1], dims[0], 8)[:, :, 4:6, 0] = 0.001
    vol.upload(thin)
    lo, hi = vol.mesh_box()
    assert lo[0] == 4 and hi[0] == 5 and len(vol.mesh()[0]) == 0


@pytest.mark.parametrize("model", ["SH1", "LED"])
def test_writers_match_oracle_files(model, tmp_path):
    """The files of Optimizer::extract_mesh / save_white_mesh / saveSDF / save_pointcloud and VolumetricGradSdf::extract_mesh /
    extract_pc / saveSDF, byte for byte."""
    from oracle import pyoracle
    from psgradientsdf_b200 import synth
    G = 40
    sc = scene(model)
    vs, _ = synth.grid_for_scene(G)
    gv, go = make_gpu(model, G, sc)
    handoff(go, sc); go.init_albedo()
    aos = gv.download()
    dims = [G, G, G]
    seen = (gv.vis() != 0).any(axis=0).astype(np.uint8)
    p = lambda n: str(tmp_path / n)
    assert go.extract_mesh(p("g_mesh.ply")) and go.save_white_mesh(p("g_obj.ply")) and go.saveSDF(p("g.sdf")) and go.save_pointcloud(p("g_spc.ply"))
    assert gv.extract_mesh(p("gv_mesh.ply")) and gv.extract_pc(p("gv_pc.ply")) and gv.saveSDF(p("gv.sdf"))
    v, c = pyoracle.extract_mesh(aos, dims, vs)
    vw, cw = pyoracle.extract_mesh(aos, dims, vs, white=True)
    assert pyoracle.write_mesh_ply(p("o_mesh.ply"), v, c) and pyoracle.write_mesh_ply(p("o_obj.ply"), vw, cw)
    assert pyoracle.write_sdf(p("o.sdf"), aos, dims, vs)
    assert pyoracle.write_volume_pc(p("o_pc.ply"), aos, seen, dims, vs)
    assert pyoracle.write_surface_pc(p("o_spc.ply"), aos, go.surface(), dims, vs)
    rd = lambda n: open(tmp_path / n, "rb").read()
    assert rd("g_mesh.ply") == rd("o_mesh.ply") and rd("gv_mesh.ply") == rd("o_mesh.ply")
    assert rd("g_obj.ply") == rd("o_obj.ply")
    assert rd("g.sdf") == rd("o.sdf") and rd("gv.sdf") == rd("o.sdf")
    assert rd("gv_pc.ply") == rd("o_pc.ply")
    assert rd("g_spc.ply") == rd("o_spc.ply")


def test_outer_loop_writes_meshes(tmp_path):
    """The loop's dumps: after_iter_3_mesh.ply is the mesh of the volume as it stands when the loop returns."""
    from oracle import pyoracle
    from psgradientsdf_b200 import synth
    model, G = "SH1", 32
    sc = scene(model)
    vs, _ = synth.grid_for_scene(G)
    gv, go = make_gpu(model, G, sc)
    handoff(go, sc)
    res = go.alternatingOptimize(True, True, True, True, max_it=3, conv_threshold=1e-9, damping=1.0, save_path=str(tmp_path) + "/")
    assert res["iterations"] == 3 and (tmp_path / "after_iter_3_mesh.ply").exists()
    v, c = pyoracle.extract_mesh(gv.download(), [G, G, G], vs)
    assert pyoracle.write_mesh_ply(str(tmp_path / "o.ply"), v, c)
    assert (tmp_path / "after_iter_3_mesh.ply").read_bytes() == (tmp_path / "o.ply").read_bytes()


def test_mesh_256():
    """256^3 (BASELINE configs[2] size): the device mesh against properties that need no CPU pass over 16.7 M voxels -- every
    vertex lies on a cell edge of the crop box, triangle count is stable across two runs, white and coloured meshes share vertices."""
    import psgradientsdf_b200 as psg
    from psgradientsdf_b200 import synth
    G = 256
    sc = scene("SH1")
    gv, go = make_gpu("SH1", G, sc)
    lo, hi = gv.mesh_box()
    v1, c1 = gv.mesh()
    v2, c2 = gv.mesh(white=True)
    assert len(v1) > 100000 and len(v1) % 3 == 0
    assert np.array_equal(v1, v2) and (c2 == 255).all()
    v3, c3 = gv.mesh()
    assert np.array_equal(v1, v3) and np.array_equal(c1, c3)
    vs = gv.geometry()[1]
    g = v1 / vs                                             # box-relative grid coordinates + lo
    frac = np.abs(g - np.round(g))
    on_grid = (frac < 1e-3).sum(axis=1)
    assert (on_grid >= 2).all()                             # a point on a cell edge has at most one free coordinate
    assert (g.min(axis=0) >= lo - 1e-3).all() and (g.max(axis=0) <= hi + 1e-3).all()
    t = v1.reshape(-1, 3, 3)
    assert (np.abs(t[:, 0] - t[:, 1]).max(axis=1) <= 1.5 * vs).all()      # the three vertices of a triangle share a cell
    assert not (t[:, 0] == t[:, 1]).all(axis=1).any()       # no degenerate triangles (computeTriangles drops them)
