"""CPU: surface extraction.  The C restatement (oracle/psg_oracle_mesh.c) against the reference's own
marching cubes and writers run here (oracle/_ref), the committed golden mesh, and the packed triangle table against a fresh probe
of all 256 cases.  Tests that need the reference build skip where it is absent."""
import os

import numpy as np
import pytest

from tests.common import assert_bit_equal, handoff, make_oracle, make_ref, scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "mesh.npz")


def _need_ref():
    from oracle import pyref
    if not pyref.available():
        pytest.skip("oracle/_ref/libpsgref.so not present (needs /root/reference at build time)")


def random_box(seed, dims=(9, 8, 7)):
    """A cropped box the way extract_mesh fills one: signed distances with exact zeros, near-zeros, equal neighbours and holes."""
    rng = np.random.default_rng(seed)
    n = int(np.prod(dims))
    t = rng.normal(0, 0.01, n).astype(np.float32)
    t[rng.random(n) < 0.05] = 0.0
    t[rng.random(n) < 0.05] = np.float32(5e-8)
    t[rng.random(n) < 0.05] = np.float32(-5e-8)
    t[rng.random(n) < 0.10] = np.float32(0.01)
    t[rng.random(n) < 0.10] = np.float32(-0.01)
    w = (rng.random(n) > 0.03).astype(np.float32) * rng.integers(1, 9, n).astype(np.float32)
    rgb = rng.integers(0, 256, (3, n)).astype(np.uint8)
    return np.asarray(dims, np.int32), t, w, rgb


def test_packed_table_matches_probe():
    """The committed header is what probing the reference's marching cubes on all 256 sign configurations gives."""
    _need_ref()
    import tools.make_mc_table as mk
    tab = mk.probe_all()
    assert sum(len(t) // 3 for t in tab) == 820
    assert open(os.path.join(ROOT, "psgradientsdf_b200", "csrc", "mc_table.h")).read() == mk.render(tab)


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_marching_cubes_vs_reference(seed):
    _need_ref()
    from oracle import pyoracle, pyref
    dims, t, w, rgb = random_box(seed, (9, 8, 7) if seed % 2 else (5, 11, 6))
    vs = np.float32(0.004)
    size = (vs * dims.astype(np.float32)).astype(np.float32)
    org = (-vs * np.array([3, 5, 7], np.float32)).astype(np.float32)
    iso = 0.0 if seed < 4 else 0.002
    vr, cr, fr = pyref.marching_cubes(dims, size, org, t, w, rgb[0], rgb[1], rgb[2], iso)
    vo, co = pyoracle.marching_cubes(dims, size, org, t, w, rgb[0], rgb[1], rgb[2], iso)
    assert len(vr) > 100
    assert_bit_equal(vr, vo, "vertices"); assert_bit_equal(cr, co, "colours")
    assert np.array_equal(fr.reshape(-1), np.arange(len(vr)))


def test_degenerate_boxes():
    from oracle import pyoracle
    for dims in [(2, 5, 5), (5, 2, 5), (1, 1, 1), (3, 3, 3)]:
        n = int(np.prod(dims))
        v, c = pyoracle.marching_cubes(dims, [1, 1, 1], [0, 0, 0], -np.ones(n, np.float32), np.ones(n, np.float32),
                                       np.zeros(n, np.uint8), np.zeros(n, np.uint8), np.zeros(n, np.uint8))
        assert len(v) == 0
    aos = np.zeros((27, 8), np.float32); aos[:, 0] = 1.0      # nothing near the surface: empty crop box
    assert pyoracle.mesh_bbox(aos, [3, 3, 3], 0.01) is None and pyoracle.extract_mesh(aos, [3, 3, 3], 0.01) is None


@pytest.mark.parametrize("model", ["SH1", "LED"])
def test_writers_vs_reference_files(model, tmp_path):
    """The files: Optimizer::extract_mesh / save_white_mesh / saveSDF and VolumetricGradSdf::extract_mesh / extract_pc of the
    reference build against the restatement, byte for byte, on a fused and initialised scene."""
    _need_ref()
    from oracle import pyoracle, pyref
    from psgradientsdf_b200 import synth
    G = 28
    sc = scene(model)
    vs, _ = synth.grid_for_scene(G)
    rv, _unused = make_ref(model, G, sc)
    ro = pyref.RefOptimizer(rv, model, vs, sc.K, str(tmp_path) + "/", reg_n=10.0)
    handoff(ro, sc); ro.init_albedo()
    aos = rv.download()
    dims = [G, G, G]
    words, lengths = rv.vis(1)
    seen = (words.reshape(-1) != 0).astype(np.uint8)
    assert ro.extract_mesh("a") and ro.save_white_mesh("a") and ro.save_sdf("a.sdf") and ro.save_pointcloud("a")
    assert rv.extract_mesh(str(tmp_path / "v_mesh.ply")) and rv.extract_pc(str(tmp_path / "v_pc.ply"))
    v, c = pyoracle.extract_mesh(aos, dims, vs)
    vw, cw = pyoracle.extract_mesh(aos, dims, vs, white=True)
    assert len(v) > 1000 and np.array_equal(v, vw) and (cw == 255).all()
    assert pyoracle.write_mesh_ply(str(tmp_path / "o_mesh.ply"), v, c)
    assert pyoracle.write_mesh_ply(str(tmp_path / "o_obj.ply"), vw, cw)
    assert pyoracle.write_sdf(str(tmp_path / "o.sdf"), aos, dims, vs)
    assert pyoracle.write_volume_pc(str(tmp_path / "o_pc.ply"), aos, seen, dims, vs)
    assert pyoracle.write_surface_pc(str(tmp_path / "o_spc.ply"), aos, ro.surface(), dims, vs)
    rd = lambda n: open(tmp_path / n, "rb").read()
    assert rd("a_pointcloud.ply") == rd("o_spc.ply")
    assert rd("a_mesh.ply") == rd("o_mesh.ply")
    assert rd("a_obj.ply") == rd("o_obj.ply")
    assert rd("v_mesh.ply") == rd("o_mesh.ply")
    assert rd("a.sdf") == rd("o.sdf")
    assert rd("v_pc.ply") == rd("o_pc.ply")


def test_golden_mesh():
    """Committed fixture (tests/golden/make_golden_mesh.py ran the reference build): crop box, vertices, colours."""
    from oracle import pyoracle
    g = np.load(GOLD)
    got = pyoracle.extract_mesh(g["aos"], g["dims"], float(g["vs"]))
    lo, hi = pyoracle.mesh_bbox(g["aos"], g["dims"], float(g["vs"]))
    assert_bit_equal(np.concatenate([lo, hi]), g["box"], "crop box")
    assert_bit_equal(got[0], g["vertices"], "vertices"); assert_bit_equal(got[1], g["colors"], "colours")
