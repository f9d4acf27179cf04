"""Device front end (normals.cu): the FALS normal estimator, the loader's pixel conversions and raw-frame fusion against the
OpenCV fixture and the CPU restatement -- bit for bit, through the C ABI."""
import os

import numpy as np
import pytest

from tests.common import handoff, make_gpu, scene

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "normals.npz")


def test_tables_and_normals_match_opencv_fixture():
    import psgradientsdf_b200 as psg
    g = np.load(GOLD)
    ne = psg.NormalEstimator(int(g["W"]), int(g["H"]), g["K"], int(g["window"]))
    t = ne.tables()
    assert np.array_equal(t[0], g["x0_n"]) and np.array_equal(t[1], g["y0_n"]) and np.array_equal(t[2], g["n_sq_inv"])
    assert np.array_equal(t[3:], g["Q"])
    nx, ny, nz = ne.compute(g["depth"])
    for a, name in ((nx, "nx"), (ny, "ny"), (nz, "nz")):
        assert np.array_equal(a, g[name], equal_nan=True), name


@pytest.mark.parametrize("shape", [(7, 9, 5), (12, 12, 11), (6, 20, 3), (5, 5, 1), (120, 160, 11), (480, 640, 11)])
def test_normals_match_oracle(shape):
    import psgradientsdf_b200 as psg
    from oracle import pyoracle
    H, W, win = shape
    rng = np.random.default_rng(H * 1000 + W)
    K = np.array([[0.82 * W, 0, W / 2 - 0.5], [0, 0.82 * W, H / 2 - 0.5], [0, 0, 1]], np.float32)
    d16 = rng.integers(2500, 17500, (H, W)).astype(np.uint16)          # 0.5 .. 3.5 m at 1/5000 m
    d16[rng.random((H, W)) < 0.25] = 0
    depth = pyoracle.convert_depth_u16(d16, np.float32(1.0 / 5000.0))
    on = pyoracle.OrNormalEstimator(W, H, K, win)
    gn = psg.NormalEstimator(W, H, K, win)
    # (a 1 x 1 window makes the 3 x 3 system singular: Q is inf / NaN there, in both)
    assert np.array_equal(gn.tables()[3:], on.Q, equal_nan=True) and np.array_equal(gn.tables()[2], on.n_sq_inv)
    for a, b in zip(gn.compute(depth), on.compute(depth)):
        assert np.array_equal(a, b, equal_nan=True)


def test_raw_frame_fusion_equals_host_front_end():
    """psg_volume_integrate_raw (u16 depth + u8 colour) against psg_volume_integrate fed by the CPU front end."""
    import psgradientsdf_b200 as psg
    from oracle import pyoracle
    from psgradientsdf_b200 import synth
    sc = scene("SH1")
    F, H, W = sc.depth.shape
    G = 40
    vs, T = synth.grid_for_scene(G)
    unit = np.float32(1.0 / 5000.0)
    d16 = np.round(sc.depth / unit).astype(np.uint16)
    c8 = np.round(sc.color * 255.0).astype(np.uint8)
    on = pyoracle.OrNormalEstimator(W, H, sc.K, 11)
    gn = psg.NormalEstimator(W, H, sc.K, 11)
    va = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64); va.set_zmin(0.5); va.set_zmax(3.5)
    vb = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64); vb.set_zmin(0.5); vb.set_zmax(3.5)
    for f in range(F):
        R, t = sc.poses_gt[f][:3, :3].copy(), sc.poses_gt[f][:3, 3].copy()
        depth = pyoracle.convert_depth_u16(d16[f], unit)
        color = pyoracle.convert_color_u8(c8[f])
        nx, ny, nz = on.compute(depth)
        if f:
            va.increase_counter(); vb.increase_counter()
        va.update(color, depth, sc.K, R, t, nx, ny, nz, on.n_sq_inv)
        vb.update_raw(c8[f], d16[f], unit, sc.K, R, t, gn)
    a, b = va.download(), vb.download()
    assert np.count_nonzero(a[:, 4]) > 1000
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert np.array_equal(va.vis(), vb.vis())


def test_u8_keyframes_equal_float_keyframes():
    import psgradientsdf_b200 as psg
    sc = scene("SH1")
    c8 = np.round(sc.color * 255.0).astype(np.uint8)
    cf = (c8.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)
    out = []
    for use_u8 in (False, True):
        gv, go = make_gpu("SH1", 40, sc)
        fi = np.arange(sc.color.shape[0], dtype=np.int32)
        if use_u8:
            go.set_keyframes_u8(fi, sc.poses_init, c8)
        else:
            go.set_keyframes(fi, sc.poses_init, cf)
        go.init(); go.init_albedo()
        out.append((go.energy(), gv.download()))
    assert out[0][0] == out[1][0]
    assert np.array_equal(out[0][1].view(np.uint32), out[1][1].view(np.uint32))


def test_overlapped_upload_gives_the_same_state():
    """psg_opt_set_upload_overlap: keyframes copied on a second stream while init builds the work list."""
    sc = scene("LED")            # the LED init sweeps the images inside psg_opt_init: the wait must come before it
    c8 = np.round(sc.color * 255.0).astype(np.uint8)
    out = []
    for overlap in (False, True):
        for use_u8 in (False, True):
            gv, go = make_gpu("LED", 40, sc)
            go.set_upload_overlap(overlap)
            fi = np.arange(sc.color.shape[0], dtype=np.int32)
            if use_u8:
                go.set_keyframes_u8(fi, sc.poses_init, c8)
            else:
                go.set_keyframes(fi, sc.poses_init, (c8.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32))
            go.init(); go.init_albedo()
            out.append((go.energy(), go.light().copy()))
    for o in out[1:]:
        assert o[0] == out[0][0] and np.array_equal(o[1], out[0][1])


def test_sharpness_matches_opencv_fixture_and_oracle():
    import psgradientsdf_b200 as psg
    from oracle import pyoracle
    g = np.load(GOLD)
    ne = psg.NormalEstimator(int(g["W"]), int(g["H"]), g["K"], int(g["window"]))
    c = g["color"]
    for img, want in zip((c, np.ascontiguousarray(c[::-1, ::-1]), g["blurred"]), g["sharpness"]):
        assert np.float32(ne.sharpness(img)) == want
    assert np.float32(ne.sharpness(g["color_u8"])) == g["sharpness"][0]          # the 8-bit entry converts like load_color
    rng = np.random.default_rng(5)
    for (H, W) in ((2, 2), (3, 7), (480, 640), (1709, 1139)):
        img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        n2 = psg.NormalEstimator(W, H, g["K"], 3)
        f = (img.astype(np.float32) * np.float32(1.0 / 255.0)).astype(np.float32)
        want = np.float32(pyoracle.modified_laplacian(f))
        assert np.float32(n2.sharpness(img)) == want and np.float32(n2.sharpness(f)) == want
