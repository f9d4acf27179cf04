"""The normal-estimator / pixel-conversion restatement (oracle/psg_oracle_normals.c) against OpenCV's own outputs for the
reference's call sequence (tests/golden/normals.npz, written by tests/golden/make_golden_normals.py with cv2 4.13)."""
import os

import numpy as np

from oracle import pyoracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "normals.npz")


def test_tables_match_opencv():
    g = np.load(GOLD)
    ne = pyoracle.OrNormalEstimator(int(g["W"]), int(g["H"]), g["K"], int(g["window"]))
    for name in ("x0", "y0", "x0_n", "y0_n", "n_sq_inv"):
        assert np.array_equal(getattr(ne, name), g[name]), name
    assert np.array_equal(ne.Q, g["Q"])


def test_conversions_and_normals_match_opencv():
    g = np.load(GOLD)
    depth = pyoracle.convert_depth_u16(g["depth_u16"], float(g["depth_unit"]))
    assert np.array_equal(depth, g["depth"])
    assert np.array_equal(pyoracle.convert_color_u8(g["color_u8"]), g["color"])
    ne = pyoracle.OrNormalEstimator(int(g["W"]), int(g["H"]), g["K"], int(g["window"]))
    nx, ny, nz = ne.compute(depth)
    for a, name in ((nx, "nx"), (ny, "ny"), (nz, "nz")):
        assert np.array_equal(a, g[name], equal_nan=True), name
    assert int(np.isnan(g["nx"]).sum()) > 0          # the fixture does contain empty windows (0 / 0)


def test_window_sizes_and_tiny_images():
    """Reflect-101 borders when the image is barely larger than the window, and other odd window sizes."""
    rng = np.random.default_rng(3)
    K = np.array([[40.0, 0, 5.5], [0, 42.0, 4.0], [0, 0, 1]], np.float32)
    for (H, W, win) in ((7, 9, 5), (12, 12, 11), (6, 20, 3), (5, 5, 1)):
        depth = (0.5 + rng.random((H, W))).astype(np.float32)
        depth[rng.random((H, W)) < 0.2] = 0
        ne = pyoracle.OrNormalEstimator(W, H, K, win)
        nx, ny, nz = ne.compute(depth)
        ok = np.isfinite(nx)
        assert np.allclose((nx * nx + ny * ny + nz * nz)[ok], 1.0, atol=1e-5)


def test_sharpness_matches_opencv():
    """modifiedLaplacian (SharpDetector.h:22-37) restated against cv2.sepFilter2D + cv2.mean on three images."""
    g = np.load(GOLD)
    c = g["color"]
    got = [pyoracle.modified_laplacian(x) for x in (c, np.ascontiguousarray(c[::-1, ::-1]), g["blurred"])]
    assert [np.float32(v) for v in got] == list(g["sharpness"])
    assert got[2] < 0.2 * got[0]                     # the blurred image is what the gate is meant to reject


def test_centroid_and_keyframe_sampling_vs_reference_and_library():
    """compute_centroid and sampleKeyFrame are host code in the library (no GPU): restatement, library and -- where the
    reference build is present -- the re-played reference statements agree bit for bit."""
    import psgradientsdf_b200 as psg
    from oracle import pyref
    g = np.load(GOLD)
    P = np.eye(4, dtype=np.float32)
    P[:3, :3] = np.array([[0.936, 0.18, 0.302], [-0.1, 0.96, -0.262], [-0.337, 0.215, 0.917]], np.float32)
    P[:3, 3] = [0.1, -0.2, 0.3]
    a = pyoracle.compute_centroid(g["K"], g["depth"], P)
    b = psg.compute_centroid(g["K"], g["depth"], P)
    assert np.array_equal(a, b)
    if pyref.available():
        assert np.array_equal(a, pyref.compute_centroid(g["K"], g["depth"], P))
    for n in (0, 1, 5, 39, 40, 41, 64, 100, 1000):
        o = pyoracle.sample_keyframes(n, 40)
        l = psg.sample_keyframes(n, 40)
        assert np.array_equal(o, l), n
        assert len(l) == (n if n < 40 else 40) and (n == 0 or l[-1] == n - 1) and np.all(np.diff(l) >= 0)
