"""Rigid depth-to-SDF tracking on the device (tracker.cu) against the C restatement and the reference's own tracker."""
import numpy as np
import pytest

from tests.common import make_gpu, make_oracle, make_ref, rel_err, scene, se3

pytestmark = pytest.mark.gpu


def test_tracker_system_vs_oracle():
    model, G = "SH1", 48
    sc = scene(model)
    ov, oo = make_oracle(model, G, sc)
    gv, go = make_gpu(model, G, sc)
    unit = np.float32(1.0 / 5000.0)
    for f, pose in ((2, sc.poses_init[2]), (0, sc.poses_gt[0]), (4, sc.poses_init[5])):
        R, t = se3(pose)
        if f == 4:   # the 16-bit entry point: same depth after the loader's conversion
            d16 = np.round(sc.depth[f] / unit).astype(np.uint16)
            assert np.array_equal(d16.astype(np.float32) * unit, sc.depth[f])
            gv.tracker_set_depth(depth_u16=d16, depth_unit=unit)
        else:
            gv.tracker_set_depth(sc.depth[f])
        for sampling in (1, 2, 3):
            Eo, go_, Ho, co = ov.tracker_system(sc.depth[f], sc.K, R, t, sampling)
            Eg, gg, Hg, cg = gv.tracker_system(sc.K, R, t, sampling)
            assert cg == co and co > 100          # the same pixels hit a fused voxel: index decisions are bit-identical
            # the reference sums in one serial float; the device in double
            assert abs(Eg - Eo) <= 2e-5 * abs(Eo)
            assert rel_err(gg, go_) < 2e-5 and rel_err(Hg, Ho) < 2e-5
            assert np.array_equal(Hg, Hg.T)
        again = gv.tracker_system(sc.K, R, t, 3)
        assert again[0] == Eg and np.array_equal(again[1], gg)      # fixed-order reduction: reproducible


def test_tracker_iterations_vs_reference():
    """Two Gauss-Newton iterations: device sums + the reference's own LLT / SE3 update against RigidPointOptimizer::optimize."""
    from oracle import pyref
    if not pyref.available():
        pytest.skip("oracle/_ref not built")
    model, G = "SH1", 48
    sc = scene(model)
    rv, ro = make_ref(model, G, sc)
    gv, go = make_gpu(model, G, sc)
    f = 2
    gv.tracker_set_depth(sc.depth[f])
    iters = 2
    conv, P_ref = rv.tracker_optimize(sc.depth[f], sc.K, sc.poses_init[f], num_iterations=iters, conv_threshold=1e-12, damping=1.0)
    P = sc.poses_init[f].copy()
    for _ in range(iters):
        R, t = se3(P)
        E, g, H, c = gv.tracker_system(sc.K, R, t, 1)
        P, xi, n2 = pyref.tracker_update(g.astype(np.float32), H.astype(np.float32), 1.0, P)
    assert not conv
    assert np.abs(P - P_ref).max() < 2e-5, np.abs(P - P_ref).max()


def test_tracker_optimize_vs_reference():
    """The whole Gauss-Newton loop (device sums + this library's float LLT / SE(3) update) against the reference's optimize()."""
    from oracle import pyref
    if not pyref.available():
        pytest.skip("oracle/_ref not built")
    model, G = "SH1", 48
    sc = scene(model)
    rv, ro = make_ref(model, G, sc)
    gv, go = make_gpu(model, G, sc)
    for f, iters in ((2, 1), (2, 3), (4, 2)):
        gv.tracker_set_depth(sc.depth[f])
        conv_r, P_ref = rv.tracker_optimize(sc.depth[f], sc.K, sc.poses_init[f], num_iterations=iters, conv_threshold=1e-12, damping=1.0)
        conv_g, it_g, P = gv.tracker_optimize(sc.K, sc.poses_init[f], num_iterations=iters, conv_threshold=1e-12, damping=1.0)
        assert conv_g == conv_r and it_g == iters
        assert np.abs(P - P_ref).max() < 3e-5 * iters, (f, iters, np.abs(P - P_ref).max())
    # a converged call: start at the fixed point of a previous run with a loose threshold
    conv_r, P_ref = rv.tracker_optimize(sc.depth[4], sc.K, sc.poses_gt[4], num_iterations=5, conv_threshold=10.0, damping=1.0)
    conv_g, it_g, P = gv.tracker_optimize(sc.K, sc.poses_gt[4], num_iterations=5, conv_threshold=10.0, damping=1.0)
    assert conv_r and conv_g and it_g == 0 and np.abs(P - P_ref).max() < 1e-6
