"""Shared helpers for the parity tests: one small synthetic scene per model, and drivers that run the same call
sequence (main_ps.cpp:183-330) on any of the three implementations (reference harness, C oracle, CUDA path)."""
from __future__ import annotations

import functools

import numpy as np

from psgradientsdf_b200 import synth

MODELS = ("SH1", "SH2", "LED")


@functools.lru_cache(maxsize=None)
def scene(model: str, F: int = 6, H: int = 120, W: int = 160):
    return synth.make_scene(F=F, H=H, W=W, model=model)


def reg_for(model: str):
    # config_skorates.json: reg norm 10; config_basket_LED.json additionally uses a Laplacian weight
    return dict(reg_n=10.0, reg_l=5.0 if model == "LED" else 0.0)


def block_order(model: str):
    return ["light", "albedo", "dist", "pose"] if model == "LED" else ["albedo", "light", "dist", "pose"]


def se3(pose):
    from oracle import pyref
    if pyref.available():
        return pyref.se3_from_matrix(pose)
    return pose[:3, :3].copy(), pose[:3, 3].copy()


def make_ref(model, G, sc):
    from oracle import pyref
    vs, T = synth.grid_for_scene(G)
    vol = pyref.RefVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5)
    opt = pyref.RefOptimizer(vol, model, vs, sc.K, "/tmp/", **reg_for(model))   # created before the first update, as main does
    for f in range(sc.color.shape[0]):
        vol.integrate(sc.color[f], sc.depth[f], sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv, sc.K, sc.poses_gt[f], f > 0)
    return vol, opt


def make_oracle(model, G, sc):
    from oracle import pyoracle
    vs, T = synth.grid_for_scene(G)
    vol = pyoracle.OrVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5, 64)
    opt = pyoracle.OrOptimizer(vol, model, vs, sc.K, **reg_for(model))
    for f in range(sc.color.shape[0]):
        R, t = se3(sc.poses_gt[f])
        vol.integrate(sc.color[f], sc.depth[f], sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv, sc.K, R, t, f > 0)
    return vol, opt


def make_gpu(model, G, sc, max_frames=64):
    import psgradientsdf_b200 as psg
    vs, T = synth.grid_for_scene(G)
    vol = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, max_frames)
    vol.set_zmin(0.5); vol.set_zmax(3.5)
    opt = psg.Optimizer(vol, vs, sc.K, model=model, **reg_for(model))
    for f in range(sc.color.shape[0]):
        R, t = se3(sc.poses_gt[f])
        if f > 0:
            vol.increase_counter()
        vol.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
    return vol, opt


def handoff(opt, sc, keys=None):
    F = sc.color.shape[0]
    keys = np.arange(F, dtype=np.int32) if keys is None else np.asarray(keys, np.int32)
    opt.set_keyframes(keys, sc.poses_init[keys], sc.color[keys])
    opt.init()


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def assert_bit_equal(a, b, what):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, f"{what}: shape {a.shape} vs {b.shape}"
    if not np.array_equal(bits(a), bits(b)):
        bad = np.flatnonzero(bits(a).ravel() != bits(b).ravel())
        raise AssertionError(f"{what}: {bad.size} of {a.size} elements differ, first at {bad[:5]}: "
                             f"{a.ravel()[bad[:5]]} vs {b.ravel()[bad[:5]]}")


def rel_err(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(a).max(), 1e-30)) if a.size else 0.0
