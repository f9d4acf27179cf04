"""Synthetic RGB-D keyframe sets for tests and bench.py (plain numpy compute, no graphics API).

Scene (BASELINE.md "Synthetic inputs"): a bumpy sphere of radius ~0.35 m
centred at the world origin with a procedural albedo in [0.2, 0.9]; F cameras on a circular orbit of
radius 1.5 m looking at the centre; pinhole intrinsics scaled from the TUM set (fx=fy=525, cx=319.5,
cy=239.5 at 640x480); depth is sphere-traced, quantised to u16 at 1/5000 m; RGB is shaded with a
ground-truth SH1 / SH2 / LED model and quantised to u8.  Camera-frame normals are analytic (inward
pointing, as the reference's estimator produces, VolumetricGradSdf.cpp:123) -- normal *estimation* is
outside the hot path.

Conventions follow the reference:
  * colour images are float32 BGR = u8 * (1/255 as float32)   (ImageLoader.h:167-188)
  * depth is float32 metres = u16 * (1/5000 as float32)        (TumrgbdLoader.h:34-44, ImageLoader.h:130-146)
  * poses are camera-to-world 4x4                              (main_ps.cpp:219)
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


@dataclass
class Scene:
    H: int
    W: int
    K: np.ndarray                 # 3x3 float32
    poses_gt: np.ndarray          # F x 4 x 4 float32 camera-to-world (ground truth)
    poses_init: np.ndarray        # F x 4 x 4 float32 perturbed (what the optimizer starts from)
    color: np.ndarray             # F x H x W x 3 float32 BGR
    depth: np.ndarray             # F x H x W float32 metres (0 = invalid)
    nx: np.ndarray                # F x H x W float32 camera-frame normals, inward pointing
    ny: np.ndarray
    nz: np.ndarray
    n_sq_inv: np.ndarray          # H x W float32: 1 / (1 + x0^2 + y0^2)   (NormalEstimator.h:77)
    light_gt: np.ndarray          # model dependent
    model: str = "SH1"
    meta: dict = field(default_factory=dict)


def _bump(d: np.ndarray) -> np.ndarray:
    """Low-frequency radial modulation on unit directions d (...,3)."""
    x, y, z = d[..., 0], d[..., 1], d[..., 2]
    return 0.06 * np.sin(3.0 * x + 0.5) * np.cos(2.0 * y - 0.3) + 0.04 * np.sin(4.0 * z + 1.0) * np.cos(3.0 * x)


def sdf(p: np.ndarray, r0: float = 0.35) -> np.ndarray:
    """Approximate signed distance (positive outside) of the bumpy sphere."""
    n = np.linalg.norm(p, axis=-1)
    d = p / np.maximum(n, 1e-9)[..., None]
    return n - r0 * (1.0 + _bump(d))


def sdf_grad(p: np.ndarray, r0: float = 0.35, h: float = 1e-4) -> np.ndarray:
    g = np.empty_like(p)
    for a in range(3):
        e = np.zeros(3)
        e[a] = h
        g[..., a] = (sdf(p + e, r0) - sdf(p - e, r0)) / (2 * h)
    return g / np.maximum(np.linalg.norm(g, axis=-1), 1e-12)[..., None]


def albedo(p: np.ndarray) -> np.ndarray:
    """Procedural 3-D sinusoid albedo, RGB in [0.2, 0.9]."""
    x, y, z = p[..., 0], p[..., 1], p[..., 2]
    r = 0.55 + 0.35 * np.sin(9.0 * x + 1.0) * np.cos(7.0 * y)
    g = 0.55 + 0.35 * np.sin(8.0 * y - 0.5) * np.cos(6.0 * z)
    b = 0.55 + 0.35 * np.sin(7.0 * z + 0.3) * np.cos(9.0 * x)
    return np.stack([r, g, b], axis=-1)


def sh_basis(n: np.ndarray, order: int) -> np.ndarray:
    """SH basis in the reference's ordering (PsOptimizerJa.cpp:17-28)."""
    one = np.ones_like(n[..., 0])
    if order == 1:
        return np.stack([one, n[..., 0], n[..., 1], n[..., 2]], axis=-1)
    x, y, z = n[..., 0], n[..., 1], n[..., 2]
    return np.stack([one, x, y, z, x * y, x * z, y * z, x * x - y * y, x * x - z * z], axis=-1)


def look_at_pose(cam_pos: np.ndarray, target: np.ndarray = np.zeros(3)) -> np.ndarray:
    """Camera-to-world with +z forward, +x right, +y down."""
    fwd = target - cam_pos
    fwd = fwd / np.linalg.norm(fwd)
    up = np.array([0.0, 0.0, 1.0])
    right = np.cross(fwd, up)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    T = np.eye(4)
    T[:3, 0], T[:3, 1], T[:3, 2], T[:3, 3] = right, down, fwd, cam_pos
    return T


def _rodrigues(w: np.ndarray) -> np.ndarray:
    th = np.linalg.norm(w)
    if th < 1e-12:
        return np.eye(3)
    k = w / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * Kx + (1 - np.cos(th)) * Kx @ Kx


def make_scene(F: int = 8, H: int = 120, W: int = 160, model: str = "SH1", seed: int = 1234,
               radius: float = 1.5, r0: float = 0.35, elev_amp: float = 0.35,
               pose_sigma_t: float = 0.002, pose_sigma_r_deg: float = 0.2) -> Scene:
    rng = np.random.default_rng(seed)            # lights
    rng_pose = np.random.default_rng(seed + 1)   # pose perturbations (same for every model)
    sx, sy = W / 640.0, H / 480.0
    fx, fy = 525.0 * sx, 525.0 * sy
    cx, cy = (319.5 + 0.5) * sx - 0.5, (239.5 + 0.5) * sy - 0.5
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=np.float32)
    fx, fy, cx, cy = (float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2]))

    u, v = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    x0 = (u - cx) / fx
    y0 = (v - cy) / fy
    n_sq_inv = (1.0 / (1.0 + x0 * x0 + y0 * y0)).astype(np.float32)
    rays_cam = np.stack([x0, y0, np.ones_like(x0)], axis=-1)           # z = 1 parametrisation

    order = 2 if model == "SH2" else 1
    poses_gt = np.zeros((F, 4, 4), dtype=np.float32)
    poses_init = np.zeros((F, 4, 4), dtype=np.float32)
    color = np.zeros((F, H, W, 3), dtype=np.float32)
    depth = np.zeros((F, H, W), dtype=np.float32)
    nxs = np.zeros((F, H, W), dtype=np.float32)
    nys = np.zeros((F, H, W), dtype=np.float32)
    nzs = np.zeros((F, H, W), dtype=np.float32)
    if model == "LED":
        light_gt = np.array([1.1, 1.0, 0.9], dtype=np.float32)
    else:
        light_gt = np.zeros((F, 4 if order == 1 else 9), dtype=np.float32)

    inv255 = np.float32(1.0) / np.float32(255.0)
    unit = np.float32(1.0) / np.float32(5000.0)
    for f in range(F):
        ang = 2.0 * np.pi * f / F
        cam = np.array([radius * np.cos(ang), radius * np.sin(ang), elev_amp * np.sin(2.0 * ang + 0.4)])
        cam *= radius / np.linalg.norm(cam)
        T = look_at_pose(cam)
        poses_gt[f] = T.astype(np.float32)
        T = poses_gt[f].astype(np.float64)                              # what the optimizer could at best recover
        R, t = T[:3, :3], T[:3, 3]
        dirs = rays_cam @ R.T                                           # world, |.| >= 1, camera z advances by s
        dn = np.linalg.norm(dirs, axis=-1)
        # only rays that pass within the bounding sphere can hit; trace those as flat arrays
        rb = r0 * 1.12
        tca = -(dirs @ t) / (dn * dn)                                   # parameter of closest approach
        d2 = np.sum((t + dirs * tca[..., None]) ** 2, axis=-1)
        cand = d2 < rb * rb
        ci = np.nonzero(cand.ravel())[0]
        dc = dirs.reshape(-1, 3)[ci]
        dnc = dn.ravel()[ci]
        sc_ = tca.ravel()[ci] - np.sqrt(np.maximum(rb * rb - d2.ravel()[ci], 0.0)) / dnc   # entry into the bounding sphere
        send = tca.ravel()[ci] + np.sqrt(np.maximum(rb * rb - d2.ravel()[ci], 0.0)) / dnc
        alive = np.ones(ci.size, dtype=bool)
        for _ in range(64):
            pc = t + dc * sc_[:, None]
            d = sdf(pc, r0)
            sc_ = np.where(alive, sc_ + 0.8 * d / dnc, sc_)
            alive &= sc_ < send
        pc = t + dc * sc_[:, None]
        hitc = alive & (np.abs(sdf(pc, r0)) < 2e-4)
        s = np.zeros(H * W)
        s[ci] = np.where(hitc, sc_, 0.0)
        s = s.reshape(H, W)
        hit = np.zeros(H * W, dtype=bool)
        hit[ci] = hitc
        hit = hit.reshape(H, W)
        p = t + dirs * s[..., None]
        z16 = np.where(hit, np.clip(np.rint(s * 5000.0), 0, 65535), 0).astype(np.uint16)
        depth[f] = z16.astype(np.float32) * unit
        n_w = np.zeros((H * W, 3))
        n_w[ci] = sdf_grad(pc, r0)                                      # outward, world
        n_w = n_w.reshape(H, W, 3)
        n_c = -(n_w @ R)                                                # inward, camera frame
        nxs[f] = np.where(hit, n_c[..., 0], 0).astype(np.float32)
        nys[f] = np.where(hit, n_c[..., 1], 0).astype(np.float32)
        nzs[f] = np.where(hit, n_c[..., 2], 0).astype(np.float32)
        rho = albedo(p)
        if model == "LED":
            p_c = rays_cam * s[..., None]
            irr = -np.sum((n_w @ R) * p_c, axis=-1) / np.maximum(np.linalg.norm(p_c, axis=-1), 1e-9) ** 3
            # scaled so that intensities land in a sensible range at ~1.2 m
            rgb = rho * (light_gt.astype(np.float64) * 1.3)[None, None, :] * irr[..., None]
        else:
            to_cam = R @ np.array([0.0, 0.0, -1.0])
            l = np.zeros(4 if order == 1 else 9)
            l[0] = 0.35 + 0.05 * np.sin(ang)
            l[1:4] = 0.6 * to_cam + 0.1 * rng.standard_normal(3)
            if order == 2:
                l[4:] = 0.08 * rng.standard_normal(5)
            light_gt[f] = l.astype(np.float32)
            irr = sh_basis(n_w, order) @ light_gt[f].astype(np.float64)
            rgb = rho * irr[..., None]
        rgb8 = np.where(hit[..., None], np.clip(np.rint(rgb * 255.0), 0, 255), 0).astype(np.uint8)
        color[f] = rgb8[..., ::-1].astype(np.float32) * inv255          # BGR
        # perturbed initial pose
        dT = np.eye(4)
        if f > 0:
            dT[:3, :3] = _rodrigues(np.deg2rad(pose_sigma_r_deg) * rng_pose.standard_normal(3))
            dT[:3, 3] = pose_sigma_t * rng_pose.standard_normal(3)
        poses_init[f] = (T @ dT).astype(np.float32)
    if model == "LED":
        light_gt = (light_gt * np.float32(1.3)).astype(np.float32)
    return Scene(H=H, W=W, K=K, poses_gt=poses_gt, poses_init=poses_init, color=color, depth=depth,
                 nx=nxs, ny=nys, nz=nzs, n_sq_inv=n_sq_inv, light_gt=light_gt, model=model,
                 meta={"seed": seed, "F": F, "radius": radius, "r0": r0})


def grid_for_scene(G: int, r0: float = 0.35, fill: float = 0.6) -> tuple[float, float]:
    """voxel size so that the object (diameter 2*r0*(1+bump)) fills `fill` of a G-cube; T = 5 * vs."""
    vs = np.float32(2.0 * r0 * 1.08 / fill / G)
    return float(vs), float(np.float32(5.0) * vs)
