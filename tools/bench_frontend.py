#!/usr/bin/env python
"""Per-frame cost of the fusion front end on one GPU, with the CPU restatement timed beside it.

    python tools/bench_frontend.py [--grid 128] [--frames 30]

Prints one JSON line: ms per frame for (a) psg_volume_integrate_raw -- 16-bit depth + 8-bit colour from pinned host memory,
conversions + FALS normals + fusion on the device, (b) psg_volume_integrate fed with float planes computed on the host,
the device time of the kernels involved (launch profiler), algorithmic bytes and the fraction of the measured HBM peak, and
the device marching cubes of the fused volume (next-4) with the CPU restatement's time beside it, and
the CPU restatement's time for the normal estimation alone (oracle/psg_oracle_normals.c, 1 thread -- what the reference
spends in OpenCV per frame before it can call update()).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=128)
    ap.add_argument("--frames", type=int, default=30)
    a = ap.parse_args()
    import torch

    import psgradientsdf_b200 as psg
    from oracle import pyoracle
    from psgradientsdf_b200 import synth

    sc = synth.make_scene(F=a.frames, H=480, W=640, model="SH1")
    F, H, W = sc.depth.shape
    G = a.grid
    vs, T = synth.grid_for_scene(G)
    unit = np.float32(1.0 / 5000.0)
    d16 = torch.from_numpy(np.round(sc.depth / unit).astype(np.uint16).view(np.int16)).pin_memory().numpy().view(np.uint16)
    c8 = torch.from_numpy(np.round(sc.color * 255.0).astype(np.uint8)).pin_memory().numpy()
    Rs = [sc.poses_gt[f][:3, :3].copy() for f in range(F)]
    ts = [sc.poses_gt[f][:3, 3].copy() for f in range(F)]

    # CPU front end (reference-equivalent arithmetic), timed per frame
    on = pyoracle.OrNormalEstimator(W, H, sc.K, 11)
    t0 = time.perf_counter()
    host = []
    for f in range(F):
        depth = pyoracle.convert_depth_u16(d16[f], unit)
        color = pyoracle.convert_color_u8(c8[f])
        nx, ny, nz = on.compute(depth)
        host.append((color, depth, nx, ny, nz))
    cpu_ms = 1e3 * (time.perf_counter() - t0) / F

    gn = psg.NormalEstimator(W, H, sc.K, 11)

    def run(raw: bool):
        v = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, max(64, F)); v.set_zmin(0.5); v.set_zmax(3.5)
        v.update_raw(c8[0], d16[0], unit, sc.K, Rs[0], ts[0], gn) if raw else v.update(*host[0][:2], sc.K, Rs[0], ts[0], *host[0][2:], on.n_sq_inv)
        torch.cuda.synchronize()
        psg.profile_reset(); psg.profile_enable(True)
        t0 = time.perf_counter()
        for f in range(1, F):
            v.increase_counter()
            if raw:
                v.update_raw(c8[f], d16[f], unit, sc.K, Rs[f], ts[f], gn)
            else:
                color, depth, nx, ny, nz = host[f]
                v.update(color, depth, sc.K, Rs[f], ts[f], nx, ny, nz, on.n_sq_inv)
        torch.cuda.synchronize()
        ms = 1e3 * (time.perf_counter() - t0) / (F - 1)
        tab = psg.profile_table(); psg.profile_enable(False)
        touched = int(np.count_nonzero(v.download()[:, 4]))
        v.close()
        return ms, {k: round(1e3 * t / n, 2) for k, (n, t) in tab.items()}, touched

    raw_ms, raw_k, touched = run(True)
    flt_ms, flt_k, _ = run(False)
    # rigid tracker: one Gauss-Newton pass over the full-resolution depth image (next-2)
    vt = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, max(64, F)); vt.set_zmin(0.5); vt.set_zmax(3.5)
    for f in range(F):
        if f:
            vt.increase_counter()
        vt.update_raw(c8[f], d16[f], unit, sc.K, Rs[f], ts[f], gn)
    vt.tracker_set_depth(depth_u16=d16[1], depth_unit=unit)
    vt.tracker_system(sc.K, Rs[1], ts[1], 1)
    torch.cuda.synchronize()
    psg.profile_reset(); psg.profile_enable(True)
    t0 = time.perf_counter()
    for _ in range(20):
        Et, gt, Ht, ct = vt.tracker_system(sc.K, Rs[1], ts[1], 1)
    trk_ms = 1e3 * (time.perf_counter() - t0) / 20
    tabt = psg.profile_table(); psg.profile_enable(False)
    trk_cpu = None
    if G <= 256:   # the CPU restatement needs the whole volume on the host
        ov = pyoracle.OrVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5, max(64, F))
        ov.upload(vt.download())
        t0 = time.perf_counter()
        Eo, go_, Ho, co = ov.tracker_system(host[1][1], sc.K, Rs[1], ts[1], 1)
        trk_cpu = 1e3 * (time.perf_counter() - t0)
        assert co == ct
    # surface extraction (next-4): crop box + marching cubes on the device, triangles to the host; the CPU restatement beside it
    torch.cuda.synchronize()
    vt.mesh()                                    # warm-up (allocations)
    psg.profile_reset(); psg.profile_enable(True)
    t0 = time.perf_counter()
    for _ in range(10):
        mv, mc_ = vt.mesh()
    mesh_ms = 1e3 * (time.perf_counter() - t0) / 10
    tabm = psg.profile_table(); psg.profile_enable(False)
    box = vt.mesh_box()
    box_voxels = int(np.prod(box[1] - box[0] + 1)) if box is not None else 0
    t0 = time.perf_counter()
    vt.extract_mesh("/tmp/psg_bench_mesh.ply")
    mesh_file_ms = 1e3 * (time.perf_counter() - t0)
    mesh_cpu = None
    if G <= 256:
        aos = vt.download()
        t0 = time.perf_counter()
        mo = pyoracle.extract_mesh(aos, [G, G, G], vs)
        mesh_cpu = 1e3 * (time.perf_counter() - t0)
        assert np.array_equal(mo[0], mv) and np.array_equal(mo[1], mc_)
    vt.close()
    peak = 6555.8
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = float(json.load(open(p))["hbm_gbs"])
    N = G ** 3
    fuse_us = raw_k.get("k_integrate", 0.0)
    # 72 B per touched voxel; every voxel is at least projected, so the per-launch figure uses the touched count of
    # the final volume as an upper bound of a single frame's
    out = {"workload": f"{F} frames 640x480 into a {G}^3 volume", "ms_per_frame_raw_u16_u8": round(raw_ms, 3),
           "ms_per_frame_host_float_planes": round(flt_ms, 3), "cpu_front_end_ms_per_frame": round(cpu_ms, 2),
           "h2d_bytes_per_frame_raw": int(H * W * 5), "h2d_bytes_per_frame_float": int(H * W * 32),
           "kernel_us_raw": raw_k, "kernel_us_float": flt_k, "voxels": N, "voxels_touched_by_any_frame": touched,
           "tracker_pass_ms": round(trk_ms, 4), "tracker_kernel_us": {k: round(1e3 * t / n, 2) for k, (n, t) in tabt.items()},
           "tracker_pixels_hit": int(ct), "tracker_cpu_pass_ms": None if trk_cpu is None else round(trk_cpu, 2),
           "mesh": {"triangles": int(len(mv) // 3), "crop_box_voxels": box_voxels, "ms_device_to_host_arrays": round(mesh_ms, 3),
                    "kernel_us_per_extraction": round(1e3 * sum(t for _, t in tabm.values()) / 10, 1),
                    "kernel_us": {k: round(1e3 * t / 10, 1) for k, (n, t) in tabm.items()},
                    "ms_with_ascii_ply_file": round(mesh_file_ms, 1), "cpu_marching_cubes_ms": None if mesh_cpu is None else round(mesh_cpu, 1),
                    # both passes read dist + weight of every cell corner once (8 B per voxel of the crop box, neighbours from L1 / L2)
                    "algorithmic_bytes": 2 * 8 * box_voxels + 45 * int(len(mv) // 3)},
           "k_integrate": {"algorithmic_bytes_upper_bound": 72 * touched, "achieved_GBps": round(72 * touched / (fuse_us * 1e-6) / 1e9, 1) if fuse_us else None,
                           "peak_GBps": peak}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
