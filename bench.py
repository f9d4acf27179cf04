#!/usr/bin/env python
"""bench.py -- photometric-optimizer hot path on synthetic RGB-D keyframes (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one process per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU code (oracle/_ref), rank 0 only

A *step* is one outer ("LM") iteration of the alternating optimizer over the whole keyframe set
(PsOptimizer.cpp:303-360): albedo, light, distance and pose blocks, each followed by the energy evaluation the
reference logs.  The reference spends 8 sweeps over all active (surface voxel, keyframe) pairs on that; here every
block's sweep also returns the energy of the state it started from, so 5 sweeps give the same 4 updates and 4 logged
energies, plus the distance PCG and the small host solves.  metric = active pair evaluations per second (pairs x sweeps executed / time);
ms_per_step = ms per LM iteration.

`value`      : inputs resident in HBM when the timed region starts (volume fused, keyframes uploaded, surface list built).
`e2e`        : the same iteration driven from HOST buffers through the C ABI: every step re-sends the keyframe images and
               poses from pinned host memory (psg_opt_set_keyframes), rebuilds the surface list (psg_opt_init), runs
               initAlbedo + one LM iteration and reads energies, poses and light back.
`roofline`   : dominant kernel by device time, timed live with CUDA events on the launching stream (psg_profile_*).
`cpu_baseline`: the reference's getPSEnergy() on this box's host (1 thread: the reference is single-threaded), on a
               bounded sample of the same workload.
At N > 1 (--multi): "shard" splits ONE workload over the ranks (strong scaling): every rank holds the volume and the work list,
the sweeps are split round-robin over 32-slot blocks of the visibility-sorted work list, and the library combines the results
with NCCL (all-reduce of energies and of the light / pose normal equations, all-gather of albedo and of the distance stencil
blocks, replicated distance solve) -- see DESIGN.md section 7.  "replicas" gives every rank its own keyframe set (weak scaling,
no data-path collective; value = pairs of all ranks / slowest rank's time).  "auto" (default) applies the library's own rule:
sharding one workload pays from 2e6 visible pairs per rank; the default 128^3 x 30 workload has 0.26e6 in total, so N GPUs
serve N keyframe sets, while --grid 1024 --frames 64 is sharded.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "active voxel-frame residual+Jacobian evaluations per second (128^3 SH1 LM iteration)"
UNIT = "evals/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--grid", type=int, default=128)
    ap.add_argument("--frames", type=int, default=30)
    ap.add_argument("--model", default="SH1", choices=["SH1", "SH2", "LED"])
    ap.add_argument("--height", type=int, default=480)
    ap.add_argument("--width", type=int, default=640)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sweep-split", type=int, default=0, choices=[0, 1, 2, 4, 8], help="psg_opt_set_sweep_split (0 = chosen per launch)")
    ap.add_argument("--pcg-mode", type=int, default=0, choices=[0, 1, 2, 3, 4, 5], help="psg_opt_set_pcg_mode (0 = default)")
    ap.add_argument("--multi", default="auto", choices=["auto", "shard", "replicas"],
                    help="N > 1: shard ONE workload over the ranks (strong scaling), run one independent keyframe set per rank "
                         "(weak scaling), or choose by the library's sharding threshold (2e6 visible pairs per rank)")
    ap.add_argument("--ref-keyframes", type=int, default=30,
                    help="keyframes in the reference-arm sample (the reference is O(S^2 F): 30 keyframes at 128^3 = ~3 s per sweep)")
    return ap.parse_args()


def workload_name(a):
    return (f"synthetic TUM-style RGB-D orbit, {a.model}, {a.grid}^3 grid, {a.frames} keyframes {a.width}x{a.height}, "
            f"light+albedo+distance+pose")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def build_scene(a, seed=1234):
    from psgradientsdf_b200 import synth
    sc = synth.make_scene(F=a.frames, H=a.height, W=a.width, model=a.model, seed=seed)
    vs, T = synth.grid_for_scene(a.grid)
    return sc, vs, T


def se3_parts(pose):
    """R, t the reference's SE3(pose) hands to update(); orthonormal input poses make this the identity up to rounding,
    and the same R, t go to every implementation."""
    from oracle import pyref
    if pyref.available():
        return pyref.se3_from_matrix(pose)
    return pose[:3, :3].copy(), pose[:3, 3].copy()


# ------------------------------------------------------------------------------------------------ reference arm
def reference_sample(a, sc, vs, T, nkey):
    """The reference's own optimizer on the same fused volume with `nkey` of the keyframes."""
    from oracle import pyref
    G = a.grid
    rv = pyref.RefVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5)
    reg_l = 5.0 if a.model == "LED" else 0.0
    ro = pyref.RefOptimizer(rv, a.model, vs, sc.K, "/tmp/", reg_n=10.0, reg_l=reg_l)
    for f in range(a.frames):
        rv.integrate(sc.color[f], sc.depth[f], sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv, sc.K, sc.poses_gt[f], f > 0)
    # the reference's membership test is a linear scan: a sweep costs ~3 s x (G/128)^4 x nkey/30 on one core;
    # keep the sample around 20 s on larger grids
    nkey = max(1, min(nkey, a.frames, int(200.0 / (G / 128.0) ** 4)))
    keys = np.unique(np.linspace(0, a.frames - 1, nkey).round().astype(np.int32))
    ro.set_keyframes(keys, sc.poses_init[keys], sc.color[keys])
    ro.init()
    ro.init_albedo()
    _, _, A = ro.energy()
    return rv, ro, int(A), keys


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyref
    if not pyref.available():
        emit({"impl": "reference", "unavailable": "oracle/_ref/libpsgref.so missing (reference sources were not built)"})
        return
    sc, vs, T = build_scene(a)
    rv, ro, A, keys = reference_sample(a, sc, vs, T, a.ref_keyframes)
    for _ in range(a.warmup):
        ro.ps_energy()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        ro.ps_energy()
    dt = time.perf_counter() - t0
    val = A * a.steps / dt
    sample = (f"getPSEnergy() (PsOptimizer.cpp:47-78) over the full fused {a.grid}^3 volume with {len(keys)} of the "
              f"{a.frames} keyframes: S={ro.S} surface voxels, {A} active pairs per step")
    out = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
           "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "impl": "reference",
           "config": {"workload": workload_name(a), "sample": sample},
           "cpu_baseline": {"value": val, "unit": UNIT, "cores": 1, "kind": "reference", "sample": sample,
                            "host_cores": os.cpu_count()},
           "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    emit(out)


# ------------------------------------------------------------------------------------------------ our arm
_ITER = {}


def lm_iteration(go, model, damping=1.0):
    """One pass of the reference's while-body (PsOptimizer.cpp:303-360 / LedOptimizer.cpp:343-403): psg_opt_iteration on structs
    that are built once, as a C++ host's loop holds them."""
    key = (id(go), model, damping)
    if key not in _ITER:
        _ITER[key] = go.iteration_config(True, True, True, True, damping=damping, normal_reg=True, laplacian_reg=(model == "LED"))
    cfg, info = _ITER[key]
    go.iteration_raw(cfg, info)
    return info.energy_end, info.n_active_end   # the energies the reference logs after each block are in info.block[...]


def sweep_bytes(A, S, F, out_block):
    """Algorithmic bytes of one sweep (SURVEY.md 8d / BASELINE.md): 48 A + 80 S + out_block S + 128 F."""
    return 48 * A + 80 * S + out_block * S + 128 * F


def ncu_traffic(a, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed `ncu --set full` capture
    (profiles/r1_traffic.json); None when the capture was taken on another workload."""
    p = os.path.join(ROOT, "profiles", "r1_traffic.json")
    try:
        with open(p) as f:
            t = json.load(f)
    except OSError:
        return None
    w = t.get("workload", {})
    if (w.get("grid"), w.get("frames"), w.get("model"), w.get("height"), w.get("width")) != (a.grid, a.frames, a.model, a.height, a.width):
        return None
    k = t.get("kernels", {}).get(kernel)
    return None if k is None else k["dram_read_bytes"] + k["dram_write_bytes"]


def run_ours(a):
    import torch
    import torch.distributed as dist

    import psgradientsdf_b200 as psg

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- this path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    G, F = a.grid, a.frames
    reg_l = 5.0 if a.model == "LED" else 0.0
    fi = np.arange(F, dtype=np.int32)
    SHARD_MIN_PAIRS = 2_000_000       # kShardMinPairsPerRank (psgradientsdf_b200/csrc/comm.cu)

    def make_job(seed):
        """One keyframe set fused into its volume, optimizer created, per-step inputs in pinned host memory."""
        sc, vs, T = build_scene(a, seed)
        gv = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, max(64, F), device=local)
        gv.set_zmin(0.5); gv.set_zmax(3.5)
        go = psg.Optimizer(gv, vs, sc.K, model=a.model, reg_n=10.0, reg_l=reg_l)
        for f in range(F):
            R, t = se3_parts(sc.poses_gt[f])
            if f > 0:
                gv.increase_counter()
            gv.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
        img_pin = torch.from_numpy(np.ascontiguousarray(sc.color)).pin_memory()
        pose_pin = torch.from_numpy(np.ascontiguousarray(sc.poses_init)).pin_memory()
        return sc, vs, T, gv, go, img_pin, pose_pin

    # N > 1: one workload sharded over the ranks, or one independent keyframe set per rank.  "auto" asks the library's own rule:
    # sharding pays from 2e6 visible (surface voxel, keyframe) pairs per rank; rank 0 builds the standard
    # workload, counts, and tells the others, which then build the standard workload (shard) or their own (replicas).
    mode = "single" if world == 1 else a.multi
    job = None
    if world > 1 and mode == "auto":
        verdict = [None]
        if rank == 0:
            job = make_job(1234)
            job[4].set_keyframes(fi, job[6].numpy(), job[5].numpy()); job[4].init()
            verdict[0] = "shard" if job[4].is_sharded()[1] // world >= SHARD_MIN_PAIRS else "replicas"
        dist.broadcast_object_list(verdict, src=0)
        mode = verdict[0]
    seed = 1234 + (17 * rank if mode == "replicas" else 0)
    if job is None:
        job = make_job(seed)
    sc, vs, T, gv, go, img_pin, pose_pin = job
    if a.pcg_mode:
        go.set_pcg_mode(a.pcg_mode)
    if a.sweep_split:
        go.set_sweep_split(a.sweep_split)
    images_np, poses_np = img_pin.numpy(), pose_pin.numpy()

    def handoff():
        go.set_keyframes(fi, poses_np, images_np)
        go.init()
        go.init_albedo()

    if mode == "shard":
        ids = [psg.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        go.comm_init(ids[0], rank, world)
        go.set_shard_policy(1)
    handoff()
    S = go.S
    E0, A = go.energy()
    # units of one step over the whole job: the shared work list when sharded, the sum of the ranks' own lists otherwise
    A_job, S_job = A, S
    vis_pairs = go.is_sharded()[1]
    if mode == "replicas":
        tot = torch.tensor([A, S], dtype=torch.float64, device=f"cuda:{local}")
        dist.all_reduce(tot)
        A_job, S_job = int(tot[0].item()), int(tot[1].item())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")   # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sweeps_per_step = 5   # 4 block sweeps (each also yields the energy of the state it starts from) + 1 energy sweep
    # ---- device-resident leg
    for _ in range(a.warmup):
        lm_iteration(go, a.model)
    sampler = ClockSampler(local); sampler.start()
    launches0 = psg.kernel_launches()
    barrier()
    psg.host_times(reset=True)
    t_total = 0.0
    E = None
    for _ in range(a.steps):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        E = lm_iteration(go, a.model)
        torch.cuda.synchronize()
        t_total += time.perf_counter() - t0
    host_us = psg.host_times(reset=True)
    barrier()
    launches = psg.kernel_launches() - launches0
    clocks = sampler.stop()
    # same K steps once more with the library's CUDA-event launch timer on (kept out of the timed region above: two event
    # records per launch would be charged to the step)
    psg.profile_reset(); psg.profile_enable(True)
    for _ in range(a.steps):
        flush.zero_()
        torch.cuda.synchronize()
        lm_iteration(go, a.model)
    torch.cuda.synchronize()
    prof = psg.profile_table()
    psg.profile_enable(False)
    tt = torch.tensor([t_total], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_max = float(tt.item())
    value = A_job * sweeps_per_step * a.steps / t_max      # whole job: the shared list once, or the sum of the ranks' own lists

    # ---- e2e leg: host buffers -> C ABI every step (keyframe upload overlapped with the work-list build: the pinned image
    # buffers stay valid until psg_opt_init returns, as psg_opt_set_upload_overlap requires)
    go.set_upload_overlap(True)
    h2d = images_np.nbytes + poses_np.nbytes
    d2h = poses_np.nbytes + 4 * go.light().size + 8 * 6
    for _ in range(min(a.warmup, 2)):
        handoff(); lm_iteration(go, a.model)
    barrier()
    te = 0.0
    for _ in range(a.steps):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        handoff()
        lm_iteration(go, a.model)
        go.poses(); go.light()
        torch.cuda.synchronize()
        te += time.perf_counter() - t0
    barrier()
    tte = torch.tensor([te], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(tte, op=dist.ReduceOp.MAX)
    e2e_value = A_job * (sweeps_per_step + 1) * a.steps / float(tte.item())   # + the initAlbedo sweep

    # same hand-off with the keyframes as the loader read them (8-bit BGR, psg_opt_set_keyframes_u8): reported beside e2e
    c8_pin = torch.from_numpy(np.ascontiguousarray(np.round(sc.color * 255.0).astype(np.uint8))).pin_memory()
    c8_np = c8_pin.numpy()

    def handoff_u8():
        go.set_keyframes_u8(fi, poses_np, c8_np)
        go.init()
        go.init_albedo()

    handoff_u8(); lm_iteration(go, a.model)
    barrier()
    te8 = 0.0
    for _ in range(a.steps):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        handoff_u8()
        lm_iteration(go, a.model)
        go.poses(); go.light()
        torch.cuda.synchronize()
        te8 += time.perf_counter() - t0
    barrier()
    tte8 = torch.tensor([te8], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(tte8, op=dist.ReduceOp.MAX)
    e2e_u8 = {"value": A_job * (sweeps_per_step + 1) * a.steps / float(tte8.item()), "unit": UNIT,
              "h2d_bytes_per_step": int(c8_np.nbytes + poses_np.nbytes), "d2h_bytes_per_step": int(d2h),
              "ms_per_step": 1e3 * float(tte8.item()) / a.steps,
              "step": "as e2e, keyframes handed over as 8-bit BGR (psg_opt_set_keyframes_u8) and converted on the device"}

    if rank != 0:
        if world > 1:
            dist.barrier()
            if mode == "shard":
                go.comm_destroy()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel
    peak, peak_src = peaks()
    out_block = {"k_sweep<ALBEDO>": 24, "k_sweep<DIST>": 56}
    best = None
    table = {}
    for name, (n, ms) in prof.items():
        table[name] = {"launches": n, "ms_total": round(ms, 4), "ms_avg": round(ms / n, 5)}
        if name.startswith("k_sweep<") and (best is None or ms > prof[best][1]):
            best = name
    roof = None
    if best:
        n, ms = prof[best]
        # per launch = per rank: a rank sweeps its share of the work list
        shards = world if mode == "shard" else 1
        bytes_launch = sweep_bytes(A // shards, S // shards, F, out_block.get(best, 0))
        ach = bytes_launch / (ms / n * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": best, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": ncu_traffic(a, best) if world == 1 else None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": bytes_launch, "avg_launch_ms": ms / n,
                "note": "instruction-issue bound, not HBM bound: see DESIGN.md section 3 / profiles/r1_summary.md"}

    # ---- CPU baseline (reference, 1 thread) on a bounded sample
    cpu = None
    if not a.no_cpu_baseline:
        from oracle import pyoracle, pyref
        if pyref.available():
            rv, ro, A_s, keys = reference_sample(a, sc, vs, T, a.ref_keyframes)
            t0 = time.perf_counter(); ro.ps_energy(); dt = time.perf_counter() - t0
            cpu = {"value": A_s / dt, "unit": UNIT, "cores": 1, "kind": "reference", "host_cores": os.cpu_count(),
                   "sample": f"one getPSEnergy() sweep, reference sources, {len(keys)} of {F} keyframes on the full fused "
                             f"{G}^3 volume: S={ro.S}, {A_s} active pairs, {dt:.2f} s"}
            ro.close(); rv.close()
        # the O(1)-membership C port on the full workload, for the non-pathological comparison (BASELINE.md CPU-B)
        ov = pyoracle.OrVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5, max(64, F))
        oo = pyoracle.OrOptimizer(ov, a.model, vs, sc.K, reg_n=10.0, reg_l=reg_l)
        for f in range(F):
            R, t = se3_parts(sc.poses_gt[f])
            ov.integrate(sc.color[f], sc.depth[f], sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv, sc.K, R, t, f > 0)
        oo.set_keyframes(fi, sc.poses_init, sc.color); oo.init(); oo.init_albedo()
        t0 = time.perf_counter(); _, _, A_o = oo.energy(); dt = time.perf_counter() - t0
        port = {"value": A_o / dt, "unit": UNIT, "cores": 1, "kind": "port",
                "sample": f"one energy sweep of oracle/psg_oracle.c (dense row map) on the full workload: {A_o} pairs, {dt:.2f} s"}
        if cpu is None:
            cpu = port
        else:
            cpu["port"] = port

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
           "ms_per_step": 1e3 * t_max / a.steps, "higher_is_better": True, "scaling": "strong" if mode == "shard" else "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": workload_name(a), "surface_voxels": S_job, "active_pairs": A_job, "keyframes": F,
                      "sweeps_per_step": sweeps_per_step, "energy_after_last_step": E[0] if E else None,
                      "l2": "flushed between steps (256 MiB device write); per-step host timing bracketed by device syncs, summed",
                      "multi_gpu": ("single GPU" if world == 1 else
                                    f"ONE workload, work list sharded over {world} ranks, NCCL all-reduce / all-gather, replicated distance solve"
                                    if mode == "shard" else
                                    f"{world} independent keyframe sets, one per GPU (same geometry and camera orbit, lights and pose "
                                    f"perturbations seeded per rank), no data-path collective; surface_voxels / active_pairs are sums "
                                    f"over the ranks.  One such set has {vis_pairs} visible pairs, "
                                    f"below the {SHARD_MIN_PAIRS} per rank from which sharding one set pays "
                                    f"(DESIGN.md section 7; --multi shard forces it)")},
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "ms_per_step": 1e3 * float(tte.item()) / a.steps,
                   "step": "psg_opt_set_keyframes (pinned host images + poses, upload overlapped with the work-list build) + psg_opt_init + psg_opt_init_albedo + one LM iteration + read-back"},
           "e2e_u8": e2e_u8,
           "host_thread_us_per_step": {k: round(v, 1) for k, v in host_us.items() if k != "iterations"},
           "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu, "kernels": table}
    emit(out)
    if world > 1:
        dist.barrier()
        if mode == "shard":
            go.comm_destroy()
        dist.destroy_process_group()


_JSON_OUT = None


def emit(obj):
    """The one JSON line goes to the real stdout; everything else written to fd 1 was diverted to stderr in main()."""
    line = json.dumps(obj) + "\n"
    if _JSON_OUT is None:
        sys.stdout.write(line); sys.stdout.flush()
    else:
        os.write(_JSON_OUT, line.encode())


def main():
    # native libraries (NCCL prints its version banner) write to fd 1 directly; the contract is ONE JSON line on stdout
    global _JSON_OUT
    a = parse()
    sys.stdout.flush()
    _JSON_OUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
