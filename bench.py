#!/usr/bin/env python
"""bench.py -- photometric-optimizer hot path on synthetic RGB-D keyframes (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one process per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU code (oracle/_ref), rank 0 only

A *step* is one outer ("LM") iteration of the alternating optimizer over the whole keyframe set
(PsOptimizer.cpp:303-360): albedo, light, distance and pose blocks, each followed by the energy evaluation the
reference logs.  The reference spends 8 sweeps over all active (surface voxel, keyframe) pairs on that (computeResidual
inside every optimize*All + getPSEnergy after it); here every block's sweep also returns the energy of the state it
started from, so 5 sweeps give the same 4 updates and 4 logged energies, plus the distance PCG and the small solves.
metric = active pair evaluations per second (pairs x sweeps executed / time); ms_per_step = ms per LM iteration.

`value`      : inputs resident in HBM when the timed region starts (volume fused, keyframes uploaded, surface list built);
               timed with CUDA events on the library's stream (psg_debug_timer_*), L2 flushed between steps.
`e2e`        : the same iteration driven from HOST buffers through the C ABI: every step re-sends the keyframe images and
               poses from pinned host memory (psg_opt_set_keyframes), rebuilds the surface list (psg_opt_init), runs
               initAlbedo + one LM iteration and reads energies, poses and light back.
`roofline`   : the dominant kernel of the step BY DEVICE TIME over all kernels (sweeps and the distance solve alike), timed live
               with CUDA events on the launching stream (psg_profile_*); `kernels` has every kernel's bytes and fraction.
`config.grids`: the same step on the larger BASELINE grids (256^3 LED, 512^3 SH1), a few steps each, with per-kernel fractions.
`cpu_baseline`: the reference's own optimizer (oracle/_ref, 1 thread: the reference is single-threaded) on this box's host:
               one getPSEnergy() sweep and one full LM iteration (4 optimize*All + 4 getPSEnergy) of the same workload, plus
               the O(1)-membership C port (oracle/psg_oracle.c) for the non-pathological comparison.
`--impl reference`: a step is one full LM iteration of the reference on a bounded keyframe subset of the same fused volume.
At N > 1 every rank joins ONE communicator; the headline workload follows --multi and
`config.strong_512` reports one 512^3 x 30 job sharded over the N ranks (z-slab row ranges, distributed distance solve).
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
SWEEPS_PER_STEP = 5        # ours: 4 block sweeps (each also yields the energy of the state it starts from) + 1 energy sweep
REF_SWEEPS_PER_STEP = 8    # reference: computeResidual inside each optimize*All + getPSEnergy after it


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
    ap.add_argument("--no-grids", action="store_true", help="skip the 256^3 LED / 512^3 SH1 legs (config.grids, config.strong_512)")
    ap.add_argument("--sweep-split", type=int, default=0, choices=[0, 1, 2, 4, 8], help="psg_opt_set_sweep_split (0 = chosen per launch)")
    ap.add_argument("--pcg-mode", type=int, default=0, help="psg_opt_set_pcg_mode (0 = default)")
    ap.add_argument("--multi", default="auto", choices=["auto", "shard", "replicas"],
                    help="N > 1 headline workload: shard ONE workload over the ranks (strong scaling), run one independent keyframe "
                         "set per rank (weak scaling), or choose by the library's sharding threshold")
    ap.add_argument("--ref-keyframes", type=int, default=4,
                    help="keyframes in a step of the reference arm (one reference LM iteration costs ~2 s per keyframe at 128^3)")
    ap.add_argument("--cpu-iteration-keyframes", type=int, default=30,
                    help="keyframes in cpu_baseline's single reference LM iteration (30 = the full workload, ~40-60 s once)")
    return ap.parse_args()


def workload_name(a, grid=None, model=None, frames=None):
    return (f"synthetic TUM-style RGB-D orbit, {model or a.model}, {grid or a.grid}^3 grid, {frames or a.frames} keyframes "
            f"{a.width}x{a.height}, light+albedo+distance+pose")


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


def build_scene(a, seed=1234, model=None):
    from psgradientsdf_b200 import synth
    return synth.make_scene(F=a.frames, H=a.height, W=a.width, model=model or a.model, seed=seed)


def pose_parts(pose):
    """R, t of a camera-to-world 4x4 for the fusion call.  The synthetic poses are orthonormal by construction, so this is what
    the reference's SE3(pose) hands to update() up to rounding; no implementation is consulted to prepare the inputs."""
    return np.ascontiguousarray(pose[:3, :3], np.float32), np.ascontiguousarray(pose[:3, 3], np.float32)


def reg_for(model):
    """(reg_n, reg_l, damping): config_skorates.json for SH, config_basket_LED.json for LED."""
    return (0.1, 5.0, 3.0) if model == "LED" else (10.0, 0.0, 1.0)


# ------------------------------------------------------------------------------------------------ reference arm
def reference_job(a, sc, nkey):
    """The reference's own optimizer (oracle/_ref) on the same fused volume with `nkey` of the keyframes, after init + initAlbedo."""
    from oracle import pyref
    from psgradientsdf_b200 import synth
    G = a.grid
    vs, T = synth.grid_for_scene(G)
    reg_n, reg_l, damping = reg_for(a.model)
    rv = pyref.RefVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5)
    ro = pyref.RefOptimizer(rv, a.model, vs, sc.K, "/tmp/", reg_n=reg_n, reg_l=reg_l)
    for f in range(a.frames):
        rv.integrate(sc.color[f], sc.depth[f], sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv, sc.K, sc.poses_gt[f], f > 0)
    nkey = max(1, min(nkey, a.frames))
    keys = np.unique(np.linspace(0, a.frames - 1, nkey).round().astype(np.int32))
    ro.set_keyframes(keys, sc.poses_init[keys], sc.color[keys])
    ro.init()
    ro.init_albedo()
    _, _, A = ro.energy()
    return rv, ro, int(A), keys


def reference_iteration(ro, model):
    """One pass of the reference's while-body through its own optimize*All + getPSEnergy (PsOptimizer.cpp:303-360,
    LedOptimizer.cpp:343-403); returns per-block wall seconds the way the reference's Timer would log them."""
    reg_n, reg_l, damping = reg_for(model)
    order = ["light", "albedo", "dist", "pose"] if model == "LED" else ["albedo", "light", "dist", "pose"]
    times = {}
    for blk in order:
        t0 = time.perf_counter()
        ro.block_reference(blk, damping, False, reg_n != 0.0, reg_l != 0.0)
        t1 = time.perf_counter()
        ro.ps_energy()
        t2 = time.perf_counter()
        times[blk] = round(t1 - t0, 3); times[blk + "_energy"] = round(t2 - t1, 3)
    return times


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pyref
    if not pyref.available():
        emit({"impl": "reference", "unavailable": "oracle/_ref/libpsgref.so missing (reference sources were not built)"})
        return
    sc = build_scene(a)
    # the reference's membership test is a linear scan: an LM iteration costs ~2 s x (G/128)^4 per keyframe on one
    # core; keep a step around 5-10 s
    nkey = max(1, min(a.ref_keyframes, int(round(a.ref_keyframes / (a.grid / 128.0) ** 4)) or 1))
    rv, ro, A, keys = reference_job(a, sc, nkey)
    for _ in range(a.warmup):
        reference_iteration(ro, a.model)
    t0 = time.perf_counter()
    last = None
    for _ in range(a.steps):
        last = reference_iteration(ro, a.model)
    dt = time.perf_counter() - t0
    val = A * REF_SWEEPS_PER_STEP * a.steps / dt
    sample = (f"one LM iteration per step (optimizeAlbedoAll/LightAll/DistAll/PosesAll + getPSEnergy after each, "
              f"PsOptimizer.cpp:303-360) on the full fused {a.grid}^3 volume with {len(keys)} of the {a.frames} keyframes: "
              f"S={ro.S} surface voxels, {A} active pairs x {REF_SWEEPS_PER_STEP} sweeps per step")
    out = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
           "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic", "impl": "reference",
           "config": {"workload": workload_name(a), "sample": sample, "block_seconds_last_step": last},
           "cpu_baseline": {"value": val, "unit": UNIT, "cores": 1, "kind": "reference", "sample": sample,
                            "host_cores": os.cpu_count()},
           "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    emit(out)


# ------------------------------------------------------------------------------------------------ our arm
def sweep_bytes(A, S, F, out_block):
    """Algorithmic bytes of one sweep (SURVEY.md 8d / BASELINE.md): 48 A + 80 S + out_block S + 128 F."""
    return 48 * A + 80 * S + out_block * S + 128 * F


def kernel_bytes(name, A, S, F, pcg_iters):
    """Algorithmic bytes per launch; None for kernels without a byte model."""
    out_block = {"k_sweep<ALBEDO>": 24, "k_sweep<DIST>": 56}
    if name.startswith("k_sweep<"):
        return sweep_bytes(A, S, F, out_block.get(name, 0))
    if name == "k_dist_pcg":       # 76 S per CG iteration (+1 for the start-up pass) + 56 S to read the stencil blocks once
        return 76 * S * (pcg_iters + 1) + 56 * S
    if name in ("k_dist_assemble",):
        return 56 * S + 12 * S
    if name in ("k_dist_update", "k_grad_update", "k_reg_energy"):
        return 64 * S
    return None


def ncu_traffic(a, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed `ncu --set full` capture
    (profiles/*_traffic.json, newest round first); None when the capture was taken on another workload."""
    for fn in ("r2_traffic.json", "r1_traffic.json"):
        p = os.path.join(ROOT, "profiles", fn)
        try:
            with open(p) as f:
                t = json.load(f)
        except OSError:
            continue
        w = t.get("workload", {})
        if (w.get("grid"), w.get("frames"), w.get("model"), w.get("height"), w.get("width")) != (a.grid, a.frames, a.model, a.height, a.width):
            continue
        k = t.get("kernels", {}).get(kernel)
        if k is not None:
            return k["dram_read_bytes"] + k["dram_write_bytes"]
    return None


class Job:
    """One keyframe set fused into its volume, optimizer created, per-step inputs in pinned host memory."""

    def __init__(self, a, local, sc, G, model, max_frames=64):
        import torch

        import psgradientsdf_b200 as psg
        from psgradientsdf_b200 import synth
        self.a, self.model, self.G = a, model, G
        self.F = F = sc.color.shape[0]
        vs, T = synth.grid_for_scene(G)
        reg_n, reg_l, self.damping = reg_for(model)
        self.reg_n, self.reg_l = reg_n, reg_l
        self.gv = gv = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, max(max_frames, F), device=local)
        gv.set_zmin(0.5); gv.set_zmax(3.5)
        self.go = psg.Optimizer(gv, vs, sc.K, model=model, reg_n=reg_n, reg_l=reg_l)
        for f in range(F):
            R, t = pose_parts(sc.poses_gt[f])
            if f > 0:
                gv.increase_counter()
            gv.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
        self.img_pin = torch.from_numpy(np.ascontiguousarray(sc.color)).pin_memory()
        self.pose_pin = torch.from_numpy(np.ascontiguousarray(sc.poses_init)).pin_memory()
        self.images_np, self.poses_np = self.img_pin.numpy(), self.pose_pin.numpy()
        self.fi = np.arange(F, dtype=np.int32)
        self.cfg = None

    def handoff(self):
        self.go.set_keyframes(self.fi, self.poses_np, self.images_np)
        self.go.init()
        self.go.init_albedo()

    def iteration(self):
        """One pass of the reference's while-body: psg_opt_iteration on structs built once, as a C++ host's loop holds them."""
        if self.cfg is None:
            self.cfg = self.go.iteration_config(True, True, True, True, damping=self.damping, normal_reg=self.reg_n != 0.0,
                                                laplacian_reg=self.reg_l != 0.0)
        cfg, info = self.cfg
        self.go.iteration_raw(cfg, info)
        return info

    def close(self):
        self.go.close(); self.gv.close()


def timed_steps(job, steps, flush, sync, barrier=None):
    """K iterations, each bracketed by CUDA events on the library's stream (L2 flushed in between); returns
    (device seconds, host wall seconds, last info)."""
    dev_ms, wall = 0.0, 0.0
    info = None
    for _ in range(steps):
        flush.zero_()
        sync()
        if barrier is not None:
            barrier()
        t0 = time.perf_counter()
        job.go.timer_begin()
        info = job.iteration()
        dev_ms += job.go.timer_end()
        wall += time.perf_counter() - t0
    return dev_ms * 1e-3, wall, info


def kernel_table(prof, A, S, F, pcg_iters, peak):
    table = {}
    for name, (n, ms) in prof.items():
        row = {"launches": n, "ms_total": round(ms, 4), "ms_avg": round(ms / n, 5)}
        kb = kernel_bytes(name, A, S, F, pcg_iters)
        if kb is not None:
            row["algorithmic_bytes"] = int(kb)
            row["frac"] = round(kb / (ms / n * 1e-3) / 1e9 / peak, 4)
        table[name] = row
    return table


def profiled_steps(job, steps, flush, sync, peak):
    """The same steps once more with the library's CUDA-event launch timer on (kept out of the timed region: two event
    records per launch would be charged to the step).  Returns the per-kernel table."""
    import psgradientsdf_b200 as psg
    psg.profile_reset(); psg.profile_enable(True)
    info = None
    for _ in range(steps):
        flush.zero_(); sync()
        info = job.iteration()
    sync()
    prof = psg.profile_table()
    psg.profile_enable(False)
    return prof, info


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
    dev = f"cuda:{local}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    G, F = a.grid, a.frames
    peak, peak_src = peaks()
    SHARD_MIN_PAIRS = 3_000_000       # kShardMinPairs (psgradientsdf_b200/csrc/comm.cu)

    def sync():
        torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # N > 1: one workload sharded over the ranks, or one independent keyframe set per rank.  "auto" asks the library's own rule
    # (DESIGN.md section 7); rank 0 builds the standard workload, counts, and tells the others.
    mode = "single" if world == 1 else a.multi
    job = None
    if world > 1 and mode == "auto":
        verdict = [None]
        if rank == 0:
            job = Job(a, local, build_scene(a, 1234), G, a.model)
            job.go.set_keyframes(job.fi, job.poses_np, job.images_np); job.go.init()
            verdict[0] = "shard" if job.go.is_sharded()[1] >= SHARD_MIN_PAIRS else "replicas"
        dist.broadcast_object_list(verdict, src=0)
        mode = verdict[0]
    seed = 1234 + (17 * rank if mode == "replicas" else 0)
    if job is None:
        job = Job(a, local, build_scene(a, seed), G, a.model)
    go = job.go
    if a.pcg_mode:
        go.set_pcg_mode(a.pcg_mode)
    if a.sweep_split:
        go.set_sweep_split(a.sweep_split)

    def join_comm(g):
        ids = [psg.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        g.comm_init(ids[0], rank, world)
        g.set_shard_policy(1)

    if mode == "shard":
        join_comm(go)
    job.handoff()
    S = go.S
    E0, A = go.energy()
    # units of one step over the whole job: the shared work list when sharded, the sum of the ranks' own lists otherwise
    A_job, S_job = A, S
    vis_pairs = go.is_sharded()[1]
    if mode == "replicas":
        tot = torch.tensor([A, S], dtype=torch.float64, device=dev)
        dist.all_reduce(tot)
        A_job, S_job = int(tot[0].item()), int(tot[1].item())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    # ---- device-resident leg
    for _ in range(a.warmup):
        job.iteration()
    sampler = ClockSampler(local); sampler.start()
    launches0 = psg.kernel_launches()
    barrier()
    psg.host_times(reset=True)
    t_dev, t_wall, info = timed_steps(job, a.steps, flush, sync, barrier if mode == "shard" else None)
    host_us = psg.host_times(reset=True)
    barrier()
    launches = psg.kernel_launches() - launches0
    clocks = sampler.stop()
    prof, pinfo = profiled_steps(job, a.steps, flush, sync, peak)
    pcg_iters = int(pinfo.block[2].cg_iterations)
    pcg_kernel = {0: "general cooperative grid", 1: "8-CTA cluster", 2: "fused grid kernel", 3: "16-CTA cluster, vectors in distributed shared memory",
                  4: "distributed (sharded job)", 5: "fused grid kernel, Chronopoulos-Gear arrangement"}.get(go.last_pcg_kernel(), "?")
    t_max = allmax(t_dev)
    t_wall_max = allmax(t_wall)
    value = A_job * SWEEPS_PER_STEP * a.steps / t_max      # whole job: the shared list once, or the sum of the ranks' own lists

    # ---- e2e leg: host buffers -> C ABI every step (keyframe upload overlapped with the work-list build: the pinned image
    # buffers stay valid until psg_opt_init returns, as psg_opt_set_upload_overlap requires)
    go.set_upload_overlap(True)
    h2d = job.images_np.nbytes + job.poses_np.nbytes
    d2h = job.poses_np.nbytes + 4 * go.light().size + 8 * 6

    def e2e_leg(handoff):
        for _ in range(min(a.warmup, 2)):
            handoff(); job.iteration()
        barrier()
        te = 0.0
        for _ in range(a.steps):
            flush.zero_()
            sync()
            t0 = time.perf_counter()
            handoff()
            job.iteration()
            go.poses(); go.light()
            sync()
            te += time.perf_counter() - t0
        barrier()
        return allmax(te)

    te = e2e_leg(job.handoff)
    e2e_value = A_job * (SWEEPS_PER_STEP + 1) * a.steps / te   # + the initAlbedo sweep

    # same hand-off with the keyframes as the loader read them (8-bit BGR, psg_opt_set_keyframes_u8): reported beside e2e
    c8_pin = torch.from_numpy(np.ascontiguousarray(np.round(job.images_np * 255.0).astype(np.uint8))).pin_memory()
    c8_np = c8_pin.numpy()

    def handoff_u8():
        go.set_keyframes_u8(job.fi, job.poses_np, c8_np)
        go.init()
        go.init_albedo()

    te8 = e2e_leg(handoff_u8)
    e2e_u8 = {"value": A_job * (SWEEPS_PER_STEP + 1) * a.steps / te8, "unit": UNIT,
              "h2d_bytes_per_step": int(c8_np.nbytes + job.poses_np.nbytes), "d2h_bytes_per_step": int(d2h),
              "ms_per_step": 1e3 * te8 / a.steps,
              "step": "as e2e, keyframes handed over as 8-bit BGR (psg_opt_set_keyframes_u8) and converted on the device"}
    go.set_upload_overlap(False)
    energy_end = float(info.energy_end) if info is not None else None
    if mode == "shard":
        go.comm_destroy()
    job.close()
    del job

    # ---- the larger BASELINE grids: a few steps each, per-kernel fractions (N = 1), and ONE 512^3 job sharded over the ranks (N > 1)
    grids, strong = None, None
    if not a.no_grids:
        sc_big = {}

        def big_job(model, Gb):
            if model not in sc_big:
                sc_big[model] = build_scene(a, 1234, model=model)
            return Job(a, local, sc_big[model], Gb, model)

        def run_grid(model, Gb, nsteps=3, shard=False, solo=False):
            """solo: only the calling rank runs (no collective inside)."""
            j = big_job(model, Gb)
            if shard:
                join_comm(j.go)
            j.handoff()
            Eb, Ab = j.go.energy()
            for _ in range(2):
                j.iteration()
            if not solo:
                barrier()
            td, tw, inf = timed_steps(j, nsteps, flush, sync, barrier if shard else None)
            if not solo:
                td = allmax(td)
            prof_b, pinf = profiled_steps(j, nsteps, flush, sync, peak)
            tab = kernel_table(prof_b, Ab // (world if shard else 1), j.go.S // (world if shard else 1), j.F,
                               int(pinf.block[2].cg_iterations), peak)
            out = {"workload": workload_name(a, Gb, model, j.F), "surface_voxels": int(j.go.S), "active_pairs": int(Ab),
                   "ms_per_step": round(1e3 * td / nsteps, 4), "steps": nsteps, "pcg_iterations": int(pinf.block[2].cg_iterations),
                   "evals_per_s": Ab * SWEEPS_PER_STEP * nsteps / td, "kernels": tab}
            if shard:
                out["mode"] = f"one job sharded over {world} ranks"
                j.go.comm_destroy()
            j.close()
            return out

        if world == 1:
            grids = {"256^3 LED": run_grid("LED", 256), "512^3 SH1": run_grid("SH1", 512)}
        else:
            strong = run_grid("SH1", 512, shard=True)
            # the same job on ONE of these GPUs in the same run, for the ratio (the other ranks wait)
            if rank == 0:
                one = run_grid("SH1", 512, solo=True)
                strong["one_gpu_same_run_ms_per_step"] = one["ms_per_step"]
                strong["speedup_vs_one_gpu_same_run"] = round(one["ms_per_step"] / strong["ms_per_step"], 3)
            barrier()

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel: over ALL kernels of the step by device time
    shards = world if mode == "shard" else 1
    table = kernel_table(prof, A // shards, S // shards, F, pcg_iters, peak)
    best = max(prof, key=lambda k: prof[k][1]) if prof else None
    roof = None
    if best:
        n, ms = prof[best]
        row = table[best]
        kb = row.get("algorithmic_bytes")
        ach = (kb / (ms / n * 1e-3) / 1e9) if kb else None
        roof = {"bound": "hbm", "kernel": best, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": (ach / peak) if ach else None,
                "traffic": ncu_traffic(a, best) if world == 1 else None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": kb, "avg_launch_ms": ms / n,
                "share_of_kernel_time": round(ms / sum(v[1] for v in prof.values()), 3),
                "note": ("the distance solve is a chain of grid-wide barriers, the sweeps are instruction-issue bound: neither is "
                         "HBM bound at this size (DESIGN.md section 3, profiles/r2_summary.md); config.grids has the 512^3 fractions")}

    # ---- CPU baseline (reference sources, 1 thread): one energy sweep + one full LM iteration; and the O(1)-membership port
    cpu = None
    if not a.no_cpu_baseline and world == 1:
        from oracle import pyoracle, pyref
        from psgradientsdf_b200 import synth
        sc = build_scene(a)
        vs, T = synth.grid_for_scene(G)
        reg_n, reg_l, damping = reg_for(a.model)
        if pyref.available():
            rv, ro, A_s, keys = reference_job(a, sc, a.cpu_iteration_keyframes)
            t0 = time.perf_counter(); ro.ps_energy(); dt = time.perf_counter() - t0
            t0 = time.perf_counter(); blk_s = reference_iteration(ro, a.model); dti = time.perf_counter() - t0
            cpu = {"value": A_s / dt, "unit": UNIT, "cores": 1, "kind": "reference", "host_cores": os.cpu_count(),
                   "sample": f"one getPSEnergy() sweep, reference sources, {len(keys)} of {F} keyframes on the full fused "
                             f"{G}^3 volume: S={ro.S}, {A_s} active pairs, {dt:.2f} s",
                   "lm_iteration_ms": round(1e3 * dti, 1),
                   "lm_iteration_evals_per_s": A_s * REF_SWEEPS_PER_STEP / dti,
                   "lm_iteration_block_seconds": blk_s,
                   "lm_iteration_sample": f"one pass of the while-body (PsOptimizer.cpp:303-360) through the reference's own "
                                          f"optimize*All + getPSEnergy, {len(keys)} of {F} keyframes (BASELINE.md CPU-A)"}
            ro.close(); rv.close()
        # the O(1)-membership C port on the full workload, for the non-pathological comparison (BASELINE.md CPU-B)
        ov = pyoracle.OrVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5, max(64, F))
        oo = pyoracle.OrOptimizer(ov, a.model, vs, sc.K, reg_n=reg_n, reg_l=reg_l)
        for f in range(F):
            R, t = pose_parts(sc.poses_gt[f])
            ov.integrate(sc.color[f], sc.depth[f], sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv, sc.K, R, t, f > 0)
        fi = np.arange(F, dtype=np.int32)
        oo.set_keyframes(fi, sc.poses_init, sc.color); oo.init(); oo.init_albedo()
        t0 = time.perf_counter(); _, _, A_o = oo.energy(); dt = time.perf_counter() - t0
        order = ["light", "albedo", "dist", "pose"] if a.model == "LED" else ["albedo", "light", "dist", "pose"]
        t0 = time.perf_counter()
        for blk in order:
            oo.block(blk, apply=True, damping=damping, normal_reg=reg_n != 0.0, laplacian_reg=reg_l != 0.0)
            oo.energy()
        dti = time.perf_counter() - t0
        port = {"value": A_o / dt, "unit": UNIT, "cores": 1, "kind": "port",
                "sample": f"one energy sweep of oracle/psg_oracle.c (dense row map) on the full workload: {A_o} pairs, {dt:.2f} s",
                "lm_iteration_ms": round(1e3 * dti, 1), "lm_iteration_evals_per_s": A_o * REF_SWEEPS_PER_STEP / dti,
                "lm_iteration_sample": "the same pass through the C port's block solves + energies, all keyframes (BASELINE.md CPU-B)"}
        if cpu is None:
            cpu = port
        else:
            cpu["port"] = port

    config = {"workload": workload_name(a), "surface_voxels": S_job, "active_pairs": A_job, "keyframes": F,
              "sweeps_per_step": SWEEPS_PER_STEP, "pcg_iterations": pcg_iters, "distance_solve_kernel": pcg_kernel, "energy_after_last_step": energy_end,
              "timing": "CUDA events on the library's stream around each step (psg_debug_timer_*), max over ranks; host wall "
                        "clock per step in host_ms_per_step",
              "l2": "flushed between steps (256 MiB device write)",
              "multi_gpu": ("single GPU" if world == 1 else
                            f"ONE workload sharded over {world} ranks (DESIGN.md section 7)"
                            if mode == "shard" else
                            f"{world} independent keyframe sets, one per GPU (same geometry and camera orbit, lights and pose "
                            f"perturbations seeded per rank), no data-path collective; surface_voxels / active_pairs are sums "
                            f"over the ranks.  One such set has {vis_pairs} visible pairs, "
                            f"below the {SHARD_MIN_PAIRS} from which sharding one set pays "
                            f"(DESIGN.md section 7; --multi shard forces it); config.strong_512 is the sharded job")}
    if grids is not None:
        config["grids"] = grids
    if strong is not None:
        config["strong_512"] = strong
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
           "ms_per_step": 1e3 * t_max / a.steps, "host_ms_per_step": 1e3 * t_wall_max / a.steps, "higher_is_better": True,
           "scaling": "strong" if mode == "shard" else "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "ms_per_step": 1e3 * te / a.steps,
                   "step": "psg_opt_set_keyframes (pinned host images + poses, upload overlapped with the work-list build) + psg_opt_init + psg_opt_init_albedo + one LM iteration + read-back"},
           "e2e_u8": e2e_u8,
           "host_thread_us_per_step": {k: round(v, 1) for k, v in host_us.items() if k != "iterations"},
           "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu, "kernels": table}
    emit(out)
    if world > 1:
        dist.barrier()
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
    if os.environ.get("PSG_BENCH_WATCHDOG"):      # debugging aid: dump every thread's Python stack if the run is still going by then
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["PSG_BENCH_WATCHDOG"]), repeat=True, file=sys.stderr)
    sys.stdout.flush()
    _JSON_OUT = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
