"""Several ranks: the row-range ("z-slab") sharded path and the replicated path against the single-GPU run.

* test_local_ranks_*: the ranks of a job as threads of ONE process on ONE GPU (psg_opt_comm_init_local): the same sharded code
  path -- per-rank slot lists, halo pushes, cross-rank flag barrier and mailbox reductions inside the persistent distance
  solve, all-reduce of the small sums, psg_opt_sync_volume -- with the solve launched as one cooperative kernel whose blocks
  play all ranks.  Runs wherever there is one GPU.
* test_two_ranks_match_one_gpu: two processes on two GPUs under torchrun (NCCL + CUDA IPC), tools/multi_gpu_check.py;
  skipped on a one-GPU box.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

from tests.common import block_order, handoff, make_gpu, rel_err, scene

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run_single(model, G, sc, iters, lap):
    gv, go = make_gpu(model, G, sc)
    handoff(go, sc)
    go.init_albedo()
    log = [go.energy()]
    for it in range(iters):
        for blk in block_order(model):
            info = go.step(blk, damping=1.0, normal_reg=True, laplacian_reg=lap)
            log.append((info["energy_before"], info["n_active_before"], info["iters"], info["cg_success"], info["n_accepted"]))
    log.append(go.energy())
    it_info = go.iteration(damping=1.0, normal_reg=True, laplacian_reg=lap)
    log.append((it_info["energy_end"], it_info["n_active_end"]))
    regs = (it_info["reg_normal_after_dist"], it_info["reg_laplacian_after_dist"])
    return gv, go, log, regs


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("model", ["SH1", "LED"])
def test_local_ranks_match_one_gpu(model, world):
    import psgradientsdf_b200 as psg
    G, iters = 40, 2
    lap = model == "LED"
    sc = scene(model)
    gv1, go1, log1, regs1 = _run_single(model, G, sc, iters, lap)
    vols, opts = [], []
    for r in range(world):
        gv, go = make_gpu(model, G, sc)
        handoff(go, sc)
        vols.append(gv); opts.append(go)
    psg.comm_init_local(opts)
    psg.run_ranks(lambda r: opts[r].set_shard_policy(1), world)

    def body(r):
        o = opts[r]
        o.init_albedo()
        log = [o.energy()]
        for it in range(iters):
            for blk in block_order(model):
                info = o.step(blk, damping=1.0, normal_reg=True, laplacian_reg=lap)
                log.append((info["energy_before"], info["n_active_before"], info["iters"], info["cg_success"], info["n_accepted"]))
        log.append(o.energy())
        it_info = o.iteration(damping=1.0, normal_reg=True, laplacian_reg=lap)
        log.append((it_info["energy_end"], it_info["n_active_end"]))
        regs = (it_info["reg_normal_after_dist"], it_info["reg_laplacian_after_dist"])
        rng_ = o.row_range()
        o.sync_volume()
        return log, regs, rng_, o.light(), o.poses()

    res = psg.run_ranks(body, world)
    # the ranges tile the list, in rank order
    S = go1.S
    bounds = [res[0][2][0]] + [r[2][1] for r in res]
    assert bounds[0] == 0 and bounds[-1] == S and all(res[q][2][0] == bounds[q] for q in range(world))
    assert all(opts[q].is_sharded()[0] for q in range(world))
    # every rank saw the same energies, verdicts, light and poses ...
    for q in range(1, world):
        assert res[q][0] == res[0][0], (q, res[q][0], res[0][0])
        assert res[q][1] == res[0][1]
        assert np.array_equal(res[q][3], res[0][3]) and np.array_equal(res[q][4], res[0][4])
    # ... and after psg_opt_sync_volume holds the same volume
    states = [v.download() for v in vols]
    for q in range(1, world):
        assert np.array_equal(states[q].view(np.uint32), states[0].view(np.uint32)), f"rank {q} holds a different volume"
    # against one GPU: same pair counts and CG verdicts, iteration counts within one, energies to rounding (the dot products and
    # the per-keyframe sums are associated per rank), same final state
    worst = 0.0
    for a, b in zip(res[0][0], log1):
        assert a[1] == b[1], (a, b)
        worst = max(worst, abs(a[0] - b[0]) / abs(b[0]))
        if len(a) == 5:
            assert abs(a[2] - b[2]) <= 1 and a[3] == b[3], (a, b)
            assert abs(a[4] - b[4]) <= max(2, 1e-3 * b[4]), (a, b)
    assert worst < 2e-7, worst
    for x, y in zip(res[0][1], regs1):
        assert abs(x - y) <= 1e-6 * abs(y) + 1e-12, (res[0][1], regs1)
    s1 = gv1.download()
    surf = go1.surface()
    assert rel_err(states[0][surf][:, [0, 5, 6, 7]], s1[surf][:, [0, 5, 6, 7]]) < 1e-5
    assert rel_err(res[0][3], go1.light()) < 1e-5 and rel_err(res[0][4], go1.poses()) < 1e-6
    psg.run_ranks(lambda r: opts[r].comm_destroy(), world)


def test_local_ranks_halo_and_replicated_policy():
    """After a sharded distance block the distance plane of every rank agrees with the owners' on its own rows AND on its halo
    (the pushes inside the solve kernel); and a job switched back to the replicated policy continues bit-identically on all
    ranks."""
    import psgradientsdf_b200 as psg
    model, G, world = "SH1", 40, 2
    sc = scene(model)
    vols, opts = [], []
    for r in range(world):
        gv, go = make_gpu(model, G, sc)
        handoff(go, sc)
        vols.append(gv); opts.append(go)
    psg.comm_init_local(opts)
    psg.run_ranks(lambda r: opts[r].set_shard_policy(1), world)

    def body(r):
        o = opts[r]
        o.init_albedo()
        o.step("albedo", damping=1.0); o.step("light", damping=1.0)
        info = o.step("dist", damping=1.0, normal_reg=True)
        return info, o.row_range()

    res = psg.run_ranks(body, world)
    assert res[0][0]["iters"] == res[1][0]["iters"] > 3 and res[0][0]["n_accepted"] == res[1][0]["n_accepted"] > 0
    surf = opts[0].surface()
    d = [v.download()[surf, 0] for v in vols]          # no sync: each rank's own view
    truth = np.empty_like(d[0])
    for q in range(world):
        lo, hi, hlo, hhi = res[q][1]
        truth[lo:hi] = d[q][lo:hi]
    for q in range(world):
        lo, hi, hlo, hhi = res[q][1]
        assert hlo < lo or hhi > hi, "the test scene must give every rank a halo"
        assert np.array_equal(d[q][hlo:hhi], truth[hlo:hhi]), f"rank {q}: halo distances differ from their owners'"
    # back to the replicated policy: the whole state is synchronised first, then every rank sweeps everything
    psg.run_ranks(lambda r: opts[r].set_shard_policy(2), world)
    out = psg.run_ranks(lambda r: (opts[r].iteration(damping=1.0, normal_reg=True)["energy_end"], opts[r].is_sharded()[0]), world)
    assert out[0] == out[1] and out[0][1] == 0
    a, b = vols[0].download(), vols[1].download()
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    psg.run_ranks(lambda r: opts[r].comm_destroy(), world)


@pytest.mark.parametrize("policy", [1, 0])          # 1 = always shard, 0 = by work (replicated for this small scene)
@pytest.mark.parametrize("model", ["SH1", "LED"])
def test_two_ranks_match_one_gpu(model, policy):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    port = 29700 + 2 * policy + (1 if model == "LED" else 0)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tools", "multi_gpu_check.py"), model, "48",
                        str(policy)], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
    assert "multi-gpu check" in r.stdout


def test_local_ranks_second_handoff():
    """A sharded job that receives new keyframes (psg_opt_set_keyframes + psg_opt_init again: the work list, the partition and the
    peers' mappings are rebuilt) continues like a fresh one."""
    import psgradientsdf_b200 as psg
    model, G, world = "SH1", 40, 2
    sc = scene(model)
    vols, opts = [], []
    for r in range(world):
        gv, go = make_gpu(model, G, sc)
        vols.append(gv); opts.append(go)
    psg.comm_init_local(opts)
    psg.run_ranks(lambda r: opts[r].set_shard_policy(1), world)

    def body(r):
        o = opts[r]
        out = []
        for rep in range(3):
            handoff(o, sc)
            o.init_albedo()
            info = o.iteration(damping=1.0, normal_reg=True)
            out.append((info["energy_end"], info["n_active_end"], info["dist"]["iters"], o.row_range()))
        return out

    res = psg.run_ranks(body, world)
    for rep in range(3):   # the volume moves between the hand-offs, so the lists differ from one to the next -- but never between ranks
        assert res[0][rep][:3] == res[1][rep][:3], (rep, res[0][rep], res[1][rep])
        assert res[0][rep][2] > 3 and res[0][rep][1] > 0
        assert res[0][rep][3][0] == 0 and res[0][rep][3][1] == res[1][rep][3][0]
    a, b = vols[0].download(), vols[1].download()     # the last psg_opt_init synchronised them; the last iteration did not
    surf = opts[0].surface()
    assert np.array_equal(surf, opts[1].surface())
    psg.run_ranks(lambda r: opts[r].sync_volume(), world)
    a, b = vols[0].download(), vols[1].download()
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    psg.run_ranks(lambda r: opts[r].comm_destroy(), world)
