"""CPU: host-side logic that needs no device -- scene generator determinism and conventions, bench.py's byte model,
and the world_size-2 (gloo) rank plumbing bench.py relies on."""
import os
import subprocess
import sys

import numpy as np
import pytest

from psgradientsdf_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_scene_is_deterministic_and_quantised():
    a = synth.make_scene(F=3, H=60, W=80, model="SH1")
    b = synth.make_scene(F=3, H=60, W=80, model="SH1")
    assert np.array_equal(a.color, b.color) and np.array_equal(a.depth, b.depth) and np.array_equal(a.poses_init, b.poses_init)
    inv255 = np.float32(1.0) / np.float32(255.0)
    u8 = np.rint(a.color * 255.0).astype(np.uint8)
    assert np.array_equal(u8.astype(np.float32) * inv255, a.color)            # u8 * (1/255 as float), ImageLoader.h:181
    u16 = np.rint(a.depth * 5000.0).astype(np.uint16)
    assert np.array_equal(u16.astype(np.float32) * (np.float32(1.0) / np.float32(5000.0)), a.depth)
    assert (a.depth > 0).mean() > 0.05
    # poses are rigid, keyframe 0 is unperturbed
    for P in a.poses_init:
        R = P[:3, :3].astype(np.float64)
        assert np.abs(R @ R.T - np.eye(3)).max() < 1e-5 and abs(np.linalg.det(R) - 1) < 1e-5
    assert np.array_equal(a.poses_init[0], a.poses_gt[0])
    # normals are inward pointing (away from the camera) where depth is valid
    assert (a.nz[a.depth > 0] > 0).mean() > 0.99


def test_models_share_geometry():
    a = synth.make_scene(F=2, H=48, W=64, model="SH1")
    b = synth.make_scene(F=2, H=48, W=64, model="LED")
    assert np.array_equal(a.depth, b.depth) and np.array_equal(a.poses_init, b.poses_init) and not np.array_equal(a.color, b.color)


def test_sweep_byte_model():
    sys.path.insert(0, ROOT)
    import bench
    # 48 A + 80 S + out_block S + 128 F
    assert bench.sweep_bytes(1000, 100, 10, 24) == 48000 + 8000 + 2400 + 1280
    assert "128^3" in bench.METRIC


def test_shard_blocks_tile_the_work_list():
    import psgradientsdf_b200 as psg
    for S in (0, 1, 31, 32, 33, 1000, 28384, 458398, 2 ** 21 + 5):
        nblocks = (S + 31) // 32
        for world in (1, 2, 3, 4, 8):
            seen = []
            counts = []
            for rank in range(world):
                first, stride, count = psg.shard_blocks(S, rank, world)
                assert stride == world and first == rank and count >= 0
                seen += [first + i * stride for i in range(count)]
                counts.append(count)
            assert sorted(seen) == list(range(nblocks)), (S, world)
            assert max(counts) - min(counts) <= 1


def test_row_partition_covers_balances_and_haloes():
    """psg_partition_rows (the z-slab split of a sharded job): the ranges tile the surface list, carry equal visible-pair counts,
    and a rank's halo contains every row of another rank that its stencils can touch: the finite-difference neighbours (+-1, +-W,
    +-W*H), the rows whose stencil contains one of its rows, and the 6-point Laplacian."""
    import psgradientsdf_b200 as psg
    rng = np.random.default_rng(5)
    for G, S in ((16, 700), (32, 5000), (48, 20000)):
        W = H = G
        plane = W * H
        # a shell-like surface: voxels near a sphere, so that slabs have ragged first / last planes
        idx = np.arange(G ** 3)
        i, j, k = idx % W, (idx // W) % H, idx // plane
        r = np.sqrt((i - G / 2) ** 2 + (j - G / 2) ** 2 + (k - G / 2) ** 2)
        cand = np.flatnonzero(np.abs(r - 0.35 * G) < 1.5)
        lin = np.sort(rng.choice(cand, min(S, cand.size), replace=False)).astype(np.int32)
        vis = rng.integers(1, 2 ** 30, lin.size).astype(np.uint64)
        rowmap = -np.ones(G ** 3, np.int64); rowmap[lin] = np.arange(lin.size)
        w = np.array([bin(int(v)).count("1") + 1 for v in vis], np.int64)
        for world in (1, 2, 3, 8):
            b, lo, hi = psg.partition_rows(lin, vis, plane, world)
            assert b[0] == 0 and b[-1] == lin.size and np.all(np.diff(b) >= 0)
            loads = np.array([w[b[q]:b[q + 1]].sum() for q in range(world)])
            assert loads.max() - loads.min() <= 2 * w.max() + 1, (world, loads)
            for q in range(world):
                own = np.arange(b[q], b[q + 1])
                assert lo[q] <= b[q] and hi[q] >= b[q + 1]
                if own.size == 0:
                    continue
                for off in (1, -1, W, -W, plane, -plane):
                    nl = lin[own].astype(np.int64) + off
                    ok = (nl >= 0) & (nl < G ** 3)
                    rows = rowmap[nl[ok]]
                    rows = rows[rows >= 0]
                    assert np.all((rows >= lo[q]) & (rows < hi[q])), (G, world, q, off)
    # degenerate inputs: empty list, more ranks than rows
    b, lo, hi = psg.partition_rows(np.zeros(0, np.int32), np.zeros(0, np.uint64), 64, 4)
    assert list(b) == [0, 0, 0, 0, 0]
    b, lo, hi = psg.partition_rows(np.array([3, 9], np.int32), np.array([1, 1], np.uint64), 64, 4)
    assert b[0] == 0 and b[-1] == 2 and np.all(np.diff(b) >= 0)


_WORKER = r'''
import os, sys, torch, torch.distributed as dist
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
t = torch.tensor([1.0 + rank], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)          # bench.py: max-over-ranks timing
u = torch.tensor([float(rank + 1)], dtype=torch.float64)
dist.all_reduce(u, op=dist.ReduceOp.SUM)          # small normal-equation blocks are summed over ranks
assert t.item() == world and u.item() == world * (world + 1) / 2
# the blocks each rank would sweep tile the work list (library function, no GPU needed)
sys.path.insert(0, os.environ["PSG_ROOT"])
import psgradientsdf_b200 as psg
S = 28384
first, stride, count = psg.shard_blocks(S, rank, world)
mine = torch.zeros((S + 31) // 32, dtype=torch.int64)
mine[first::stride][:count] += 1
dist.all_reduce(mine, op=dist.ReduceOp.SUM)       # every block is swept by exactly one rank
assert bool((mine == 1).all())
dist.barrier(); dist.destroy_process_group()
print("rank", rank, "ok")
'''


def test_two_rank_gloo_plumbing(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER)
    env = dict(os.environ, PSG_ROOT=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29611", str(script)], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.count("ok") == 2


@pytest.mark.parametrize("model", ["SH1", "SH2", "LED"])
def test_small_solve_is_the_reference_cg(model):
    """The host solve of the light and pose blocks (host_solve.cpp, no device involved) against the reference's own
    Eigen::ConjugateGradient on the committed golden systems: same iterates bit for bit, same iteration count and verdict --
    including the 9 x 9 SH2 light systems, whose verdict depends on the summation order of the inner products."""
    import psgradientsdf_b200 as psg
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", f"ref_{model}.npz"))
    for name in ("light", "pose"):
        bs = 6 if name == "pose" else (3 if model == "LED" else (9 if model == "SH2" else 4))
        H = np.asarray(g[name + "_H"], np.float32)
        b = np.asarray(g[name + "_b"], np.float32)
        F = b.size // bs
        assert H.shape == (F * bs, F * bs)
        blocks = np.stack([H[f * bs:(f + 1) * bs, f * bs:(f + 1) * bs] for f in range(F)])
        x, ok, it, err = psg.small_cg(blocks, b)
        assert it == int(g[name + "_iters"]) and ok == (int(g[name + "_info"]) == 0), (model, name, it, ok)
        assert np.array_equal(x.view(np.uint32), np.asarray(g[name + "_delta"], np.float32).view(np.uint32)), (model, name)
