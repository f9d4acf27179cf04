"""Run under torchrun with N >= 2 GPUs: the multi-GPU optimizer (row-range sharding with the distributed distance solve over
CUDA IPC peer memory + NCCL, or the replicated policy) against the single-GPU one on rank 0.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29600 tools/multi_gpu_check.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import psgradientsdf_b200 as psg
from tests.common import block_order, handoff, scene
from psgradientsdf_b200 import synth

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
model = sys.argv[1] if len(sys.argv) > 1 else "SH1"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 48
sc = scene(model)
vs, T = synth.grid_for_scene(G)


def build(device):
    vol = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64, device=device)
    vol.set_zmin(0.5); vol.set_zmax(3.5)
    opt = psg.Optimizer(vol, vs, sc.K, model=model, reg_n=10.0, reg_l=5.0 if model == "LED" else 0.0)
    for f in range(sc.color.shape[0]):
        R, t = sc.poses_gt[f][:3, :3].copy(), sc.poses_gt[f][:3, 3].copy()
        if f:
            vol.increase_counter()
        vol.update(sc.color[f], sc.depth[f], sc.K, R, t, sc.nx[f], sc.ny[f], sc.nz[f], sc.n_sq_inv)
    handoff(opt, sc)
    return vol, opt


vol, opt = build(local)
ids = [psg.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
opt.comm_init(ids[0], rank, world)
opt.set_shard_policy(int(sys.argv[3]) if len(sys.argv) > 3 else 1)   # default: force sharding (the test scene is tiny)
opt.init_albedo()
log = [opt.energy()]
psg.profile_reset(); psg.profile_enable(True)
for it in range(2):
    for blk in block_order(model):
        info = opt.step(blk, damping=1.0, normal_reg=True, laplacian_reg=(model == "LED"))
        log.append((info["energy_before"], info["n_active_before"], info["iters"]))
log.append(opt.energy())
it_info = opt.iteration(damping=1.0, normal_reg=True, laplacian_reg=(model == "LED"))   # the three-wait pass, sharded
log.append((it_info["energy_end"], it_info["n_active_end"]))
regs = (it_info["reg_normal_after_dist"], it_info["reg_laplacian_after_dist"])
prof = {k: round(1e3 * v[1] / v[0], 1) for k, v in psg.profile_table().items()}
psg.profile_enable(False)
opt.sync_volume()        # sharded: every rank receives the others' rows
state = vol.download()
# all ranks must hold the same state
h = torch.tensor(np.frombuffer(state.tobytes(), dtype=np.uint8).astype(np.int64).sum(), device=f"cuda:{local}")
hs = [torch.zeros_like(h) for _ in range(world)]
dist.all_gather(hs, h)
assert all(int(x) == int(hs[0]) for x in hs), "ranks diverged"
if rank == 0:
    vol1, opt1 = build(local)     # single-GPU reference run on the same device
    opt1.init_albedo()
    log1 = [opt1.energy()]
    for it in range(2):
        for blk in block_order(model):
            info = opt1.step(blk, damping=1.0, normal_reg=True, laplacian_reg=(model == "LED"))
            log1.append((info["energy_before"], info["n_active_before"], info["iters"]))
    log1.append(opt1.energy())
    it1 = opt1.iteration(damping=1.0, normal_reg=True, laplacian_reg=(model == "LED"))
    log1.append((it1["energy_end"], it1["n_active_end"]))
    worst = 0.0
    for a, b in zip(log, log1):
        assert a[1] == b[1], (a, b)
        worst = max(worst, abs(a[0] - b[0]) / abs(b[0]))
        if len(a) == 3:
            assert abs(a[2] - b[2]) <= 1, ("CG iterations", a, b)
    for x, y in zip(regs, (it1["reg_normal_after_dist"], it1["reg_laplacian_after_dist"])):
        assert abs(x - y) <= 1e-6 * abs(y) + 1e-12, (regs, it1)
    s1 = vol1.download()
    surf = opt1.surface()
    d_rms = float(np.sqrt(np.mean((state[surf, 0].astype(np.float64) - s1[surf, 0]) ** 2))) / vs
    a_rms = float(np.sqrt(np.mean((state[surf, 5:8].astype(np.float64) - s1[surf, 5:8]) ** 2)))
    print(f"multi-gpu check: world={world} model={model} G={G} S={opt1.S} sharded={opt.is_sharded()[0]} rows={opt.row_range()} "
          f"kernel us {prof} CG iterations {[x[2] for x in log if len(x) == 3]} "
          f"worst relative energy difference {worst:.3e}, "
          f"final RMS distance {d_rms:.3e} voxels, albedo {a_rms:.3e}")
    assert worst < 5e-6 and d_rms < 1e-4 and a_rms < 1e-4
dist.barrier()
opt.comm_destroy()
dist.destroy_process_group()
