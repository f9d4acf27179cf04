"""Two ranks on two GPUs (when the box has them): the sharded and the replicated multi-GPU paths against the single-GPU run.
Runs tools/multi_gpu_check.py under torchrun; skipped on a one-GPU box."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


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
