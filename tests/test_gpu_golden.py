"""GPU: the CUDA path, called through the C ABI (ctypes over libpsg_b200.so), against the committed reference fixtures."""
import numpy as np
import pytest

from tests import golden_common
from tests.common import MODELS, reg_for

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("split", [1, 0])
@pytest.mark.parametrize("model", MODELS)
def test_cuda_matches_reference_fixtures(model, split):
    """split = 1: one warp per 32 voxels walks the keyframes in order (the reference's summation order, bit-exact
    per-voxel sums); split = 0: the launch-time choice (8 warps at this size)."""
    import psgradientsdf_b200 as psg

    def mkvol(G, vs, T):
        v = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64)
        v.set_zmin(0.5); v.set_zmax(3.5)
        return v

    def mkopt(vol, vs, K):
        o = psg.Optimizer(vol, vs, K, model=model, **reg_for(model))
        o.set_sweep_split(split)
        return o

    def integ(vol, f, inp, R, t):
        if f > 0:
            vol.increase_counter()
        vol.update(inp["color"][f], inp["depth"][f], inp["K"], R, t, inp["nx"][f], inp["ny"][f], inp["nz"][f], inp["n_sq_inv"])

    before = psg.kernel_launches()
    golden_common.replay(model, mkvol, mkopt, integ, gpu=True, exact_order=(split == 1))
    assert psg.kernel_launches() > before, "the CUDA library launched no kernels"
