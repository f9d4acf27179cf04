"""CPU: the plain-C restatement (oracle/psg_oracle.c) against the committed reference fixtures."""
import numpy as np
import pytest

from tests import golden_common
from tests.common import MODELS, reg_for


@pytest.mark.parametrize("model", MODELS)
def test_oracle_matches_reference_fixtures(model):
    from oracle import pyoracle

    def mkvol(G, vs, T):
        return pyoracle.OrVolume([G, G, G], vs, [0, 0, 0], T, 0.5, 3.5, 64)

    def mkopt(vol, vs, K):
        return pyoracle.OrOptimizer(vol, model, vs, K, **reg_for(model))

    def integ(vol, f, inp, R, t):
        vol.integrate(inp["color"][f], inp["depth"][f], inp["nx"][f], inp["ny"][f], inp["nz"][f], inp["n_sq_inv"],
                      inp["K"], R, t, f > 0)

    golden_common.replay(model, mkvol, mkopt, integ, gpu=False)
