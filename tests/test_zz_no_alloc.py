"""GPU: the allocation-free energy entry point (psg_opt_energy_noalloc) against psg_opt_energy.

Who needs it: a host whose GPU memory is already fully handed out when the optimizer runs -- a framework's caching
allocator, a per-tenant partition, a process that pre-allocates everything before going live.  There a cudaMalloc from
inside the optimisation loop may fail or stall the device, so this variant keeps its scratch in an array that is part
of the loaded library instead.  The file is named so that pytest collects it last: if it fails, every other test has
already run.  It builds its own volume and optimizer, so it passes alone in a fresh process as well as after the rest.
"""
import ctypes as C

import numpy as np
import pytest

from tests.common import block_order, handoff, make_gpu, scene

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model", ["SH1", "LED"])
def test_noalloc_energy_matches_allocating_entry_point(model):
    import psgradientsdf_b200 as psg
    sc = scene(model)
    gv, go = make_gpu(model, 32, sc)
    handoff(go, sc)
    go.init_albedo()
    L = psg.lib()
    for call in range(2):                      # two calls in a row share the built-in array: stale contents must not leak
        e_ref, n_ref = go.energy()             # existing entry point (may allocate)
        e = C.c_double(); n = C.c_longlong()
        st = L.psg_opt_energy_noalloc(go.h, C.byref(e), C.byref(n))   # straight through the C ABI
        assert st == 0, L.psg_last_error()
        assert n.value == n_ref, (call, n.value, n_ref)
        # both run the same kernel with a fixed-order reduction; allow the tolerance the other energy tests use
        assert abs(e.value - e_ref) <= 1e-6 * abs(e_ref), (call, e.value, e_ref)
        # change the state between the two calls so the second answer differs from the first
        for blk in block_order(model)[:2]:
            go.step(blk, damping=1.0, normal_reg=True, laplacian_reg=(model == "LED"))
    e1, n1 = go.energy_noalloc()
    e2, n2 = go.energy()
    assert n1 == n2 and abs(e1 - e2) <= 1e-6 * abs(e2)


def test_noalloc_rejects_what_it_cannot_hold():
    import psgradientsdf_b200 as psg
    L = psg.lib()
    e = C.c_double(); n = C.c_longlong()
    assert L.psg_opt_energy_noalloc(None, C.byref(e), C.byref(n)) == 1          # null handle
    sc = scene("SH1")
    gv, go = make_gpu("SH1", 16, sc)
    go.set_keyframes(np.arange(2, dtype=np.int32), sc.poses_init[:2], sc.color[:2])
    assert L.psg_opt_energy_noalloc(go.h, C.byref(e), C.byref(n)) == 3          # before psg_opt_init
    assert b"psg_opt_init" in L.psg_last_error()
    header = open(psg.HEADER_PATH).read()
    assert "2 097 152 surface voxels" in header and "psg_opt_energy_noalloc" in header
