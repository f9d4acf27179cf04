"""GPU: the allocation-free energy entry point (psg_opt_energy_noalloc) against psg_opt_energy and against the oracle.

Who needs it: a host whose GPU memory is already fully handed out when the optimizer runs -- a framework's caching
allocator, a per-tenant partition, a process that pre-allocates everything before going live.  There a cudaMalloc from
inside the optimisation loop may fail or stall the device, so this variant keeps its scratch in an array that is part
of the loaded library instead.  The file is named so that pytest collects it last: if it fails, every other test has
already run.  It builds its own volume and optimizer, so it passes alone in a fresh process as well as after the rest.
"""
import ctypes as C

import numpy as np
import pytest

from tests.common import block_order, handoff, make_gpu, make_oracle, scene

pytestmark = pytest.mark.gpu


def _free_bytes():
    """Free device memory as the driver reports it (device-wide, so allocations made by the library show up)."""
    import torch
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info(0)[0]


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


def test_noalloc_energy_matches_the_oracle_and_allocates_nothing():
    """A fresh optimizer on which psg_opt_energy has never run: the entry point must not lean on buffers another call
    left behind.  One warm-up call (address look-up, kernel load), then ten calls between two readings of the driver's
    free-memory counter; the value is checked against the CPU restatement, not against the same kernel."""
    import psgradientsdf_b200 as psg
    model = "SH1"
    sc = scene(model)
    ov, oo = make_oracle(model, 32, sc)
    gv, go = make_gpu(model, 32, sc)
    handoff(oo, sc); handoff(go, sc)
    oo.init_albedo(); go.init_albedo()
    _, e_or, n_or = oo.energy()
    L = psg.lib()
    e = C.c_double(); n = C.c_longlong()
    assert L.psg_opt_energy_noalloc(go.h, C.byref(e), C.byref(n)) == 0, L.psg_last_error()   # warm-up
    free0 = _free_bytes()
    for _ in range(10):
        assert L.psg_opt_energy_noalloc(go.h, C.byref(e), C.byref(n)) == 0, L.psg_last_error()
        assert n.value == n_or
        assert abs(e.value - e_or) <= 1e-6 * abs(e_or), (e.value, e_or)
    assert _free_bytes() == free0, "device memory was allocated or freed across psg_opt_energy_noalloc calls"


def test_noalloc_rejects_what_it_cannot_hold():
    import psgradientsdf_b200 as psg
    from psgradientsdf_b200 import synth
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
    # a work list larger than the built-in array: every voxel of a 130^3 volume on the surface (dist = 0, seen in frame 0)
    G = 130
    vs, T = synth.grid_for_scene(G)
    big = psg.VolumetricGradSdf([G, G, G], vs, [0, 0, 0], T, 64)
    aos = np.zeros((G ** 3, 8), np.float32)
    aos[:, 3] = 1.0; aos[:, 4] = 1.0; aos[:, 5:8] = 0.5
    big.upload(aos)
    big.set_vis(np.ones((1, G ** 3), np.uint64), 0)
    opt = psg.Optimizer(big, vs, sc.K, model="SH1")
    opt.set_keyframes(np.arange(1, dtype=np.int32), sc.poses_init[:1], sc.color[:1])
    opt.init()
    assert opt.S == G ** 3 > 2097152
    assert L.psg_opt_energy_noalloc(opt.h, C.byref(e), C.byref(n)) == 1
    assert b"exceed the built-in scratch" in L.psg_last_error()
    assert opt.energy()[1] >= 0                                                  # the allocating entry point still serves it
