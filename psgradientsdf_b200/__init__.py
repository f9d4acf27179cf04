"""psgradientsdf_b200 -- B200-native hot path of PSgradientSDF's photometric optimizer.

The product is the sm_100a shared library ``libpsg_b200.so`` (C ABI in ``include/psg.h``, sources in ``csrc/``).
This module is a thin ctypes binding of that ABI plus Python look-alikes of the reference's two classes
(``VolumetricGradSdf`` and ``PsOptimizer`` / ``LedOptimizer``, cpp/include/sdf_tracker/VolumetricGradSdf.h and
cpp/include/ps_optimizer/*.h) used by the tests and bench.py.  There is no CPU fallback: loading fails loudly when
the library has not been built, and every compute call fails with the library's error text when no CUDA device is
usable.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpsg_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "psg.h")

MODEL = {"SH1": 0, "SH2": 1, "LED": 2}
LOSS = {"l2": 0, "cauchy": 1, "huber": 2, "tukey": 3, "trunc_l2": 4}
BLOCK = {"albedo": 0, "light": 1, "dist": 2, "pose": 3}
REG_ALBEDO, REG_NORMAL, REG_LAPLACIAN = 1, 2, 4


class PsgError(RuntimeError):
    pass


class Settings(C.Structure):
    _fields_ = [("model", C.c_int), ("loss", C.c_int), ("lam", C.c_float), ("reg_weight_rho", C.c_float),
                ("reg_weight_n", C.c_float), ("reg_weight_l", C.c_float)]


class StepInfo(C.Structure):
    _fields_ = [("n", C.c_int), ("cg_success", C.c_int), ("cg_iterations", C.c_int), ("cg_error", C.c_float),
                ("applied", C.c_int), ("n_accepted", C.c_longlong), ("energy_before", C.c_double),
                ("n_active_before", C.c_longlong)]


class IterationConfig(C.Structure):
    _fields_ = [("light", C.c_int), ("albedo", C.c_int), ("distance", C.c_int), ("pose", C.c_int), ("damping", C.c_float),
                ("reg_flags", C.c_int), ("closing_energy", C.c_int)]


class IterationInfo(C.Structure):
    _fields_ = [("ran", C.c_int * 4), ("block", StepInfo * 4), ("energy_after", C.c_double * 4),
                ("reg_normal_after_dist", C.c_double), ("reg_laplacian_after_dist", C.c_double), ("energy_end", C.c_double),
                ("n_active_end", C.c_longlong), ("reg_albedo_after_albedo", C.c_double)]


class TrackerConfig(C.Structure):
    _fields_ = [("num_iterations", C.c_int), ("conv_threshold", C.c_float), ("damping", C.c_float), ("sampling", C.c_int)]


class LoopConfig(C.Structure):
    _fields_ = [("light", C.c_int), ("albedo", C.c_int), ("distance", C.c_int), ("pose", C.c_int),
                ("max_it", C.c_int), ("conv_threshold", C.c_float), ("damping", C.c_float), ("upsample", C.c_int),
                ("write_artifacts", C.c_int), ("verbose", C.c_int)]


class LoopResult(C.Structure):
    _fields_ = [("iterations", C.c_int), ("converged", C.c_int), ("diverged", C.c_int), ("last_rel_diff", C.c_float),
                ("E", C.c_double), ("E_normal", C.c_double), ("E_laplacian", C.c_double), ("E_albedo", C.c_double),
                ("E_total", C.c_double), ("ms_per_iteration", C.c_double), ("iteration_ms", C.c_float * 64)]


_lib = None
f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")


def lib():
    """Load libpsg_b200.so (built by psgradientsdf_b200.build / __graft_entry__.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PsgError(f"{LIB_PATH} is missing: run `python -m psgradientsdf_b200.build` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.psg_last_error.restype = C.c_char_p
        L.psg_kernel_launches.restype = C.c_longlong
        vp = C.c_void_p
        L.psg_volume_create.argtypes = [C.POINTER(vp), i32p, C.c_float, f32p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int]
        L.psg_volume_destroy.argtypes = [vp]
        L.psg_volume_geometry.argtypes = [vp, i32p, C.POINTER(C.c_float), f32p]
        L.psg_volume_integrate.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, f32p, f32p, f32p, C.c_int]
        L.psg_volume_download.argtypes = [vp, f32p]
        L.psg_volume_upload.argtypes = [vp, f32p]
        L.psg_volume_vis_words.argtypes = [vp, C.POINTER(C.c_int)]
        L.psg_volume_download_vis.argtypes = [vp, u64p]
        L.psg_volume_upload_vis.argtypes = [vp, u64p, C.c_int]
        L.psg_opt_create.argtypes = [C.POINTER(vp), vp, C.c_float, f32p, C.POINTER(Settings)]
        L.psg_opt_destroy.argtypes = [vp]
        L.psg_opt_set_keyframes.argtypes = [vp, C.c_int, i32p, f32p, vp, C.c_int, C.c_int]
        L.psg_opt_set_keytimestamps.argtypes = [vp, C.c_int, C.POINTER(C.c_char_p)]
        L.psg_opt_init.argtypes = [vp]
        L.psg_opt_num_surface.argtypes = [vp, C.POINTER(C.c_int)]
        L.psg_opt_get_surface.argtypes = [vp, vp, vp]
        L.psg_opt_light_len.argtypes = [vp, C.POINTER(C.c_int)]
        L.psg_opt_get_light.argtypes = [vp, f32p]
        L.psg_opt_set_light.argtypes = [vp, f32p]
        L.psg_opt_get_poses.argtypes = [vp, f32p]
        L.psg_opt_set_poses.argtypes = [vp, f32p]
        L.psg_opt_get_reg.argtypes = [vp, f32p]
        L.psg_opt_set_reg.argtypes = [vp, f32p]
        L.psg_opt_init_albedo.argtypes = [vp]
        L.psg_opt_set_sweep_split.argtypes = [vp, C.c_int]
        L.psg_opt_set_pcg_mode.argtypes = [vp, C.c_int]
        L.psg_opt_energy.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
        L.psg_opt_energy_noalloc.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
        L.psg_opt_reg_energies.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.psg_opt_residual.argtypes = [vp, f32p, f32p]
        L.psg_opt_probe_pairs.argtypes = [vp, f32p, f32p, vp, f32p, f32p]
        L.psg_opt_step.argtypes = [vp, C.c_int, C.c_float, C.c_int, C.POINTER(StepInfo)]
        L.psg_normals_create.argtypes = [C.POINTER(vp), C.c_int, C.c_int, f32p, C.c_int, C.c_int]
        L.psg_normals_destroy.argtypes = [vp]
        L.psg_normals_tables.argtypes = [vp, f32p]
        L.psg_normals_compute.argtypes = [vp, f32p, f32p, f32p, f32p]
        L.psg_volume_integrate_raw.argtypes = [vp, vp, vp, vp, C.c_float, f32p, f32p, f32p, C.c_int]
        L.psg_opt_set_keyframes_u8.argtypes = [vp, C.c_int, i32p, f32p, vp, C.c_int, C.c_int]
        L.psg_tracker_set_depth.argtypes = [vp, f32p, C.c_int, C.c_int]
        L.psg_tracker_set_depth_u16.argtypes = [vp, vp, C.c_float, C.c_int, C.c_int]
        L.psg_tracker_system.argtypes = [vp, f32p, f32p, f32p, C.c_int, C.POINTER(C.c_double), vp, vp, C.POINTER(C.c_longlong)]
        L.psg_opt_get_surface_voxels.argtypes = [vp, f32p]
        L.psg_opt_set_shard_policy.argtypes = [vp, C.c_int]
        L.psg_opt_is_sharded.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_longlong)]
        L.psg_opt_set_upload_overlap.argtypes = [vp, C.c_int]
        L.psg_tracker_optimize.argtypes = [vp, f32p, f32p, C.POINTER(TrackerConfig), f32p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.psg_normals_sharpness.argtypes = [vp, f32p, C.POINTER(C.c_float)]
        L.psg_normals_sharpness_u8.argtypes = [vp, vp, C.POINTER(C.c_float)]
        L.psg_compute_centroid.argtypes = [f32p, f32p, C.c_int, C.c_int, f32p, f32p]
        L.psg_sample_keyframes.argtypes = [C.c_int, C.c_int, i32p, C.POINTER(C.c_int)]
        L.psg_opt_iteration.argtypes = [vp, C.POINTER(IterationConfig), C.POINTER(IterationInfo)]
        LL = C.POINTER(C.c_longlong)
        L.psg_volume_mesh_box.argtypes = [vp, C.c_float, i32p, i32p, C.POINTER(C.c_int)]
        L.psg_volume_mesh_extract.argtypes = [vp, C.c_float, C.c_int, LL]
        L.psg_volume_mesh_download.argtypes = [vp, vp, vp]
        L.psg_volume_extract_mesh.argtypes = [vp, C.c_char_p, C.c_int, LL]
        L.psg_volume_extract_pc.argtypes = [vp, C.c_char_p, LL]
        L.psg_volume_save_sdf.argtypes = [vp, C.c_char_p]
        L.psg_opt_extract_mesh.argtypes = [vp, C.c_char_p, C.c_int, LL]
        L.psg_opt_save_sdf.argtypes = [vp, C.c_char_p]
        L.psg_opt_save_pointcloud.argtypes = [vp, C.c_char_p]
        L.psg_opt_last_system.argtypes = [vp, vp, vp, vp]
        L.psg_opt_last_vectors.argtypes = [vp, vp, vp]
        L.psg_opt_upsample.argtypes = [vp]
        L.psg_debug_timer_begin.argtypes = [vp]
        L.psg_debug_timer_end.argtypes = [vp, C.POINTER(C.c_float)]
        L.psg_opt_alternating_optimize.argtypes = [vp, C.POINTER(LoopConfig), C.c_char_p, C.POINTER(LoopResult)]
        L.psg_comm_unique_id.argtypes = [C.c_char_p]
        L.psg_opt_comm_init.argtypes = [vp, C.c_char_p, C.c_int, C.c_int]
        L.psg_opt_comm_destroy.argtypes = [vp]
        L.psg_opt_comm_init_local.argtypes = [C.POINTER(vp), C.c_int]
        L.psg_opt_row_range.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.psg_opt_sync_volume.argtypes = [vp]
        L.psg_partition_rows.argtypes = [C.c_int, i32p, u64p, C.c_longlong, C.c_int, i32p, i32p, i32p]
        L.psg_shard_blocks.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.psg_opt_energy_partial.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
        L.psg_profile_enable.argtypes = [C.c_int]
        L.psg_profile_name.restype = C.c_char_p
        L.psg_profile_name.argtypes = [C.c_int]
        L.psg_profile_get.argtypes = [C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_double)]
        _lib = L
    return _lib


def _check(status: int) -> None:
    if status != 0:
        raise PsgError(f"psg status {status}: {lib().psg_last_error().decode(errors='replace')}")


def kernel_launches() -> int:
    return int(lib().psg_kernel_launches())


def comm_unique_id() -> bytes:
    """NCCL unique id (128 bytes) created on this rank; broadcast it to the other ranks and pass it to comm_init."""
    buf = C.create_string_buffer(128)
    _check(lib().psg_comm_unique_id(buf))
    return buf.raw


def comm_init_local(opts):
    """The given optimizers (one device) become ranks 0 .. n-1 of one job whose collectives are met by n threads, one per optimizer
    (psg_opt_comm_init_local)."""
    arr = (C.c_void_p * len(opts))(*[o.h for o in opts])
    _check(lib().psg_opt_comm_init_local(arr, len(opts)))


def partition_rows(lin, vis, plane: int, world: int):
    """(bounds[world + 1], halo_lo[world], halo_hi[world]) of psg_partition_rows."""
    lin = np.ascontiguousarray(lin, np.int32); vis = np.ascontiguousarray(vis, np.uint64)
    b = np.zeros(world + 1, np.int32); lo = np.zeros(world, np.int32); hi = np.zeros(world, np.int32)
    _check(lib().psg_partition_rows(int(lin.size), lin, vis, int(plane), int(world), b, lo, hi))
    return b, lo, hi


def run_ranks(fn, n: int):
    """fn(rank) on n threads at once (the collective calls of a local rank group must be made concurrently); re-raises the first
    exception; returns the list of results."""
    import threading
    out, err = [None] * n, [None] * n

    def work(r):
        try:
            out[r] = fn(r)
        except BaseException as e:   # noqa: BLE001
            err[r] = e
    th = [threading.Thread(target=work, args=(r,)) for r in range(n)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for e in err:
        if e is not None:
            raise e
    return out


def shard_blocks(S: int, rank: int, world: int):
    """(first, stride, count): the 32-slot blocks of the work list that `rank` of `world` sweeps."""
    f, st, c = C.c_int(), C.c_int(), C.c_int()
    _check(lib().psg_shard_blocks(int(S), int(rank), int(world), C.byref(f), C.byref(st), C.byref(c)))
    return f.value, st.value, c.value


def compute_centroid(K, depth, pose):
    """compute_centroid (main_ps.cpp:346-375): host code, no GPU needed."""
    d = np.ascontiguousarray(depth, np.float32)
    out = np.zeros(3, np.float32)
    _check(lib().psg_compute_centroid(np.ascontiguousarray(K, np.float32).reshape(-1), d, d.shape[0], d.shape[1],
                                      np.ascontiguousarray(pose, np.float32).reshape(-1), out))
    return out


def sample_keyframes(n: int, max_num: int = 40):
    """sampleKeyFrame (main_ps.cpp:392-421): positions of the keyframes that are kept."""
    kept = np.zeros(max(n, max_num) + 1, np.int32)
    m = C.c_int()
    _check(lib().psg_sample_keyframes(int(n), int(max_num), kept, C.byref(m)))
    return kept[:m.value].copy()


def host_times(reset: bool = True) -> dict:
    """Where the host thread's time went inside psg_opt_iteration (microseconds per iteration since the last reset)."""
    us = (C.c_double * 3)()
    it = C.c_longlong()
    lib().psg_debug_host_times.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.c_int]
    _check(lib().psg_debug_host_times(us, C.byref(it), int(reset)))
    n = max(it.value, 1)
    return {"iterations": it.value, "queue_us": us[0] / n, "wait_us": us[1] / n, "host_solve_us": us[2] / n}


def small_cg(H_blocks, b):
    """The host solve of the light / pose blocks (psg_debug_small_cg) on (nblocks, bs, bs) dense symmetric blocks: host code only,
    no device needed.  Returns (x, success, iterations, error)."""
    Hb = np.ascontiguousarray(H_blocks, np.float32)
    bb = np.ascontiguousarray(b, np.float32).reshape(-1)
    nblocks, bs = Hb.shape[0], Hb.shape[1]
    assert Hb.shape == (nblocks, bs, bs) and bb.size == nblocks * bs
    x = np.zeros(bb.size, np.float32)
    ok, it, err = C.c_int(), C.c_int(), C.c_float()
    L = lib()
    L.psg_debug_small_cg.argtypes = [C.c_int, C.c_int, f32p, f32p, f32p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_float)]
    _check(L.psg_debug_small_cg(nblocks, bs, Hb.reshape(-1), bb, x, C.byref(ok), C.byref(it), C.byref(err)))
    return x, bool(ok.value), it.value, err.value


def profile_enable(on: bool = True) -> None:
    lib().psg_profile_enable(int(on))


def profile_reset() -> None:
    lib().psg_profile_reset()


def profile_table() -> dict:
    """{kernel name: (launches, total device ms)} from the built-in CUDA-event launch timer."""
    out = {}
    for i in range(lib().psg_profile_kernels()):
        n, ms = C.c_longlong(), C.c_double()
        _check(lib().psg_profile_get(i, C.byref(n), C.byref(ms)))
        if n.value:
            out[lib().psg_profile_name(i).decode()] = (n.value, ms.value)
    return out


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class VolumetricGradSdf:
    """Look-alike of the reference's volume class (VolumetricGradSdf.h:65-122) over psg_volume_*."""

    def __init__(self, grid_dim, voxel_size, shift, T, max_frames=64, device=0):
        self.h = C.c_void_p()
        self.grid_dim = np.ascontiguousarray(grid_dim, np.int32)
        self._zmin, self._zmax = 0.5, 10.0   # Sdf.h:39-40 defaults
        self._args = (self.grid_dim, float(voxel_size), np.ascontiguousarray(shift, np.float32), float(T), int(max_frames), int(device))
        self._created = False
        self._pending_counter = False

    def set_zmin(self, z):
        self._zmin = float(z)

    def set_zmax(self, z):
        self._zmax = float(z)

    def _ensure(self):
        if not self._created:
            gd, vs, sh, T, mf, dev = self._args
            _check(lib().psg_volume_create(C.byref(self.h), gd, vs, sh, T, self._zmin, self._zmax, mf, dev))
            self._created = True

    def close(self):
        if self._created:
            lib().psg_volume_destroy(self.h)
            self._created = False

    def increase_counter(self):
        self._pending_counter = True

    def update(self, color, depth, K, R, t, nx, ny, nz, n_sq_inv):
        """Sdf::update(color, depth, K, pose, NEst) with the estimator's outputs passed explicitly."""
        self._ensure()
        H, W = depth.shape
        c = lambda a: np.ascontiguousarray(a, np.float32)
        arrs = [c(color), c(depth), c(nx), c(ny), c(nz), c(n_sq_inv)]
        _check(lib().psg_volume_integrate(self.h, *[_ptr(a) for a in arrs], H, W, c(K).reshape(9), c(R).reshape(9),
                                          c(t).reshape(3), int(self._pending_counter)))
        self._pending_counter = False

    def tracker_set_depth(self, depth=None, depth_u16=None, depth_unit=None):
        """The frame RigidPointOptimizer::optimize tracks: float depth in metres, or 16-bit depth + unit."""
        self._ensure()
        if depth_u16 is not None:
            d = np.ascontiguousarray(depth_u16, np.uint16)
            _check(lib().psg_tracker_set_depth_u16(self.h, d.ctypes.data, float(depth_unit), d.shape[0], d.shape[1]))
        else:
            d = np.ascontiguousarray(depth, np.float32)
            _check(lib().psg_tracker_set_depth(self.h, d, d.shape[0], d.shape[1]))

    def tracker_system(self, K, R, t, sampling=1):
        """One pass of optimize_sampled's inner loop at pose (R, t): (E_sum, g[6], H[6,6], count), sums in float64."""
        E, cnt = C.c_double(), C.c_longlong()
        g, Hm = np.zeros(6), np.zeros((6, 6))
        c = lambda a: np.ascontiguousarray(a, np.float32).reshape(-1)
        _check(lib().psg_tracker_system(self.h, c(K), c(R), c(t), int(sampling), C.byref(E), g.ctypes.data, Hm.ctypes.data, C.byref(cnt)))
        return E.value, g, Hm, cnt.value

    def tracker_optimize(self, K, pose, num_iterations=50, conv_threshold=1e-3, damping=1.0, sampling=1):
        """RigidPointOptimizer::optimize on the frame given to tracker_set_depth: (converged, iterations, pose 4x4)."""
        cfg = TrackerConfig(int(num_iterations), float(conv_threshold), float(damping), int(sampling))
        out = np.zeros(16, np.float32)
        conv, it = C.c_int(), C.c_int()
        c = lambda a: np.ascontiguousarray(a, np.float32).reshape(-1)
        _check(lib().psg_tracker_optimize(self.h, c(K), c(pose), C.byref(cfg), out, C.byref(conv), C.byref(it)))
        return bool(conv.value), it.value, out.reshape(4, 4)

    # ---- surface extraction
    def mesh_box(self, voxel_size=0.0):
        """Crop box of extract_mesh / saveSDF: (lo[3], hi[3]) or None when no voxel lies within sqrt(3) * vs of the surface."""
        self._ensure()
        lo, hi, any_ = np.zeros(3, np.int32), np.zeros(3, np.int32), C.c_int()
        _check(lib().psg_volume_mesh_box(self.h, float(voxel_size), lo, hi, C.byref(any_)))
        return (lo, hi) if any_.value else None

    def mesh(self, white=False, voxel_size=0.0):
        """Marching cubes on the device: (vertices 3T x 3 float32, colours 3T x 3 uint8); face t is (3t, 3t+1, 3t+2)."""
        self._ensure()
        nt = C.c_longlong()
        _check(lib().psg_volume_mesh_extract(self.h, float(voxel_size), int(white), C.byref(nt)))
        v, c = np.zeros((3 * nt.value, 3), np.float32), np.zeros((3 * nt.value, 3), np.uint8)
        if nt.value:
            _check(lib().psg_volume_mesh_download(self.h, v.ctypes.data, c.ctypes.data))
        return v, c

    def extract_mesh(self, filename="mesh.ply", white=False):
        """VolumetricGradSdf::extract_mesh: the reference's PLY file; False (and no file) when there is no triangle."""
        self._ensure()
        nt = C.c_longlong()
        _check(lib().psg_volume_extract_mesh(self.h, str(filename).encode(), int(white), C.byref(nt)))
        return nt.value > 0

    def extract_pc(self, filename="cloud.ply"):
        self._ensure()
        n = C.c_longlong()
        _check(lib().psg_volume_extract_pc(self.h, str(filename).encode(), C.byref(n)))
        return True

    def saveSDF(self, filename="sdf.sdf"):
        self._ensure()
        _check(lib().psg_volume_save_sdf(self.h, str(filename).encode()))
        return True

    def update_raw(self, color_u8, depth_u16, depth_unit, K, R, t, NEst: "NormalEstimator"):
        """load_color + load_depth + NEst->compute + update() from the loader's raw pixels (psg_volume_integrate_raw)."""
        self._ensure()
        color_u8 = np.ascontiguousarray(color_u8, np.uint8)
        depth_u16 = np.ascontiguousarray(depth_u16, np.uint16)
        assert depth_u16.shape == (NEst.H, NEst.W) and color_u8.shape == (NEst.H, NEst.W, 3)
        _check(lib().psg_volume_integrate_raw(self.h, NEst.h, color_u8.ctypes.data, depth_u16.ctypes.data, float(depth_unit),
                                              np.ascontiguousarray(K, np.float32).reshape(-1),
                                              np.ascontiguousarray(R, np.float32).reshape(-1),
                                              np.ascontiguousarray(t, np.float32).reshape(-1), int(self._pending_counter)))
        self._pending_counter = False

    @property
    def N(self):
        dims, _, _ = self.geometry()
        return int(dims[0]) * int(dims[1]) * int(dims[2])

    def geometry(self):
        self._ensure()
        dims = np.zeros(3, np.int32)
        vs = C.c_float()
        org = np.zeros(3, np.float32)
        _check(lib().psg_volume_geometry(self.h, dims, C.byref(vs), org))
        return dims, np.float32(vs.value), org

    def download(self):
        self._ensure()
        a = np.zeros((self.N, 8), np.float32)
        _check(lib().psg_volume_download(self.h, a))
        return a

    def upload(self, aos):
        self._ensure()
        _check(lib().psg_volume_upload(self.h, np.ascontiguousarray(aos, np.float32)))

    def vis(self):
        self._ensure()
        nw = C.c_int()
        _check(lib().psg_volume_vis_words(self.h, C.byref(nw)))
        w = np.zeros((nw.value, self.N), np.uint64)
        _check(lib().psg_volume_download_vis(self.h, w))
        return w

    def set_vis(self, words, frame_counter):
        self._ensure()
        _check(lib().psg_volume_upload_vis(self.h, np.ascontiguousarray(words, np.uint64), int(frame_counter)))


class NormalEstimator:
    """cv::NormalEstimator<float> (normals/NormalEstimator.h:125-176) on the device."""

    def __init__(self, width, height, K, window=11, device=0):
        self.W, self.H, self.window = int(width), int(height), int(window)
        self.h = C.c_void_p()
        _check(lib().psg_normals_create(C.byref(self.h), self.W, self.H, np.ascontiguousarray(K, np.float32).reshape(-1),
                                        self.window, int(device)))

    def close(self):
        if self.h:
            lib().psg_normals_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def tables(self):
        """(9, H, W): x0_n_sq_inv, y0_n_sq_inv, n_sq_inv, Q11, Q12, Q13, Q22, Q23, Q33."""
        t = np.zeros((9, self.H, self.W), np.float32)
        _check(lib().psg_normals_tables(self.h, t))
        return t

    def n_sq_inv(self):
        return self.tables()[2]

    def sharpness(self, color):
        """modifiedLaplacian (SharpDetector.h:22-37) of a float BGR image, or of the 8-bit image the loader read."""
        m = C.c_float()
        if color.dtype == np.uint8:
            c = np.ascontiguousarray(color)
            assert c.shape == (self.H, self.W, 3)
            _check(lib().psg_normals_sharpness_u8(self.h, c.ctypes.data, C.byref(m)))
        else:
            c = np.ascontiguousarray(color, np.float32)
            assert c.shape == (self.H, self.W, 3)
            _check(lib().psg_normals_sharpness(self.h, c, C.byref(m)))
        return m.value

    def compute(self, depth):
        depth = np.ascontiguousarray(depth, np.float32)
        assert depth.shape == (self.H, self.W)
        nx, ny, nz = (np.zeros((self.H, self.W), np.float32) for _ in range(3))
        _check(lib().psg_normals_compute(self.h, depth, nx, ny, nz))
        return nx, ny, nz


class Optimizer:
    """Look-alike of Optimizer / PsOptimizer / LedOptimizer (Optimizer.h:24-182) over psg_opt_*."""

    def __init__(self, tSDF: VolumetricGradSdf, voxel_size, K, model="SH1", loss="cauchy", lam=0.2, reg_rho=0.0,
                 reg_n=0.0, reg_l=0.0):
        tSDF._ensure()
        self.vol = tSDF
        self.model = model
        self.h = C.c_void_p()
        st = Settings(MODEL[model], LOSS[loss], lam, reg_rho, reg_n, reg_l)
        _check(lib().psg_opt_create(C.byref(self.h), tSDF.h, float(voxel_size), np.ascontiguousarray(K, np.float32).reshape(9),
                                    C.byref(st)))
        self._F = 0

    def close(self):
        if self.h:
            lib().psg_opt_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):   # an optimizer holds one of the device's few keyframe-table slots (psg_opt_create): give it back
        try:
            self.close()
        except Exception:
            pass

    # setImages + setKeyframes + setPoses (main_ps.cpp:323-326)
    def set_keyframes(self, frame_idx, poses, images):
        F, H, W, _ = images.shape
        self._F = F
        img = np.ascontiguousarray(images, np.float32)
        _check(lib().psg_opt_set_keyframes(self.h, F, np.ascontiguousarray(frame_idx, np.int32),
                                           np.ascontiguousarray(poses, np.float32).reshape(-1), _ptr(img), H, W))
        self._pending_images = img      # with set_upload_overlap the copy may still be running: keep the buffer until init()

    def set_keyframes_u8(self, frame_idx, poses, images_u8):
        """setImages with the 8-bit BGR pixels the loader read (psg_opt_set_keyframes_u8)."""
        F, H, W, _ = images_u8.shape
        self._F = F
        img = np.ascontiguousarray(images_u8, np.uint8)
        _check(lib().psg_opt_set_keyframes_u8(self.h, F, np.ascontiguousarray(frame_idx, np.int32),
                                              np.ascontiguousarray(poses, np.float32).reshape(-1), img.ctypes.data, H, W))
        self._pending_images = img

    # setKeytimestamps (main_ps.cpp:325): the strings savePoses prints in front of every pose
    def set_keytimestamps(self, stamps):
        arr = (C.c_char_p * len(stamps))(*[str(s).encode() for s in stamps])
        _check(lib().psg_opt_set_keytimestamps(self.h, len(stamps), arr))

    def init(self):
        _check(lib().psg_opt_init(self.h))
        self._pending_images = None

    @property
    def S(self):
        s = C.c_int()
        _check(lib().psg_opt_num_surface(self.h, C.byref(s)))
        return s.value

    @property
    def F(self):
        return self._F

    def surface(self):
        s = np.zeros(self.S, np.int32)
        _check(lib().psg_opt_get_surface(self.h, _ptr(s), None))
        return s

    def surface_voxels(self):
        """(S, 8) SdfVoxel records of the surface voxels, in surface-list order (psg_opt_get_surface_voxels)."""
        a = np.zeros((self.S, 8), np.float32)
        _check(lib().psg_opt_get_surface_voxels(self.h, a))
        return a

    # Optimizer::extract_mesh / save_white_mesh / save_pointcloud / saveSDF (OptimizerAux.cpp:278-577); the reference prefixes
    # save_path_ and appends _mesh.ply / _obj.ply / _pointcloud.ply itself, here the caller passes the whole file name
    def extract_mesh(self, filename, white=False):
        nt = C.c_longlong()
        _check(lib().psg_opt_extract_mesh(self.h, str(filename).encode(), int(white), C.byref(nt)))
        return nt.value > 0

    def save_white_mesh(self, filename):
        return self.extract_mesh(filename, white=True)

    def save_pointcloud(self, filename):
        _check(lib().psg_opt_save_pointcloud(self.h, str(filename).encode()))
        return True

    def saveSDF(self, filename):
        _check(lib().psg_opt_save_sdf(self.h, str(filename).encode()))
        return True

    def vis_keyframes(self):
        v = np.zeros(self.S, np.uint64)
        _check(lib().psg_opt_get_surface(self.h, None, _ptr(v)))
        return v

    def light(self):
        n = C.c_int()
        _check(lib().psg_opt_light_len(self.h, C.byref(n)))
        l = np.zeros(n.value, np.float32)
        _check(lib().psg_opt_get_light(self.h, l))
        return l

    def set_light(self, l):
        _check(lib().psg_opt_set_light(self.h, np.ascontiguousarray(l, np.float32).reshape(-1)))

    def poses(self):
        p = np.zeros(self.F * 16, np.float32)
        _check(lib().psg_opt_get_poses(self.h, p))
        return p.reshape(self.F, 4, 4)

    def set_poses(self, p):
        _check(lib().psg_opt_set_poses(self.h, np.ascontiguousarray(p, np.float32).reshape(-1)))

    def reg(self):
        r = np.zeros(3, np.float32)
        _check(lib().psg_opt_get_reg(self.h, r))
        return r

    def set_reg(self, rho, n, l):
        _check(lib().psg_opt_set_reg(self.h, np.array([rho, n, l], np.float32)))

    def init_albedo(self):
        _check(lib().psg_opt_init_albedo(self.h))

    def set_upload_overlap(self, on: bool):
        """Queue keyframe uploads behind psg_opt_init's work-list build; the arrays given to set_keyframes* must then stay alive
        (and pinned, to actually overlap) until init() has returned."""
        _check(lib().psg_opt_set_upload_overlap(self.h, int(on)))
        self._overlap = bool(on)

    def set_pcg_mode(self, mode: int):
        _check(lib().psg_opt_set_pcg_mode(self.h, int(mode)))

    def set_sweep_split(self, warps: int):
        _check(lib().psg_opt_set_sweep_split(self.h, int(warps)))

    def energy(self):
        e = C.c_double()
        na = C.c_longlong()
        _check(lib().psg_opt_energy(self.h, C.byref(e), C.byref(na)))
        return e.value, na.value

    def comm_init(self, unique_id: bytes, rank: int, world: int):
        """Join the NCCL communicator: from here on init_albedo / energy / step / alternatingOptimize are collective."""
        _check(lib().psg_opt_comm_init(self.h, C.create_string_buffer(unique_id, 128), int(rank), int(world)))

    def set_shard_policy(self, policy: int):
        """0 = shard the sweeps when there is enough work per rank, 1 = always, 2 = never (psg_opt_set_shard_policy)."""
        _check(lib().psg_opt_set_shard_policy(self.h, int(policy)))

    def is_sharded(self):
        sh, vp_ = C.c_int(), C.c_longlong()
        _check(lib().psg_opt_is_sharded(self.h, C.byref(sh), C.byref(vp_)))
        return bool(sh.value), vp_.value

    def row_range(self):
        """(row_lo, row_hi, halo_lo, halo_hi) of this rank (psg_opt_row_range)."""
        a, b, c, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        _check(lib().psg_opt_row_range(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return a.value, b.value, c.value, d.value

    def sync_volume(self):
        _check(lib().psg_opt_sync_volume(self.h))

    def comm_destroy(self):
        _check(lib().psg_opt_comm_destroy(self.h))

    def energy_partial(self, rank: int, world: int):
        s = C.c_double()
        n = C.c_longlong()
        _check(lib().psg_opt_energy_partial(self.h, int(rank), int(world), C.byref(s), C.byref(n)))
        return s.value, n.value

    def energy_noalloc(self):
        """psg_opt_energy_noalloc: same value as energy(), computed without any device allocation (see include/psg.h)."""
        e = C.c_double()
        na = C.c_longlong()
        _check(lib().psg_opt_energy_noalloc(self.h, C.byref(e), C.byref(na)))
        return e.value, na.value

    def reg_energies(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        _check(lib().psg_opt_reg_energies(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def residual(self):
        r = np.zeros((self.S, self.F, 3), np.float32)
        w = np.zeros((self.S, self.F, 3), np.float32)
        _check(lib().psg_opt_residual(self.h, r, w))
        return r, w

    def probe_pairs(self):
        S, F = self.S, self.F
        m = np.zeros((S, F), np.float32)
        n = np.zeros((S, F), np.float32)
        ok = np.zeros((S, F), np.uint8)
        I = np.zeros((S, F, 3), np.float32)
        rend = np.zeros((S, F, 3), np.float32)
        _check(lib().psg_opt_probe_pairs(self.h, m, n, _ptr(ok), I, rend))
        return m, n, ok, I, rend

    def image_probe(self, frame, rows, cols):
        """(rgb, g0, g1) of interpolateImage / computeImageGradient at the given positions of keyframe `frame` (psg_debug_image_probe)."""
        rows = np.ascontiguousarray(rows, np.float32); cols = np.ascontiguousarray(cols, np.float32)
        n = rows.size
        rgb = np.zeros((n, 3), np.float32); g0 = np.zeros((n, 3), np.float32); g1 = np.zeros((n, 3), np.float32)
        L = lib()
        L.psg_debug_image_probe.argtypes = [C.c_void_p, C.c_int, C.c_int, f32p, f32p, f32p, f32p, f32p]
        _check(L.psg_debug_image_probe(self.h, int(frame), n, rows, cols, rgb.reshape(-1), g0.reshape(-1), g1.reshape(-1)))
        return rgb, g0, g1

    def step(self, which, damping=1.0, albedo_reg=False, normal_reg=False, laplacian_reg=False):
        flags = (REG_ALBEDO if albedo_reg else 0) | (REG_NORMAL if normal_reg else 0) | (REG_LAPLACIAN if laplacian_reg else 0)
        info = StepInfo()
        _check(lib().psg_opt_step(self.h, BLOCK[which], damping, flags, C.byref(info)))
        return dict(n=info.n, cg_success=bool(info.cg_success), iters=info.cg_iterations, err=info.cg_error,
                    applied=bool(info.applied), n_accepted=info.n_accepted, energy_before=info.energy_before,
                    n_active_before=info.n_active_before)

    def iteration(self, light=True, albedo=True, distance=True, pose=True, damping=1.0, normal_reg=False, laplacian_reg=False,
                  albedo_reg=False):
        """One pass of alternatingOptimize's while-body (psg_opt_iteration): dict with per-block step records, the energy
        after each block, the regulariser energies after the distance block and the closing energy."""
        cfg = IterationConfig(int(light), int(albedo), int(distance), int(pose), float(damping),
                              (REG_NORMAL if normal_reg else 0) | (REG_LAPLACIAN if laplacian_reg else 0) |
                              (REG_ALBEDO if albedo_reg else 0), 0)
        info = IterationInfo()
        _check(lib().psg_opt_iteration(self.h, C.byref(cfg), C.byref(info)))
        out = {"energy_end": info.energy_end, "n_active_end": info.n_active_end,
               "reg_normal_after_dist": info.reg_normal_after_dist, "reg_laplacian_after_dist": info.reg_laplacian_after_dist,
               "reg_albedo_after_albedo": info.reg_albedo_after_albedo}
        for name, b in BLOCK.items():
            if info.ran[b]:
                st = info.block[b]
                out[name] = dict(n=st.n, cg_success=bool(st.cg_success), iters=st.cg_iterations, err=st.cg_error,
                                 applied=bool(st.applied), n_accepted=st.n_accepted, energy_before=st.energy_before,
                                 n_active_before=st.n_active_before, energy_after=info.energy_after[b])
        return out

    def iteration_config(self, light=True, albedo=True, distance=True, pose=True, damping=1.0, normal_reg=False, laplacian_reg=False):
        """(psg_iteration_config, psg_iteration_info) to reuse with iteration_raw."""
        return (IterationConfig(int(light), int(albedo), int(distance), int(pose), float(damping),
                                (REG_NORMAL if normal_reg else 0) | (REG_LAPLACIAN if laplacian_reg else 0), 0), IterationInfo())

    def iteration_raw(self, cfg, info):
        """psg_opt_iteration on caller-held structs: what a C / C++ host's loop does per pass, without building Python dicts."""
        _check(lib().psg_opt_iteration(self.h, C.byref(cfg), C.byref(info)))
        return info

    def last_system(self, which):
        """(H_blocks, b, delta) of the last light / pose step as float64 arrays."""
        led = self.model == "LED"
        bs = 6 if which == "pose" else (3 if led else (9 if self.model == "SH2" else 4))
        nblk = 1 if (which == "light" and led) else self.F
        H = np.zeros((nblk, bs, bs))
        b = np.zeros(nblk * bs)
        d = np.zeros(nblk * bs)
        _check(lib().psg_opt_last_system(self.h, _ptr(H), _ptr(b), _ptr(d)))
        return H, b, d

    def last_vectors(self, which):
        n = 3 * self.S if which == "albedo" else self.S
        b = np.zeros(n, np.float32)
        d = np.zeros(n, np.float32)
        _check(lib().psg_opt_last_vectors(self.h, _ptr(b), _ptr(d)))
        return b, d

    def upsample(self):
        _check(lib().psg_opt_upsample(self.h))

    def last_pcg_kernel(self) -> int:
        k = C.c_int()
        lib().psg_debug_last_pcg_kernel.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        _check(lib().psg_debug_last_pcg_kernel(self.h, C.byref(k)))
        return k.value

    def timer_begin(self):
        """CUDA event on the optimizer's stream (psg_debug_timer_begin)."""
        _check(lib().psg_debug_timer_begin(self.h))

    def timer_end(self) -> float:
        """Second event, wait, device milliseconds since timer_begin."""
        ms = C.c_float()
        _check(lib().psg_debug_timer_end(self.h, C.byref(ms)))
        return ms.value

    def alternatingOptimize(self, light=True, albedo=True, distance=True, pose=True, max_it=100, conv_threshold=1e-4,
                            damping=1.0, upsample=False, save_path=None, verbose=False):
        """{Ps,Led}Optimizer::alternatingOptimize; returns the LoopResult as a dict."""
        cfg = LoopConfig(int(light), int(albedo), int(distance), int(pose), int(max_it), float(conv_threshold),
                         float(damping), int(upsample), int(save_path is not None), int(verbose))
        res = LoopResult()
        _check(lib().psg_opt_alternating_optimize(self.h, C.byref(cfg), save_path.encode() if save_path else None,
                                                  C.byref(res)))
        d = {k: getattr(res, k) for k, _ in LoopResult._fields_ if k != "iteration_ms"}
        d["iteration_ms"] = [res.iteration_ms[i] for i in range(min(res.iterations, 64))]
        return d
