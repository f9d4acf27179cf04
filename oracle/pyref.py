"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/_ref/libpsgref.so (the reference's own sources +
oracle/ref_harness.cpp).  Imported by tests/, bench.py's cpu_baseline / --impl reference legs and
__graft_entry__.smoke() only -- never by the product package."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PSG_REF_LIB") or os.path.join(_HERE, "_ref", "libpsgref.so")   # PSG_REF_LIB: an alternative build (e.g. -DNDEBUG)

_lib = None

f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref_vol_create.restype = C.c_void_p
        L.ref_vol_create.argtypes = [i32p, C.c_float, f32p, C.c_float, C.c_float, C.c_float]
        L.ref_vol_destroy.argtypes = [C.c_void_p]
        L.ref_se3_from_matrix.argtypes = [f32p, f32p, f32p]
        L.ref_compute_centroid.argtypes = [f32p, f32p, C.c_int, C.c_int, f32p, f32p]
        L.ref_tracker_system.argtypes = [C.c_void_p, f32p, C.c_int, C.c_int, f32p, f32p, C.c_int, C.POINTER(C.c_float), f32p, f32p,
                                         C.POINTER(C.c_longlong)]
        L.ref_tracker_update.restype = C.c_float
        L.ref_tracker_update.argtypes = [f32p, f32p, C.c_float, f32p, f32p, f32p]
        L.ref_tracker_optimize.argtypes = [C.c_void_p, f32p, C.c_int, C.c_int, f32p, f32p, C.c_int, C.c_float, C.c_float, f32p]
        L.ref_vol_integrate.argtypes = [C.c_void_p, f32p, f32p, f32p, f32p, f32p, f32p, C.c_int, C.c_int, f32p, f32p, C.c_int]
        L.ref_vol_num_voxels.restype = C.c_longlong
        L.ref_vol_num_voxels.argtypes = [C.c_void_p]
        L.ref_vol_counter.argtypes = [C.c_void_p]
        L.ref_vol_dims.argtypes = [C.c_void_p, i32p, C.POINTER(C.c_float), f32p]
        L.ref_vol_download.argtypes = [C.c_void_p, f32p]
        L.ref_vol_upload.argtypes = [C.c_void_p, f32p]
        L.ref_vol_vis.argtypes = [C.c_void_p, u64p, C.c_int, i32p]
        L.ref_vol_set_vis.argtypes = [C.c_void_p, u64p, C.c_int, i32p]
        L.ref_opt_create.restype = C.c_void_p
        L.ref_opt_create.argtypes = [C.c_void_p, C.c_int, C.c_float, f32p, C.c_char_p, C.c_int, C.c_float, C.c_float,
                                     C.c_float, C.c_float, C.c_float, C.c_int, C.c_float, C.c_int]
        L.ref_opt_destroy.argtypes = [C.c_void_p]
        L.ref_opt_set_keyframes.argtypes = [C.c_void_p, C.c_int, i32p, f32p, f32p, C.c_int, C.c_int]
        L.ref_opt_init.argtypes = [C.c_void_p]
        L.ref_opt_set_keytimestamps.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p)]
        L.ref_opt_num_surface.argtypes = [C.c_void_p]
        L.ref_opt_surface.argtypes = [C.c_void_p, i32p]
        L.ref_opt_num_frames.argtypes = [C.c_void_p]
        L.ref_opt_voxel_size.restype = C.c_float
        L.ref_opt_voxel_size.argtypes = [C.c_void_p]
        L.ref_opt_light_len.argtypes = [C.c_void_p]
        L.ref_opt_get_light.argtypes = [C.c_void_p, f32p]
        L.ref_opt_set_light.argtypes = [C.c_void_p, f32p]
        L.ref_opt_get_poses.argtypes = [C.c_void_p, f32p]
        L.ref_opt_set_poses.argtypes = [C.c_void_p, f32p]
        L.ref_opt_get_reg.argtypes = [C.c_void_p, f32p]
        L.ref_opt_set_reg.argtypes = [C.c_void_p, f32p]
        L.ref_opt_init_albedo.argtypes = [C.c_void_p]
        L.ref_opt_energy.restype = C.c_float
        L.ref_opt_energy.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
        for n in ("ref_opt_ps_energy", "ref_opt_normal_energy", "ref_opt_laplacian_energy", "ref_opt_albedo_reg_energy"):
            getattr(L, n).restype = C.c_float
            getattr(L, n).argtypes = [C.c_void_p]
        L.ref_opt_probe_pairs.argtypes = [C.c_void_p, f32p, f32p, u8p, f32p, f32p]
        L.ref_opt_residual.argtypes = [C.c_void_p, f32p, f32p]
        L.ref_opt_dist_grad.argtypes = [C.c_void_p, f32p]
        L.ref_opt_jacobian.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ref_last_jacobian_fetch.argtypes = [i64p, i32p, f32p]
        L.ref_opt_block.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_opt_block_result.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_float)]
        L.ref_opt_block_H.argtypes = [C.c_void_p, C.POINTER(C.c_longlong), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                      C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_opt_block_reference.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_opt_alternating.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)]
        L.ref_last_stdout.restype = C.c_char_p
        L.ref_opt_subsampling.argtypes = [C.c_void_p]
        L.ref_set_quiet.argtypes = [C.c_int]
        L.ref_marching_cubes.argtypes = [i32p, f32p, f32p, f32p, f32p, u8p, u8p, u8p, C.c_float, C.POINTER(C.c_longlong),
                                         C.POINTER(C.c_longlong), C.c_void_p, C.c_void_p, C.c_void_p]
        for n in ("ref_vol_extract_mesh", "ref_vol_extract_pc", "ref_vol_save_sdf", "ref_opt_extract_mesh", "ref_opt_save_white_mesh",
                  "ref_opt_save_pointcloud", "ref_opt_save_sdf"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_char_p]
        _lib = L
    return _lib


LOSS = {"l2": 0, "cauchy": 1, "huber": 2, "tukey": 3, "trunc_l2": 4}
MODEL = {"SH1": 0, "SH2": 1, "LED": 2}
BLOCK = {"albedo": 0, "light": 1, "dist": 2, "pose": 3}


def se3_from_matrix(pose: np.ndarray):
    R = np.zeros(9, np.float32)
    t = np.zeros(3, np.float32)
    lib().ref_se3_from_matrix(np.ascontiguousarray(pose, np.float32).reshape(16), R, t)
    return R.reshape(3, 3), t


def compute_centroid(K, depth, pose):
    """main_ps.cpp:346-375 re-played with the reference's Eigen types."""
    d = np.ascontiguousarray(depth, np.float32)
    out = np.zeros(3, np.float32)
    lib().ref_compute_centroid(np.ascontiguousarray(K, np.float32).reshape(-1), d, d.shape[0], d.shape[1],
                               np.ascontiguousarray(pose, np.float32).reshape(-1), out)
    return out


def tracker_update(g, H, damping, pose):
    """xi = damping * H.llt().solve(g); pose <- SE3::exp(-xi) * pose (RigidPointOptimizer.cpp:67-75): (new pose, xi, |xi|^2)."""
    out = np.zeros(16, np.float32)
    xi = np.zeros(6, np.float32)
    c = lambda a: np.ascontiguousarray(a, np.float32).reshape(-1)
    n2 = lib().ref_tracker_update(c(g), c(H), float(damping), c(pose), out, xi)
    return out.reshape(4, 4), xi, float(n2)


def marching_cubes(dims, size, origin, tsdf, weights, red, green, blue, iso=0.0):
    """third/mesh/MarchingCubes.cpp, unmodified: (vertices T*3 x 3 float32, colours T*3 x 3 uint8, faces T x 3 int32)."""
    dims = np.asarray(dims, np.int32)
    a = (dims, np.asarray(size, np.float32), np.asarray(origin, np.float32), np.ascontiguousarray(tsdf, np.float32).reshape(-1),
         np.ascontiguousarray(weights, np.float32).reshape(-1), np.ascontiguousarray(red, np.uint8).reshape(-1),
         np.ascontiguousarray(green, np.uint8).reshape(-1), np.ascontiguousarray(blue, np.uint8).reshape(-1), float(iso))
    nv, nf = C.c_longlong(), C.c_longlong()
    lib().ref_marching_cubes(*a, C.byref(nv), C.byref(nf), None, None, None)
    v = np.zeros((max(nv.value, 1), 3), np.float32); c = np.zeros((max(nv.value, 1), 3), np.uint8); f = np.zeros((max(nf.value, 1), 3), np.int32)
    lib().ref_marching_cubes(*a, C.byref(nv), C.byref(nf), v.ctypes.data, c.ctypes.data, f.ctypes.data)
    return v[:nv.value], c[:nv.value], f[:nf.value]


class RefVolume:
    def __init__(self, dims, voxel_size, shift, T, zmin=0.5, zmax=3.5):
        self.dims = np.asarray(dims, np.int32)
        self.h = lib().ref_vol_create(self.dims, voxel_size, np.asarray(shift, np.float32), T, zmin, zmax)

    def close(self):
        if self.h:
            lib().ref_vol_destroy(self.h)
            self.h = None

    @property
    def N(self):
        return int(lib().ref_vol_num_voxels(self.h))

    def geometry(self):
        dims = np.zeros(3, np.int32)
        vs = C.c_float()
        org = np.zeros(3, np.float32)
        lib().ref_vol_dims(self.h, dims, C.byref(vs), org)
        return dims, np.float32(vs.value), org

    def integrate(self, color, depth, nx, ny, nz, n_sq_inv, K, pose, increase_counter):
        H, W = depth.shape
        c = lambda a: np.ascontiguousarray(a, np.float32)
        lib().ref_vol_integrate(self.h, c(color), c(depth), c(nx), c(ny), c(nz), c(n_sq_inv), H, W,
                                c(K).reshape(9), c(pose).reshape(16), int(increase_counter))

    def tracker_system(self, depth, K, pose, sampling=1):
        """One pass of RigidPointOptimizer::optimize_sampled's inner loop at `pose`, by the reference's own Sdf lookups."""
        depth = np.ascontiguousarray(depth, np.float32)
        E, cnt = C.c_float(), C.c_longlong()
        g, Hm = np.zeros(6, np.float32), np.zeros((6, 6), np.float32)
        c = lambda a: np.ascontiguousarray(a, np.float32).reshape(-1)
        lib().ref_tracker_system(self.h, depth, depth.shape[0], depth.shape[1], c(K), c(pose), int(sampling), C.byref(E), g, Hm,
                                 C.byref(cnt))
        return E.value, g, Hm, cnt.value

    def tracker_optimize(self, depth, K, pose, num_iterations=50, conv_threshold=1e-3, damping=1.0):
        """RigidPointOptimizer(num_iterations, conv_threshold, damping, tSDF).optimize(depth, K) from `pose`: (converged, pose)."""
        depth = np.ascontiguousarray(depth, np.float32)
        out = np.zeros(16, np.float32)
        c = lambda a: np.ascontiguousarray(a, np.float32).reshape(-1)
        conv = lib().ref_tracker_optimize(self.h, depth, depth.shape[0], depth.shape[1], c(K), c(pose), int(num_iterations),
                                          float(conv_threshold), float(damping), out)
        return bool(conv), out.reshape(4, 4)

    def extract_mesh(self, filename):
        """VolumetricGradSdf::extract_mesh (VolumetricGradSdf.cpp:234-318), the file it writes."""
        return bool(lib().ref_vol_extract_mesh(self.h, filename.encode()))

    def extract_pc(self, filename):
        return bool(lib().ref_vol_extract_pc(self.h, filename.encode()))

    def save_sdf(self, filename):
        return bool(lib().ref_vol_save_sdf(self.h, filename.encode()))

    def download(self):
        a = np.zeros((self.N, 8), np.float32)
        lib().ref_vol_download(self.h, a)
        return a

    def upload(self, aos):
        lib().ref_vol_upload(self.h, np.ascontiguousarray(aos, np.float32))

    def vis(self, nwords=1):
        w = np.zeros((nwords, self.N), np.uint64)
        ln = np.zeros(self.N, np.int32)
        lib().ref_vol_vis(self.h, w, nwords, ln)
        return w, ln

    def set_vis(self, words, lengths):
        words = np.ascontiguousarray(words, np.uint64)
        lib().ref_vol_set_vis(self.h, words, words.shape[0], np.ascontiguousarray(lengths, np.int32))


class RefOptimizer:
    def __init__(self, vol: RefVolume, model, voxel_size, K, save_path="/tmp/", loss="cauchy", lam=0.2, damping=1.0,
                 reg_rho=0.0, reg_n=0.0, reg_l=0.0, max_it=100, conv=5e-3, upsample=False):
        self.vol = vol
        self.model = model
        self.h = lib().ref_opt_create(vol.h, MODEL[model], voxel_size, np.ascontiguousarray(K, np.float32).reshape(9),
                                      save_path.encode(), LOSS[loss], lam, damping, reg_rho, reg_n, reg_l, max_it, conv,
                                      int(upsample))

    def close(self):
        if self.h:
            lib().ref_opt_destroy(self.h)
            self.h = None

    def set_keyframes(self, frame_idx, poses, images):
        F, H, W, _ = images.shape
        lib().ref_opt_set_keyframes(self.h, F, np.ascontiguousarray(frame_idx, np.int32),
                                    np.ascontiguousarray(poses, np.float32).reshape(-1),
                                    np.ascontiguousarray(images, np.float32).reshape(-1), H, W)

    def set_keytimestamps(self, stamps):
        arr = (C.c_char_p * len(stamps))(*[str(x).encode() for x in stamps])
        lib().ref_opt_set_keytimestamps(self.h, len(stamps), arr)

    def init(self):
        lib().ref_opt_init(self.h)

    @property
    def S(self):
        return lib().ref_opt_num_surface(self.h)

    @property
    def F(self):
        return lib().ref_opt_num_frames(self.h)

    def surface(self):
        s = np.zeros(self.S, np.int32)
        lib().ref_opt_surface(self.h, s)
        return s

    def light(self):
        l = np.zeros(lib().ref_opt_light_len(self.h), np.float32)
        lib().ref_opt_get_light(self.h, l)
        return l

    def set_light(self, l):
        lib().ref_opt_set_light(self.h, np.ascontiguousarray(l, np.float32).reshape(-1))

    def poses(self):
        p = np.zeros(self.F * 16, np.float32)
        lib().ref_opt_get_poses(self.h, p)
        return p.reshape(self.F, 4, 4)

    def set_poses(self, p):
        lib().ref_opt_set_poses(self.h, np.ascontiguousarray(p, np.float32).reshape(-1))

    def reg(self):
        r = np.zeros(3, np.float32)
        lib().ref_opt_get_reg(self.h, r)
        return r

    def set_reg(self, rho, n, l):
        lib().ref_opt_set_reg(self.h, np.array([rho, n, l], np.float32))

    def voxel_size(self):
        return np.float32(lib().ref_opt_voxel_size(self.h))

    def init_albedo(self):
        lib().ref_opt_init_albedo(self.h)

    def energy(self):
        e64 = C.c_double()
        na = C.c_longlong()
        e = lib().ref_opt_energy(self.h, C.byref(e64), C.byref(na))
        return np.float32(e), e64.value, na.value

    def ps_energy(self):
        """getPSEnergy() only (what bench.py times)."""
        return np.float32(lib().ref_opt_ps_energy(self.h))

    def normal_energy(self):
        return np.float32(lib().ref_opt_normal_energy(self.h))

    def laplacian_energy(self):
        return np.float32(lib().ref_opt_laplacian_energy(self.h))

    def albedo_reg_energy(self):
        return np.float32(lib().ref_opt_albedo_reg_energy(self.h))

    def probe_pairs(self):
        S, F = self.S, self.F
        m = np.zeros((S, F), np.float32)
        n = np.zeros((S, F), np.float32)
        ok = np.zeros((S, F), np.uint8)
        I = np.zeros((S, F, 3), np.float32)
        rend = np.zeros((S, F, 3), np.float32)
        lib().ref_opt_probe_pairs(self.h, m, n, ok, I, rend)
        return m, n, ok, I, rend

    def residual(self):
        n = 3 * self.S * self.F
        r = np.zeros(n, np.float32)
        w = np.zeros(n, np.float32)
        lib().ref_opt_residual(self.h, r, w)
        # -> (S, F, 3)
        return (r.reshape(self.F, self.S, 3).transpose(1, 0, 2).copy(),
                w.reshape(self.F, self.S, 3).transpose(1, 0, 2).copy())

    def image_probe(self, frame, rows, cols):
        rows = np.ascontiguousarray(rows, np.float32); cols = np.ascontiguousarray(cols, np.float32)
        n = rows.size
        rgb = np.zeros((n, 3), np.float32); g0 = np.zeros((n, 3), np.float32); g1 = np.zeros((n, 3), np.float32)
        L = lib()
        L.ref_opt_image_probe.argtypes = [C.c_void_p, C.c_int, C.c_int, f32p, f32p, f32p, f32p, f32p]
        L.ref_opt_image_probe(self.h, int(frame), n, rows, cols, rgb.reshape(-1), g0.reshape(-1), g1.reshape(-1))
        return rgb, g0, g1

    def dist_grad(self):
        g = np.zeros((self.S, 6), np.float32)
        lib().ref_opt_dist_grad(self.h, g)
        return g[:, :3].copy(), g[:, 3:].copy()

    def jacobian(self, which):
        import scipy.sparse as sp
        idx = which if isinstance(which, int) else BLOCK[which]
        nnz, rows, cols = C.c_longlong(), C.c_int(), C.c_int()
        lib().ref_opt_jacobian(self.h, idx, C.byref(nnz), C.byref(rows), C.byref(cols))
        colptr = np.zeros(cols.value + 1, np.int64)
        rowidx = np.zeros(nnz.value, np.int32)
        vals = np.zeros(nnz.value, np.float32)
        lib().ref_last_jacobian_fetch(colptr, rowidx, vals)
        return sp.csc_matrix((vals, rowidx, colptr), shape=(rows.value, cols.value))

    def block(self, which, apply=True, damping=1.0, albedo_reg=False, normal_reg=False, laplacian_reg=False):
        """Re-played block step; returns dict(H (scipy csc), b, delta, info, iters, err)."""
        import scipy.sparse as sp
        n = lib().ref_opt_block(self.h, BLOCK[which], int(apply), damping, int(albedo_reg), int(normal_reg),
                                int(laplacian_reg))
        b = np.zeros(n, np.float32)
        d = np.zeros(n, np.float32)
        info, iters, err = C.c_int(), C.c_int(), C.c_float()
        lib().ref_opt_block_result(self.h, b.ctypes.data, d.ctypes.data, C.byref(info), C.byref(iters), C.byref(err))
        nnz, rows, cols = C.c_longlong(), C.c_int(), C.c_int()
        lib().ref_opt_block_H(self.h, C.byref(nnz), C.byref(rows), C.byref(cols), None, None, None)
        colptr = np.zeros(cols.value + 1, np.int64)
        rowidx = np.zeros(nnz.value, np.int32)
        vals = np.zeros(nnz.value, np.float32)
        lib().ref_opt_block_H(self.h, C.byref(nnz), C.byref(rows), C.byref(cols), colptr.ctypes.data,
                              rowidx.ctypes.data, vals.ctypes.data)
        H = sp.csc_matrix((vals, rowidx, colptr), shape=(rows.value, cols.value))
        return dict(H=H, b=b, delta=d, info=info.value, iters=iters.value, err=err.value)

    def block_reference(self, which, damping=1.0, albedo_reg=False, normal_reg=False, laplacian_reg=False):
        lib().ref_opt_block_reference(self.h, BLOCK[which], damping, int(albedo_reg), int(normal_reg), int(laplacian_reg))

    def alternating(self, light=True, albedo=True, distance=True, pose=True):
        sec = C.c_double()
        r = lib().ref_opt_alternating(self.h, int(light), int(albedo), int(distance), int(pose), C.byref(sec))
        out = lib().ref_last_stdout().decode(errors="replace")
        return bool(r), sec.value, out

    def subsampling(self):
        lib().ref_opt_subsampling(self.h)

    # the reference's writers: <save_path><name>_mesh.ply, _obj.ply, _pointcloud.ply, <save_path><name>
    def extract_mesh(self, name):
        return bool(lib().ref_opt_extract_mesh(self.h, name.encode()))

    def save_white_mesh(self, name):
        return bool(lib().ref_opt_save_white_mesh(self.h, name.encode()))

    def save_pointcloud(self, name):
        return bool(lib().ref_opt_save_pointcloud(self.h, name.encode()))

    def save_sdf(self, name):
        return bool(lib().ref_opt_save_sdf(self.h, name.encode()))
