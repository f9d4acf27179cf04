"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/libpsgoracle.so (the plain-C restatement,
psg_oracle.c).  Same method names as oracle/pyref.py so tests can run either against the other.
Imported by tests/, bench.py's cpu_baseline / --impl reference legs and __graft_entry__.smoke() only."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpsgoracle.so")
_lib = None

f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")

LOSS = {"l2": 0, "cauchy": 1, "huber": 2, "tukey": 3, "trunc_l2": 4}
MODEL = {"SH1": 0, "SH2": 1, "LED": 2}
BLOCK = {"albedo": 0, "light": 1, "dist": 2, "pose": 3}


def build(force: bool = False) -> None:
    srcs = [os.path.join(_HERE, n) for n in ("psg_oracle.c", "psg_oracle_normals.c", "psg_oracle_mesh.c")]
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(x) for x in srcs):
        subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.or_volume_create.restype = vp
        L.or_volume_create.argtypes = [i32p, C.c_float, f32p, C.c_float, C.c_float, C.c_float, C.c_int]
        L.or_volume_destroy.argtypes = [vp]
        L.or_volume_num_voxels.restype = C.c_longlong
        L.or_volume_num_voxels.argtypes = [vp]
        L.or_volume_geometry.argtypes = [vp, i32p, C.POINTER(C.c_float), f32p]
        L.or_volume_integrate.argtypes = [vp, f32p, f32p, f32p, f32p, f32p, f32p, C.c_int, C.c_int, f32p, f32p, f32p, C.c_int]
        L.or_volume_download.argtypes = [vp, f32p]
        L.or_volume_upload.argtypes = [vp, f32p]
        L.or_volume_vis_words.argtypes = [vp]
        L.or_volume_vis.argtypes = [vp, u64p, i32p]
        L.or_volume_set_vis.argtypes = [vp, u64p, i32p]
        L.or_opt_create.restype = vp
        L.or_opt_create.argtypes = [vp, C.c_int, C.c_float, f32p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float]
        L.or_opt_destroy.argtypes = [vp]
        L.or_opt_set_keyframes.argtypes = [vp, C.c_int, i32p, f32p, f32p, C.c_int, C.c_int]
        L.or_opt_init.argtypes = [vp]
        L.or_opt_num_surface.argtypes = [vp]
        L.or_opt_surface.argtypes = [vp, i32p]
        L.or_opt_vis_keyframes.argtypes = [vp, u64p]
        L.or_opt_light_len.argtypes = [vp]
        L.or_opt_get_light.argtypes = [vp, f32p]
        L.or_opt_set_light.argtypes = [vp, f32p]
        L.or_opt_get_poses.argtypes = [vp, f32p]
        L.or_opt_set_poses.argtypes = [vp, f32p]
        L.or_opt_get_reg.argtypes = [vp, f32p]
        L.or_opt_set_reg.argtypes = [vp, f32p]
        L.or_opt_voxel_size.restype = C.c_float
        L.or_opt_voxel_size.argtypes = [vp]
        L.or_opt_init_albedo.argtypes = [vp]
        L.or_opt_energy.restype = C.c_float
        L.or_opt_energy.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
        for n in ("or_opt_normal_energy", "or_opt_laplacian_energy", "or_opt_albedo_reg_energy"):
            getattr(L, n).restype = C.c_float
            getattr(L, n).argtypes = [vp, C.POINTER(C.c_double)]
        L.or_opt_probe_pairs.argtypes = [vp, f32p, f32p, u8p, f32p, f32p]
        L.or_opt_residual.argtypes = [vp, f32p, f32p]
        L.or_opt_dist_grad.argtypes = [vp, f32p]
        L.or_opt_block.argtypes = [vp, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.or_opt_block_result.argtypes = [vp, vp, vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_float)]
        L.or_opt_block_H.restype = C.c_longlong
        L.or_opt_block_H.argtypes = [vp, vp, vp, vp]
        L.or_opt_subsampling.argtypes = [vp]
        L.or_opt_set_accum64.argtypes = [vp, C.c_int]
        L.or_tracker_system.argtypes = [vp, f32p, C.c_int, C.c_int, f32p, f32p, f32p, C.c_int, C.POINTER(C.c_float), f32p, f32p,
                                        C.POINTER(C.c_longlong)]
        L.or_normals_cache.argtypes = [C.c_int, C.c_int, f32p, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p]
        L.or_normals_compute.argtypes = [C.c_int, C.c_int, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, f32p, f32p]
        L.or_convert_depth_u16.argtypes = [vp, C.c_size_t, C.c_float, f32p]
        L.or_convert_color_u8.argtypes = [vp, C.c_size_t, f32p]
        L.or_modified_laplacian.restype = C.c_float
        L.or_modified_laplacian.argtypes = [f32p, C.c_int, C.c_int]
        L.or_compute_centroid.argtypes = [f32p, f32p, C.c_int, C.c_int, f32p, f32p]
        L.or_sample_keyframes.argtypes = [C.c_int, C.c_int, i32p]
        L.or_mesh_bbox.argtypes = [f32p, i32p, C.c_float, i32p, i32p]
        L.or_marching_cubes.restype = C.c_longlong
        L.or_marching_cubes.argtypes = [i32p, f32p, f32p, f32p, f32p, u8p, u8p, u8p, C.c_float, vp, vp]
        L.or_extract_mesh.restype = C.c_longlong
        L.or_extract_mesh.argtypes = [f32p, i32p, C.c_float, C.c_int, vp, vp, vp, vp]
        L.or_write_mesh_ply.argtypes = [C.c_char_p, C.c_longlong, f32p, u8p]
        L.or_write_sdf.argtypes = [C.c_char_p, f32p, i32p, C.c_float]
        L.or_write_volume_pc.argtypes = [C.c_char_p, f32p, u8p, i32p, C.c_float]
        L.or_write_surface_pc.argtypes = [C.c_char_p, f32p, i32p, C.c_int, i32p, C.c_float]
        _lib = L
    return _lib


class OrVolume:
    def __init__(self, dims, voxel_size, shift, T, zmin=0.5, zmax=3.5, max_frames=64):
        self.dims = np.asarray(dims, np.int32)
        self.h = lib().or_volume_create(self.dims, voxel_size, np.asarray(shift, np.float32), T, zmin, zmax, max_frames)

    def close(self):
        if self.h:
            lib().or_volume_destroy(self.h)
            self.h = None

    @property
    def N(self):
        return int(lib().or_volume_num_voxels(self.h))

    def geometry(self):
        dims = np.zeros(3, np.int32)
        vs = C.c_float()
        org = np.zeros(3, np.float32)
        lib().or_volume_geometry(self.h, dims, C.byref(vs), org)
        return dims, np.float32(vs.value), org

    def integrate(self, color, depth, nx, ny, nz, n_sq_inv, K, R, t, increase_counter):
        H, W = depth.shape
        c = lambda a: np.ascontiguousarray(a, np.float32)
        lib().or_volume_integrate(self.h, c(color), c(depth), c(nx), c(ny), c(nz), c(n_sq_inv), H, W,
                                  c(K).reshape(9), c(R).reshape(9), c(t).reshape(3), int(increase_counter))

    def tracker_system(self, depth, K, R, t, sampling=1):
        """(E, g[6], H[6,6], count) of one RigidPointOptimizer::optimize_sampled pass at pose (R, t)."""
        depth = np.ascontiguousarray(depth, np.float32)
        E, cnt = C.c_float(), C.c_longlong()
        g, Hm = np.zeros(6, np.float32), np.zeros((6, 6), np.float32)
        c = lambda a: np.ascontiguousarray(a, np.float32).reshape(-1)
        lib().or_tracker_system(self.h, depth, depth.shape[0], depth.shape[1], c(K), c(R), c(t), int(sampling), C.byref(E), g, Hm,
                                C.byref(cnt))
        return E.value, g, Hm, cnt.value

    def download(self):
        a = np.zeros((self.N, 8), np.float32)
        lib().or_volume_download(self.h, a)
        return a

    def upload(self, aos):
        lib().or_volume_upload(self.h, np.ascontiguousarray(aos, np.float32))

    def vis(self):
        nw = lib().or_volume_vis_words(self.h)
        w = np.zeros((nw, self.N), np.uint64)
        ln = np.zeros(self.N, np.int32)
        lib().or_volume_vis(self.h, w, ln)
        return w, ln

    def set_vis(self, words, lengths):
        lib().or_volume_set_vis(self.h, np.ascontiguousarray(words, np.uint64), np.ascontiguousarray(lengths, np.int32))


class OrOptimizer:
    def __init__(self, vol: OrVolume, model, voxel_size, K, loss="cauchy", lam=0.2, reg_rho=0.0, reg_n=0.0, reg_l=0.0):
        self.vol = vol
        self.model = model
        self.h = lib().or_opt_create(vol.h, MODEL[model], voxel_size, np.ascontiguousarray(K, np.float32).reshape(9),
                                     LOSS[loss], lam, reg_rho, reg_n, reg_l)

    def close(self):
        if self.h:
            lib().or_opt_destroy(self.h)
            self.h = None

    def set_keyframes(self, frame_idx, poses, images):
        F, H, W, _ = images.shape
        self._F = F
        lib().or_opt_set_keyframes(self.h, F, np.ascontiguousarray(frame_idx, np.int32),
                                   np.ascontiguousarray(poses, np.float32).reshape(-1),
                                   np.ascontiguousarray(images, np.float32).reshape(-1), H, W)

    def init(self):
        lib().or_opt_init(self.h)

    @property
    def S(self):
        return lib().or_opt_num_surface(self.h)

    @property
    def F(self):
        return self._F

    def surface(self):
        s = np.zeros(self.S, np.int32)
        lib().or_opt_surface(self.h, s)
        return s

    def vis_keyframes(self):
        v = np.zeros(self.S, np.uint64)
        lib().or_opt_vis_keyframes(self.h, v)
        return v

    def light(self):
        l = np.zeros(lib().or_opt_light_len(self.h), np.float32)
        lib().or_opt_get_light(self.h, l)
        return l

    def set_light(self, l):
        lib().or_opt_set_light(self.h, np.ascontiguousarray(l, np.float32).reshape(-1))

    def poses(self):
        p = np.zeros(self.F * 16, np.float32)
        lib().or_opt_get_poses(self.h, p)
        return p.reshape(self.F, 4, 4)

    def set_poses(self, p):
        lib().or_opt_set_poses(self.h, np.ascontiguousarray(p, np.float32).reshape(-1))

    def reg(self):
        r = np.zeros(3, np.float32)
        lib().or_opt_get_reg(self.h, r)
        return r

    def set_reg(self, rho, n, l):
        lib().or_opt_set_reg(self.h, np.array([rho, n, l], np.float32))

    def voxel_size(self):
        return np.float32(lib().or_opt_voxel_size(self.h))

    def init_albedo(self):
        lib().or_opt_init_albedo(self.h)

    def energy(self):
        e64 = C.c_double()
        na = C.c_longlong()
        e = lib().or_opt_energy(self.h, C.byref(e64), C.byref(na))
        return np.float32(e), e64.value, na.value

    def _reg_energy(self, name):
        e64 = C.c_double()
        e = getattr(lib(), name)(self.h, C.byref(e64))
        return np.float32(e), e64.value

    def normal_energy(self):
        return self._reg_energy("or_opt_normal_energy")[0]

    def laplacian_energy(self):
        return self._reg_energy("or_opt_laplacian_energy")[0]

    def albedo_reg_energy(self):
        return self._reg_energy("or_opt_albedo_reg_energy")[0]

    def probe_pairs(self):
        S, F = self.S, self.F
        m = np.zeros((S, F), np.float32)
        n = np.zeros((S, F), np.float32)
        ok = np.zeros((S, F), np.uint8)
        I = np.zeros((S, F, 3), np.float32)
        rend = np.zeros((S, F, 3), np.float32)
        lib().or_opt_probe_pairs(self.h, m, n, ok, I, rend)
        return m, n, ok, I, rend

    def residual(self):
        r = np.zeros((self.S, self.F, 3), np.float32)
        w = np.zeros((self.S, self.F, 3), np.float32)
        lib().or_opt_residual(self.h, r, w)
        return r, w

    def image_probe(self, frame, rows, cols):
        rows = np.ascontiguousarray(rows, np.float32); cols = np.ascontiguousarray(cols, np.float32)
        n = rows.size
        rgb = np.zeros((n, 3), np.float32); g0 = np.zeros((n, 3), np.float32); g1 = np.zeros((n, 3), np.float32)
        L = lib()
        L.or_opt_image_probe.argtypes = [C.c_void_p, C.c_int, C.c_int, f32p, f32p, f32p, f32p, f32p]
        L.or_opt_image_probe(self.h, int(frame), n, rows, cols, rgb.reshape(-1), g0.reshape(-1), g1.reshape(-1))
        return rgb, g0, g1

    def dist_grad(self):
        g = np.zeros((self.S, 6), np.float32)
        lib().or_opt_dist_grad(self.h, g)
        return g[:, :3].copy(), g[:, 3:].copy()

    def block(self, which, apply=True, damping=1.0, albedo_reg=False, normal_reg=False, laplacian_reg=False):
        import scipy.sparse as sp
        n = lib().or_opt_block(self.h, BLOCK[which], int(apply), damping, int(albedo_reg), int(normal_reg),
                               int(laplacian_reg))
        b = np.zeros(n, np.float32)
        d = np.zeros(n, np.float32)
        ok, iters, err = C.c_int(), C.c_int(), C.c_float()
        lib().or_opt_block_result(self.h, b.ctypes.data, d.ctypes.data, C.byref(ok), C.byref(iters), C.byref(err))
        nnz = lib().or_opt_block_H(self.h, None, None, None)
        rows = np.zeros(nnz, np.int32)
        cols = np.zeros(nnz, np.int32)
        vals = np.zeros(nnz, np.float32)
        lib().or_opt_block_H(self.h, rows.ctypes.data, cols.ctypes.data, vals.ctypes.data)
        H = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsc()
        # info follows Eigen's ComputationInfo: 0 = Success
        return dict(H=H, b=b, delta=d, info=0 if ok.value else 2, iters=iters.value, err=err.value)

    def set_accum64(self, on=True):
        """Diagnostic: light / pose normal equations summed in double (not what the reference does)."""
        lib().or_opt_set_accum64(self.h, int(on))

    def subsampling(self):
        lib().or_opt_subsampling(self.h)


class OrNormalEstimator:
    """cv::NormalEstimator<float> (NormalEstimator.h:125-176) restated in oracle/psg_oracle_normals.c."""

    def __init__(self, width, height, K, window=11):
        self.W, self.H, self.window = int(width), int(height), int(window)
        shp = (self.H, self.W)
        self.x0, self.y0, self.x0_n, self.y0_n, self.n_sq_inv = (np.zeros(shp, np.float32) for _ in range(5))
        self.Q = np.zeros((6,) + shp, np.float32)
        lib().or_normals_cache(self.W, self.H, np.ascontiguousarray(K, np.float32).reshape(-1), self.window, self.x0, self.y0,
                               self.x0_n, self.y0_n, self.n_sq_inv, self.Q)

    def compute(self, depth):
        depth = np.ascontiguousarray(depth, np.float32)
        nx, ny, nz = (np.zeros((self.H, self.W), np.float32) for _ in range(3))
        lib().or_normals_compute(self.W, self.H, self.window, depth, self.x0_n, self.y0_n, self.n_sq_inv, self.Q, nx, ny, nz)
        return nx, ny, nz


def convert_depth_u16(d16, unit):
    d16 = np.ascontiguousarray(d16, np.uint16)
    out = np.zeros(d16.shape, np.float32)
    lib().or_convert_depth_u16(d16.ctypes.data, d16.size, float(unit), out)
    return out


def convert_color_u8(c8):
    c8 = np.ascontiguousarray(c8, np.uint8)
    out = np.zeros(c8.shape, np.float32)
    lib().or_convert_color_u8(c8.ctypes.data, c8.size, out)
    return out


def modified_laplacian(color_bgr):
    """SharpDetector.h:22-37 on a float BGR image."""
    c = np.ascontiguousarray(color_bgr, np.float32)
    return float(lib().or_modified_laplacian(c, c.shape[0], c.shape[1]))


def compute_centroid(K, depth, pose):
    """main_ps.cpp:346-375."""
    d = np.ascontiguousarray(depth, np.float32)
    out = np.zeros(3, np.float32)
    lib().or_compute_centroid(np.ascontiguousarray(K, np.float32).reshape(-1), d, d.shape[0], d.shape[1],
                              np.ascontiguousarray(pose, np.float32).reshape(-1), out)
    return out


def sample_keyframes(n, max_num):
    """main_ps.cpp:392-421: indices of the keyframes that are kept."""
    idx = np.zeros(max(n, max_num) + 1, np.int32)
    m = lib().or_sample_keyframes(int(n), int(max_num), idx)
    return idx[:m].copy()


# ---- surface extraction (psg_oracle_mesh.c) --------------------------------------------------------------------------------
def mesh_bbox(aos, dims, vs):
    """Crop box of extract_mesh / saveSDF: (lo[3], hi[3]) or None when no voxel lies within sqrt(3) * vs of the surface."""
    lo, hi = np.zeros(3, np.int32), np.zeros(3, np.int32)
    any_ = lib().or_mesh_bbox(np.ascontiguousarray(aos, np.float32), np.asarray(dims, np.int32), float(vs), lo, hi)
    return (lo, hi) if any_ else None


def marching_cubes(dims, size, origin, tsdf, weights, red, green, blue, iso=0.0):
    """MarchingCubes::computeIsoSurface on cropped arrays: (vertices 3T x 3 float32, colours 3T x 3 uint8)."""
    a = (np.asarray(dims, np.int32), np.asarray(size, np.float32), np.asarray(origin, np.float32),
         np.ascontiguousarray(tsdf, np.float32).reshape(-1), np.ascontiguousarray(weights, np.float32).reshape(-1),
         np.ascontiguousarray(red, np.uint8).reshape(-1), np.ascontiguousarray(green, np.uint8).reshape(-1),
         np.ascontiguousarray(blue, np.uint8).reshape(-1), float(iso))
    nt = lib().or_marching_cubes(*a, None, None)
    v, c = np.zeros((max(3 * nt, 1), 3), np.float32), np.zeros((max(3 * nt, 1), 3), np.uint8)
    lib().or_marching_cubes(*a, v.ctypes.data, c.ctypes.data)
    return v[:3 * nt], c[:3 * nt]


def extract_mesh(aos, dims, vs, white=False):
    """Optimizer::extract_mesh / save_white_mesh on an N x 8 voxel array: (vertices, colours) or None (empty crop box)."""
    aos = np.ascontiguousarray(aos, np.float32)
    dims = np.asarray(dims, np.int32)
    nt = lib().or_extract_mesh(aos, dims, float(vs), int(white), None, None, None, None)
    if nt < 0:
        return None
    v, c = np.zeros((max(3 * nt, 1), 3), np.float32), np.zeros((max(3 * nt, 1), 3), np.uint8)
    lib().or_extract_mesh(aos, dims, float(vs), int(white), None, None, v.ctypes.data, c.ctypes.data)
    return v[:3 * nt], c[:3 * nt]


def write_mesh_ply(path, vertices, colors):
    v, c = np.ascontiguousarray(vertices, np.float32), np.ascontiguousarray(colors, np.uint8)
    return bool(lib().or_write_mesh_ply(path.encode(), len(v) // 3, v.reshape(-1) if len(v) else np.zeros(1, np.float32),
                                        c.reshape(-1) if len(c) else np.zeros(1, np.uint8)))


def write_sdf(path, aos, dims, vs):
    return bool(lib().or_write_sdf(path.encode(), np.ascontiguousarray(aos, np.float32), np.asarray(dims, np.int32), float(vs)))


def write_volume_pc(path, aos, seen, dims, vs):
    return bool(lib().or_write_volume_pc(path.encode(), np.ascontiguousarray(aos, np.float32), np.ascontiguousarray(seen, np.uint8),
                                         np.asarray(dims, np.int32), float(vs)))


def write_surface_pc(path, aos, lin, dims, vs):
    """Optimizer::save_pointcloud over the surface list `lin`."""
    lin = np.ascontiguousarray(lin, np.int32)
    return bool(lib().or_write_surface_pc(path.encode(), np.ascontiguousarray(aos, np.float32), lin if len(lin) else np.zeros(1, np.int32),
                                          len(lin), np.asarray(dims, np.int32), float(vs)))
