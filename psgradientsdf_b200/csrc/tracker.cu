// Rigid depth-to-SDF tracking, the per-pixel part.
//
// Replaces the double loop of RigidPointOptimizer::optimize_sampled (cpp/include/sdf_tracker/RigidPointOptimizer.cpp:38-62):
// every sampling-th pixel with zmin < z < zmax is back-projected, moved by the current pose, looked up in the gradient-SDF
// volume (Sdf::weights / Sdf::tsdf, VolumetricGradSdf.h:76-93; VoxelGrid::nearest_index, VoxelGrid.cpp:57-72) and
// contributes phi^2, phi * J and J J^T with J = (g_hat, point x g_hat) to the cost, gradient and approximate Hessian.
// The 6 x 6 solve and the SE(3) update (RigidPointOptimizer.cpp:67-75: Eigen LLT, Sophus exp) stay with the host, which
// calls this once per Gauss-Newton iteration; the depth image stays on the device between iterations.
//
// Per-pixel arithmetic is the reference's float arithmetic operation by operation, so which pixels count (voxel index,
// weight test) is decided identically; the sums are formed in double (the reference: one serial float each), reduced
// warp -> block -> grid in a fixed order, hence reproducible.
#include <cmath>

#include "psg_host.h"

namespace psg {

constexpr int kTrackVals = 29;   // E | g[6] | H lower triangle (21) | count

struct TrackParams {
    float fx_inv, fy_inv, cx, cy;
    float R[9], t[3];
    int H, W, sampling, nx, ny;
};

__global__ void __launch_bounds__(256) k_track_system(VolView v, const float* __restrict__ depth, TrackParams P, double* __restrict__ partials) {
    __shared__ double s_w[8][kTrackVals];
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    double acc[kTrackVals];
#pragma unroll
    for (int i = 0; i < kTrackVals; ++i) acc[i] = 0.0;
    if (s < P.nx * P.ny) {
        const int sy = s / P.nx, sx = s - sy * P.nx;
        const int x = sx * P.sampling, y = sy * P.sampling;
        const float z = __ldg(depth + (long long)y * P.W + x);
        if (!(z <= v.zmin || z >= v.zmax)) {
            const float x0 = ((float)x - P.cx) * P.fx_inv;
            const float y0 = ((float)y - P.cy) * P.fy_inv;
            const float pc[3] = { x0 * z, y0 * z, z };
            float Rp[3], point[3], fi[3];
            mat_vec(P.R, pc, Rp);
#pragma unroll
            for (int a = 0; a < 3; ++a) { point[a] = Rp[a] + P.t[a]; fi[a] = (point[a] - v.origin[a]) / v.vs; }
            // VoxelGrid::nearest_index: strictly inside, then round half up through a double add
            const bool inside = !(fi[0] <= 0 || fi[1] <= 0 || fi[2] <= 0 || fi[0] >= (float)(v.dims[0] - 1) ||
                                  fi[1] >= (float)(v.dims[1] - 1) || fi[2] >= (float)(v.dims[2] - 1));
            if (inside) {
                const int im = (int)((double)fi[0] + 0.5), jm = (int)((double)fi[1] + 0.5), km = (int)((double)fi[2] + 0.5);
                const long long I = (long long)im + (long long)jm * v.dims[0] + (long long)km * v.dims[0] * v.dims[1];
                if (v.w[I] > 0.f) {
                    const float grad[3] = { v.gx[I], v.gy[I], v.gz[I] };
                    float gn[3], d[3];
                    normalized3(grad, gn);
                    // voxel2world(world2voxel(point)) - point: world2voxel adds a float 0.5 and truncates (VoxelGrid.h:63-66)
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        const int idx = (int)(fi[a] + 0.5f);
                        d[a] = (v.origin[a] + v.vs * (float)idx) - point[a];
                    }
                    const float phi0 = v.dist[I] + dot3(gn, d);
                    const float J[6] = { gn[0], gn[1], gn[2], point[1] * gn[2] - point[2] * gn[1], point[2] * gn[0] - point[0] * gn[2],
                                         point[0] * gn[1] - point[1] * gn[0] };
                    acc[0] = (double)(phi0 * phi0);
                    int t = 7;
#pragma unroll
                    for (int a = 0; a < 6; ++a) {
                        acc[1 + a] = (double)(phi0 * J[a]);
#pragma unroll
                        for (int b = 0; b <= a; ++b) acc[t++] = (double)(J[a] * J[b]);
                    }
                    acc[28] = 1.0;
                }
            }
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < kTrackVals; ++i) { const double w = warp_sum(acc[i]); if (lane == 0) s_w[warp][i] = w; }
    __syncthreads();
    if (threadIdx.x < kTrackVals) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += s_w[w][threadIdx.x];
        partials[(long long)blockIdx.x * 32 + threadIdx.x] = t;
    }
}

// fixed-order sum of the block partials: 8 interleaved chains per value, combined in chain order
__global__ void __launch_bounds__(256) k_track_reduce(const double* __restrict__ partials, int nblocks, double* __restrict__ out) {
    __shared__ double s_c[8][32];
    const int v = threadIdx.x & 31, c = threadIdx.x >> 5;
    double t = 0.0;
    if (v < kTrackVals) for (int b = c; b < nblocks; b += 8) t += partials[(long long)b * 32 + v];
    s_c[c][v] = t;
    __syncthreads();
    if (threadIdx.x < kTrackVals) {
        double r = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) r += s_c[k][threadIdx.x];
        out[threadIdx.x] = r;
    }
}

__global__ void __launch_bounds__(256) k_track_depth_u16(const unsigned short* __restrict__ src, float unit, float* __restrict__ dst, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = (float)src[i] * unit;   // ImageLoader::load_depth: convertTo(CV_32FC1, unit_)
}

}  // namespace psg

using namespace psg;

extern "C" {

int psg_tracker_set_depth(psg_volume* v, const float* depth, int H, int W) {
    if (!v || !depth || H <= 0 || W <= 0) { set_error("psg_tracker_set_depth: null or empty argument"); return PSG_ERR_INVALID; }
    try {
        PSG_CUDA(cudaSetDevice(v->device));
        v->track_depth.alloc((size_t)H * W);
        PSG_CUDA(cudaMemcpyAsync(v->track_depth.p, depth, (size_t)H * W * sizeof(float), cudaMemcpyHostToDevice, v->stream));
        PSG_CUDA(cudaStreamSynchronize(v->stream));   // the host buffer may be reused as soon as this returns
        v->track_H = H; v->track_W = W;
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

int psg_tracker_set_depth_u16(psg_volume* v, const unsigned short* depth_u16, float depth_unit, int H, int W) {
    if (!v || !depth_u16 || H <= 0 || W <= 0) { set_error("psg_tracker_set_depth_u16: null or empty argument"); return PSG_ERR_INVALID; }
    try {
        PSG_CUDA(cudaSetDevice(v->device));
        const size_t n = (size_t)H * W;
        v->track_depth.alloc(n);
        v->staging.alloc((n + 1) / 2);
        unsigned short* raw = reinterpret_cast<unsigned short*>(v->staging.p);
        PSG_CUDA(cudaMemcpyAsync(raw, depth_u16, n * 2, cudaMemcpyHostToDevice, v->stream));
        { Prof pf(v->stream, KID_TRACK); k_track_depth_u16<<<(unsigned)((n + 255) / 256), 256, 0, v->stream>>>(raw, depth_unit, v->track_depth.p, (int)n); }
        PSG_LAUNCH_CHECK();
        PSG_CUDA(cudaStreamSynchronize(v->stream));
        v->track_H = H; v->track_W = W;
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

int psg_tracker_system(psg_volume* v, const float K[9], const float R[9], const float t[3], int sampling, double* E_sum,
                       double g6[6], double H36[36], long long* count) {
    if (!v || !K || !R || !t || sampling < 1) { set_error("psg_tracker_system: bad argument"); return PSG_ERR_INVALID; }
    if (!v->track_depth.p || v->track_H <= 0) { set_error("psg_tracker_system: call psg_tracker_set_depth first"); return PSG_ERR_STATE; }
    try {
        PSG_CUDA(cudaSetDevice(v->device));
        TrackParams P;
        P.fx_inv = 1.f / K[0]; P.fy_inv = 1.f / K[4]; P.cx = K[2]; P.cy = K[5];
        for (int a = 0; a < 9; ++a) P.R[a] = R[a];
        for (int a = 0; a < 3; ++a) P.t[a] = t[a];
        P.H = v->track_H; P.W = v->track_W; P.sampling = sampling;
        P.nx = (P.W + sampling - 1) / sampling; P.ny = (P.H + sampling - 1) / sampling;
        const int n = P.nx * P.ny;
        const int nblocks = (n + 255) / 256;
        v->track_partials.alloc((size_t)nblocks * 32 + 32);
        double* out = v->track_partials.p + (size_t)nblocks * 32;
        { Prof pf(v->stream, KID_TRACK); k_track_system<<<nblocks, 256, 0, v->stream>>>(v->view(), v->track_depth.p, P, v->track_partials.p); }
        PSG_LAUNCH_CHECK();
        { Prof pf(v->stream, KID_TRACK); k_track_reduce<<<1, 256, 0, v->stream>>>(v->track_partials.p, nblocks, out); }
        PSG_LAUNCH_CHECK();
        double h[kTrackVals];
        PSG_CUDA(cudaMemcpyAsync(h, out, sizeof(h), cudaMemcpyDeviceToHost, v->stream));
        PSG_CUDA(cudaStreamSynchronize(v->stream));
        if (E_sum) *E_sum = h[0];
        if (g6) for (int a = 0; a < 6; ++a) g6[a] = h[1 + a];
        if (H36) {
            int k = 7;
            for (int a = 0; a < 6; ++a) for (int b = 0; b <= a; ++b) { H36[6 * a + b] = h[k]; H36[6 * b + a] = h[k]; ++k; }
        }
        if (count) *count = (long long)(h[28] + 0.5);
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}


// ---- host side of RigidPointOptimizer::optimize_sampled (RigidPointOptimizer.cpp:26-77): float 6 x 6 Cholesky solve and the
// SE(3) update pose <- exp(-xi) * pose, written after Eigen's LLT and Sophus' SE3 / SO3 (third/Sophus/sophus/se3.hpp:761-782,
// so3.hpp:325-339, 583-619).  The pose is carried as unit quaternion + translation like Sophus does.
namespace {

struct Se3f { float q[4]; /* x y z w */ float t[3]; };

// Eigen::Quaternion<float>(Matrix3f)
void quat_from_rot(const float* R /* row-major */, float q[4]) {
    auto m = [&](int r, int c) { return R[3 * r + c]; };
    float tr = m(0, 0) + m(1, 1) + m(2, 2);
    if (tr > 0.f) {
        tr = std::sqrt(tr + 1.f);
        q[3] = 0.5f * tr;
        tr = 0.5f / tr;
        q[0] = (m(2, 1) - m(1, 2)) * tr; q[1] = (m(0, 2) - m(2, 0)) * tr; q[2] = (m(1, 0) - m(0, 1)) * tr;
    } else {
        int i = 0;
        if (m(1, 1) > m(0, 0)) i = 1;
        if (m(2, 2) > m(i, i)) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        tr = std::sqrt(m(i, i) - m(j, j) - m(k, k) + 1.f);
        q[i] = 0.5f * tr;
        tr = 0.5f / tr;
        q[3] = (m(k, j) - m(j, k)) * tr; q[j] = (m(j, i) + m(i, j)) * tr; q[k] = (m(k, i) + m(i, k)) * tr;
    }
}

// Eigen::QuaternionBase::toRotationMatrix
void rot_from_quat(const float q[4], float* R /* row-major */) {
    const float tx = 2.f * q[0], ty = 2.f * q[1], tz = 2.f * q[2];
    const float twx = tx * q[3], twy = ty * q[3], twz = tz * q[3], txx = tx * q[0], txy = ty * q[0], txz = tz * q[0];
    const float tyy = ty * q[1], tyz = tz * q[1], tzz = tz * q[2];
    R[0] = 1.f - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
    R[3] = txy + twz; R[4] = 1.f - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1.f - (txx + tyy);
}

inline float sum3h(float a, float b, float c) { return a + (b + c); }

// xi = H.llt().solve(g) for a 6 x 6 float system; false if H is not positive definite
bool llt_solve6(const float* H /* row-major, symmetric */, const float* g, float* x) {
    float L[36];
    for (int i = 0; i < 36; ++i) L[i] = H[i];
    for (int k = 0; k < 6; ++k) {
        float d = L[6 * k + k];
        for (int j = 0; j < k; ++j) d -= L[6 * k + j] * L[6 * k + j];
        if (!(d > 0.f)) return false;
        d = std::sqrt(d);
        L[6 * k + k] = d;
        for (int i = k + 1; i < 6; ++i) {
            float v = L[6 * i + k];
            for (int j = 0; j < k; ++j) v -= L[6 * i + j] * L[6 * k + j];
            L[6 * i + k] = v / d;
        }
    }
    float y[6];
    for (int i = 0; i < 6; ++i) {
        float v = g[i];
        for (int j = 0; j < i; ++j) v -= L[6 * i + j] * y[j];
        y[i] = v / L[6 * i + i];
    }
    for (int i = 5; i >= 0; --i) {
        float v = y[i];
        for (int j = i + 1; j < 6; ++j) v -= L[6 * j + i] * x[j];
        x[i] = v / L[6 * i + i];
    }
    return true;
}

// pose <- SE3::exp(a) * pose
void se3_exp_left(const float a[6], Se3f& P) {
    const float w[3] = { a[3], a[4], a[5] };
    const float theta_sq = sum3h(w[0] * w[0], w[1] * w[1], w[2] * w[2]);
    const float eps = 1e-10f;   // Sophus::Constants<float>::epsilon()
    float theta, imag, real;
    if (theta_sq < eps * eps) {
        theta = 0.f;
        const float t4 = theta_sq * theta_sq;
        imag = 0.5f - (float)(1.0 / 48.0) * theta_sq + (float)(1.0 / 3840.0) * t4;
        real = 1.f - (float)(1.0 / 8.0) * theta_sq + (float)(1.0 / 384.0) * t4;
    } else {
        theta = std::sqrt(theta_sq);
        const float half = 0.5f * theta;
        imag = std::sin(half) / theta;
        real = std::cos(half);
    }
    const float qe[4] = { imag * w[0], imag * w[1], imag * w[2], real };
    // V (se3.hpp:768-780)
    const float Om[9] = { 0.f, -w[2], w[1], w[2], 0.f, -w[0], -w[1], w[0], 0.f };
    float V[9];
    if (theta < eps) {
        rot_from_quat(qe, V);
    } else {
        float Om2[9];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) Om2[3 * r + c] = sum3h(Om[3 * r] * Om[c], Om[3 * r + 1] * Om[3 + c], Om[3 * r + 2] * Om[6 + c]);
        const float th2 = theta * theta;
        const float c1 = (1.f - std::cos(theta)) / th2, c2 = (theta - std::sin(theta)) / (th2 * theta);
        for (int i = 0; i < 9; ++i) V[i] = ((i % 4 == 0 ? 1.f : 0.f) + c1 * Om[i]) + c2 * Om2[i];
    }
    const float te[3] = { sum3h(V[0] * a[0], V[1] * a[1], V[2] * a[2]), sum3h(V[3] * a[0], V[4] * a[1], V[5] * a[2]),
                          sum3h(V[6] * a[0], V[7] * a[1], V[8] * a[2]) };
    // translation: te + qe * t (Eigen's _transformVector)
    {
        const float* v = P.t;
        float uv[3] = { qe[1] * v[2] - qe[2] * v[1], qe[2] * v[0] - qe[0] * v[2], qe[0] * v[1] - qe[1] * v[0] };
        for (int i = 0; i < 3; ++i) uv[i] += uv[i];
        const float c[3] = { qe[1] * uv[2] - qe[2] * uv[1], qe[2] * uv[0] - qe[0] * uv[2], qe[0] * uv[1] - qe[1] * uv[0] };
        float rt[3];
        for (int i = 0; i < 3; ++i) rt[i] = v[i] + qe[3] * uv[i] + c[i];
        for (int i = 0; i < 3; ++i) P.t[i] = te[i] + rt[i];
    }
    // rotation: quaternion product (so3.hpp:334-338), then SO3(quaternion) normalises
    const float* b = P.q;
    float r[4];
    r[3] = qe[3] * b[3] - qe[0] * b[0] - qe[1] * b[1] - qe[2] * b[2];
    r[0] = qe[3] * b[0] + qe[0] * b[3] + qe[1] * b[2] - qe[2] * b[1];
    r[1] = qe[3] * b[1] + qe[1] * b[3] + qe[2] * b[0] - qe[0] * b[2];
    r[2] = qe[3] * b[2] + qe[2] * b[3] + qe[0] * b[1] - qe[1] * b[0];
    const float len = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3]);
    for (int i = 0; i < 4; ++i) P.q[i] = r[i] / len;
}

}  // namespace

int psg_tracker_optimize(psg_volume* v, const float K[9], const float pose_in[16], const psg_tracker_config* cfg, float pose_out[16],
                         int* converged, int* iterations) {
    if (!v || !K || !pose_in || !cfg || !pose_out) { set_error("psg_tracker_optimize: null argument"); return PSG_ERR_INVALID; }
    if (cfg->num_iterations < 0 || cfg->sampling < 1) { set_error("psg_tracker_optimize: bad configuration"); return PSG_ERR_INVALID; }
    Se3f P;
    float R[9];
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) R[3 * r + c] = pose_in[4 * r + c]; P.t[r] = pose_in[4 * r + 3]; }
    quat_from_rot(R, P.q);
    int conv = 0, it = 0;
    const float thr_sq = cfg->conv_threshold * cfg->conv_threshold;
    for (; it < cfg->num_iterations; ++it) {
        rot_from_quat(P.q, R);
        double E = 0.0, g6[6], H36[36];
        long long count = 0;
        const int st = psg_tracker_system(v, K, R, P.t, cfg->sampling, &E, g6, H36, &count);
        if (st != PSG_OK) return st;
        if (count == 0) break;                                   // "return false" (RigidPointOptimizer.cpp:64-66)
        float Hf[36], gf[6], xi[6];
        for (int i = 0; i < 36; ++i) Hf[i] = (float)H36[i];
        for (int i = 0; i < 6; ++i) gf[i] = (float)g6[i];
        if (!llt_solve6(Hf, gf, xi)) break;
        float n2 = 0.f;
        for (int i = 0; i < 6; ++i) { xi[i] = cfg->damping * xi[i]; }
        for (int i = 0; i < 6; ++i) n2 += xi[i] * xi[i];
        if (n2 < thr_sq) { conv = 1; break; }
        float neg[6];
        for (int i = 0; i < 6; ++i) neg[i] = -xi[i];
        se3_exp_left(neg, P);
    }
    rot_from_quat(P.q, R);
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) pose_out[4 * r + c] = R[3 * r + c]; pose_out[4 * r + 3] = P.t[r]; }
    pose_out[12] = 0.f; pose_out[13] = 0.f; pose_out[14] = 0.f; pose_out[15] = 1.f;
    if (converged) *converged = conv;
    if (iterations) *iterations = it;
    return PSG_OK;
}

}  // extern "C"
