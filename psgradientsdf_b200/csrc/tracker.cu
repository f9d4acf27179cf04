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

}  // extern "C"
