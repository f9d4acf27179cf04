// Gradient-SDF volume: structure-of-arrays store in HBM and depth-frame fusion.
//
// Replaces VolumetricGradSdf::init / update (cpp/include/sdf_tracker/VolumetricGradSdf.cpp:14-38, 51-138).
// Layout: eight float planes {dist, gx, gy, gz, weight, r, g, b} of N = W*H*D voxels, x fastest, plus
// ceil(max_frames/64) planes of 64-bit visibility words (bit c of a voxel = "fused in input frame c",
// the reference's std::vector<bool> vis_, VolumetricGradSdf.h:42).
#include "psg_host.h"

namespace psg {

__global__ void __launch_bounds__(256) k_fill_volume(VolView v) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < v.N; i += stride) {
        v.dist[i] = v.T; v.gx[i] = 0.f; v.gy[i] = 0.f; v.gz[i] = 0.f;
        v.w[i] = 0.f; v.r[i] = 0.f; v.g[i] = 0.f; v.b[i] = 0.f;
        for (int w = 0; w < v.nwords; ++w) v.vis[(long long)w * v.N + i] = 0ull;
    }
}

// planar staging (colour HxWx3 BGR | depth | nx | ny | nz | n_sq_inv) -> one 32-byte record per pixel
// {depth, nx, ny, nz, n_sq_inv, b, g, r}: the fusion kernel then gathers one sector per voxel.
__global__ void __launch_bounds__(256) k_pack_frame(const float* __restrict__ planar, float* __restrict__ packed, int HW) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= HW) return;
    const float* color = planar;
    const float* depth = planar + 3ll * HW;
    const float* nx = depth + HW;
    const float* ny = nx + HW;
    const float* nz = ny + HW;
    const float* nsq = nz + HW;
    float4 a = make_float4(depth[p], nx[p], ny[p], nz[p]);
    float4 b = make_float4(nsq[p], color[3 * p], color[3 * p + 1], color[3 * p + 2]);
    reinterpret_cast<float4*>(packed)[2 * p] = a;
    reinterpret_cast<float4*>(packed)[2 * p + 1] = b;
}

struct FuseParams {
    float fx, fy, cx, cy;
    float R[9], t[3];
    int H, W, counter;
};

// One thread per voxel, x fastest (coalesced plane accesses).  Gates in the reference's order
// (VolumetricGradSdf.cpp:84-116); voxels that fail a gate touch no volume memory.
// 3-D launch: (x-run, j, k) come from the block / thread indices -- a linear index would cost two 64-bit divisions per voxel,
// more than the projection itself (the kernel is instruction bound: most voxels are rejected before they touch memory).
__global__ void __launch_bounds__(256) k_integrate(VolView v, const float* __restrict__ frame, FuseParams P) {
    const int W0 = v.dims[0], W1 = v.dims[1];
    const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (i >= W0) return;
    const int j = (int)blockIdx.y, k = (int)blockIdx.z;
    const long long lin = (long long)i + (long long)W0 * ((long long)j + (long long)W1 * k);

    const float p[3] = { v.origin[0] + v.vs * (float)i, v.origin[1] + v.vs * (float)j, v.origin[2] + v.vs * (float)k };
    const float q[3] = { p[0] - P.t[0], p[1] - P.t[1], p[2] - P.t[2] };
    float pt[3];
    matT_vec(P.R, q, pt);
    if (pt[2] < 0.f) return;
    // static_cast<int>(c + f*x/z + 0.5): float expression, +0.5 in double, truncation toward zero; values that
    // do not fit an int (or NaN) become INT_MIN on x86 and are rejected by the n<0 test.
    const double xn = (double)(P.cx + P.fx * pt[0] / pt[2]) + 0.5;
    const double xm = (double)(P.cy + P.fy * pt[1] / pt[2]) + 0.5;
    if (!(xn > -2147483649.0 && xn < 2147483648.0) || !(xm > -2147483649.0 && xm < 2147483648.0)) return;
    const int n = __double2int_rz(xn);
    const int m = __double2int_rz(xm);
    if (n < 0 || n >= P.W || m < 0 || m >= P.H) return;

    const float4* px = reinterpret_cast<const float4*>(frame) + 2ll * ((long long)m * P.W + n);
    const float4 a = __ldg(px);
    const float z = a.x;
    if (z <= v.zmin || z >= v.zmax) return;
    const float sdf = z - pt[2];
    float w = 0.f;
    if (sdf >= 0.f) w = 1.f; else if (sdf >= -v.T) w = 1.f + sdf * v.invT;   // Sdf::weight (Sdf.h:48-66)
    if (w == 0.f) return;
    const float nrm[3] = { a.y, a.z, a.w };
    if ((double)sqnorm3(nrm) < .1) return;
    const float4 b = __ldg(px + 1);
    const float zi = (float)(1.0 / (double)pt[2]);
    const float xy[3] = { zi * pt[0], zi * pt[1], zi * pt[2] };
    const float d = dot3(nrm, xy);
    if ((double)(d * d * b.x) < .25 * .25) return;

    const float wsum = v.w[lin] + w;
    v.w[lin] = wsum;
    const float tr = fmaxf(-v.T, fminf(v.T, sdf));   // Sdf::truncate (Sdf.h:44-46)
    float cur = v.dist[lin];
    v.dist[lin] = cur + (tr - cur) * w / wsum;
    v.gx[lin] -= sum3((w * P.R[0]) * nrm[0], (w * P.R[1]) * nrm[1], (w * P.R[2]) * nrm[2]);
    v.gy[lin] -= sum3((w * P.R[3]) * nrm[0], (w * P.R[4]) * nrm[1], (w * P.R[5]) * nrm[2]);
    v.gz[lin] -= sum3((w * P.R[6]) * nrm[0], (w * P.R[7]) * nrm[1], (w * P.R[8]) * nrm[2]);
    cur = v.r[lin]; v.r[lin] = cur + (b.w - cur) * w / wsum;   // z_color[2] = red
    cur = v.g[lin]; v.g[lin] = cur + (b.z - cur) * w / wsum;
    cur = v.b[lin]; v.b[lin] = cur + (b.y - cur) * w / wsum;
    // vis.resize(counter_); vis.push_back(true)
    v.vis[(long long)(P.counter >> 6) * v.N + lin] |= 1ull << (P.counter & 63);
}

__global__ void __launch_bounds__(256) k_aos_to_soa(VolView v, const float* __restrict__ aos, long long begin, long long count) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float4 a = reinterpret_cast<const float4*>(aos)[2 * i];
    const float4 b = reinterpret_cast<const float4*>(aos)[2 * i + 1];
    const long long l = begin + i;
    v.dist[l] = a.x; v.gx[l] = a.y; v.gy[l] = a.z; v.gz[l] = a.w;
    v.w[l] = b.x; v.r[l] = b.y; v.g[l] = b.z; v.b[l] = b.w;
}
__global__ void __launch_bounds__(256) k_soa_to_aos(VolView v, float* __restrict__ aos, long long begin, long long count) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const long long l = begin + i;
    reinterpret_cast<float4*>(aos)[2 * i] = make_float4(v.dist[l], v.gx[l], v.gy[l], v.gz[l]);
    reinterpret_cast<float4*>(aos)[2 * i + 1] = make_float4(v.w[l], v.r[l], v.g[l], v.b[l]);
}

static inline unsigned blocks_for(long long n, int bs) { return (unsigned)((n + bs - 1) / bs); }

void launch_fill_volume(psg_volume* v) {
    const VolView view = v->view();
    { Prof pf(v->stream, KID_FILL); k_fill_volume<<<148 * 8, 256, 0, v->stream>>>(view); }
    PSG_LAUNCH_CHECK();
}

void launch_pack_frame(psg_volume* v, const float* planar_dev, int H, int W) {
    v->frame.alloc((size_t)H * W * 8);
    { Prof pf(v->stream, KID_PACK_FRAME); k_pack_frame<<<blocks_for((long long)H * W, 256), 256, 0, v->stream>>>(planar_dev, v->frame.p, H * W); }
    PSG_LAUNCH_CHECK();
}

void launch_integrate(psg_volume* v, int H, int W, const float* K, const float* R, const float* t, int counter) {
    FuseParams P;
    P.fx = K[0]; P.fy = K[4]; P.cx = K[2]; P.cy = K[5];
    for (int a = 0; a < 9; ++a) P.R[a] = R[a];
    for (int a = 0; a < 3; ++a) P.t[a] = t[a];
    P.H = H; P.W = W; P.counter = counter;
    const VolView view = v->view();
    int bx = (v->dims[0] + 31) / 32 * 32;
    if (bx > 256) bx = 256;
    const dim3 grid((unsigned)((v->dims[0] + bx - 1) / bx), (unsigned)v->dims[1], (unsigned)v->dims[2]);
    { Prof pf(v->stream, KID_FUSE); k_integrate<<<grid, bx, 0, v->stream>>>(view, v->frame.p, P); }
    PSG_LAUNCH_CHECK();
}

void launch_aos_to_soa(psg_volume* v, const float* aos_dev, long long begin, long long count) {
    { Prof pf(v->stream, KID_AOS); k_aos_to_soa<<<blocks_for(count, 256), 256, 0, v->stream>>>(v->view(), aos_dev, begin, count); }
    PSG_LAUNCH_CHECK();
}
void launch_soa_to_aos(psg_volume* v, float* aos_dev, long long begin, long long count) {
    { Prof pf(v->stream, KID_AOS); k_soa_to_aos<<<blocks_for(count, 256), 256, 0, v->stream>>>(v->view(), aos_dev, begin, count); }
    PSG_LAUNCH_CHECK();
}

}  // namespace psg

psg::VolView psg_volume::view() const {
    psg::VolView w;
    for (int a = 0; a < 3; ++a) { w.dims[a] = dims[a]; w.origin[a] = origin[a]; }
    w.N = N; w.vs = vs; w.T = T; w.invT = invT; w.zmin = zmin; w.zmax = zmax;
    w.dist = f[0].p; w.gx = f[1].p; w.gy = f[2].p; w.gz = f[3].p; w.w = f[4].p; w.r = f[5].p; w.g = f[6].p; w.b = f[7].p;
    w.vis = vis.p; w.nwords = nwords;
    return w;
}
