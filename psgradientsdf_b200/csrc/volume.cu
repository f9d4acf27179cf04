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

// One voxel of VolumetricGradSdf::update.  Gates in the reference's order (VolumetricGradSdf.cpp:84-116); a voxel that fails a
// gate touches no volume memory.
__device__ __forceinline__ void fuse_voxel(const VolView& v, const float* __restrict__ frame, const FuseParams& P, int i, int j, int k,
                                           long long lin) {
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

// Depth range of the valid pixels (zmin < depth < zmax, the reference's gate) per 8 x 8 pixel tile (0 = none).
constexpr int kTile = 8;
__global__ void __launch_bounds__(256) k_depth_tiles(const float* __restrict__ frame, int H, int W, float zmin, float zmax, int tx,
                                                    int ntiles, float* __restrict__ tile_max) {
    // one warp per tile: lane -> 2 pixels of the 8 x 8 tile
    const int tile = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (tile >= ntiles) return;
    const int ty0 = (tile / tx) * kTile, tx0 = (tile % tx) * kTile;
    float m = 0.f;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int q = lane + 32 * h, px = tx0 + (q & 7), py = ty0 + (q >> 3);
        if (px < W && py < H) {
            const float d = __ldg(frame + 8ll * ((long long)py * W + px));
            if (d > zmin && d < zmax) m = fmaxf(m, d);
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) tile_max[tile] = m;
}

// Small volumes: one thread per voxel, 3-D launch ((i, j, k) from the block / thread indices -- a linear index would cost two
// 64-bit divisions per voxel, more than the projection itself).
__global__ void __launch_bounds__(256) k_integrate(VolView v, const float* __restrict__ frame, FuseParams P) {
    const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
    if (i >= v.dims[0]) return;
    const int j = (int)blockIdx.y, k = (int)blockIdx.z;
    const long long lin = (long long)i + (long long)v.dims[0] * ((long long)j + (long long)v.dims[1] * k);
    fuse_voxel(v, frame, P, i, j, k, lin);
}

// Fusion over bricks of 8 x 8 x 8 voxels.  Most voxels of a frame are rejected only after they have been projected -- they lie
// behind the surface by more than T or on pixels without depth -- and the projection is what a per-voxel kernel spends its
// time on (instruction bound; ~4 % of the voxels are written).  So the bricks are tested first, one thread per brick
// (k_cull_bricks): the 8 corners are projected with the voxel formulas; if all lie in front of the camera, the brick's voxels
// can only hit pixels inside the corners' bounding box (+2 px), and if that box misses the image, holds no valid depth, or
// every valid depth in it is more than T in front of the brick's nearest corner (all weights 0), no voxel of the brick can
// pass the reference's gates.  The surviving bricks are appended to a list and k_integrate_list runs the unchanged per-voxel
// code on them (persistent blocks striding over the list), so the result is the per-voxel result bit for bit.
constexpr int kBrick = 8;
__global__ void __launch_bounds__(128) k_cull_bricks(VolView v, FuseParams P, const float* __restrict__ tile_max, int tx, int nbx, int nby,
                                                    int nbricks, int* __restrict__ list, int* __restrict__ count) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    bool skip = true;
    if (b < nbricks) {
    const int bx = b % nbx, by = (b / nbx) % nby, bz = b / (nbx * nby);
    const int i0 = bx * kBrick, j0 = by * kBrick, k0 = bz * kBrick;
    float zlo = 3.4e38f, zhi = -3.4e38f, nlo = 3.4e38f, nhi = -3.4e38f, mlo = 3.4e38f, mhi = -3.4e38f;
    bool finite = true;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const int ci = min(i0 + ((c & 1) ? kBrick - 1 : 0), v.dims[0] - 1);
        const int cj = min(j0 + ((c & 2) ? kBrick - 1 : 0), v.dims[1] - 1);
        const int ck = min(k0 + ((c & 4) ? kBrick - 1 : 0), v.dims[2] - 1);
        const float p[3] = { v.origin[0] + v.vs * (float)ci, v.origin[1] + v.vs * (float)cj, v.origin[2] + v.vs * (float)ck };
        const float q[3] = { p[0] - P.t[0], p[1] - P.t[1], p[2] - P.t[2] };
        float pt[3];
        matT_vec(P.R, q, pt);
        const float n = P.cx + P.fx * pt[0] / pt[2], m = P.cy + P.fy * pt[1] / pt[2];
        finite = finite && (n == n) && (m == m);
        zlo = fminf(zlo, pt[2]); zhi = fmaxf(zhi, pt[2]);
        nlo = fminf(nlo, n); nhi = fmaxf(nhi, n); mlo = fminf(mlo, m); mhi = fmaxf(mhi, m);
    }
    skip = false;
    const float eps = 1e-4f * fmaxf(1.f, fabsf(zhi)) + v.vs * 1e-3f;
    if (zhi < -eps) {
        skip = true;                                    // the whole brick is behind the camera (gate pt.z < 0)
    } else if (zlo > eps && finite) {
        const float n0f = nlo - 2.f, n1f = nhi + 2.f, m0f = mlo - 2.f, m1f = mhi + 2.f;
        if (n1f < 0.f || n0f >= (float)P.W || m1f < 0.f || m0f >= (float)P.H) {
            skip = true;                                // projects outside the image
        } else {
            const int a0 = max(0, (int)floorf(n0f)) / kTile, a1 = min(P.W - 1, (int)ceilf(n1f)) / kTile;
            const int b0 = max(0, (int)floorf(m0f)) / kTile, b1 = min(P.H - 1, (int)ceilf(m1f)) / kTile;
            if ((a1 - a0 + 1) * (b1 - b0 + 1) <= 256) {   // (huge footprints: not worth the scan, keep the brick)
                float dmax = 0.f;
                for (int bb = b0; bb <= b1; ++bb)
                    for (int aa = a0; aa <= a1; ++aa) dmax = fmaxf(dmax, __ldg(tile_max + bb * tx + aa));
                // no valid depth under the brick, or even its nearest corner is more than T behind every valid depth
                if (dmax == 0.f || dmax - zlo < -(v.T * 1.0001f + eps)) skip = true;
            }
        }
    }
    }
    // warp-aggregated append: one atomic per warp (every survivor adding to the same counter would serialise in L2)
    const unsigned keep = __ballot_sync(0xffffffffu, !skip);
    if (!skip) {
        const int lane = threadIdx.x & 31;
        const int leader = __ffs(keep) - 1;
        int base = 0;
        if (lane == leader) base = atomicAdd(count, __popc(keep));
        base = __shfl_sync(keep, base, leader);
        list[base + __popc(keep & ((1u << lane) - 1u))] = b;
    }
}

__global__ void __launch_bounds__(256) k_integrate_list(VolView v, const float* __restrict__ frame, FuseParams P, const int* __restrict__ list,
                                                       const int* __restrict__ count, int nbx, int nby) {
    const int n = *count;
    const int ii = threadIdx.x & 7, jj = (threadIdx.x >> 3) & 7, kk = threadIdx.x >> 6;   // kk in 0..3: two z layers each
    for (int e = blockIdx.x; e < n; e += gridDim.x) {
        const int b = list[e];
        const int bx = b % nbx, by = (b / nbx) % nby, bz = b / (nbx * nby);
        const int i = bx * kBrick + ii, j = by * kBrick + jj;
        if (i >= v.dims[0] || j >= v.dims[1]) continue;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int k = bz * kBrick + kk + 4 * h;
            if (k >= v.dims[2]) continue;
            const long long lin = (long long)i + (long long)v.dims[0] * ((long long)j + (long long)v.dims[1] * k);
            fuse_voxel(v, frame, P, i, j, k, lin);
        }
    }
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
    const int nbx = (v->dims[0] + kBrick - 1) / kBrick, nby = (v->dims[1] + kBrick - 1) / kBrick, nbz = (v->dims[2] + kBrick - 1) / kBrick;
    const long long nbricks_ll = (long long)nbx * nby * nbz;
    if (nbricks_ll < 32768) {
        // below ~256^3 the three launches of the culled path cost more than projecting every voxel
        int bx = (v->dims[0] + 31) / 32 * 32;
        if (bx > 256) bx = 256;
        const dim3 grid3((unsigned)((v->dims[0] + bx - 1) / bx), (unsigned)v->dims[1], (unsigned)v->dims[2]);
        { Prof pf(v->stream, KID_FUSE); k_integrate<<<grid3, bx, 0, v->stream>>>(view, v->frame.p, P); }
        PSG_LAUNCH_CHECK();
        return;
    }
    const int nbricks = (int)nbricks_ll;
    const int tx = (W + kTile - 1) / kTile, ty = (H + kTile - 1) / kTile;
    v->tiles.alloc((size_t)tx * ty);
    { Prof pf(v->stream, KID_FUSE); k_depth_tiles<<<(tx * ty + 7) / 8, 256, 0, v->stream>>>(v->frame.p, H, W, v->zmin, v->zmax, tx, tx * ty, v->tiles.p); }
    PSG_LAUNCH_CHECK();
    v->brick_list.alloc((size_t)nbricks + 1);
    int* count = v->brick_list.p + nbricks;
    PSG_CUDA(cudaMemsetAsync(count, 0, sizeof(int), v->stream));
    { Prof pf(v->stream, KID_FUSE);
      k_cull_bricks<<<(nbricks + 127) / 128, 128, 0, v->stream>>>(view, P, v->tiles.p, tx, nbx, nby, nbricks, v->brick_list.p, count); }
    PSG_LAUNCH_CHECK();
    const int grid = nbricks < 148 * 8 ? nbricks : 148 * 8;
    { Prof pf(v->stream, KID_FUSE); k_integrate_list<<<grid, 256, 0, v->stream>>>(view, v->frame.p, P, v->brick_list.p, count, nbx, nby); }
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
