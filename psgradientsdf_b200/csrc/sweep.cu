// The per-(surface voxel, keyframe) photometric sweep: projection, bilinear sampling, shading, robust loss and the
// Jacobian contribution of the active block, plus the reductions that turn them into normal equations.
//
// One kernel template, k_sweep<MODEL, MODE>.  A thread owns one surface voxel (its record lives in registers for
// the whole sweep); the block walks the keyframes any of its voxels sees, in ascending order, all lanes on the
// same keyframe (uniform pose / light, gathers into one image).  Per-voxel sums (albedo, distance stencil) are
// therefore formed in the reference's own order (ascending keyframe, then channel).  Per-keyframe sums (light,
// pose) are reduced with a warp butterfly reduce-scatter, then per-warp shared-memory slices, then one
// double-precision atomic per block and entry.
//
// Reference functions covered: Optimizer::getIntensity + interpolateImage + computeImageGradient
// (OptimizerAux.cpp:207-231, Auxilary.h:41-123), {Ps,Led}Optimizer::renderedIntensity, computeResidual,
// getPSEnergy, rhoJacobian/albedoJacobian, lightJacobian, poseJacobian, distJacobian
// (PsOptimizerJa.cpp, LedOptimizerJa.cpp), Optimizer::computeDistGrad / normalJacobian / distRegJacobian /
// computeDistLaplacian / distLaplacianJacobian (Optimizer.cpp), Optimizer::initAlbedo (Optimizer.cpp:50-81),
// LedOptimizer::computeLightIntensive (LedOptimizer.cpp:76-112), Optimizer::updateAlbedo (OptimizerAux.cpp:120-150).
#include <mutex>

#include "psg_host.h"

namespace psg {

enum Mode { M_ENERGY = 0, M_INIT_ALBEDO, M_LED_INIT, M_PROBE, M_ALBEDO, M_LIGHT, M_POSE, M_DIST };

// A block is 32 surface voxels (one per lane) times W warps; the warps split the keyframes the 32 voxels see
// between them (W = 1, 2, 4 or 8, chosen at launch so that small surfaces still fill the 148 SMs).
constexpr int kMaxSplit = 8;
constexpr int kBlock = 128;   // block size of the auxiliary per-voxel kernels in this file
constexpr int kWarps = kBlock / 32;

struct SweepArgs {
    VolView vol;
    SurfView surf;
    Camera cam;
    LossParams loss;
    const float4* images;
    int F;
    int blk_first, blk_stride, blk_count;   // 32-slot blocks of the work list this launch walks: blk_first + i * blk_stride, i < blk_count
    float vs_inv;
    int cslot;          // which copy of the keyframe table in constant memory (c_frames) belongs to this optimizer
    float damping, reg_n, reg_l;
    int flags;
    double* red;        // per-keyframe reductions (zeroed by the launcher)
    int red_copies;     // blocks spread their atomics over this many copies of `red` (kRedDoubles apart), combined afterwards
    double* scal;       // 8 scalar results: [0..5] mode specific, [6] sum of robust losses, [7] active pairs
    double* partials;   // [gridDim.x][8] per-block scalar partials, combined in block order by the last block
    unsigned* counter;  // arrival counter for that (self-resetting)
    float* out0;        // mode dependent per-voxel outputs
    float* out1;
    // probe outputs
    float* pm; float* pn; unsigned char* pok; float* pI; float* prend; float* pr; float* pw;
};

struct Footprint { int x, y; float4 t00, t01, t10, t11; };

// Per-keyframe R, t, light of up to kConstSlots optimizers on this device, read through the constant bank with the warp-uniform
// keyframe index (operands of the arithmetic instructions, no load instructions, no registers).  The authoritative table lives
// in device memory (psg_optimizer::frames_dev), where the small solves update it; a 6 KB device-to-device copy refreshes the
// optimizer's slot here whenever it changed (sync_const_table).
__constant__ FrameParams c_frames[kConstSlots][kMaxFrames + 1];

__device__ __forceinline__ const float4* texel_ptr(const SweepArgs& A, int f, int row, int col) {
    return A.images + ((long long)f * A.cam.H + row) * A.cam.W + col;
}

// interpolateImage (Auxilary.h:41-61) called as interpolateImage(row = n, col = m): weights partly double, each
// weight * texel rounded to float, four float adds left to right; nearest texel in the last row / column.
__device__ __forceinline__ void sample_image(const SweepArgs& A, int f, float rowf, float colf, Footprint& fp, float* rgb) {
    const int x = (int)floorf(rowf), y = (int)floorf(colf);
    fp.x = x; fp.y = y;
    float bgr[3];
    if (x + 1 < A.cam.H && y + 1 < A.cam.W) {
        const float4* p0 = texel_ptr(A, f, x, y);
        const float4* p1 = p0 + A.cam.W;
        fp.t00 = __ldg(p0); fp.t01 = __ldg(p0 + 1); fp.t10 = __ldg(p1); fp.t11 = __ldg(p1 + 1);
        const double wy1 = (double)y + 1.0 - (double)colf;
        const double wx1 = (double)x + 1.0 - (double)rowf;
        const float fx_ = rowf - (float)x;
        const float fy_ = colf - (float)y;
        const double a1 = wy1 * (double)fx_;
        const double a2 = wy1 * wx1;
        const float a3 = fy_ * fx_;
        const double a4 = (double)fy_ * wx1;
        const float c10[3] = { fp.t10.x, fp.t10.y, fp.t10.z }, c00[3] = { fp.t00.x, fp.t00.y, fp.t00.z };
        const float c11[3] = { fp.t11.x, fp.t11.y, fp.t11.z }, c01[3] = { fp.t01.x, fp.t01.y, fp.t01.z };
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float s = (float)((double)c10[c] * a1) + (float)((double)c00[c] * a2);
            s = s + c11[c] * a3;
            s = s + (float)((double)c01[c] * a4);
            bgr[c] = s;
        }
    } else {
        fp.t00 = __ldg(texel_ptr(A, f, x, y));
        fp.t01 = fp.t00; fp.t10 = fp.t00; fp.t11 = fp.t00;
        bgr[0] = fp.t00.x; bgr[1] = fp.t00.y; bgr[2] = fp.t00.z;
    }
    rgb[0] = bgr[2]; rgb[1] = bgr[1]; rgb[2] = bgr[0];
}

__device__ __forceinline__ void sub_bgr(const float4& a, const float4& b, float* o) { o[0] = a.x - b.x; o[1] = a.y - b.y; o[2] = a.z - b.z; }
__device__ __forceinline__ void negadd_bgr(const float4& a, const float4& b, float* o) { o[0] = -a.x + b.x; o[1] = -a.y + b.y; o[2] = -a.z + b.z; }

// computeImageGradient (Auxilary.h:64-123), both directions at once: g0 = along columns, g1 = along rows (RGB).
// `fp` holds the 2x2 footprint at (x, y) when it is interior.  The reference's reads one element past the image at
// the bottom-right corner cell are clamped to the centre texel.
__device__ __forceinline__ void image_gradient(const SweepArgs& A, int f, float rowf, float colf, const Footprint& fp,
                                               float* g0, float* g1) {
    const int x = fp.x, y = fp.y, H = A.cam.H, W = A.cam.W;
    const float w01 = rowf - (float)x, w11 = colf - (float)y;
    const float w00 = (float)(1.0 - (double)w01), w10 = (float)(1.0 - (double)w11);
    float v0[3], v1[3];
    if (x + 1 < H && y + 1 < W) {
        sub_bgr(fp.t01, fp.t00, v0); sub_bgr(fp.t11, fp.t10, v1);
#pragma unroll
        for (int c = 0; c < 3; ++c) g0[c] = w00 * v0[2 - c] + w01 * v1[2 - c];
        sub_bgr(fp.t10, fp.t00, v0); sub_bgr(fp.t11, fp.t01, v1);
#pragma unroll
        for (int c = 0; c < 3; ++c) g1[c] = w10 * v0[2 - c] + w11 * v1[2 - c];
        return;
    }
    const float4 c00 = __ldg(texel_ptr(A, f, x, y));
    // direction 0
    if (x + 1 >= H) {
        const float4 c01 = (y + 1 < W) ? __ldg(texel_ptr(A, f, x, y + 1)) : c00;
        sub_bgr(c01, c00, v0);
#pragma unroll
        for (int c = 0; c < 3; ++c) g0[c] = v0[2 - c];
    } else {   // x + 1 < H && y + 1 >= W : backward differences
        negadd_bgr(__ldg(texel_ptr(A, f, x, y - 1)), c00, v0);
        negadd_bgr(__ldg(texel_ptr(A, f, x + 1, y - 1)), __ldg(texel_ptr(A, f, x + 1, y)), v1);
#pragma unroll
        for (int c = 0; c < 3; ++c) g0[c] = w00 * v0[2 - c] + w01 * v1[2 - c];
    }
    // direction 1
    if (x + 1 >= H && y + 1 < W) {
        negadd_bgr(__ldg(texel_ptr(A, f, x - 1, y)), c00, v0);
        negadd_bgr(__ldg(texel_ptr(A, f, x - 1, y + 1)), __ldg(texel_ptr(A, f, x, y + 1)), v1);
#pragma unroll
        for (int c = 0; c < 3; ++c) g1[c] = w10 * v0[2 - c] + w11 * v1[2 - c];
    } else {   // y + 1 >= W
        const float4 c10 = (x + 1 < H) ? __ldg(texel_ptr(A, f, x + 1, y)) : c00;
        sub_bgr(c10, c00, v0);
#pragma unroll
        for (int c = 0; c < 3; ++c) g1[c] = v0[2 - c];
    }
}

// Optimizer::normalJacobian(grad, direction, lag = false) (Optimizer.cpp:269-284)
__device__ __forceinline__ void normal_jacobian(float vs_inv, const float* grad, const float* direction, float* Jn) {
    float nd[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) nd[a] = -vs_inv * direction[a];
    const float N_inv = (float)(1.0 / (double)fmaxf(norm3(grad), 0.001f));
    const float dN = (float)(cube_d(N_inv) * (double)dot3(nd, grad));
#pragma unroll
    for (int a = 0; a < 3; ++a) Jn[a] = N_inv * nd[a] - dN * grad[a];
}

// V values per lane -> lane l adds the warp sum of value l to dst[l] (one double atomic per lane and keyframe).
template <int V>
__device__ __forceinline__ void warp_accumulate(float* vals, int lane, double* dst /* >= V doubles */) {
    const float s = warp_reduce_scatter<V>(vals, lane);
    if (lane < V && s != 0.f) atomicAdd(&dst[lane], (double)s);
}

// Deterministic grid-wide sum of 8 per-block doubles, two levels so that no single block serialises the whole grid:
// blocks are grouped by kRedGroup; the last block of a group to arrive adds the group's partials (fixed association:
// thread t takes elements t, t + blockDim, ... -- independent, coalesced loads -- then the threads holding the same slot
// are combined in thread order) into a group partial; the last group to finish adds the group partials the same way and
// writes out[0..8).  Counters reset themselves.
constexpr int kRedGroup = 128;

__device__ __forceinline__ void fixed_order_sum8(const double* src, size_t n /* multiple of 8 */, double* dst, double* s_scratch) {
    const int t = threadIdx.x;
    double acc = 0.0;
    size_t e = t;
    for (; e + 3 * (size_t)blockDim.x < n; e += 4 * (size_t)blockDim.x) {
        const double a0 = __ldcg(src + e), a1 = __ldcg(src + e + blockDim.x);
        const double a2 = __ldcg(src + e + 2 * (size_t)blockDim.x), a3 = __ldcg(src + e + 3 * (size_t)blockDim.x);
        acc += a0; acc += a1; acc += a2; acc += a3;
    }
    for (; e < n; e += blockDim.x) acc += __ldcg(src + e);
    __syncthreads();
    s_scratch[t] = acc;
    __syncthreads();
    if (t < 8) {
        double v = 0.0;
        for (int j = t; j < (int)blockDim.x; j += 8) v += s_scratch[j];
        dst[t] = v;
    }
}

__device__ __forceinline__ void grid_sum_scalars(const double* block_vals, double* partials, unsigned* counter, double* out,
                                                 bool* s_last, double* s_scratch /* blockDim.x doubles */) {
    const int t = threadIdx.x;
    const unsigned ngroups = (gridDim.x + kRedGroup - 1) / kRedGroup;
    const unsigned group = blockIdx.x / kRedGroup;
    const unsigned gfirst = group * kRedGroup;
    const unsigned gsize = min((unsigned)kRedGroup, gridDim.x - gfirst);
    double* group_partials = partials + (size_t)gridDim.x * 8;   // [ngroups][8] behind the block partials
    if (t < 8) partials[(size_t)blockIdx.x * 8 + t] = block_vals[t];
    __threadfence();
    __syncthreads();
    if (t == 0) *s_last = (atomicAdd(counter + 1 + group, 1u) == gsize - 1);
    __syncthreads();
    if (!*s_last) return;
    __threadfence();
    fixed_order_sum8(partials + (size_t)gfirst * 8, (size_t)gsize * 8, group_partials + (size_t)group * 8, s_scratch);
    if (t == 0) counter[1 + group] = 0u;
    __threadfence();
    __syncthreads();
    if (t == 0) *s_last = (atomicAdd(counter, 1u) == ngroups - 1);
    __syncthreads();
    if (!*s_last) return;
    __threadfence();
    fixed_order_sum8(group_partials, (size_t)ngroups * 8, out, s_scratch);
    if (t == 0) *counter = 0u;
}

template <int MODEL, int MODE>
__global__ void __launch_bounds__(32 * kMaxSplit, (MODE == M_DIST) ? 2 : ((MODE == M_POSE) ? 3 : 4)) k_sweep(const __grid_constant__ SweepArgs A) {
    constexpr int NB = (MODEL == PSG_MODEL_SH2) ? 9 : 4;
    constexpr bool LED = (MODEL == PSG_MODEL_LED);
    // per-keyframe reduction width (padded to a power of two / multiple of 32 lanes)
    constexpr int VLIGHT = LED ? 0 : (NB == 4 ? 16 : 64);
    constexpr int VRED = (MODE == M_POSE) ? 32 : ((MODE == M_LIGHT) ? VLIGHT : 0);
    // per-voxel accumulators that the keyframe-splitting warps have to combine
    constexpr int NACC = (MODE == M_DIST) ? 14 : ((MODE == M_ALBEDO) ? 6 : ((MODE == M_INIT_ALBEDO) ? 4 : 0));

    extern __shared__ float s_dyn[];   // [W][NACC][32]
    __shared__ double s_red[kMaxSplit][8];
    __shared__ double s_tot[8];
    __shared__ double s_scratch[32 * kMaxSplit];
    __shared__ bool s_last;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, W = blockDim.x >> 5;
    const int S = A.surf.S;
    const int slot = (A.blk_first + (int)blockIdx.x * A.blk_stride) * 32 + lane;   // position in the visibility-sorted work list
    const bool live = slot < A.surf.nslots;

    // ---- voxel record (five coalesced 128-bit loads; everything per-voxel is precomputed by k_gather_records)
    int lin = 0, s = 0;
    unsigned long long vis = 0ull;
    float dist = 0.f, gh[3] = {0, 0, 0}, rho[3] = {0, 0, 0}, q0[3] = {0, 0, 0};
    float fd[3] = {0, 0, 0}, posf[3] = {1, 1, 1}, nfd[3] = {0, 0, 0};
    float rec1w = 0.f;
    if (live) {
        vis = A.surf.pvis[slot];
        const float4 r0 = A.surf.rec0[slot], r1 = A.surf.rec1[slot], r2 = A.surf.rec2[slot], r3 = A.surf.rec3[slot],
                     r4 = A.surf.rec4[slot];
        dist = r0.x; gh[0] = r0.y; gh[1] = r0.z; gh[2] = r0.w;
        rho[0] = r1.x; rho[1] = r1.y; rho[2] = r1.z; rec1w = r1.w;
        fd[0] = r2.x; fd[1] = r2.y; fd[2] = r2.z; s = __float_as_int(r2.w);
        lin = __float_as_int(r3.w);
        nfd[0] = r4.x; nfd[1] = r4.y; nfd[2] = r4.z;
        const int pb = __float_as_int(r1.w);
        const float pw[3] = { r3.x, r3.y, r3.z };
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            posf[a] = ((pb >> a) & 1) ? 1.f : -1.f;
            q0[a] = pw[a] - dist * gh[a];
        }
    }
    float sh_n[NB];   // SH basis at the finite-difference normal (shading in the residual)
    float sh_g[NB];   // SH basis at the stored gradient (albedo / light Jacobians)
    sh_eval<NB>(nfd, sh_n);
    sh_eval<NB>(gh, sh_g);

    // ---- keyframes any of the 32 voxels sees (every warp of the block holds the same 32 voxels)
    unsigned long long uni;
    {
        const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)vis);
        const unsigned hi = __reduce_or_sync(0xffffffffu, (unsigned)(vis >> 32));
        uni = ((unsigned long long)hi << 32) | lo;
    }

    // ---- mode-specific per-voxel state
    double acc_d[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // ENERGY: [0] loss; LED_INIT: I[3], rend[3]; LED light: H[3], b[3]
    long long cnt = 0;
    double eloss = 0.0;                           // robust loss of the current state (every mode that forms residuals)
    int accepted = 0;
    float vacc[14];                               // ALBEDO: H[3], b[3]; INIT_ALBEDO: sum[3]; DIST: A(10) + b(4)
#pragma unroll
    for (int i = 0; i < 14; ++i) vacc[i] = 0.f;
    int icount = 0;

    // DIST: frame independent parts of distJacobian (PsOptimizerJa.cpp:194-221)
    float dn[4][3], dx[4][3];
    int cols[4] = { s, -1, -1, -1 };
    if constexpr (MODE == M_DIST) {
        if (live) {
            normal_jacobian(A.vs_inv, fd, posf, dn[0]);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                float dir[3] = {0.f, 0.f, 0.f};
                dir[k] = LED ? (0.f + posf[k]) : (0.f - posf[k]);   // LedOptimizerJa.cpp:158 "+=" vs PsOptimizerJa.cpp:201 "-="
                normal_jacobian(A.vs_inv, fd, dir, dn[k + 1]);
                cols[k + 1] = A.surf.nbr_row[k * S + s];
            }
#pragma unroll
            for (int a = 0; a < 3; ++a) dx[0][a] = -gh[a] - dist * dn[0][a];
#pragma unroll
            for (int k = 1; k < 4; ++k)
#pragma unroll
                for (int a = 0; a < 3; ++a) dx[k][a] = -dist * dn[k][a];
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int a = 0; a < 3; ++a) { dn[k][a] = 0.f; dx[k][a] = 0.f; }
        }
    }

    const float fx = A.cam.fx, fy = A.cam.fy, cx = A.cam.cx, cy = A.cam.cy;
    float led[3] = {0.f, 0.f, 0.f};
    if constexpr (LED) { led[0] = c_frames[A.cslot][kMaxFrames].light[0]; led[1] = c_frames[A.cslot][kMaxFrames].light[1]; led[2] = c_frames[A.cslot][kMaxFrames].light[2]; }

    // ---- keyframe loop: ascending, every lane of a warp on the same keyframe; warp w takes every W-th visible keyframe
    if (W > 1) {   // keep every W-th set bit, starting with the warp-th (one pass over the bits, outside the hot loop)
        unsigned long long mine = 0ull, rest = uni;
        int k = 0;
        while (rest) {
            const unsigned long long low = rest & (0ull - rest);
            if ((k & (W - 1)) == warp) mine |= low;
            rest ^= low;
            ++k;
        }
        uni = mine;
    }
    while (uni) {
        const int f = __ffsll((long long)uni) - 1;
        uni &= uni - 1;
        const FrameParams& P = c_frames[A.cslot][f];   // constant bank: no load at the head of the iteration
        const bool act = live && ((vis >> f) & 1ull);

        bool ok = false;
        float pt[3] = {0, 0, 1}, I[3] = {0, 0, 0}, rend[3] = {0, 0, 0}, r[3] = {0, 0, 0}, w[3] = {0, 0, 0};
        float m = 0.f, n = 0.f, pair_loss = 0.f;
        Footprint fp; fp.x = 0; fp.y = 0;
        if (act) {
            const float q[3] = { q0[0] - P.t[0], q0[1] - P.t[1], q0[2] - P.t[2] };
            matT_vec(P.R, q, pt);
            const float z_inv = (float)(1.0 / (double)pt[2]);
            m = fx * pt[0] * z_inv + cx;
            n = fy * pt[1] * z_inv + cy;
            ok = !(m < 0.f || m >= (float)A.cam.W || n < 0.f || n >= (float)A.cam.H) && (m == m) && (n == n);
            if (ok) {
                sample_image(A, f, n, m, fp, I);
                float irr;
                if constexpr (LED) {   // LedOptimizerJa.cpp:15-29
                    float nR[3];
#pragma unroll
                    for (int a = 0; a < 3; ++a) nR[a] = sum3((-nfd[0]) * P.R[a], (-nfd[1]) * P.R[3 + a], (-nfd[2]) * P.R[6 + a]);
                    irr = dot3(nR, pt);
                    irr /= (float)cube_d(norm3(pt));
#pragma unroll
                    for (int c = 0; c < 3; ++c) rend[c] = rho[c] * led[c] * irr;
                } else {     // PsOptimizerJa.cpp:30-40
                    irr = dot_sh<NB>(P.light, sh_n);
#pragma unroll
                    for (int c = 0; c < 3; ++c) rend[c] = rho[c] * irr;
                }
#pragma unroll
                for (int c = 0; c < 3; ++c) r[c] = I[c] - rend[c];
                if (A.loss.loss == PSG_LOSS_CAUCHY) {
                    // Cauchy (every shipped config): w = 1 / (1 + (r/lambda)^2), loss = sum_c log(1 + (r_c/lambda)^2).
                    // One division per channel serves both; the three logarithms are taken as one log of the product
                    // (|r| <= 1 and lambda > 0 keep it far from overflow), which changes a per-pair loss by ~1 ulp.
                    float tq[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) { const float q = r[c] / A.loss.lambda; tq[c] = 1.f + q * q; w[c] = 1.f / tq[c]; }
                    if constexpr (MODE != M_PROBE && MODE != M_INIT_ALBEDO && MODE != M_LED_INIT) pair_loss = logf((tq[0] * tq[1]) * tq[2]);
                } else {
#pragma unroll
                    for (int c = 0; c < 3; ++c) w[c] = irls_weight(A.loss, r[c]);
                    if constexpr (MODE != M_PROBE && MODE != M_INIT_ALBEDO && MODE != M_LED_INIT) pair_loss = robust_loss3(A.loss, r);
                }
            }
        }

        if constexpr (MODE != M_PROBE && MODE != M_INIT_ALBEDO && MODE != M_LED_INIT) {
            if (ok) { eloss += (double)pair_loss; ++cnt; }
        }
        if constexpr (MODE == M_ENERGY) {
        } else if constexpr (MODE == M_INIT_ALBEDO) {
            if (ok) { vacc[0] += I[0]; vacc[1] += I[1]; vacc[2] += I[2]; ++icount; }
        } else if constexpr (MODE == M_LED_INIT) {
            if (ok) {
#pragma unroll
                for (int c = 0; c < 3; ++c) { acc_d[c] += (double)I[c]; acc_d[3 + c] += (double)rend[c]; }
            }
        } else if constexpr (MODE == M_PROBE) {
            if (act) {
                const long long k = (long long)s * A.F + f;
                A.pm[k] = m; A.pn[k] = n; A.pok[k] = ok ? 1 : 0;
                if (ok) {
#pragma unroll
                    for (int c = 0; c < 3; ++c) { A.pI[3 * k + c] = I[c]; A.prend[3 * k + c] = rend[c]; A.pr[3 * k + c] = r[c]; A.pw[3 * k + c] = w[c]; }
                }
            }
        } else if constexpr (MODE == M_ALBEDO) {
            if (ok) {
                float Jr[3];
                if constexpr (LED) {   // LedOptimizerJa.cpp:85-99
                    float gR[3];
#pragma unroll
                    for (int a = 0; a < 3; ++a) gR[a] = sum3(gh[0] * P.R[a], gh[1] * P.R[3 + a], gh[2] * P.R[6 + a]);
                    float refl = dot3(gR, pt);
                    refl /= (float)cube_d(norm3(pt));
#pragma unroll
                    for (int c = 0; c < 3; ++c) Jr[c] = refl * led[c];
                } else {     // PsOptimizerJa.cpp:118-122
                    const float j = -dot_sh<NB>(P.light, sh_g);
                    Jr[0] = j; Jr[1] = j; Jr[2] = j;
                }
#pragma unroll
                for (int c = 0; c < 3; ++c) { vacc[c] += (Jr[c] * w[c]) * Jr[c]; vacc[3 + c] += (Jr[c] * w[c]) * r[c]; }
            }
        } else if constexpr (MODE == M_LIGHT) {
            if constexpr (LED) {       // LedOptimizerJa.cpp:101-115: one global 3-vector, diagonal system
                if (ok) {
                    float gR[3];
#pragma unroll
                    for (int a = 0; a < 3; ++a) gR[a] = sum3(gh[0] * P.R[a], gh[1] * P.R[3 + a], gh[2] * P.R[6 + a]);
                    float refl = dot3(gR, pt);
                    refl /= (float)cube_d(norm3(pt));
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const float j = refl * rho[c];
                        acc_d[c] += (double)((j * w[c]) * j);
                        acc_d[3 + c] += (double)((j * w[c]) * r[c]);
                    }
                }
            } else {         // PsOptimizerJa.cpp:132-143: rows -rho_c * SH(g_hat), lower triangle + rhs
                constexpr int NT = NB * (NB + 1) / 2;
                float vals[VLIGHT > 0 ? VLIGHT : 1];
#pragma unroll
                for (int i = 0; i < VLIGHT; ++i) vals[i] = 0.f;
                if (ok) {
                    // sum_c w_c (rho_c SH_i)(rho_c SH_k) = (sum_c w_c rho_c^2) SH_i SH_k, sum_c w_c (-rho_c SH_i) r_c likewise:
                    // the channel sums are two scalars per pair.  Contracted arithmetic: these sums are accumulated in fp64
                    // and compared with the reference's float-serial sums to 2e-5, not bit for bit.
                    float s2 = 0.f, sr = 0.f;
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const float wr = w[c] * rho[c];
                        s2 = __fmaf_rn(wr, rho[c], s2);
                        sr = __fmaf_rn(-wr, r[c], sr);
                    }
                    int t = 0;
#pragma unroll
                    for (int i = 0; i < NB; ++i) {
                        const float si = s2 * sh_g[i];
#pragma unroll
                        for (int k = 0; k <= i; ++k) { vals[t] = si * sh_g[k]; ++t; }
                        vals[NT + i] = sr * sh_g[i];
                    }
                }
                double* dst = A.red + (size_t)(blockIdx.x % (unsigned)A.red_copies) * kRedDoubles + (long long)f * VRED;
                if constexpr (VLIGHT == 16) {
                    warp_accumulate<16>(vals, lane, dst);
                } else {
                    warp_accumulate<32>(vals, lane, dst);
                    warp_accumulate<32>(vals + 32, lane, dst + 32);
                }
            }
        } else if constexpr (MODE == M_POSE || MODE == M_DIST) {
            // shared front part of poseJacobian / distJacobian: division-form projection and its own bounds test
            bool okj = false;
            float Aim[9];
#pragma unroll
            for (int i = 0; i < 9; ++i) Aim[i] = 0.f;
            if (ok) {
                const float mj = fx * pt[0] / pt[2] + cx;
                const float nj = fy * pt[1] / pt[2] + cy;
                okj = !(mj < 0.f || mj >= (float)A.cam.W || nj < 0.f || nj >= (float)A.cam.H);
                if (okj) {
                    const int xj = (int)floorf(nj), yj = (int)floorf(mj);
                    if (xj != fp.x || yj != fp.y) { float dummy[3]; sample_image(A, f, nj, mj, fp, dummy); }
                    float g0[3], g1[3];
                    image_gradient(A, f, nj, mj, fp, g0, g1);
                    const float z_inv = (float)(1.0 / (double)pt[2]);
                    const float z_inv_sq = z_inv * z_inv;
                    const float pi[6] = { fx * z_inv, 0.f, -fx * pt[0] * z_inv_sq, 0.f, fy * z_inv, -fy * pt[1] * z_inv_sq };
                    // G = grad I (3x2) * d pi (2x3); pi[1] = pi[3] = 0.  From here on the Jacobian algebra is contracted (FMA):
                    // it feeds normal equations that are accumulated in fp64 / compared to 2e-5, not the bit-exact residual path.
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        Aim[3 * c] = g0[c] * pi[0];
                        Aim[3 * c + 1] = g1[c] * pi[4];
                        Aim[3 * c + 2] = __fmaf_rn(g0[c], pi[2], g1[c] * pi[5]);
                    }
                }
            }
            if constexpr (MODE == M_POSE) {
                float vals[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) vals[i] = 0.f;
                if (okj) {
                    const float Sk[9] = { 0.f, -pt[2], pt[1], pt[2], 0.f, -pt[0], -pt[1], pt[0], 0.f };
                    float J[3][6];
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const float a0 = Aim[3 * c], a1 = Aim[3 * c + 1], a2 = Aim[3 * c + 2];
#pragma unroll
                        for (int j = 0; j < 3; ++j)   // -G R^T
                            J[c][j] = -__fmaf_rn(a0, P.R[3 * j], __fmaf_rn(a1, P.R[3 * j + 1], a2 * P.R[3 * j + 2]));
                        // G [p]x
                        J[c][3] = __fmaf_rn(a1, pt[2], -(a2 * pt[1]));
                        J[c][4] = __fmaf_rn(a2, pt[0], -(a0 * pt[2]));
                        J[c][5] = __fmaf_rn(a0, pt[1], -(a1 * pt[0]));
                    }
                    if constexpr (LED) {   // LedOptimizerJa.cpp:63-78
                        const float l3 = (float)cube_d(norm3(pt));
                        const float l5 = (float)pow5_d(norm3(pt));
                        float skp[3], nR[3];
#pragma unroll
                        for (int a = 0; a < 3; ++a) skp[a] = sum3(Sk[3 * a] * pt[0], Sk[3 * a + 1] * pt[1], Sk[3 * a + 2] * pt[2]);
#pragma unroll
                        for (int a = 0; a < 3; ++a) nR[a] = sum3(gh[0] * P.R[a], gh[1] * P.R[3 + a], gh[2] * P.R[6 + a]);
                        const float nRp = dot3(nR, pt);
#pragma unroll
                        for (int c = 0; c < 3; ++c) {
                            const float kt = -(rho[c] * led[c] / l3);
                            const float kr = -3 * rho[c] * led[c] / l5 * nRp;
#pragma unroll
                            for (int j = 0; j < 3; ++j) { J[c][j] = J[c][j] + kt * gh[j]; J[c][3 + j] = J[c][3 + j] + kr * skp[j]; }
                        }
                    }
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        int t = 0;
#pragma unroll
                        for (int i = 0; i < 6; ++i) {
                            const float jw = J[c][i] * w[c];
#pragma unroll
                            for (int k = 0; k <= i; ++k) { vals[t] = __fmaf_rn(jw, J[c][k], vals[t]); ++t; }
                            vals[21 + i] = __fmaf_rn(jw, r[c], vals[21 + i]);
                        }
                    }
                }
                warp_accumulate<32>(vals, lane, A.red + (size_t)(blockIdx.x % (unsigned)A.red_copies) * kRedDoubles + (long long)f * VRED);
            } else if (okj) {   // M_DIST
                // M = A * R^T, dI_k = M * dx_k
                float Mx[9];
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int j = 0; j < 3; ++j)
                        Mx[3 * c + j] = __fmaf_rn(Aim[3 * c], P.R[3 * j], __fmaf_rn(Aim[3 * c + 1], P.R[3 * j + 1], Aim[3 * c + 2] * P.R[3 * j + 2]));
                float Jd[4][3];
#pragma unroll
                for (int k = 0; k < 4; ++k)
#pragma unroll
                    for (int c = 0; c < 3; ++c)   // dI_k for now
                        Jd[k][c] = __fmaf_rn(Mx[3 * c], dx[k][0], __fmaf_rn(Mx[3 * c + 1], dx[k][1], Mx[3 * c + 2] * dx[k][2]));
                if constexpr (LED) {   // LedOptimizerJa.cpp:178-207
                    float gR[3], pRt[3];
#pragma unroll
                    for (int a = 0; a < 3; ++a) gR[a] = sum3(nfd[0] * P.R[a], nfd[1] * P.R[3 + a], nfd[2] * P.R[6 + a]);
                    const float gRp = dot3(gR, pt);
#pragma unroll
                    for (int a = 0; a < 3; ++a) pRt[a] = sum3(pt[0] * P.R[3 * a], pt[1] * P.R[3 * a + 1], pt[2] * P.R[3 * a + 2]);
                    const float radius = (float)cube_d(norm3(pt));
                    const double r5 = pow5_d(norm3(pt));
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        float dR_[3];
#pragma unroll
                        for (int a = 0; a < 3; ++a) dR_[a] = sum3(dn[k][0] * P.R[a], dn[k][1] * P.R[3 + a], dn[k][2] * P.R[6 + a]);
                        float dm = dot3(dR_, pt) + dot3(nfd, dx[k]);
                        const float dm2 = (float)((double)(-3 * dot3(pRt, dx[k])) / r5);
                        dm = dm / radius + dm2 * gRp;
#pragma unroll
                        for (int c = 0; c < 3; ++c) Jd[k][c] = Jd[k][c] + (rho[c] * led[c]) * dm;
                    }
                } else if constexpr (NB == 4) {   // PsOptimizerJa.cpp:225-240
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float ld = __fmaf_rn(P.light[1], dn[k][0], __fmaf_rn(P.light[2], dn[k][1], P.light[3] * dn[k][2]));
#pragma unroll
                        for (int c = 0; c < 3; ++c) Jd[k][c] = __fmaf_rn(-rho[c], ld, Jd[k][c]);
                    }
                } else {                // PsOptimizerJa.cpp:242-278
                    const float D[27] = { 0, 1, 0, 0, nfd[1], nfd[2], 0, 2 * nfd[0], 2 * nfd[0],
                                          0, 0, 1, 0, nfd[0], 0, nfd[2], -2 * nfd[1], 0,
                                          0, 0, 0, 1, 0, nfd[0], nfd[1], 0, -2 * nfd[2] };
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        float ld = 0.f;   // l^T (dSH/dn)^T dn_k, common to the three channels
#pragma unroll
                        for (int i = 1; i < 9; ++i) {   // column 0 of D is zero
                            const float dsh = __fmaf_rn(D[i], dn[k][0], __fmaf_rn(D[9 + i], dn[k][1], D[18 + i] * dn[k][2]));
                            ld = __fmaf_rn(P.light[i], dsh, ld);
                        }
#pragma unroll
                        for (int c = 0; c < 3; ++c) Jd[k][c] = __fmaf_rn(-rho[c], ld, Jd[k][c]);
                    }
                }
                // stencil normal equations, lower triangle in slot order: (0,0),(1,0),(1,1),(2,0)...(3,3), then b[4]
                // (entries of neighbours that are not surface voxels are dropped when the block is written out)
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    int t = 0;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float jw = Jd[i][c] * w[c];
#pragma unroll
                        for (int k = 0; k <= i; ++k) { vacc[t] = __fmaf_rn(jw, Jd[k][c], vacc[t]); ++t; }
                        vacc[10 + i] = __fmaf_rn(jw, r[c], vacc[10 + i]);
                    }
                }
            }
        }
    }

    // ---- epilogues
    // (a) per-voxel partial sums of the keyframe-splitting warps -> warp 0, in warp order
    if constexpr (NACC > 0) {
        if (W > 1) {
            if constexpr (MODE == M_INIT_ALBEDO) vacc[3] = (float)icount;
#pragma unroll
            for (int i = 0; i < NACC; ++i) s_dyn[(warp * NACC + i) * 32 + lane] = vacc[i];
            __syncthreads();
            if (warp == 0) {
#pragma unroll
                for (int i = 0; i < NACC; ++i) {
                    float t = s_dyn[i * 32 + lane];
                    for (int wv = 1; wv < W; ++wv) t += s_dyn[(wv * NACC + i) * 32 + lane];
                    vacc[i] = t;
                }
                if constexpr (MODE == M_INIT_ALBEDO) icount = (int)(vacc[3] + 0.5f);
            }
        }
    }
    if constexpr (MODE == M_INIT_ALBEDO) {
        if (warp == 0 && live && icount) {
            const float nr = vacc[0] / (float)icount, ng = vacc[1] / (float)icount, nbl = vacc[2] / (float)icount;
            A.vol.r[lin] = nr; A.vol.g[lin] = ng; A.vol.b[lin] = nbl;
            A.surf.rec1[slot] = make_float4(nr, ng, nbl, rec1w);
            rho[0] = nr; rho[1] = ng; rho[2] = nbl;
        }
    }
    if constexpr (MODE == M_ALBEDO) {
        // H.diagonal() += damping * H.diagonal(); Jacobi-preconditioned CG on a diagonal system = b * (1 / H);
        // Optimizer::updateAlbedo gate (OptimizerAux.cpp:120-147): accept a channel only if it stays strictly in (0, 1)
        if (warp == 0 && live && (A.flags & PSG_REG_ALBEDO)) {
            // reg albedo != 0: the system is no longer diagonal; hand J^T W J and J^T W r to the regularised solve (pcg.cu)
#pragma unroll
            for (int c = 0; c < 3; ++c) { A.out0[3 * (long long)s + c] = vacc[3 + c]; A.out1[3 * (long long)s + c] = vacc[c]; }
        } else if (warp == 0 && live) {
            float* chan[3] = { A.vol.r, A.vol.g, A.vol.b };
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                float Hd = vacc[c];
                if (A.damping != 0.f) Hd = Hd + A.damping * Hd;
                const float inv = (Hd != 0.f) ? 1.f / Hd : 1.f;
                const float delta = inv * vacc[3 + c];
                const float nv = rho[c] - delta;
                if ((double)nv > 0.0 && (double)nv < 1.0) { chan[c][lin] = nv; rho[c] = nv; ++accepted; }
                if (A.out0) { A.out0[3 * (long long)s + c] = vacc[3 + c]; A.out1[3 * (long long)s + c] = delta; }
            }
            if (accepted) A.surf.rec1[slot] = make_float4(rho[0], rho[1], rho[2], rec1w);
        }
    }
    if constexpr (MODE == M_DIST) {
        if (warp == 0 && live) {
            // columns of neighbours that are not surface voxels do not exist (PsOptimizerJa.cpp:536-552)
            {
                int t = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
#pragma unroll
                    for (int k = 0; k <= i; ++k) { if (!(cols[i] >= 0 && cols[k] >= 0)) vacc[t] = 0.f; ++t; }
                    if (cols[i] < 0) vacc[10 + i] = 0.f;
                }
            }
            // Eikonal rows (Optimizer.cpp:196-218, 477-537), scaled by reg_weight_n (PsOptimizer.cpp:136-142)
            if (A.flags & PSG_REG_NORMAL) {
                float nd[3], Jr[4];
#pragma unroll
                for (int a = 0; a < 3; ++a) nd[a] = -A.vs_inv * posf[a];
                Jr[0] = dot3(fd, nd);
#pragma unroll
                for (int a = 0; a < 3; ++a) Jr[a + 1] = fd[a] * (0.f + A.vs_inv * posf[a]);
                const float gn = norm3(fd);
                if (gn > 0.f) {
#pragma unroll
                    for (int a = 0; a < 4; ++a) Jr[a] /= gn;
                }
                const float rr = gn - 1;
                int t = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
#pragma unroll
                    for (int k = 0; k <= i; ++k) { if (cols[i] >= 0 && cols[k] >= 0) vacc[t] += A.reg_n * (Jr[i] * Jr[k]); ++t; }
                    if (cols[i] >= 0) vacc[10 + i] += A.reg_n * (Jr[i] * rr);
                }
            }
            // Laplacian: only the diagonal entry reaches the matrix (Optimizer.cpp:540-590; PsOptimizer.cpp:144-150)
            if (A.flags & PSG_REG_LAPLACIAN) {
                const long long stride[3] = { 1, A.vol.dims[0], (long long)A.vol.dims[0] * A.vol.dims[1] };
                float dd[3];
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const long long l1 = (long long)lin + stride[a], l0 = (long long)lin - stride[a];
                    const float d1 = (l1 < A.vol.N) ? A.vol.dist[l1] : dist;
                    const float d0 = (l0 >= 0) ? A.vol.dist[l0] : dist;
                    dd[a] = d1 + d0 - 2 * dist;
                }
                const float lap = (dd[0] + dd[1] + dd[2]) * A.vs_inv * A.vs_inv;
                const float jd = -6 * (A.vs_inv * A.vs_inv);
                vacc[0] += A.reg_l * (jd * jd);
                vacc[10] += A.reg_l * (jd * lap);
            }
#pragma unroll
            for (int i = 0; i < 14; ++i) A.out0[(long long)i * S + s] = vacc[i];
        }
    }
    // (b) scalar sums: warp -> block (warp order) -> grid (fixed order, last block): [0..5] mode specific, [6] loss, [7] pairs
    if constexpr (MODE != M_PROBE && MODE != M_INIT_ALBEDO) {
        double v8[8] = { acc_d[0], acc_d[1], acc_d[2], acc_d[3], acc_d[4], acc_d[5], eloss, (double)cnt };
        if constexpr (MODE == M_ALBEDO) v8[5] = (double)accepted;
        __syncthreads();   // s_dyn / s_red reuse
#pragma unroll
        for (int i = 0; i < 8; ++i) { const double v = warp_sum(v8[i]); if (lane == 0) s_red[warp][i] = v; }
        __syncthreads();
        if (tid < 8) {
            double v = 0.0;
            for (int i = 0; i < W; ++i) v += s_red[i][tid];
            s_tot[tid] = v;
        }
        __syncthreads();
        grid_sum_scalars(s_tot, A.partials, A.counter, A.scal, &s_last, s_scratch);
    }
}

// copy 0 += copies 1.. of the per-keyframe sums, in copy order
__global__ void __launch_bounds__(256) k_combine_red(double* __restrict__ red, int n, int copies) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double t = red[i];
    for (int c = 1; c < copies; ++c) t += red[(size_t)c * kRedDoubles + i];
    red[i] = t;
}

// ------------------------------------------------------------------------------------------------ launchers
static SweepArgs base_args(psg_optimizer* o, int slot = 0) {
    refresh_frame_table(o);
    SweepArgs A;
    memset(&A, 0, sizeof(A));
    A.vol = o->vol->view();
    A.surf = surf_view(o);
    A.cam.fx = o->K[0]; A.cam.fy = o->K[4]; A.cam.cx = o->K[2]; A.cam.cy = o->K[5]; A.cam.H = o->H; A.cam.W = o->W;
    A.loss.loss = o->set.loss; A.loss.lambda = o->set.lambda; A.loss.lambda_sq = o->set.lambda * o->set.lambda;
    A.images = o->images.p; A.F = o->F;
    shard_blocks(A.surf.nslots, 0, 1, &A.blk_first, &A.blk_stride, &A.blk_count);   // a sharded rank's slot list holds its own rows only
    A.vs_inv = o->vs_inv;
    A.cslot = o->cslot;
    A.reg_n = o->set.reg_weight_n; A.reg_l = o->set.reg_weight_l;
    o->red.alloc((size_t)kRedDoubles * kRedCopies + kMailDoubles);
    A.red = red_dev(o);
    A.red_copies = 1;
    A.scal = mail_dev(o) + 8 * slot;
    const size_t grid = (size_t)(o->S + 31) / 32 + 1;
    const size_t ngroups = grid / kRedGroup + 2;
    o->sweep_partials.alloc((grid + ngroups) * 8);
    if (o->sweep_counter.n < ngroups + 1) {
        o->sweep_counter.alloc(ngroups + 1);
        PSG_CUDA(cudaMemsetAsync(o->sweep_counter.p, 0, (ngroups + 1) * sizeof(unsigned), o->stream));
    }
    A.partials = o->sweep_partials.p;
    A.counter = o->sweep_counter.p;
    return A;
}

// ---- the keyframe table's copies in constant memory
namespace {
std::mutex g_slot_mutex;
unsigned g_slots_used[16] = {0};          // per device: bit s set = slot s of c_frames taken
FrameParams* g_const_base[16] = {nullptr};   // per device: address of c_frames
}
int acquire_const_slot(int device) {
    std::lock_guard<std::mutex> lk(g_slot_mutex);
    if (device < 0 || device >= 16) { set_error("device ordinal %d not supported", device); throw CudaError{-1}; }
    if (!g_const_base[device]) {
        void* p = nullptr;
        PSG_CUDA(cudaGetSymbolAddress(&p, c_frames));   // the current device's instance of the symbol
        g_const_base[device] = static_cast<FrameParams*>(p);
    }
    for (int s = 0; s < kConstSlots; ++s)
        if (!(g_slots_used[device] & (1u << s))) { g_slots_used[device] |= 1u << s; return s; }
    set_error("at most %d optimizers per device can exist at a time (their keyframe tables live in constant memory)", kConstSlots);
    throw CudaError{-1};
}
void release_const_slot(int device, int slot) {
    std::lock_guard<std::mutex> lk(g_slot_mutex);
    if (device >= 0 && device < 16 && slot >= 0) g_slots_used[device] &= ~(1u << slot);
}

// per-keyframe R, t, light -> the table the sweeps read.  The host's copy is uploaded when an API call changed it; the device
// table (which the small solves update in place; fetch_mail brings their result back into o->poses / o->light, so the host copy
// is current whenever an API call returns) is then copied into the optimizer's slot of constant memory, stream ordered.
void refresh_frame_table(psg_optimizer* o) {
    if (o->frames_dirty) {
        FrameParams* fr = o->frame_table.f;
        memset(fr, 0, sizeof(FrameTable));
        for (int f = 0; f < o->F; ++f) {
            const float* P = o->poses.data() + 16 * f;
            for (int a = 0; a < 3; ++a) { for (int b = 0; b < 3; ++b) fr[f].R[3 * a + b] = P[4 * a + b]; fr[f].t[a] = P[4 * a + 3]; }
            if (o->set.model != PSG_MODEL_LED) for (int i = 0; i < o->nb; ++i) fr[f].light[i] = o->light[(size_t)f * o->nb + i];
        }
        if (o->set.model == PSG_MODEL_LED) for (int c = 0; c < 3 && c < (int)o->light.size(); ++c) fr[kMaxFrames].light[c] = o->light[c];
        // pageable source: the driver stages the 6 KB before the call returns, the host copy may change right after
        PSG_CUDA(cudaMemcpyAsync(o->frames_dev.p, fr, sizeof(FrameTable), cudaMemcpyHostToDevice, o->stream));
        o->frames_dirty = false;
        o->const_dirty = true;
    }
    if (o->const_dirty) {
        PSG_CUDA(cudaMemcpyAsync(g_const_base[o->vol->device] + (size_t)o->cslot * (kMaxFrames + 1), o->frames_dev.p, sizeof(FrameTable),
                                 cudaMemcpyDeviceToDevice, o->stream));
        o->const_dirty = false;
    }
}

template <int MODE>
static void launch_sweep(psg_optimizer* o, SweepArgs& A, size_t red_doubles, size_t dyn_smem) {
    // many blocks adding into the same few hundred doubles serialise in L2: from a few thousand blocks on, spread them over
    // private copies and add the copies up afterwards (k_combine_red)
    if (red_doubles && A.blk_count >= 2048) A.red_copies = kRedCopies;
    if (red_doubles) PSG_CUDA(cudaMemsetAsync(A.red, 0, ((size_t)(A.red_copies - 1) * kRedDoubles + red_doubles) * sizeof(double), o->stream));
    if (o->sharded || A.blk_count <= 0)   // ranks without work still take part in the all-reduce
        PSG_CUDA(cudaMemsetAsync(A.scal, 0, 8 * sizeof(double), o->stream));
    if (o->S == 0 || A.blk_count <= 0) return;
    const int grid = A.blk_count;
    // keyframe split: double W while the launch stays within ~32 warps per SM (beyond that the replicated prologue costs more than it
    // hides).  The distance sweep (128 registers, 16 warps per SM) stops one doubling earlier than the 64-register modes: measured at
    // 128^3 (888 blocks), W = 8 against 4: albedo 34.0 -> 32.2 us, light 25.8 -> 24.3, pose 31.7 -> 28.9, but distance 33.8 -> 41.1.
    int W = 1;
    const long long budget = (MODE == M_DIST) ? 148ll * 32 : 2 * 148ll * 32;
    while (W < kMaxSplit && (long long)grid * W * 2 <= budget) W *= 2;
    if (o->sweep_split > 0) W = o->sweep_split;
    const int kBlockDim = 32 * W;
    if (W > 1 && dyn_smem < (size_t)W * 14 * 32 * sizeof(float)) dyn_smem = (size_t)W * 14 * 32 * sizeof(float);
#define PSG_SWEEP_CASE(MODEL)                                                                                          \
    case MODEL: {                                                                                                      \
        if (dyn_smem > 48 * 1024)                                                                                      \
            PSG_CUDA(cudaFuncSetAttribute(k_sweep<MODEL, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_smem)); \
        Prof pf(o->stream, KID_SWEEP_ENERGY + MODE);                                                                   \
        k_sweep<MODEL, MODE><<<grid, kBlockDim, dyn_smem, o->stream>>>(A);                                                \
    } break;
    switch (o->set.model) {
        PSG_SWEEP_CASE(PSG_MODEL_SH1)
        PSG_SWEEP_CASE(PSG_MODEL_SH2)
        PSG_SWEEP_CASE(PSG_MODEL_LED)
        default: set_error("unknown model %d", o->set.model); throw CudaError{-1};
    }
#undef PSG_SWEEP_CASE
    PSG_LAUNCH_CHECK();
    if (red_doubles && A.red_copies > 1) {
        k_combine_red<<<(unsigned)((red_doubles + 255) / 256), 256, 0, o->stream>>>(A.red, (int)red_doubles, A.red_copies);
        count_launch(1);
        PSG_LAUNCH_CHECK();
    }
}

// Results travel through a small "mailbox" behind the per-keyframe reduction buffer: 8 slots of 8 doubles.  A launch
// function queues kernels (and, across GPUs, the collectives on their results) without waiting; fetch_mail brings n doubles
// of per-keyframe sums plus the whole mailbox to pinned host memory in one round trip.  Several blocks of an outer
// iteration can therefore be in flight before the host waits once (psg_opt_iteration).
// device / pinned-host layout: mailbox first, per-keyframe sums behind it, so that one copy brings both
double* mail_dev(psg_optimizer* o) { return o->red.p; }
double* red_dev(psg_optimizer* o) { return o->red.p + kMailDoubles; }

// Across ranks: what the kernels queued so far left as per-rank partial sums becomes the job's sums (stream ordered; called
// before a kernel that consumes them on the device -- the small solves -- and by fetch_mail).
void reduce_pending(psg_optimizer* o) {
    if (o->world <= 1 || !(o->reduce_hi >= o->reduce_lo || o->reduce_reg)) return;
    if (o->sharded) {
        // one all-reduce for everything queued since the last one: per-keyframe sums, the scalar slots in flight and the
        // regulariser energies / accepted-update count of the distance block (partial sums over each rank's rows)
        double* ptr[3]; size_t cnt[3]; int k = 0;
        if (o->reduce_red) { ptr[k] = red_dev(o); cnt[k++] = o->reduce_red; }
        if (o->reduce_hi >= o->reduce_lo) { ptr[k] = mail_dev(o) + 8 * o->reduce_lo; cnt[k++] = (size_t)8 * (o->reduce_hi - o->reduce_lo + 1); }
        if (o->reduce_reg) { ptr[k] = mail_dev(o) + 8 * kSlotReg; cnt[k++] = 8; }
        comm_allreduce_doubles(o, ptr, cnt, k);
    } else {
        // replicated sweeps: every rank holds complete sums, equal up to the order of the fp64 atomics; rank 0's copy
        // becomes everybody's, so that the small solves that follow are identical on all ranks
        comm_broadcast_doubles(o, o->red.p, kMailDoubles + o->reduce_red);
    }
    o->reduce_lo = 1 << 30; o->reduce_hi = -1; o->reduce_red = 0; o->reduce_reg = false;
}

const double* fetch_mail(psg_optimizer* o, size_t n) {
    o->ensure_host_staging();
    double* h = o->host_staging;
    trace("[rank %d] fetch_mail: red %zu, slots %d..%d, reg %d, sharded %d", o->rank, o->reduce_red, o->reduce_lo, o->reduce_hi, (int)o->reduce_reg, (int)o->sharded);
    reduce_pending(o);
    PSG_CUDA(cudaMemcpyAsync(h, o->red.p, (kMailDoubles + n) * sizeof(double), cudaMemcpyDeviceToHost, o->stream));
    // a light / pose block solved on the device since the last round trip: its update lives in the keyframe table
    if (o->small_pending) PSG_CUDA(cudaMemcpyAsync(o->host_frames, o->frames_dev.p, sizeof(FrameTable), cudaMemcpyDeviceToHost, o->stream));
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    if (o->small_pending) adopt_device_frames(o);
    trace("[rank %d] fetch_mail done", o->rank);
    return h;
}
const double* mail_slot(const psg_optimizer* o, int slot) { return o->host_staging + 8 * slot; }
const double* red_host(const psg_optimizer* o) { return o->host_staging + kMailDoubles; }

// partial sums of this rank's share: summed over the ranks by the next fetch_mail (slots in flight are contiguous)
static void reduce_over_ranks(psg_optimizer* o, size_t n, int slot) {
    if (o->world <= 1) return;
    if (slot < o->reduce_lo) o->reduce_lo = slot;
    if (slot > o->reduce_hi) o->reduce_hi = slot;
    if (n > o->reduce_red) o->reduce_red = n;
}

// legacy one-block round trip: slot 0
static const double* fetch_red(psg_optimizer* o, size_t n, bool over_ranks = true) {
    if (over_ranks) reduce_over_ranks(o, n, 0);
    const double* h = fetch_mail(o, n);
    o->last_energy_sum = mail_slot(o, 0)[6];
    o->last_active = (long long)(mail_slot(o, 0)[7] + 0.5);
    return h;
}
static inline const double* scalars(const double* h) { return h; }   // slot 0

void sweep_energy(psg_optimizer* o, double* E_sum, long long* n_active) {
    SweepArgs A = base_args(o);
    launch_sweep<M_ENERGY>(o, A, 0, 0);
    fetch_red(o, 0);
    *E_sum = o->last_energy_sum;
    *n_active = o->last_active;
}

void sweep_energy_partial(psg_optimizer* o, int rank, int world, double* sum, long long* n_active) {
    SweepArgs A = base_args(o);
    shard_blocks(A.surf.nslots, rank, world, &A.blk_first, &A.blk_stride, &A.blk_count);
    PSG_CUDA(cudaMemsetAsync(A.scal, 0, 8 * sizeof(double), o->stream));   // an empty range launches nothing
    launch_sweep<M_ENERGY>(o, A, 0, 0);
    fetch_red(o, 0, /*reduce_over_ranks=*/false);
    *sum = o->last_energy_sum;
    *n_active = o->last_active;
}

void sweep_init_albedo(psg_optimizer* o) {
    SweepArgs A = base_args(o);
    launch_sweep<M_INIT_ALBEDO>(o, A, 0, 0);
    if (o->sharded) o->vol_unsynced = true;   // other ranks' rows of the albedo planes are theirs to write
}

void sweep_led_light_init(psg_optimizer* o, double sums[6]) {
    SweepArgs A = base_args(o);
    launch_sweep<M_LED_INIT>(o, A, 0, 0);
    const double* sc = scalars(fetch_red(o, 0));
    for (int i = 0; i < 6; ++i) sums[i] = sc[i];
}

void sweep_probe(psg_optimizer* o, float* m, float* n, unsigned char* ok, float* I3, float* rend3, float* r3, float* w3) {
    SweepArgs A = base_args(o);
    A.pm = m; A.pn = n; A.pok = ok; A.pI = I3; A.prend = rend3; A.pr = r3; A.pw = w3;
    launch_sweep<M_PROBE>(o, A, 0, 0);   // a sharded rank probes the pairs of its own rows
}

void launch_albedo(psg_optimizer* o, float damping, float* b_dev, float* delta_dev, int slot, bool regularised) {
    SweepArgs A = base_args(o, slot);
    A.damping = damping; A.out0 = b_dev; A.out1 = delta_dev;
    if (regularised) A.flags = PSG_REG_ALBEDO;   // out0 = J^T W r, out1 = diagonal of J^T W J, no update
    launch_sweep<M_ALBEDO>(o, A, 0, 0);
    if (o->sharded) o->vol_unsynced = true;
    reduce_over_ranks(o, 0, slot);
}

// light / pose: queues the sweep; returns how many doubles of per-keyframe sums it leaves in o->red (0 for the LED light)
size_t launch_small_block(psg_optimizer* o, int block, int slot) {
    SweepArgs A = base_args(o, slot);
    const bool led = o->set.model == PSG_MODEL_LED;
    size_t nred = 0;
    if (block == PSG_BLOCK_LIGHT && led) {
        launch_sweep<M_LIGHT>(o, A, 0, 0);
    } else {
        const int V = block == PSG_BLOCK_POSE ? 32 : (o->nb == 4 ? 16 : 64);
        nred = (size_t)o->F * V;
        if (block == PSG_BLOCK_POSE) launch_sweep<M_POSE>(o, A, nred, 0);
        else launch_sweep<M_LIGHT>(o, A, nred, 0);
    }
    reduce_over_ranks(o, nred, slot);
    return nred;
}

// distance: leaves the stencil blocks in o->vecs; the scalars (energy before the update) are fetched with the PCG result
void launch_dist_sweep(psg_optimizer* o, int flags, int slot) {
    SweepArgs A = base_args(o, slot);
    A.flags = flags;
    o->vecs.alloc((size_t)(o->S > 0 ? o->S : 1) * 25 + 8);
    A.out0 = o->vecs.p;   // planes 0..13: stencil normal equations
    launch_sweep<M_DIST>(o, A, 0, 0);   // sharded: the stencil blocks of the rows this rank owns, consumed where they are produced
    reduce_over_ranks(o, 0, slot);
}

void launch_energy(psg_optimizer* o, int slot) {
    SweepArgs A = base_args(o, slot);
    launch_sweep<M_ENERGY>(o, A, 0, 0);
    reduce_over_ranks(o, 0, slot);
}

// ---- regulariser energies (Optimizer.cpp:86-136) -------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_reg_energy(VolView vol, SurfView surf, float vs_inv, double* red, int row0, int row1) {
    __shared__ double s_red[kWarps][3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int S = surf.S;
    const int s = row0 + blockIdx.x * kBlock + tid;
    double en = 0.0, el = 0.0, er = 0.0;
    if (s < row1) {
        const int lin = surf.lin[s];
        const float dist = vol.dist[lin];
        const float rho[3] = { vol.r[lin], vol.g[lin], vol.b[lin] };
        const unsigned char pb = surf.pos[s];
        const long long stride[3] = { 1, vol.dims[0], (long long)vol.dims[0] * vol.dims[1] };
        float fd[3], dd[3], G[9];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const int nl = surf.nbr_lin[a * S + s];
            const float p = ((pb >> a) & 1) ? 1.f : -1.f;
            const float dn = nl >= 0 ? vol.dist[nl] : dist;
            fd[a] = (p * (dn - dist)) * vs_inv;
            const long long l1 = (long long)lin + stride[a], l0 = (long long)lin - stride[a];
            const float d1 = (l1 < vol.N) ? vol.dist[l1] : dist;
            const float d0 = (l0 >= 0) ? vol.dist[l0] : dist;
            dd[a] = d1 + d0 - 2 * dist;
            // Optimizer::computeAlbedoGrad (Optimizer.cpp:396-460)
            const float nr = nl >= 0 ? vol.r[nl] : rho[0], ng = nl >= 0 ? vol.g[nl] : rho[1], nbb = nl >= 0 ? vol.b[nl] : rho[2];
            G[a] = (p * (nr - rho[0])) * vs_inv; G[3 + a] = (p * (ng - rho[1])) * vs_inv; G[6 + a] = (p * (nbb - rho[2])) * vs_inv;
        }
        const float e = norm3(fd) - 1;
        en = (double)(e * e);
        const float lap = (dd[0] + dd[1] + dd[2]) * vs_inv * vs_inv;
        el = (double)(lap * lap);
        er = (double)((norm3(G) + norm3(G + 3)) + norm3(G + 6));
    }
    en = warp_sum(en); el = warp_sum(el); er = warp_sum(er);
    if (lane == 0) { s_red[warp][0] = en; s_red[warp][1] = el; s_red[warp][2] = er; }
    __syncthreads();
    if (tid < 3) {
        double v = 0.0;
#pragma unroll
        for (int i = 0; i < kWarps; ++i) v += s_red[i][tid];
        atomicAdd(&red[tid], v);
    }
}

void launch_reg_energies(psg_optimizer* o, int slot) {
    SweepArgs A = base_args(o, slot);   // replicated: every rank evaluates all rows; sharded: its own rows, summed with the next flush
    PSG_CUDA(cudaMemsetAsync(A.scal, 0, 8 * sizeof(double), o->stream));
    const int r0 = o->sharded ? o->row_lo : 0, r1 = o->sharded ? o->row_hi : o->S;
    if (r1 > r0) {
        { Prof pf(o->stream, KID_REG_ENERGY); k_reg_energy<<<(r1 - r0 + kBlock - 1) / kBlock, kBlock, 0, o->stream>>>(A.vol, A.surf, o->vs_inv, A.scal, r0, r1); }
        PSG_LAUNCH_CHECK();
    }
    if (o->sharded) o->reduce_reg = true;
}

void reg_energies(psg_optimizer* o, double* En, double* El, double* Er) {
    launch_reg_energies(o, kSlotReg);
    if (o->world > 1 && !o->sharded) reduce_over_ranks(o, 0, kSlotReg);   // replicated: rank 0's sums become everybody's
    fetch_mail(o, 0);
    const double* r = mail_slot(o, kSlotReg);
    *En = r[0]; *El = r[1]; *Er = r[2];
}

// ---- diagnostic: interpolateImage + computeImageGradient at caller-chosen positions (the sweeps' own device functions)
__global__ void __launch_bounds__(128) k_image_probe(SweepArgs A, int f, int n, const float* __restrict__ rows, const float* __restrict__ cols,
                                                    float* __restrict__ rgb3, float* __restrict__ g0o, float* __restrict__ g1o) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Footprint fp;
    float rgb[3], g0[3], g1[3];
    sample_image(A, f, rows[i], cols[i], fp, rgb);
    image_gradient(A, f, rows[i], cols[i], fp, g0, g1);
#pragma unroll
    for (int c = 0; c < 3; ++c) { rgb3[3 * i + c] = rgb[c]; g0o[3 * i + c] = g0[c]; g1o[3 * i + c] = g1[c]; }
}

void image_probe(psg_optimizer* o, int f, int n, const float* rows_dev, const float* cols_dev, float* rgb3, float* g0, float* g1) {
    SweepArgs A = base_args(o);
    k_image_probe<<<(n + 127) / 128, 128, 0, o->stream>>>(A, f, n, rows_dev, cols_dev, rgb3, g0, g1);
    count_launch();
    PSG_LAUNCH_CHECK();
}

// ------------------------------------------------------------------------------------------------------------------
// Energy evaluation that performs no device allocation.
//
// Some hosts hand the whole GPU memory to one owner at start-up (a framework's caching allocator, a fixed partition
// per tenant, a real-time process that pre-allocates everything before it goes live).  For them a cudaMalloc issued
// from inside the optimisation loop is not acceptable: it can fail although "enough" memory exists, and it stalls the
// device while the driver services it.  psg_opt_energy() may allocate (reduction buffers and the per-keyframe table
// grow on demand); this variant runs the same kernel with all of its scratch in one array that is part of the loaded
// module, so the only device memory it touches beyond the caller's volume / keyframes is that array.
//
// Layout of the array: 8 doubles of results | arrival counter | per-block
// partial sums for up to kNoAllocMaxSurface / 32 blocks.  About 4.2 MB, held by every GPU that loads the library until
// the process exits.
constexpr int kNoAllocMaxSurface = 1 << 21;   // 2 097 152 surface voxels
constexpr int kNoAllocMaxDevices = 64;
struct NoAllocScratch {
    double red[8];
    unsigned counter[kNoAllocMaxSurface / 32 / kRedGroup + 8];   // [0] top level, [1 + g] per group of blocks
    double partials[((size_t)kNoAllocMaxSurface / 32 + kNoAllocMaxSurface / 32 / kRedGroup + 8) * 8];   // blocks, then groups
};
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];
// device address of the array per device ordinal (the symbol name itself is only a host-side handle)
static NoAllocScratch* g_noalloc_dev[kNoAllocMaxDevices] = { nullptr };

}  // namespace psg

extern "C" int psg_opt_energy_noalloc(psg_optimizer* o, double* E, long long* n_active) {
    using namespace psg;
    if (!o) { set_error("psg_opt_energy_noalloc: null optimizer"); return PSG_ERR_INVALID; }
    if (!o->inited) { set_error("psg_opt_init must be called first"); return PSG_ERR_STATE; }
    if (o->S > kNoAllocMaxSurface) {
        set_error("psg_opt_energy_noalloc: %d surface voxels exceed the built-in scratch (at most %d); use psg_opt_energy",
                  o->S, kNoAllocMaxSurface);
        return PSG_ERR_INVALID;
    }
    const int dev = o->vol->device;
    if (dev < 0 || dev >= kNoAllocMaxDevices) { set_error("psg_opt_energy_noalloc: device ordinal %d not supported", dev); return PSG_ERR_INVALID; }
    try {
        PSG_CUDA(cudaSetDevice(dev));
        if (!g_noalloc_dev[dev]) {
            // one-time set-up for this device, with the device current: resolve the array's device address
            void* ptr = nullptr;
            PSG_CUDA(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
            g_noalloc_dev[dev] = static_cast<NoAllocScratch*>(ptr);
        }
        NoAllocScratch* sc = g_noalloc_dev[dev];
        if (o->S == 0) { const double zero = 0.0; if (E) *E = zero / zero; if (n_active) *n_active = 0; return PSG_OK; }   // E / S with S = 0, as psg_opt_energy

        refresh_frame_table(o);   // uploads the keyframe table if the host changed it (a copy into memory allocated at creation)
        // the array keeps whatever the previous call left: results and the arrival counter must start at zero
        PSG_CUDA(cudaMemsetAsync(sc, 0, sizeof(sc->red) + sizeof(sc->counter), o->stream));

        SweepArgs A;
        memset(&A, 0, sizeof(A));
        A.vol = o->vol->view();
        A.surf = surf_view(o);   // re-gathers the compact records if the volume changed; launches a kernel, allocates nothing
        A.cam.fx = o->K[0]; A.cam.fy = o->K[4]; A.cam.cx = o->K[2]; A.cam.cy = o->K[5]; A.cam.H = o->H; A.cam.W = o->W;
        A.loss.loss = o->set.loss; A.loss.lambda = o->set.lambda; A.loss.lambda_sq = o->set.lambda * o->set.lambda;
        A.images = o->images.p; A.F = o->F;
        shard_blocks(A.surf.nslots, 0, 1, &A.blk_first, &A.blk_stride, &A.blk_count);   // this rank's whole slot list (no communication on this path)
        A.vs_inv = o->vs_inv;
        A.cslot = o->cslot;
        A.reg_n = o->set.reg_weight_n; A.reg_l = o->set.reg_weight_l;
        A.red = sc->red; A.red_copies = 1; A.scal = sc->red; A.partials = sc->partials; A.counter = sc->counter;

        const int grid = A.blk_count;
        int W = 1;
        while (W < kMaxSplit && (long long)grid * W * 2 <= 148ll * 32) W *= 2;
        if (o->sweep_split > 0) W = o->sweep_split;
        const size_t smem = W > 1 ? (size_t)W * 14 * 32 * sizeof(float) : 0;
        if (grid > 0) switch (o->set.model) {
            case PSG_MODEL_SH1: k_sweep<PSG_MODEL_SH1, M_ENERGY><<<grid, 32 * W, smem, o->stream>>>(A); break;
            case PSG_MODEL_SH2: k_sweep<PSG_MODEL_SH2, M_ENERGY><<<grid, 32 * W, smem, o->stream>>>(A); break;
            default: k_sweep<PSG_MODEL_LED, M_ENERGY><<<grid, 32 * W, smem, o->stream>>>(A); break;
        }
        count_launch();
        PSG_LAUNCH_CHECK();
        double r[8];
        PSG_CUDA(cudaMemcpyAsync(r, sc->red, sizeof(r), cudaMemcpyDeviceToHost, o->stream));
        PSG_CUDA(cudaStreamSynchronize(o->stream));
        if (E) *E = r[6] / (double)o->S;
        if (n_active) *n_active = (long long)(r[7] + 0.5);
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

