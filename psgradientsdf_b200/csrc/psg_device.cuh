// Device-side building blocks shared by the sm_100a kernels.
//
// Arithmetic contract (DESIGN.md "Numerics"): this translation unit set is compiled with -fmad=false and
// IEEE division / square root, and every expression that feeds a voxel / pixel INDEX or a visibility
// decision is written in the operation order the reference executes on x86-64 (no FMA contraction):
// Eigen's fully unrolled 3-term reductions associate as a0 + (a1 + a2)
// (third/eigen/Eigen/src/Core/Redux.h:100-116); its dynamic 4- and 9-term inner products use the SSE packet
// order (Redux.h:236-281).  That is what makes visibility masks, surface lists and pixel indices bit-exact.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace psg {

constexpr int kMaxFrames = 64;   // keyframes per optimizer (one 64-bit visibility word per surface voxel)

// ---- plain views passed to kernels by value -------------------------------------------------------------
struct VolView {
    int dims[3];
    long long N;
    float vs;          // voxel size (VoxelGrid.h:32)
    float origin[3];   // position of voxel (0,0,0) (VoxelGrid.h:34)
    float T, invT, zmin, zmax;
    float* dist; float* gx; float* gy; float* gz; float* w; float* r; float* g; float* b;   // SoA, N each
    unsigned long long* vis;   // frame-indexed visibility bits, word-major: vis[word * N + lin]
    int nwords;
};

struct FramePose { float R[9]; float t[3]; };   // camera-to-world rotation (row-major) and translation

// Per-keyframe parameters, one entry per keyframe, read uniformly by a whole warp.
struct __align__(16) FrameParams {
    float R[9];
    float t[3];
    float light[9];   // SH coefficients of this keyframe (4 or 9 used); unused for LED
    float pad[3];
};
// All keyframes' parameters: one table in device memory (psg_optimizer::frames_dev) that the small solves (small_solve.cu) update
// in place, so that an outer iteration runs without a host round trip per block; the sweeps read a copy in constant memory
// (sweep.cu: c_frames).  Entry kMaxFrames holds the LED model's global light colour in light[0..2].
struct FrameTable { FrameParams f[kMaxFrames + 1]; };
constexpr int kConstSlots = 10;  // copies of the table in constant memory: optimizers alive at a time per device (10 x 6240 B of the 64 KB)

struct SurfView {
    int S;                              // number of surface voxels (rows)
    int nslots;                         // slots of this rank's sweep work list: S, or the rows it owns when the list is sharded
    const int* lin;                     // S   linear voxel index, ascending (Optimizer::surface_points_)
    const unsigned long long* vis;      // S   keyframe-indexed visibility (Optimizer::select_vis)
    const int* nbr_lin;                 // 3*S linear index of the finite-difference neighbour per axis (or -1)
    const int* nbr_row;                 // 3*S surface row of that neighbour, -1 if it is not a surface voxel
    const unsigned char* pos;           // S   bit a set = forward difference along axis a
    // Slot-ordered work list of the sweeps.  Slot t holds surface row perm[t]; slots are sorted by keyframe visibility so
    // that the 32 voxels of a warp see (nearly) the same keyframes (in lin order a warp straddles both sides of the
    // object and half of its lanes idle).  Each field is one coalesced 128-bit load per lane:
    const unsigned long long* pvis;     // S   keyframe visibility of the slot
    float4* rec0;                       // S   {dist, g_hat.xyz}            g_hat = normalised stored gradient
    float4* rec1;                       // S   {r, g, b, bits(pos)}
    float4* rec2;                       // S   {fd.xyz, bits(row)}          fd = finite-difference gradient, row = perm[t]
    float4* rec3;                       // S   {voxel centre in world xyz, bits(lin)}
    float4* rec4;                       // S   {normalised fd xyz, -}
};

struct Camera { float fx, fy, cx, cy; int H, W; };

struct LossParams { int loss; float lambda, lambda_sq; };

// ---- 3-vector helpers in the reference's association order -----------------------------------------------
__device__ __forceinline__ float sum3(float a0, float a1, float a2) { return a0 + (a1 + a2); }
__device__ __forceinline__ float dot3(const float* a, const float* b) { return sum3(a[0] * b[0], a[1] * b[1], a[2] * b[2]); }
__device__ __forceinline__ float sqnorm3(const float* a) { return sum3(a[0] * a[0], a[1] * a[1], a[2] * a[2]); }
__device__ __forceinline__ float norm3(const float* a) { return sqrtf(sqnorm3(a)); }
// Eigen normalized(): returned unchanged when the squared norm is 0 (Core/Dot.h:121-131)
__device__ __forceinline__ void normalized3(const float* a, float* o) {
    const float z = sqnorm3(a);
    if (z > 0.f) { const float s = sqrtf(z); o[0] = a[0] / s; o[1] = a[1] / s; o[2] = a[2] / s; }
    else { o[0] = a[0]; o[1] = a[1]; o[2] = a[2]; }
}
// o = M^T v / o = M v, M row-major
__device__ __forceinline__ void matT_vec(const float* M, const float* v, float* o) {
#pragma unroll
    for (int a = 0; a < 3; ++a) o[a] = sum3(M[a] * v[0], M[3 + a] * v[1], M[6 + a] * v[2]);
}
__device__ __forceinline__ void mat_vec(const float* M, const float* v, float* o) {
#pragma unroll
    for (int a = 0; a < 3; ++a) o[a] = sum3(M[3 * a] * v[0], M[3 * a + 1] * v[1], M[3 * a + 2] * v[2]);
}
// dynamic-size Eigen inner product of length NB (4 or 9): SSE packet order
template <int NB>
__device__ __forceinline__ float dot_sh(const float* a, const float* b) {
    if (NB == 4) return (a[0] * b[0] + a[2] * b[2]) + (a[1] * b[1] + a[3] * b[3]);
    const float p0 = a[0] * b[0] + a[4] * b[4], p1 = a[1] * b[1] + a[5] * b[5];
    const float p2 = a[2] * b[2] + a[6] * b[6], p3 = a[3] * b[3] + a[7] * b[7];
    return ((p0 + p2) + (p1 + p3)) + a[8] * b[8];
}
// (float) std::pow(x, 3 | 5): the float argument is promoted to double (C++11 <cmath>)
__device__ __forceinline__ double cube_d(float x) { const double d = (double)x; return d * d * d; }
__device__ __forceinline__ double pow5_d(float x) { const double d = (double)x; const double d2 = d * d; return d2 * d2 * d; }

// SH basis in the reference's order (PsOptimizerJa.cpp:17-28)
template <int NB>
__device__ __forceinline__ void sh_eval(const float* n, float* sh) {
    sh[0] = 1.f; sh[1] = n[0]; sh[2] = n[1]; sh[3] = n[2];
    if (NB == 9) {
        sh[4] = n[0] * n[1]; sh[5] = n[0] * n[2]; sh[6] = n[1] * n[2];
        sh[7] = n[0] * n[0] - n[1] * n[1]; sh[8] = n[0] * n[0] - n[2] * n[2];
    }
}

// VoxelGrid::line2idx (VoxelGrid.h:88-97)
__device__ __forceinline__ void line2idx(int lin, const int* dims, int* idx) {
    const int WH = dims[0] * dims[1];
    int rest = lin;
    idx[2] = rest / WH; rest -= idx[2] * WH;
    idx[1] = rest / dims[0]; rest -= idx[1] * dims[0];
    idx[0] = rest;
}

// ---- robust loss (Optimizer.cpp:140-186), per channel ----------------------------------------------------
__device__ __forceinline__ float irls_weight(const LossParams& L, float r) {
    const float q = r / L.lambda;
    switch (L.loss) {
        case 1: return 1.f / (1.f + q * q);                                          // CAUCHY
        case 3: { const float u = 1.f - q * q; return (r * r < L.lambda_sq) ? u * u : 0.f; }   // TUKEY
        case 2: return (r * r < L.lambda_sq) ? 1.f : L.lambda * fabsf(1.f / r);      // HUBER
        case 4: return (r * r < L.lambda_sq) ? 1.f : 0.f;                            // TRUNC_L2
        default: return 1.f;                                                          // L2
    }
}
__device__ __forceinline__ float robust_loss1(const LossParams& L, float r) {
    const float q = r / L.lambda;
    switch (L.loss) {
        case 1: return logf(1.f + q * q);
        case 3: { const float u = 1.f - q * q; return (r * r < L.lambda_sq) ? 1.f - u * u * u : 1.f; }
        case 2: return (r * r < L.lambda_sq) ? .5f * (r * r) : L.lambda * (fabsf(r) - 0.5f * L.lambda);
        case 4: { const float c = fminf(fmaxf(r, -L.lambda), L.lambda); return c * c; }
        default: return r * r;
    }
}
__device__ __forceinline__ float robust_loss3(const LossParams& L, const float* r) {
    return (robust_loss1(L, r[0]) + robust_loss1(L, r[1])) + robust_loss1(L, r[2]);
}

// ---- warp / block reductions -----------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Butterfly reduce-scatter of V (power of two, <= 32) per-lane values across the 32 lanes of a warp:
// afterwards lane l holds the warp-wide sum of vals[l % V] in vals[0] (V + log-many shuffles instead of 5 V).
template <int V>
__device__ __forceinline__ float warp_reduce_scatter(float* vals, int lane) {
    static_assert(V == 1 || V == 2 || V == 4 || V == 8 || V == 16 || V == 32, "V must be a power of two");
    // halve the number of live values per lane while exchanging the other half with the partner lane
#pragma unroll
    for (int half = V / 2; half >= 1; half >>= 1) {
        const bool upper = (lane & half) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            const float send = upper ? vals[i] : vals[i + half];
            const float keep = upper ? vals[i + half] : vals[i];
            vals[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
        }
    }
    float v = vals[0];
    // lanes that differ only in bits >= log2(V) hold partial sums of the same index
#pragma unroll
    for (int o = 16; o >= V; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace psg
