// The light and pose blocks' small solves on the device: one thread block runs Eigen's preconditioned conjugate gradient on the
// F dense bs x bs blocks the sweep accumulated, operation by operation in the order the reference's x86-64 build executes it
// (host_solve.cpp restates the same order for the host; psg_debug_small_cg exposes that one as the checker of this kernel), and
// applies the update to the keyframe table in device memory.  Nothing waits for the host: an outer iteration queues all four
// blocks and the closing energy and reads everything back in one round trip.
//
//   reference: PsOptimizer.cpp:175-203 (light), 206-234 (pose), LedOptimizer.cpp:143-163, 246-275; OptimizerAux.cpp:190-205
//   Eigen:     IterativeLinearSolvers/ConjugateGradient.h:29-98, SparseCore/SparseSelfAdjointView.h:299-331, Core/Redux.h:236-281
#include <cstring>

#include "psg_device.cuh"
#include "psg_host.h"

namespace psg {

namespace {

constexpr int kSolveMaxUnknowns = 576;   // 64 keyframes x 9 SH coefficients at most
constexpr int kSolveMaxThreads = 704;    // 22 warps of three 9 x 9 blocks
// One product array in shared memory: [0, n & ~7) the products of the strided range in place; kSumPad x -0.0f behind them (the
// identity of a float add: loads come in chunks of 8 x 8 and need no remainder handling); then eight floats for the products
// past the last full stride of eight -- [0..3] one more packet of four, [4..6] the tail -- which hold -0.0f where the length has
// no such element, so that they can be added unconditionally.
constexpr int kSumPad = 64;
constexpr int kSumExtra = kSolveMaxUnknowns + kSumPad;
constexpr int kSumFloats = kSumExtra + 8;

__device__ __forceinline__ int sum_slot(int e, int n) {
    const int aligned = n & ~3, aligned2 = n & ~7;
    return e < aligned2 ? e : (e < aligned ? kSumExtra + (e - aligned2) : kSumExtra + 4 + (e - aligned));
}
__device__ __forceinline__ float lds(unsigned a) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ void sts(unsigned a, float v) { asm volatile("st.shared.f32 [%0], %1;" :: "r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ unsigned keep(unsigned v) { unsigned r; asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v)); return r; }   // held in a register, not recomputed

// What a lane needs to add up a product array, computed once (byte addresses in the shared window, relative to array 0).
struct SumPlan {
    unsigned chain;     // first element of accumulation chain c = lane & 7
    unsigned extra;     // this lane's element of the extra packet
    unsigned tail;      // first tail element
    int chunks;         // chunks of eight strides after the first element of the chain (>= 1: a chunk of pad is harmless)
    int n;              // < 8: no strided range at all
};

// sum_i a[i] * b[i] (products formed by the callers, stored with sum_slot) as Eigen's vectorised reduction associates it: SSE2
// packets of four floats, two packet accumulators p0 / p1 over strides of eight, their sum, one more packet if the length
// allows, the horizontal sum (l0 + l2) + (l1 + l3), then the tail element by element (Core/Redux.h:236-281; eigen_dot in
// host_solve.cpp is the same order on the host).  Called by all lanes of a warp: lane L carries accumulation chain L & 7 (p0's
// four lanes, then p1's; four groups of eight lanes do the same work), the merges are butterfly shuffles whose operands
// commute, so every lane returns the sum.  A chunk of eight loads per chain is issued while the previous chunk is added up.
__device__ __forceinline__ float eigen_sum(const SumPlan& P, unsigned array_off) {
    const unsigned full = 0xffffffffu;
    float a;
    if (P.n >= 8) {
        unsigned q = P.chain + array_off;
        a = lds(q);
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = lds(q + 32 * (k + 1));
        for (int ch = 1; ; ++ch) {
            const bool more = ch < P.chunks;
            float w[8];
            if (more) {
                q += 256;
#pragma unroll
                for (int k = 0; k < 8; ++k) w[k] = lds(q + 32 * (k + 1));
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) a += v[k];
            if (!more) break;
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = w[k];
        }
        a = a + __shfl_xor_sync(full, a, 4);                 // p0 + p1
        a += lds(P.extra + array_off);                       // one more packet (or -0)
    } else if (P.n >= 4) {
        a = lds(P.extra + array_off);                        // a single packet
    } else {                                                 // Eigen's scalar path
        a = lds(P.tail + array_off);
        for (int i = 1; i < P.n; ++i) a += lds(P.tail + array_off + 4 * i);
        return a;
    }
    const float t = a + __shfl_xor_sync(full, a, 2);         // l0 + l2 | l1 + l3
    float r = t + __shfl_xor_sync(full, t, 1);
    r += lds(P.tail + array_off);
    r += lds(P.tail + array_off + 4);
    r += lds(P.tail + array_off + 8);
    return r;
}

struct SmallArgs {
    const double* red;      // per-keyframe sums of the sweep: V doubles per keyframe, lower triangle row by row, then the rhs
    const double* scal;     // LED light: the sweep's scalar slot (diagonal [0..2], rhs [3..5])
    int nblk, bs, V;
    int led_light;          // one 3 x 3 diagonal block read from `scal`
    int kind;               // PSG_BLOCK_LIGHT / PSG_BLOCK_POSE
    int damp;               // H.diagonal() += damping * H.diagonal()
    int pose_needs_success; // LED: updatePose only on Success
    float damping;
    FrameParams* frames;    // updated in place
    float* sys;             // H | b | x for psg_opt_last_system
    float* verdict;         // {success, iterations, error, applied}
};

// Sophus SO3::exp followed by Eigen's Quaternion::toRotationMatrix, as capi.cu's host version (so3.hpp:583-619).  sin / cos are
// evaluated in double and rounded: the correctly rounded float, which is what the host's libm returns for these arguments.
__device__ void so3_exp_matrix_dev(const float* w, float* Rm) {
    const float theta_sq = w[0] * w[0] + (w[1] * w[1] + w[2] * w[2]);
    float imag, real;
    if (theta_sq < 1e-10f * 1e-10f) {
        const float t4 = theta_sq * theta_sq;
        imag = 0.5f - (float)(1.0 / 48.0) * theta_sq + (float)(1.0 / 3840.0) * t4;
        real = 1.f - (float)(1.0 / 8.0) * theta_sq + (float)(1.0 / 384.0) * t4;
    } else {
        const float theta = sqrtf(theta_sq);
        const float half = 0.5f * theta;
        imag = (float)sin((double)half) / theta;
        real = (float)cos((double)half);
    }
    const float qw = real, qx = imag * w[0], qy = imag * w[1], qz = imag * w[2];
    const float tx = 2.f * qx, ty = 2.f * qy, tz = 2.f * qz;
    const float twx = tx * qw, twy = ty * qw, twz = tz * qw, txx = tx * qx, txy = ty * qx, txz = tz * qx;
    const float tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
    Rm[0] = 1.f - (tyy + tzz); Rm[1] = txy - twz; Rm[2] = txz + twy;
    Rm[3] = txy + twz; Rm[4] = 1.f - (txx + tzz); Rm[5] = tyz - twx;
    Rm[6] = txz - twy; Rm[7] = tyz + twx; Rm[8] = 1.f - (txx + tyy);
}

// One thread per unknown e = (block f, row i), a block never straddling two warps (a warp takes 32 / BS blocks): the thread keeps
// row i of its symmetric block, its entries of x, res, p, z and the preconditioner in registers; the search direction of a block
// travels between its rows by warp shuffles; shared memory only carries the element products of the three inner products, which
// EVERY warp then adds up in Eigen's order (no broadcast).  Two block barriers per CG iteration.
template <int BS>
__global__ void __launch_bounds__(kSolveMaxThreads) k_small_solve(const SmallArgs A) {
    __shared__ float s_sum[3 * kSumFloats];           // three product arrays
    constexpr int bb = BS * BS;
    constexpr int BPW = 32 / BS;                      // blocks per warp
    constexpr unsigned kArr = kSumFloats * 4;         // bytes from one product array to the next
    const int nblk = A.nblk, n = nblk * BS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int f = warp * BPW + lane / BS, i = lane % BS;
    const int base = lane - i;                        // first lane of this block's rows
    const bool live = lane < BPW * BS && f < nblk;
    const int e = live ? f * BS + i : 0;
    const unsigned s0 = (unsigned)__cvta_generic_to_shared(s_sum);
    const unsigned slot = keep(s0 + 4u * (live ? sum_slot(e, n) : kSumExtra + 7));   // where this unknown's products go in array 0 (lanes without one: an unused float)
    SumPlan P;
    P.chain = keep(s0 + 4u * (lane & 7));
    P.extra = keep(s0 + 4u * (kSumExtra + (lane & 3)));
    P.tail = keep(s0 + 4u * (kSumExtra + 4));
    P.chunks = (int)keep((unsigned)max(1, ((n & ~7) / 8 - 1 + 7) / 8));
    P.n = n;
    const unsigned off_d = keep((lane & 8) ? 2 * kArr : kArr);   // second phase: half of the lanes add up array 1, half array 2
    for (int q = threadIdx.x; q < kSumPad + 8; q += blockDim.x)
#pragma unroll
        for (int arr = 0; arr < 3; ++arr) s_sum[arr * kSumFloats + (n & ~7) + (q < kSumPad ? q : kSolveMaxUnknowns - (n & ~7) + q)] = -0.f;

    // ---- the system as the host used to unpack it: doubles -> floats, then the damping on the diagonal
    float hrow[BS];   // row i of the block (the lower triangle mirrored)
    float be = 0.f;
    if (A.led_light) {
#pragma unroll
        for (int j = 0; j < BS; ++j) hrow[j] = 0.f;
        if (live) {
            const float d = (float)A.scal[i];
#pragma unroll
            for (int j = 0; j < BS; ++j) if (j == i) hrow[j] = d;
            be = (float)A.scal[3 + i];
        }
    } else {
        constexpr int NT = BS * (BS + 1) / 2;
        const double* v = A.red + (size_t)(live ? f : 0) * A.V;
#pragma unroll
        for (int j = 0; j < BS; ++j) {
            const int hi = i > j ? i : j, lo = i > j ? j : i;
            hrow[j] = live ? (float)v[hi * (hi + 1) / 2 + lo] : 0.f;
        }
        if (live) be = (float)v[NT + i];
    }
    float diag = 0.f;
#pragma unroll
    for (int j = 0; j < BS; ++j) if (j == i) diag = hrow[j];
    if (A.damp && A.damping != 0.f) {
        diag = diag + A.damping * diag;
#pragma unroll
        for (int j = 0; j < BS; ++j) if (j == i) hrow[j] = diag;
    }
    if (live) {
#pragma unroll
        for (int j = 0; j < BS; ++j) A.sys[(size_t)f * bb + i * BS + j] = hrow[j];
        A.sys[(size_t)nblk * bb + e] = be;
    }
    const float invd = diag != 0.f ? 1.f / diag : 1.f;       // DiagonalPreconditioner::factorize
    float x = 0.f, res = be, pe = 0.f, ze = 0.f;

    // ---- ConjugateGradient.h:29-98
    const float tol = 1.1920929e-07f;
    const int maxIters = 2 * n;
    int it = 0, success = 1;
    float err = 0.f;
    __syncthreads();                                  // pads before products
    sts(slot, be * be);
    __syncthreads();
    const float rhsNorm2 = eigen_sum(P, 0);
    if (rhsNorm2 != 0.f) {
        float threshold = tol * tol * rhsNorm2;
        if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
        float residualNorm2 = rhsNorm2;                       // res = b: the same products in the same order
        if (residualNorm2 >= threshold) {
            pe = invd * res;
            sts(slot + kArr, res * pe);
            __syncthreads();
            float absNew = eigen_sum(P, kArr);
            while (it < maxIters) {
                // tmp = H.selfadjointView<Lower>() * p: row i receives the scattered terms of the columns before it, its diagonal
                // term, then the sum gathered over the rows below it -- two chains, each in Eigen's order
                float t = 0.f, g = 0.f;
#pragma unroll
                for (int j = 0; j < BS; ++j) {
                    const float v = hrow[j] * __shfl_sync(0xffffffffu, pe, base + j);
                    if (j <= i) t += v; else g += v;
                }
                const float tmp = t + g;
                sts(slot, pe * tmp);
                __syncthreads();
                const float alpha = absNew / eigen_sum(P, 0);
                x += alpha * pe;
                res -= alpha * tmp;
                ze = invd * res;
                sts(slot + kArr, res * res);
                sts(slot + 2 * kArr, res * ze);
                __syncthreads();
                // both sums at once: the lower half of every group of sixteen lanes adds up res . res, the upper half res . z
                const float mine = eigen_sum(P, off_d);
                const float other = __shfl_xor_sync(0xffffffffu, mine, 8);
                residualNorm2 = (lane & 8) ? other : mine;
                if (residualNorm2 < threshold) break;
                const float absOld = absNew;
                absNew = (lane & 8) ? mine : other;
                const float beta = absNew / absOld;
                pe = ze + beta * pe;
                ++it;
            }
        }
        err = sqrtf(residualNorm2 / rhsNorm2);
        success = err <= tol ? 1 : 0;
    }
    if (live) A.sys[(size_t)nblk * bb + n + e] = x;

    // ---- the update, in the keyframe table
    int applied = 1;
    if (A.kind == PSG_BLOCK_LIGHT) {
        // light_[frame] -= delta, unconditionally (PsOptimizer.cpp:199-201, LedOptimizer.cpp:159)
        if (live) {
            float* L = A.led_light ? &A.frames[kMaxFrames].light[i] : &A.frames[f].light[i];
            *L = *L - x;
        }
    } else {
        // Optimizer::updatePose (OptimizerAux.cpp:190-205): t <- t - xi[0:3], R <- R * exp(-xi[3:6]); SH unconditionally
        // (PsOptimizer.cpp:232), LED only on Success (LedOptimizer.cpp:271-273)
        applied = (!A.pose_needs_success || success) ? 1 : 0;
        __syncthreads();
        if (live) s_sum[e] = x;
        __syncthreads();
        const int fr = threadIdx.x;                   // one thread per keyframe
        if (applied && fr < nblk) {
            FrameParams& P = A.frames[fr];
            const float* d = s_sum + 6 * fr;
            float R[9], E[9];
#pragma unroll
            for (int a = 0; a < 9; ++a) R[a] = P.R[a];
            const float w[3] = { -d[3], -d[4], -d[5] };
            so3_exp_matrix_dev(w, E);
            P.t[0] = P.t[0] - d[0]; P.t[1] = P.t[1] - d[1]; P.t[2] = P.t[2] - d[2];
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    P.R[3 * a + c] = R[3 * a] * E[c] + (R[3 * a + 1] * E[3 + c] + R[3 * a + 2] * E[6 + c]);
        }
    }
    if (threadIdx.x == 0) {
        A.verdict[0] = (float)success; A.verdict[1] = (float)it; A.verdict[2] = err; A.verdict[3] = (float)applied;
    }
}

}  // namespace

// Queues the solve of the block whose sweep (and, across ranks, whose sums' reduction) is already queued on the stream.
void launch_small_solve(psg_optimizer* o, int block, int slot, float damping) {
    const bool led = o->set.model == PSG_MODEL_LED;
    SmallArgs A;
    memset(&A, 0, sizeof(A));
    A.kind = block;
    A.led_light = (block == PSG_BLOCK_LIGHT && led) ? 1 : 0;
    A.bs = block == PSG_BLOCK_POSE ? 6 : (led ? 3 : o->nb);
    A.nblk = A.led_light ? 1 : o->F;
    A.V = block == PSG_BLOCK_POSE ? 32 : (o->nb == 4 ? 16 : 64);
    A.damp = !(block == PSG_BLOCK_LIGHT && !led);   // every block but the SH light block (PsOptimizer.cpp:175-203)
    A.damping = damping;
    A.pose_needs_success = led ? 1 : 0;
    A.red = red_dev(o);
    A.scal = mail_dev(o) + 8 * slot;
    A.frames = o->frames_dev.p;
    A.sys = o->small_sys.p;     // allocated at creation for the largest system
    A.verdict = reinterpret_cast<float*>(mail_dev(o) + 8 * kSlotSmall) + (block == PSG_BLOCK_POSE ? 4 : 0);
    const int bpw = 32 / A.bs;                         // a block's rows stay within one warp
    const int threads = (A.nblk + bpw - 1) / bpw * 32;
    {
        Prof pf(o->stream, KID_SMALL_SOLVE);
        switch (A.bs) {
            case 3: k_small_solve<3><<<1, threads, 0, o->stream>>>(A); break;
            case 4: k_small_solve<4><<<1, threads, 0, o->stream>>>(A); break;
            case 6: k_small_solve<6><<<1, threads, 0, o->stream>>>(A); break;
            default: k_small_solve<9><<<1, threads, 0, o->stream>>>(A); break;
        }
    }
    count_launch(1);
    PSG_LAUNCH_CHECK();
    o->small_pending |= 1 << block;
    o->const_dirty = true;   // the sweeps' copy of the table is refreshed before the next one is queued
}

// after fetch_mail: {success, iterations, error, applied} of the block's solve
void read_small_verdict(const psg_optimizer* o, int block, int* success, int* iterations, float* error, int* applied) {
    const float* v = reinterpret_cast<const float*>(mail_slot(o, kSlotSmall)) + (block == PSG_BLOCK_POSE ? 4 : 0);
    *success = (int)(v[0] + 0.5f); *iterations = (int)(v[1] + 0.5f); *error = v[2]; *applied = (int)(v[3] + 0.5f);
}

// the device table as fetch_mail landed it -> o->poses / o->light
void adopt_device_frames(psg_optimizer* o) {
    const FrameTable* T = o->host_frames;
    const bool led = o->set.model == PSG_MODEL_LED;
    if (o->small_pending & (1 << PSG_BLOCK_POSE))
        for (int f = 0; f < o->F; ++f) {
            float* P = o->poses.data() + 16 * f;
            for (int a = 0; a < 3; ++a) { for (int b = 0; b < 3; ++b) P[4 * a + b] = T->f[f].R[3 * a + b]; P[4 * a + 3] = T->f[f].t[a]; }
        }
    if (o->small_pending & (1 << PSG_BLOCK_LIGHT)) {
        if (led) for (int c = 0; c < 3; ++c) o->light[c] = T->f[kMaxFrames].light[c];
        else for (int f = 0; f < o->F; ++f) for (int i = 0; i < o->nb; ++i) o->light[(size_t)f * o->nb + i] = T->f[f].light[i];
    }
    o->small_pending = 0;
}

}  // namespace psg
