// Distance block: assembly of the sparse S x S normal equations from per-voxel 4-point stencil blocks, the
// Jacobi-preconditioned conjugate gradient the reference runs on it, and the gated distance / gradient update.
//
// Replaces the tail of {Ps,Led}Optimizer::optimizeDistAll (PsOptimizer.cpp:124-172, LedOptimizer.cpp:198-244):
//   H = J^T W J (+ reg terms), H.diagonal() += damping * H.diagonal(),
//   Eigen::ConjugateGradient<SparseMatrix<float>> (Lower, DiagonalPreconditioner, tolerance = epsilon, <= 2n iterations,
//   third/eigen/Eigen/src/IterativeLinearSolvers/ConjugateGradient.h:28-93), Optimizer::updateDist + updateGrad
//   (OptimizerAux.cpp:152-188).
//
// The matrix is never formed.  Every residual row of the reference touches the voxel and its three finite-difference
// neighbours, so H is the sum over surface voxels t of a 4x4 block A_t scattered to rows/cols (t, n_x(t), n_y(t), n_z(t)).
// y = H p is evaluated in two gather phases without atomics: u_t = A_t * p[stencil(t)] (4 values per voxel), then
// y_s = u_s[0] + sum over the <= 6 stencils that contain s of u_t[slot].  Which stencils contain s is static per
// surface list (surface.cu, k_neighbors).  Dot products are block partials combined in a fixed order, so a solve is
// reproducible run to run.  One persistent cooperative kernel runs the whole solve (grid.sync between phases).
#include <cooperative_groups.h>

#include <algorithm>
#include <cstring>

#include "psg_host.h"

namespace cg = cooperative_groups;

namespace psg {

constexpr int kPcgBlock = 256;

struct PcgArgs {
    int S;
    const float* blk;        // 14 planes of S: stencil lower triangle (10) + rhs (4)
    const int* nbr_row;      // 3*S
    const int* inc;          // 6*S
    float* rhs; float* invd; float* ddiag;   // assembled rhs, 1 / damped diagonal, damping * diagonal
    float* x; float* r; float* p; float* tmp; float* U;   // U: 4 planes
    double* partials;        // [3][gridDim.x]
    unsigned long long* barrier;   // grid barrier arrival counter (zeroed by k_dist_assemble)
    float* result;           // [0] iterations, [1] residualNorm2, [2] rhsNorm2, [3] threshold
    float damping;
    int max_iters;
};

// slot pair (i >= k) -> index into the lower triangle stored in planes 0..9
__device__ __forceinline__ int tri(int i, int k) { return i * (i + 1) / 2 + k; }

__global__ void __launch_bounds__(kPcgBlock) k_dist_assemble(PcgArgs A) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    const int S = A.S;
    if (s == 0) *A.barrier = 0ull;
    if (s >= S) return;
    float rhs = A.blk[(long long)10 * S + s];
    float diag = A.blk[(long long)0 * S + s];
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const int t = A.inc[(h * 3 + a) * S + s];
            if (t >= 0) {
                rhs += A.blk[(long long)(10 + a + 1) * S + t];
                diag += A.blk[(long long)tri(a + 1, a + 1) * S + t];
            }
        }
    const float dd = (A.damping != 0.f) ? A.damping * diag : 0.f;
    const float full = diag + dd;
    A.rhs[s] = rhs;
    A.ddiag[s] = dd;
    A.invd[s] = (full != 0.f) ? 1.f / full : 1.f;   // DiagonalPreconditioner (BasicPreconditioners.h:67-82)
}

template <int BS>
__device__ __forceinline__ double block_sum(double v, double* s_buf) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) s_buf[warp] = v;
    __syncthreads();
    double t = 0.0;
    if (warp == 0) {
        t = lane < (BS / 32) ? s_buf[lane] : 0.0;
        t = warp_sum(t);
    }
    return t;   // valid in warp 0
}

// every block sums all block partials of one slot in the same order -> identical value everywhere
__device__ __forceinline__ double grid_total(const double* partials, int nblocks, double* s_buf) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (warp == 0) {
        double t = 0.0;
        for (int i = lane; i < nblocks; i += 32) t += partials[i];
        t = warp_sum(t);
        if (lane == 0) s_buf[0] = t;
    }
    __syncthreads();
    return s_buf[0];
}

// two sums at once (one pair of block barriers instead of two)
template <int BS>
__device__ __forceinline__ double2 block_sum2(double a, double b, double* s_buf /* 2 * BS / 32 */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    a = warp_sum(a); b = warp_sum(b);
    __syncthreads();
    if (lane == 0) { s_buf[warp] = a; s_buf[BS / 32 + warp] = b; }
    __syncthreads();
    double2 t = make_double2(0.0, 0.0);
    if (warp == 0) {
        t.x = lane < (BS / 32) ? s_buf[lane] : 0.0;
        t.y = lane < (BS / 32) ? s_buf[BS / 32 + lane] : 0.0;
        t.x = warp_sum(t.x); t.y = warp_sum(t.y);
    }
    return t;   // valid in warp 0
}

// block partials written by other blocks: read through L2 (an SM's L1 may still hold last iteration's line)
__device__ __forceinline__ double grid_total_cg(const double* partials, int nblocks, double* s_buf) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (warp == 0) {
        double t = 0.0;
        for (int i = lane; i < nblocks; i += 32) t += __ldcg(partials + i);
        t = warp_sum(t);
        if (lane == 0) s_buf[0] = t;
    }
    __syncthreads();
    return s_buf[0];
}

__device__ __forceinline__ double2 grid_total2_cg(const double* pa, const double* pb, int nblocks, double* s_buf) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (warp == 0) {
        double ta = 0.0, tb = 0.0;
        for (int i = lane; i < nblocks; i += 32) { ta += __ldcg(pa + i); tb += __ldcg(pb + i); }
        ta = warp_sum(ta); tb = warp_sum(tb);
        if (lane == 0) { s_buf[0] = ta; s_buf[1] = tb; }
    }
    __syncthreads();
    return make_double2(s_buf[0], s_buf[1]);
}

// Grid-wide barrier of the cooperative launch (all blocks co-resident): one monotonic arrival counter, zeroed by
// k_dist_assemble; a block arrives with a gpu-scope release and its thread 0 spins with gpu-scope acquires.  Data that other
// blocks wrote before the barrier is then read with ld.global.cg (L2), never through this SM's L1.
__device__ __forceinline__ void grid_barrier(unsigned long long* ctr, unsigned long long& target, unsigned nb) {
    __syncthreads();
    target += nb;
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(ctr) : "memory");
        unsigned long long v;
        do {
            asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(ctr) : "memory");
        } while (v < target);
    }
    __syncthreads();
}

// Cooperative launch over the whole GPU, vectors in global memory, three grid barriers per iteration (the search-direction
// update is folded into the stencil phase: neighbours' new p is recomputed from the published r, 1/diag and previous p).
// A thread walks its rows in groups of RPT and issues the loads of a whole group before using any of them, so a phase costs
// two dependent memory round trips per group instead of two per row.  RPT = 1 for small surfaces (one row per thread, every
// SM busy), 4 for large ones.
template <int BS, int RPT>
__global__ void __launch_bounds__(BS) k_dist_pcg(PcgArgs A, float* pbuf /* 2 x S */) {
    __shared__ double s_buf[2 * (BS / 32)];
    const int S = A.S;
    const int nb = gridDim.x;
    const int T = nb * BS;
    const int gtid = blockIdx.x * BS + threadIdx.x;
    double* P0 = A.partials;
    double* P1 = A.partials + nb;
    double* P2 = A.partials + 2 * nb;
    unsigned long long bar_target = 0;

    double a0 = 0.0, a1 = 0.0;
    for (int s = gtid; s < S; s += T) {
        const float b = __ldg(A.rhs + s);
        const float pv = __ldg(A.invd + s) * b;
        A.x[s] = 0.f; A.r[s] = b; pbuf[s] = pv;
        a0 += (double)b * (double)b;
        a1 += (double)b * (double)pv;
    }
    {
        const double2 t = block_sum2<BS>(a0, a1, s_buf);
        if (threadIdx.x == 0) { P0[blockIdx.x] = t.x; P1[blockIdx.x] = t.y; }
    }
    grid_barrier(A.barrier, bar_target, nb);
    const double2 t01 = grid_total2_cg(P0, P1, nb, s_buf);
    const float rhsNorm2 = (float)t01.x;
    float absNew = (float)t01.y;
    const float tol = 1.1920929e-07f;   // NumTraits<float>::epsilon(), the default tolerance
    float threshold = tol * tol * rhsNorm2;
    if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
    float residualNorm2 = rhsNorm2;
    int it = 0;
    if (rhsNorm2 != 0.f && !(residualNorm2 < threshold)) {
        float gamma = 0.f, beta = 1.f;
        int cur = 0;
        while (it < A.max_iters) {
            const float* pc = pbuf + (size_t)cur * S;
            float* pn = pbuf + (size_t)(cur ^ 1) * S;
            // ---- phase 1: p <- z + beta p, u_t = A_t * p[stencil(t)]
            for (int base = gtid; base < S; base += T * RPT) {
                int row[RPT], col[RPT][3];
                float pj[RPT];
#pragma unroll
                for (int j = 0; j < RPT; ++j) {
                    row[j] = base + j * T;
                    const bool in = row[j] < S;
#pragma unroll
                    for (int a = 0; a < 3; ++a) col[j][a] = in ? __ldg(A.nbr_row + a * S + row[j]) : -1;
                    pj[j] = in ? (__ldg(A.invd + row[j]) * __ldcg(A.r + row[j])) * gamma + beta * __ldcg(pc + row[j]) : 0.f;
                }
#pragma unroll
                for (int j = 0; j < RPT; ++j) {
                    if (row[j] >= S) continue;
                    pn[row[j]] = pj[j];
                    float px[4];
                    px[0] = pj[j];
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        const int c = col[j][a];
                        px[a + 1] = c >= 0 ? (__ldg(A.invd + c) * __ldcg(A.r + c)) * gamma + beta * __ldcg(pc + c) : 0.f;
                    }
                    float Ab[10];
#pragma unroll
                    for (int i = 0; i < 10; ++i) Ab[i] = __ldg(A.blk + (long long)i * S + row[j]);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float u = 0.f;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float aik = (i >= k) ? Ab[i * (i + 1) / 2 + k] : Ab[k * (k + 1) / 2 + i];
                            if (k == 0 || col[j][k - 1] >= 0) u += aik * px[k];
                        }
                        const bool valid = (i == 0) || col[j][i - 1] >= 0;
                        A.U[(long long)i * S + row[j]] = valid ? u : 0.f;
                    }
                }
            }
            grid_barrier(A.barrier, bar_target, nb);
            // ---- phase 2: tmp_s = u_s[0] + incident u + damping * diag * p_s ; p . tmp
            double pAp = 0.0;
            for (int base = gtid; base < S; base += T * RPT) {
                int tn[RPT][6];
#pragma unroll
                for (int j = 0; j < RPT; ++j)
#pragma unroll
                    for (int h = 0; h < 6; ++h) tn[j][h] = (base + j * T < S) ? __ldg(A.inc + h * S + base + j * T) : -1;
#pragma unroll
                for (int j = 0; j < RPT; ++j) {
                    const int s = base + j * T;
                    if (s >= S) continue;
                    float y = __ldcg(A.U + s);
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int a = 0; a < 3; ++a) {
                            const int t = tn[j][h * 3 + a];
                            if (t >= 0) y += __ldcg(A.U + (long long)(a + 1) * S + t);
                        }
                    const float ps = __ldcg(pn + s);
                    y += __ldg(A.ddiag + s) * ps;
                    A.tmp[s] = y;
                    pAp += (double)ps * (double)y;
                }
            }
            pAp = block_sum<BS>(pAp, s_buf); if (threadIdx.x == 0) P0[blockIdx.x] = pAp;
            grid_barrier(A.barrier, bar_target, nb);
            const float alpha = absNew / (float)grid_total_cg(P0, nb, s_buf);
            // ---- phase 3 (own rows only): x += alpha p ; residual -= alpha tmp ; norms
            double rn = 0.0, rz = 0.0;
            for (int base = gtid; base < S; base += T * RPT) {
                float xv[RPT], pv[RPT], rv[RPT], tv[RPT], iv[RPT];
#pragma unroll
                for (int j = 0; j < RPT; ++j) {
                    const int s = base + j * T;
                    const bool in = s < S;
                    xv[j] = in ? __ldcg(A.x + s) : 0.f; pv[j] = in ? __ldcg(pn + s) : 0.f; rv[j] = in ? __ldcg(A.r + s) : 0.f;
                    tv[j] = in ? __ldcg(A.tmp + s) : 0.f; iv[j] = in ? __ldg(A.invd + s) : 0.f;
                }
#pragma unroll
                for (int j = 0; j < RPT; ++j) {
                    const int s = base + j * T;
                    if (s >= S) continue;
                    A.x[s] = xv[j] + alpha * pv[j];
                    const float r2 = rv[j] - alpha * tv[j];
                    A.r[s] = r2;
                    rn += (double)r2 * (double)r2;
                    rz += (double)r2 * (double)(iv[j] * r2);
                }
            }
            {
                const double2 t = block_sum2<BS>(rn, rz, s_buf);
                if (threadIdx.x == 0) { P1[blockIdx.x] = t.x; P2[blockIdx.x] = t.y; }
            }
            grid_barrier(A.barrier, bar_target, nb);
            const double2 t12 = grid_total2_cg(P1, P2, nb, s_buf);
            residualNorm2 = (float)t12.x;
            if (residualNorm2 < threshold) break;
            const float absOld = absNew;
            absNew = (float)t12.y;
            beta = absNew / absOld;
            gamma = 1.f;
            cur ^= 1;
            ++it;
        }
    } else if (rhsNorm2 == 0.f) {
        residualNorm2 = 0.f;
    }
    if (gtid == 0) {
        A.result[0] = (float)it;
        A.result[1] = residualNorm2;
        A.result[2] = rhsNorm2;
        A.result[3] = threshold;
    }
}

// Barrier and dot-product reduction in one: warp 0 forms the block's partials, publishes them with the arrival, waits for the
// other blocks and sums all partials (same association as block_sum2 + grid_total2_cg: identical bits); the other warps only
// pass two block barriers.
template <int BS>
__device__ __forceinline__ double2 grid_reduce2(double a, double b, double* Pa, double* Pb, unsigned long long* ctr,
                                                unsigned long long& target, int nb, double* s_buf /* 2 * BS / 32 + 2 */) {
    constexpr int NW = BS / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    a = warp_sum(a); b = warp_sum(b);
    if (lane == 0) { s_buf[warp] = a; s_buf[NW + warp] = b; }
    target += (unsigned)nb;
    __syncthreads();
    if (warp == 0) {
        double ta = lane < NW ? s_buf[lane] : 0.0, tb = lane < NW ? s_buf[NW + lane] : 0.0;
        ta = warp_sum(ta); tb = warp_sum(tb);
        if (lane == 0) {
            Pa[blockIdx.x] = ta; Pb[blockIdx.x] = tb;
            asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(ctr) : "memory");
            unsigned long long v;
            do {
                asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(ctr) : "memory");
            } while (v < target);
        }
        __syncwarp();
        double xa = 0.0, xb = 0.0;
        for (int i = lane; i < nb; i += 32) { xa += __ldcg(Pa + i); xb += __ldcg(Pb + i); }
        xa = warp_sum(xa); xb = warp_sum(xb);
        if (lane == 0) { s_buf[2 * NW] = xa; s_buf[2 * NW + 1] = xb; }
    }
    __syncthreads();
    const double2 out = make_double2(s_buf[2 * NW], s_buf[2 * NW + 1]);
    return out;
}

// ---- surfaces with at most one row per resident thread: assembly, solve, update and regulariser energies in ONE launch
// Everything a row owns (its stencil block, neighbour / incidence indices, x, r, p, z, 1/diag) stays in registers for the
// whole solve; only what other rows read crosses L2: z = r / diag and p (for the neighbours' on-the-fly direction update),
// the stencil products U, and the block partials of the dot products.  The prologue is k_dist_assemble, the epilogue is
// k_dist_update + k_grad_update (+ k_reg_energy), separated by the same grid barrier -- five kernel launches become one.
struct SolveSmallArgs {
    PcgArgs A;
    VolView vol;
    SurfView surf;
    const int* lin;
    double thr;                      // sqrt(3) * voxel size: Optimizer::updateDist's gate
    float vs_inv;
    unsigned long long* accepted;
    int only_on_success;             // SH: the update is applied only if CG reports Success
    double* reg_out;                 // {normal, laplacian, albedo} regulariser energy sums after the update, or nullptr
    float* zbuf;                     // S floats: published z = r / diag
};

template <int BS>
__global__ void __launch_bounds__(BS, (BS >= 1024) ? 1 : (BS >= 512) ? 2 : 4) k_dist_solve_small(SolveSmallArgs G, float* pbuf /* 2 x S */) {
    const PcgArgs& A = G.A;
    __shared__ double s_buf[2 * (BS / 32) + 2];
    __shared__ double s_reg[BS / 32][3];
    const int S = A.S;
    const int nb = gridDim.x;
    const int row = blockIdx.x * BS + threadIdx.x;
    const bool in = row < S;
    double* P0 = A.partials;
    double* P1 = A.partials + nb;
    double* P2 = A.partials + 2 * nb;
    double* P3 = A.partials + 3 * nb;   // unused second lane of the p.Ap reduction
    unsigned long long bar_target = 0;

    // ---- prologue: assemble the row (k_dist_assemble) and keep it
    int col[3] = { -1, -1, -1 }, tn[6] = { -1, -1, -1, -1, -1, -1 };
    float Ab[10], invd = 0.f, dd = 0.f, b = 0.f;
#pragma unroll
    for (int i = 0; i < 10; ++i) Ab[i] = 0.f;
    if (in) {
#pragma unroll
        for (int a = 0; a < 3; ++a) col[a] = __ldg(A.nbr_row + a * S + row);
#pragma unroll
        for (int h = 0; h < 6; ++h) tn[h] = __ldg(A.inc + h * S + row);
#pragma unroll
        for (int i = 0; i < 10; ++i) Ab[i] = __ldg(A.blk + (long long)i * S + row);
        float rhs = __ldg(A.blk + (long long)10 * S + row);
        float diag = Ab[0];
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int t = tn[h * 3 + a];
                if (t >= 0) {
                    rhs += __ldg(A.blk + (long long)(10 + a + 1) * S + t);
                    diag += __ldg(A.blk + (long long)tri(a + 1, a + 1) * S + t);
                }
            }
        dd = (A.damping != 0.f) ? A.damping * diag : 0.f;
        const float full = diag + dd;
        invd = (full != 0.f) ? 1.f / full : 1.f;
        b = rhs;
        A.rhs[row] = rhs;          // diagnostics (psg_opt_last_vectors)
    }
    float x = 0.f, r = b, z = invd * b, p = z, u0 = 0.f, y = 0.f;
    if (in) { pbuf[row] = p; G.zbuf[row] = z; }
    const double2 t01 = grid_reduce2<BS>((double)b * (double)b, (double)b * (double)z, P0, P1, A.barrier, bar_target, nb, s_buf);
    const float rhsNorm2 = (float)t01.x;
    float absNew = (float)t01.y;
    const float tol = 1.1920929e-07f;
    float threshold = tol * tol * rhsNorm2;
    if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
    float residualNorm2 = rhsNorm2;
    int it = 0;
    if (rhsNorm2 != 0.f && !(residualNorm2 < threshold)) {
        float gamma = 0.f, beta = 1.f;   // first pass: the published p already is the search direction
        int cur = 0;
        while (it < A.max_iters) {
            const float* pc = pbuf + (size_t)cur * S;
            float* pn = pbuf + (size_t)(cur ^ 1) * S;
            // ---- phase 1: p <- z + beta p (own row from registers, neighbours recomputed from the published z and p)
            if (in) {
                p = z * gamma + beta * p;
                pn[row] = p;
                float px[4];
                px[0] = p;
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const int c = col[a];
                    px[a + 1] = c >= 0 ? __ldcg(G.zbuf + c) * gamma + beta * __ldcg(pc + c) : 0.f;
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float u = 0.f;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float aik = (i >= k) ? Ab[i * (i + 1) / 2 + k] : Ab[k * (k + 1) / 2 + i];
                        if (k == 0 || col[k - 1] >= 0) u += aik * px[k];
                    }
                    const bool valid = (i == 0) || col[i - 1] >= 0;
                    if (i == 0) u0 = u; else A.U[(long long)i * S + row] = valid ? u : 0.f;
                }
            }
            grid_barrier(A.barrier, bar_target, nb);
            // ---- phase 2: y = u_s[0] + incident u + damping * diag * p_s ; p . y
            double pAp = 0.0;
            if (in) {
                y = u0;
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        const int t = tn[h * 3 + a];
                        if (t >= 0) y += __ldcg(A.U + (long long)(a + 1) * S + t);
                    }
                y += dd * p;
                pAp = (double)p * (double)y;
            }
            const float alpha = absNew / (float)grid_reduce2<BS>(pAp, 0.0, P0, P3, A.barrier, bar_target, nb, s_buf).x;
            // ---- phase 3 (registers only, then publish z)
            double rn = 0.0, rz = 0.0;
            if (in) {
                x += alpha * p;
                r = r - alpha * y;
                z = invd * r;
                G.zbuf[row] = z;
                rn = (double)r * (double)r;
                rz = (double)r * (double)z;
            }
            const double2 t12 = grid_reduce2<BS>(rn, rz, P1, P2, A.barrier, bar_target, nb, s_buf);
            residualNorm2 = (float)t12.x;
            if (residualNorm2 < threshold) break;
            const float absOld = absNew;
            absNew = (float)t12.y;
            beta = absNew / absOld;
            gamma = 1.f;
            cur ^= 1;
            ++it;
        }
    } else if (rhsNorm2 == 0.f) {
        residualNorm2 = 0.f;
    }
    if (in) A.x[row] = x;
    if (row == 0) {
        A.result[0] = (float)it;
        A.result[1] = residualNorm2;
        A.result[2] = rhsNorm2;
        A.result[3] = threshold;
    }
    // ---- epilogue: Optimizer::updateDist (OptimizerAux.cpp:162-188), every block knows the verdict
    const float err = rhsNorm2 != 0.f ? sqrtf(residualNorm2 / rhsNorm2) : 0.f;
    const bool apply = !G.only_on_success || err <= tol;
    if (apply) {
        int acc = 0;
        if (in && fabs((double)x) < G.thr) { G.vol.dist[G.lin[row]] -= x; acc = 1; }
        const unsigned m = __ballot_sync(0xffffffffu, acc);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(G.accepted, (unsigned long long)__popc(m));
    }
    if (!apply && !G.reg_out) return;   // uniform over the grid
    grid_barrier(A.barrier, bar_target, nb);
    // Optimizer::updateGrad (OptimizerAux.cpp:152-160) + the slot records; any one-to-one thread -> slot map will do
    if (apply && in) {
        const int t = row;
        const int s = __float_as_int(G.surf.rec2[t].w);
        const int lin = __float_as_int(G.surf.rec3[t].w);
        const float dist = __ldcg(G.vol.dist + lin);
        const int pb = __float_as_int(G.surf.rec1[t].w);
        float fd[3], nfd[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const int nl = G.surf.nbr_lin[a * S + s];
            const float pp = ((pb >> a) & 1) ? 1.f : -1.f;
            const float dn = nl >= 0 ? __ldcg(G.vol.dist + nl) : dist;
            fd[a] = (pp * (dn - dist)) * G.vs_inv;
        }
        G.vol.gx[lin] = fd[0]; G.vol.gy[lin] = fd[1]; G.vol.gz[lin] = fd[2];
        normalized3(fd, nfd);
        G.surf.rec0[t] = make_float4(dist, nfd[0], nfd[1], nfd[2]);
        G.surf.rec2[t] = make_float4(fd[0], fd[1], fd[2], __int_as_float(s));
        G.surf.rec4[t] = make_float4(nfd[0], nfd[1], nfd[2], 0.f);
    }
    // getNormalEnergy / getLaplacianEnergy / getAlbedoRegEnergy (Optimizer.cpp:86-136) on the updated distances
    if (G.reg_out) {
        double en = 0.0, el = 0.0, er = 0.0;
        if (in) {
            const int s = row;
            const int lin = G.lin[s];
            const float dist = __ldcg(G.vol.dist + lin);
            const float rho[3] = { G.vol.r[lin], G.vol.g[lin], G.vol.b[lin] };
            const unsigned char pb = G.surf.pos[s];
            const long long stride[3] = { 1, G.vol.dims[0], (long long)G.vol.dims[0] * G.vol.dims[1] };
            float fd[3], d2[3], Gr[9];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int nl = G.surf.nbr_lin[a * S + s];
                const float pp = ((pb >> a) & 1) ? 1.f : -1.f;
                const float dn = nl >= 0 ? __ldcg(G.vol.dist + nl) : dist;
                fd[a] = (pp * (dn - dist)) * G.vs_inv;
                const long long l1 = (long long)lin + stride[a], l0 = (long long)lin - stride[a];
                const float d1 = (l1 < G.vol.N) ? __ldcg(G.vol.dist + l1) : dist;
                const float d0 = (l0 >= 0) ? __ldcg(G.vol.dist + l0) : dist;
                d2[a] = d1 + d0 - 2 * dist;
                const float nr = nl >= 0 ? G.vol.r[nl] : rho[0], ng = nl >= 0 ? G.vol.g[nl] : rho[1], nbb = nl >= 0 ? G.vol.b[nl] : rho[2];
                Gr[a] = (pp * (nr - rho[0])) * G.vs_inv; Gr[3 + a] = (pp * (ng - rho[1])) * G.vs_inv; Gr[6 + a] = (pp * (nbb - rho[2])) * G.vs_inv;
            }
            const float e = norm3(fd) - 1;
            en = (double)(e * e);
            const float lap = (d2[0] + d2[1] + d2[2]) * G.vs_inv * G.vs_inv;
            el = (double)(lap * lap);
            er = (double)((norm3(Gr) + norm3(Gr + 3)) + norm3(Gr + 6));
        }
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        en = warp_sum(en); el = warp_sum(el); er = warp_sum(er);
        if (lane == 0) { s_reg[warp][0] = en; s_reg[warp][1] = el; s_reg[warp][2] = er; }
        __syncthreads();
        if (threadIdx.x < 3) {
            double v = 0.0;
#pragma unroll
            for (int i = 0; i < BS / 32; ++i) v += s_reg[i][threadIdx.x];
            atomicAdd(&G.reg_out[threadIdx.x], v);
        }
    }
}

// three sums at once: warp 0 forms the block's partials, publishes them with the arrival, waits for the other blocks and adds all
// partials (fixed association: identical bits in every block)
template <int BS>
__device__ __forceinline__ double3 grid_reduce3(double a, double b, double d, double* Pa, double* Pb, double* Pd, unsigned long long* ctr,
                                                unsigned long long& target, int nb, double* s_buf /* 3 * BS / 32 + 3 */) {
    constexpr int NW = BS / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    a = warp_sum(a); b = warp_sum(b); d = warp_sum(d);
    if (lane == 0) { s_buf[warp] = a; s_buf[NW + warp] = b; s_buf[2 * NW + warp] = d; }
    target += (unsigned)nb;
    __syncthreads();
    if (warp == 0) {
        double ta = lane < NW ? s_buf[lane] : 0.0, tb = lane < NW ? s_buf[NW + lane] : 0.0, td = lane < NW ? s_buf[2 * NW + lane] : 0.0;
        ta = warp_sum(ta); tb = warp_sum(tb); td = warp_sum(td);
        if (lane == 0) {
            Pa[blockIdx.x] = ta; Pb[blockIdx.x] = tb; Pd[blockIdx.x] = td;
            asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(ctr) : "memory");
            unsigned long long v;
            do {
                asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(ctr) : "memory");
            } while (v < target);
        }
        __syncwarp();
        double xa = 0.0, xb = 0.0, xd = 0.0;
        for (int i = lane; i < nb; i += 32) { xa += __ldcg(Pa + i); xb += __ldcg(Pb + i); xd += __ldcg(Pd + i); }
        xa = warp_sum(xa); xb = warp_sum(xb); xd = warp_sum(xd);
        if (lane == 0) { s_buf[3 * NW] = xa; s_buf[3 * NW + 1] = xb; s_buf[3 * NW + 2] = xd; }
    }
    __syncthreads();
    return make_double3(s_buf[3 * NW], s_buf[3 * NW + 1], s_buf[3 * NW + 2]);
}

// ---- the fused small-surface kernel in the Chronopoulos-Gear arrangement: TWO grid barriers per CG iteration ---------------------
// With u = M^-1 r, w = A u and the recurrences p = u + beta p, s = w + beta s (= A p), x += alpha p, r -= alpha s, ONE fused
// reduction per iteration yields gamma = (r, u), delta = (w, u) and |r|^2; beta = gamma / gamma_old, alpha = gamma / (delta - beta
// gamma / alpha_old).  Same iterates as Eigen's loop in exact arithmetic and, in float, the same iteration counts on every system
// tried.  The update of a row needs nothing from other rows, so a thread does NOT wait for its neighbours' new u: it recomputes
// it from what they published one phase earlier (r, s double-buffered; w; 1/diag) with the very operations their owner
// executes -- same bits.  An iteration is then: [update + stencil products] barrier [gather into w + three dot products] barrier.
template <int BS>
__global__ void __launch_bounds__(BS, (BS >= 1024) ? 1 : (BS >= 512) ? 2 : 4) k_dist_solve_small_cg(SolveSmallArgs G, float* pbuf /* 2 x S */) {
    const PcgArgs& A = G.A;
    __shared__ double s_buf[3 * (BS / 32) + 3];
    __shared__ double s_reg[BS / 32][3];
    const int S = A.S;
    const int nb = gridDim.x;
    const int row = blockIdx.x * BS + threadIdx.x;
    const bool in = row < S;
    double* P0 = A.partials;
    double* P1 = A.partials + nb;
    double* P2 = A.partials + 2 * nb;
    unsigned long long bar_target = 0;
    // published per row: r and s double-buffered (pbuf, planes p / tmp), w (plane r), 1/diag (plane invd), U1..U3 (plane U + 1..3)
    // (A.p and A.tmp are consecutive planes of S floats, as are the two halves of pbuf)
    float* const Rb0 = pbuf;
    float* const Sb0 = A.p;
    float* Wp = A.r;

    // ---- prologue: assemble the row (k_dist_assemble) and keep it
    int col[3] = { -1, -1, -1 }, tn[6] = { -1, -1, -1, -1, -1, -1 };
    float Ab[10], invd = 0.f, dd = 0.f, b = 0.f;
#pragma unroll
    for (int i = 0; i < 10; ++i) Ab[i] = 0.f;
    if (in) {
#pragma unroll
        for (int a = 0; a < 3; ++a) col[a] = __ldg(A.nbr_row + a * S + row);
#pragma unroll
        for (int h = 0; h < 6; ++h) tn[h] = __ldg(A.inc + h * S + row);
#pragma unroll
        for (int i = 0; i < 10; ++i) Ab[i] = __ldg(A.blk + (long long)i * S + row);
        float rhs = __ldg(A.blk + (long long)10 * S + row);
        float diag = Ab[0];
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int t = tn[h * 3 + a];
                if (t >= 0) {
                    rhs += __ldg(A.blk + (long long)(10 + a + 1) * S + t);
                    diag += __ldg(A.blk + (long long)tri(a + 1, a + 1) * S + t);
                }
            }
        dd = (A.damping != 0.f) ? A.damping * diag : 0.f;
        const float full = diag + dd;
        invd = (full != 0.f) ? 1.f / full : 1.f;
        b = rhs;
        A.rhs[row] = rhs;          // diagnostics (psg_opt_last_vectors)
        A.invd[row] = invd; Rb0[row] = rhs; Sb0[row] = 0.f; Wp[row] = 0.f;
    }
    float x = 0.f, r = b, u = invd * b, p = 0.f, sv = 0.f, w = 0.f, u0 = 0.f;
    grid_barrier(A.barrier, bar_target, nb);
    const float tol = 1.1920929e-07f;
    float rhsNorm2 = 0.f, threshold = 0.f, residualNorm2 = 0.f, gamma = 0.f, alpha = 0.f, beta = 0.f;
    int it = 0, cur = 0;
    bool first = true;
    for (;;) {
        // ---- phase 1: own update (registers), neighbours' new u recomputed, stencil products
        if (in) {
            if (!first) {
                p = u + beta * p; sv = w + beta * sv;
                x += alpha * p; r = r - alpha * sv;
                u = invd * r;
                Rb0[(size_t)(cur ^ 1) * S + row] = r; Sb0[(size_t)(cur ^ 1) * S + row] = sv;
            }
            float px[4];
            px[0] = u;
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int c = col[a];
                if (c < 0) { px[a + 1] = 0.f; continue; }
                const float rc = __ldcg(Rb0 + (size_t)cur * S + c), ic = __ldcg(A.invd + c);
                if (first) px[a + 1] = ic * rc;
                else {
                    const float sc = __ldcg(Wp + c) + beta * __ldcg(Sb0 + (size_t)cur * S + c);
                    px[a + 1] = ic * (rc - alpha * sc);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float uu = 0.f;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float aik = (i >= k) ? Ab[i * (i + 1) / 2 + k] : Ab[k * (k + 1) / 2 + i];
                    if (k == 0 || col[k - 1] >= 0) uu += aik * px[k];
                }
                const bool valid = (i == 0) || col[i - 1] >= 0;
                if (i == 0) u0 = uu; else A.U[(long long)i * S + row] = valid ? uu : 0.f;
            }
        }
        if (!first) cur ^= 1;
        grid_barrier(A.barrier, bar_target, nb);
        // ---- phase 2: w = u_s[0] + incident u + damping * diag * u_s, published ; (r, u), (w, u), (r, r)
        double g = 0.0, d = 0.0, n = 0.0;
        if (in) {
            w = u0;
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const int t = tn[h * 3 + a];
                    if (t >= 0) w += __ldcg(A.U + (long long)(a + 1) * S + t);
                }
            w += dd * u;
            Wp[row] = w;
            g = (double)r * (double)u; d = (double)w * (double)u; n = (double)r * (double)r;
        }
        const double3 t3 = grid_reduce3<BS>(g, d, n, P0, P1, P2, A.barrier, bar_target, nb, s_buf);
        if (first) {
            first = false;
            rhsNorm2 = (float)t3.z;
            threshold = tol * tol * rhsNorm2;
            if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
            residualNorm2 = rhsNorm2;
            if (rhsNorm2 == 0.f) { residualNorm2 = 0.f; break; }
            if (residualNorm2 < threshold) break;
            gamma = (float)t3.x;
            alpha = gamma / (float)t3.y; beta = 0.f;
            if (it >= A.max_iters) break;
            continue;
        }
        residualNorm2 = (float)t3.z;
        if (residualNorm2 < threshold) break;
        const float gnew = (float)t3.x;
        beta = gnew / gamma;
        alpha = gnew / ((float)t3.y - beta * gnew / alpha);
        gamma = gnew;
        ++it;
        if (it >= A.max_iters) break;
    }
    if (in) A.x[row] = x;
    if (row == 0) {
        A.result[0] = (float)it;
        A.result[1] = residualNorm2;
        A.result[2] = rhsNorm2;
        A.result[3] = threshold;
    }
    // ---- epilogue: Optimizer::updateDist (OptimizerAux.cpp:162-188), every block knows the verdict
    const float err = rhsNorm2 != 0.f ? sqrtf(residualNorm2 / rhsNorm2) : 0.f;
    const bool apply = !G.only_on_success || err <= tol;
    if (apply) {
        int acc = 0;
        if (in && fabs((double)x) < G.thr) { G.vol.dist[G.lin[row]] -= x; acc = 1; }
        const unsigned m = __ballot_sync(0xffffffffu, acc);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(G.accepted, (unsigned long long)__popc(m));
    }
    if (!apply && !G.reg_out) return;   // uniform over the grid
    grid_barrier(A.barrier, bar_target, nb);
    // Optimizer::updateGrad (OptimizerAux.cpp:152-160) + the slot records; any one-to-one thread -> slot map will do
    if (apply && in) {
        const int t = row;
        const int s = __float_as_int(G.surf.rec2[t].w);
        const int lin = __float_as_int(G.surf.rec3[t].w);
        const float dist = __ldcg(G.vol.dist + lin);
        const int pb = __float_as_int(G.surf.rec1[t].w);
        float fd[3], nfd[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const int nl = G.surf.nbr_lin[a * S + s];
            const float pp = ((pb >> a) & 1) ? 1.f : -1.f;
            const float dn = nl >= 0 ? __ldcg(G.vol.dist + nl) : dist;
            fd[a] = (pp * (dn - dist)) * G.vs_inv;
        }
        G.vol.gx[lin] = fd[0]; G.vol.gy[lin] = fd[1]; G.vol.gz[lin] = fd[2];
        normalized3(fd, nfd);
        G.surf.rec0[t] = make_float4(dist, nfd[0], nfd[1], nfd[2]);
        G.surf.rec2[t] = make_float4(fd[0], fd[1], fd[2], __int_as_float(s));
        G.surf.rec4[t] = make_float4(nfd[0], nfd[1], nfd[2], 0.f);
    }
    // getNormalEnergy / getLaplacianEnergy / getAlbedoRegEnergy (Optimizer.cpp:86-136) on the updated distances
    if (G.reg_out) {
        double en = 0.0, el = 0.0, er = 0.0;
        if (in) {
            const int s = row;
            const int lin = G.lin[s];
            const float dist = __ldcg(G.vol.dist + lin);
            const float rho[3] = { G.vol.r[lin], G.vol.g[lin], G.vol.b[lin] };
            const unsigned char pb = G.surf.pos[s];
            const long long stride[3] = { 1, G.vol.dims[0], (long long)G.vol.dims[0] * G.vol.dims[1] };
            float fd[3], d2[3], Gr[9];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int nl = G.surf.nbr_lin[a * S + s];
                const float pp = ((pb >> a) & 1) ? 1.f : -1.f;
                const float dn = nl >= 0 ? __ldcg(G.vol.dist + nl) : dist;
                fd[a] = (pp * (dn - dist)) * G.vs_inv;
                const long long l1 = (long long)lin + stride[a], l0 = (long long)lin - stride[a];
                const float d1 = (l1 < G.vol.N) ? __ldcg(G.vol.dist + l1) : dist;
                const float d0 = (l0 >= 0) ? __ldcg(G.vol.dist + l0) : dist;
                d2[a] = d1 + d0 - 2 * dist;
                const float nr = nl >= 0 ? G.vol.r[nl] : rho[0], ng = nl >= 0 ? G.vol.g[nl] : rho[1], nbb = nl >= 0 ? G.vol.b[nl] : rho[2];
                Gr[a] = (pp * (nr - rho[0])) * G.vs_inv; Gr[3 + a] = (pp * (ng - rho[1])) * G.vs_inv; Gr[6 + a] = (pp * (nbb - rho[2])) * G.vs_inv;
            }
            const float e = norm3(fd) - 1;
            en = (double)(e * e);
            const float lap = (d2[0] + d2[1] + d2[2]) * G.vs_inv * G.vs_inv;
            el = (double)(lap * lap);
            er = (double)((norm3(Gr) + norm3(Gr + 3)) + norm3(Gr + 6));
        }
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        en = warp_sum(en); el = warp_sum(el); er = warp_sum(er);
        if (lane == 0) { s_reg[warp][0] = en; s_reg[warp][1] = el; s_reg[warp][2] = er; }
        __syncthreads();
        if (threadIdx.x < 3) {
            double v = 0.0;
#pragma unroll
            for (int i = 0; i < BS / 32; ++i) v += s_reg[i][threadIdx.x];
            atomicAdd(&G.reg_out[threadIdx.x], v);
        }
    }
}

// ---- surfaces of at most 32 768 rows: the whole distance block inside ONE 16-CTA cluster, vectors in distributed shared memory --
// The fused grid kernel above is a chain of grid barriers (~2 us each through L2, three per CG iteration) around a few hundred
// nanoseconds of arithmetic.  A thread-block cluster has a hardware barrier (barrier.cluster, ~0.2 us) and can read its peers'
// shared memory directly, so here: 16 CTAs x 1024 threads on 16 SMs of one GPC (non-portable cluster size), RPT <= 2 rows per
// thread (row = cta * RPC + j * 1024 + tid, RPC = 1024 RPT).  Everything the solve re-reads lives on chip: a row's stencil block
// and incidence addresses in its CTA's shared memory, x / r / p / z / 1/diag in registers, and what OTHER rows read -- z, p
// (double-buffered) and the stencil products U1..U3 -- in shared-memory planes that neighbours gather with ld.shared::cluster
// (a neighbour row's CTA and offset are fixed for the whole solve: mapped once, kept as 32-bit cluster addresses).  Dot products:
// every CTA writes its two partial sums into a slot of EVERY CTA's shared memory before the phase's barrier and adds the 16
// slots in slot order after it, so all CTAs hold the same bits and take the same branches.  Assembly (k_dist_assemble) is the
// prologue; Optimizer::updateDist, updateGrad, the record refresh and the regulariser energies are the epilogue, as in
// k_dist_solve_small.  Global memory is touched only by prologue and epilogue.
constexpr int kClNC = 16, kClBS = 1024;

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned mapa_u32(unsigned addr, unsigned cta) {
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta));
    return r;
}
__device__ __forceinline__ float ld_dsmem(unsigned addr) {
    float v;
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void st_dsmem_f64(unsigned addr, double v) { asm volatile("st.shared::cluster.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// per-CTA partials -> every CTA, then (after the caller's cluster barrier) the totals in slot order
template <int RPT>
struct ClSmem {
    static constexpr int RPC = 1024 * RPT;
    float Ab[10][RPC];          // stencil block of the CTA's rows
    unsigned taddr[6][RPC];     // cluster address of plane U1's entry of each incident row (0 = none)
    float z[RPC];               // published z = r / diag
    float p[2][RPC];            // search direction, double-buffered
    float U[3][RPC];            // stencil products of slots 1..3
    double slot[2][kClNC][2];   // [parity][sender CTA][value]
    double red[2 * 32];
};

template <int RPT>
__device__ __forceinline__ void cl_post(ClSmem<RPT>& sm, int par, unsigned cta, double a, double b) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    a = warp_sum(a); b = warp_sum(b);
    if (lane == 0) { sm.red[warp] = a; sm.red[32 + warp] = b; }
    __syncthreads();
    if (warp == 0) {
        double ta = sm.red[lane], tb = sm.red[32 + lane];
        ta = warp_sum(ta); tb = warp_sum(tb);
        if (lane < kClNC) {
            const unsigned dst = mapa_u32(smem_u32(&sm.slot[par][cta][0]), (unsigned)lane);
            st_dsmem_f64(dst, ta); st_dsmem_f64(dst + 8, tb);
        }
    }
}
template <int RPT>
__device__ __forceinline__ double2 cl_total(const ClSmem<RPT>& sm, int par) {
    double a = 0.0, b = 0.0;
#pragma unroll
    for (int q = 0; q < kClNC; ++q) { a += sm.slot[par][q][0]; b += sm.slot[par][q][1]; }
    return make_double2(a, b);
}

template <int RPT>
__global__ void __launch_bounds__(kClBS, 1) k_dist_solve_cluster16(SolveSmallArgs G) {
    extern __shared__ __align__(16) unsigned char cl_raw[];
    ClSmem<RPT>& sm = *reinterpret_cast<ClSmem<RPT>*>(cl_raw);
    constexpr int RPC = ClSmem<RPT>::RPC;
    constexpr int LOG_RPC = (RPT == 1) ? 10 : 11;
    const PcgArgs& A = G.A;
    const int S = A.S, tid = threadIdx.x;
    unsigned cta;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta));
    const unsigned z0 = smem_u32(&sm.z[0]);
    constexpr unsigned kP0 = (unsigned)(sizeof(float) * RPC);              // z -> p[0]
    constexpr unsigned kU = (unsigned)(sizeof(float) * RPC * 3);           // z -> U[0]

    // ---- prologue: assemble the rows (k_dist_assemble), stage what the solve re-reads
    int row[RPT]; bool in[RPT];
    unsigned naddr[RPT][3];      // cluster address of z of the finite-difference neighbour per axis (0 = not a surface voxel)
    float invd[RPT], dd[RPT], x[RPT], r[RPT], z[RPT], p[RPT], u0[RPT], y[RPT];
    double s0 = 0.0, s1 = 0.0;
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
        const int li = j * kClBS + tid;
        row[j] = (int)cta * RPC + li;
        in[j] = row[j] < S;
        invd[j] = 0.f; dd[j] = 0.f; x[j] = 0.f; r[j] = 0.f; z[j] = 0.f; p[j] = 0.f; u0[j] = 0.f; y[j] = 0.f;
#pragma unroll
        for (int a = 0; a < 3; ++a) naddr[j][a] = 0u;
#pragma unroll
        for (int h = 0; h < 6; ++h) sm.taddr[h][li] = 0u;
#pragma unroll
        for (int i = 0; i < 10; ++i) sm.Ab[i][li] = 0.f;
        sm.z[li] = 0.f; sm.p[0][li] = 0.f; sm.p[1][li] = 0.f; sm.U[0][li] = 0.f; sm.U[1][li] = 0.f; sm.U[2][li] = 0.f;
        if (!in[j]) continue;
        const int s = row[j];
        float rhs = __ldg(A.blk + (long long)10 * S + s);
        float diag = 0.f;
#pragma unroll
        for (int i = 0; i < 10; ++i) { const float v = __ldg(A.blk + (long long)i * S + s); sm.Ab[i][li] = v; if (i == 0) diag = v; }
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const int c = __ldg(A.nbr_row + a * S + s);
            if (c >= 0) naddr[j][a] = mapa_u32(z0 + 4u * (unsigned)(c & (RPC - 1)), (unsigned)(c >> LOG_RPC));
        }
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int t = __ldg(A.inc + (h * 3 + a) * S + s);
                if (t >= 0) {
                    rhs += __ldg(A.blk + (long long)(10 + a + 1) * S + t);
                    diag += __ldg(A.blk + (long long)tri(a + 1, a + 1) * S + t);
                    // U plane a of row t: where phase 2 gathers the product of t's slot a + 1
                    sm.taddr[h * 3 + a][li] = mapa_u32(z0 + kU + (unsigned)(sizeof(float) * RPC) * (unsigned)a + 4u * (unsigned)(t & (RPC - 1)), (unsigned)(t >> LOG_RPC));
                }
            }
        dd[j] = (A.damping != 0.f) ? A.damping * diag : 0.f;
        const float full = diag + dd[j];
        invd[j] = (full != 0.f) ? 1.f / full : 1.f;
        A.rhs[s] = rhs;          // diagnostics (psg_opt_last_vectors)
        r[j] = rhs; z[j] = invd[j] * rhs; p[j] = z[j];
        sm.z[li] = z[j]; sm.p[0][li] = p[j];
        s0 += (double)rhs * (double)rhs; s1 += (double)rhs * (double)z[j];
    }
    int par = 0;
    cl_post<RPT>(sm, par, cta, s0, s1);
    cluster_sync_all();
    const double2 t01 = cl_total<RPT>(sm, par);
    const float rhsNorm2 = (float)t01.x;
    float absNew = (float)t01.y;
    const float tol = 1.1920929e-07f;
    float threshold = tol * tol * rhsNorm2;
    if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
    float residualNorm2 = rhsNorm2;
    int it = 0;
    if (rhsNorm2 != 0.f && !(residualNorm2 < threshold)) {
        float gamma = 0.f, beta = 1.f;   // first pass: the published p already is the search direction
        int cur = 0;
        while (it < A.max_iters) {
            const unsigned pc_off = kP0 + (unsigned)cur * (unsigned)(sizeof(float) * RPC);
            // ---- phase 1: p <- z + beta p (neighbours recomputed from the published z and p), u = A_t * p[stencil]
#pragma unroll
            for (int j = 0; j < RPT; ++j) {
                if (!in[j]) continue;
                const int li = j * kClBS + tid;
                p[j] = z[j] * gamma + beta * p[j];
                sm.p[cur ^ 1][li] = p[j];
                float px[4];
                px[0] = p[j];
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const unsigned na = naddr[j][a];
                    px[a + 1] = na ? ld_dsmem(na) * gamma + beta * ld_dsmem(na + pc_off) : 0.f;
                }
                float Ab[10];
#pragma unroll
                for (int i = 0; i < 10; ++i) Ab[i] = sm.Ab[i][li];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float u = 0.f;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float aik = (i >= k) ? Ab[i * (i + 1) / 2 + k] : Ab[k * (k + 1) / 2 + i];
                        if (k == 0 || naddr[j][k - 1]) u += aik * px[k];
                    }
                    const bool valid = (i == 0) || naddr[j][i - 1];
                    if (i == 0) u0[j] = u; else sm.U[i - 1][li] = valid ? u : 0.f;
                }
            }
            cluster_sync_all();
            // ---- phase 2: y = u_s[0] + incident u + damping * diag * p_s ; p . y
            double pAp = 0.0;
#pragma unroll
            for (int j = 0; j < RPT; ++j) {
                if (!in[j]) continue;
                const int li = j * kClBS + tid;
                float yy = u0[j];
#pragma unroll
                for (int h = 0; h < 6; ++h) {
                    const unsigned ta = sm.taddr[h][li];
                    if (ta) yy += ld_dsmem(ta);
                }
                yy += dd[j] * p[j];
                y[j] = yy;
                pAp += (double)p[j] * (double)yy;
            }
            par ^= 1;
            cl_post<RPT>(sm, par, cta, pAp, 0.0);
            cluster_sync_all();
            const float alpha = absNew / (float)cl_total<RPT>(sm, par).x;
            // ---- phase 3 (registers, then publish z)
            double rn = 0.0, rz = 0.0;
#pragma unroll
            for (int j = 0; j < RPT; ++j) {
                if (!in[j]) continue;
                x[j] += alpha * p[j];
                r[j] = r[j] - alpha * y[j];
                z[j] = invd[j] * r[j];
                sm.z[j * kClBS + tid] = z[j];
                rn += (double)r[j] * (double)r[j];
                rz += (double)r[j] * (double)z[j];
            }
            par ^= 1;
            cl_post<RPT>(sm, par, cta, rn, rz);
            cluster_sync_all();
            const double2 t12 = cl_total<RPT>(sm, par);
            residualNorm2 = (float)t12.x;
            if (residualNorm2 < threshold) break;
            const float absOld = absNew;
            absNew = (float)t12.y;
            beta = absNew / absOld;
            gamma = 1.f;
            cur ^= 1;
            ++it;
        }
    } else if (rhsNorm2 == 0.f) {
        residualNorm2 = 0.f;
    }
#pragma unroll
    for (int j = 0; j < RPT; ++j) if (in[j]) A.x[row[j]] = x[j];
    if (cta == 0 && tid == 0) {
        A.result[0] = (float)it; A.result[1] = residualNorm2; A.result[2] = rhsNorm2; A.result[3] = threshold;
    }
    // ---- epilogue: Optimizer::updateDist (OptimizerAux.cpp:162-188); every CTA knows the verdict
    const float err = rhsNorm2 != 0.f ? sqrtf(residualNorm2 / rhsNorm2) : 0.f;
    const bool apply = !G.only_on_success || err <= tol;
    if (apply) {
        int acc = 0;
#pragma unroll
        for (int j = 0; j < RPT; ++j)
            if (in[j] && fabs((double)x[j]) < G.thr) { G.vol.dist[G.lin[row[j]]] -= x[j]; ++acc; }
        acc += __shfl_xor_sync(0xffffffffu, acc, 16); acc += __shfl_xor_sync(0xffffffffu, acc, 8); acc += __shfl_xor_sync(0xffffffffu, acc, 4);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2); acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        if ((tid & 31) == 0 && acc) atomicAdd(G.accepted, (unsigned long long)acc);
    }
    if (!apply && !G.reg_out) return;   // uniform over the cluster
    __threadfence();
    cluster_sync_all();
    // Optimizer::updateGrad (OptimizerAux.cpp:152-160) + the slot records; any one-to-one thread -> slot map will do
    double en = 0.0, el = 0.0, er = 0.0;
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
        if (!in[j]) continue;
        if (apply) {
            const int t = row[j];
            const int s = __float_as_int(G.surf.rec2[t].w);
            const int lin = __float_as_int(G.surf.rec3[t].w);
            const float dist = __ldcg(G.vol.dist + lin);
            const int pb = __float_as_int(G.surf.rec1[t].w);
            float fd[3], nfd[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int nl = G.surf.nbr_lin[a * S + s];
                const float pp = ((pb >> a) & 1) ? 1.f : -1.f;
                const float dn = nl >= 0 ? __ldcg(G.vol.dist + nl) : dist;
                fd[a] = (pp * (dn - dist)) * G.vs_inv;
            }
            G.vol.gx[lin] = fd[0]; G.vol.gy[lin] = fd[1]; G.vol.gz[lin] = fd[2];
            normalized3(fd, nfd);
            G.surf.rec0[t] = make_float4(dist, nfd[0], nfd[1], nfd[2]);
            G.surf.rec2[t] = make_float4(fd[0], fd[1], fd[2], __int_as_float(s));
            G.surf.rec4[t] = make_float4(nfd[0], nfd[1], nfd[2], 0.f);
        }
        // getNormalEnergy / getLaplacianEnergy / getAlbedoRegEnergy (Optimizer.cpp:86-136) on the updated distances
        if (G.reg_out) {
            const int s = row[j];
            const int lin = G.lin[s];
            const float dist = __ldcg(G.vol.dist + lin);
            const float rho[3] = { G.vol.r[lin], G.vol.g[lin], G.vol.b[lin] };
            const unsigned char pb = G.surf.pos[s];
            const long long stride[3] = { 1, G.vol.dims[0], (long long)G.vol.dims[0] * G.vol.dims[1] };
            float fd[3], d2[3], Gr[9];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int nl = G.surf.nbr_lin[a * S + s];
                const float pp = ((pb >> a) & 1) ? 1.f : -1.f;
                const float dn = nl >= 0 ? __ldcg(G.vol.dist + nl) : dist;
                fd[a] = (pp * (dn - dist)) * G.vs_inv;
                const long long l1 = (long long)lin + stride[a], l0 = (long long)lin - stride[a];
                const float d1 = (l1 < G.vol.N) ? __ldcg(G.vol.dist + l1) : dist;
                const float d0 = (l0 >= 0) ? __ldcg(G.vol.dist + l0) : dist;
                d2[a] = d1 + d0 - 2 * dist;
                const float nr = nl >= 0 ? G.vol.r[nl] : rho[0], ng = nl >= 0 ? G.vol.g[nl] : rho[1], nbb = nl >= 0 ? G.vol.b[nl] : rho[2];
                Gr[a] = (pp * (nr - rho[0])) * G.vs_inv; Gr[3 + a] = (pp * (ng - rho[1])) * G.vs_inv; Gr[6 + a] = (pp * (nbb - rho[2])) * G.vs_inv;
            }
            const float e = norm3(fd) - 1;
            en += (double)(e * e);
            const float lap = (d2[0] + d2[1] + d2[2]) * G.vs_inv * G.vs_inv;
            el += (double)(lap * lap);
            er += (double)((norm3(Gr) + norm3(Gr + 3)) + norm3(Gr + 6));
        }
    }
    if (G.reg_out) {
        const int lane = tid & 31, warp = tid >> 5;
        en = warp_sum(en); el = warp_sum(el); er = warp_sum(er);
        __syncthreads();
        if (lane == 0) { sm.red[warp] = en; sm.red[32 + warp] = el; }
        __shared__ double s_er[32];
        if (lane == 0) s_er[warp] = er;
        __syncthreads();
        if (tid < 3) {
            double v = 0.0;
            for (int i = 0; i < 32; ++i) v += (tid == 0) ? sm.red[i] : (tid == 1) ? sm.red[32 + i] : s_er[i];
            atomicAdd(&G.reg_out[tid], v);
        }
    }
}

// ---- small surfaces: the whole solve inside ONE thread-block cluster ------------------------------------------------
// At 128^3 an iteration touches ~4 MB; the grid-wide kernel above spends its time in grid.sync() and in dependent L2
// round trips.  Here <= 8 CTAs x 1024 threads form one cluster; a thread owns RPT rows (row = gtid + j * threads), keeps
// their x, r, p, 1/diag in registers for the whole solve, issues the loads of all its rows back to back (two dependent
// levels per phase instead of 2 x RPT), synchronises with the hardware cluster barrier, and needs three barriers per
// iteration: the search-direction update p <- z + beta p is not a phase of its own -- every thread recomputes the
// neighbours' new p on the fly from the published r, 1/diag and previous p (same operands, same operations, hence the
// same bits as the owner computes).
constexpr int kClusterBlock = 1024;

template <int RPT>
__global__ void __launch_bounds__(kClusterBlock) k_dist_pcg_cluster(PcgArgs A, float* __restrict__ pbuf /* 2 x S */) {
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ double s_buf[kClusterBlock / 32];
    const int S = A.S;
    const int nb = gridDim.x;
    const int T = nb * kClusterBlock;
    const int gtid = blockIdx.x * kClusterBlock + threadIdx.x;
    double* P0 = A.partials;
    double* P1 = A.partials + nb;
    double* P2 = A.partials + 2 * nb;

    int row[RPT];
    float x[RPT], r[RPT], p[RPT], invd[RPT], dd[RPT], u0[RPT], tmp[RPT];
    double a0 = 0.0, a1 = 0.0;
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
        row[j] = gtid + j * T;
        const bool in = row[j] < S;
        const float b = in ? A.rhs[row[j]] : 0.f;
        invd[j] = in ? A.invd[row[j]] : 0.f;
        dd[j] = in ? A.ddiag[row[j]] : 0.f;
        x[j] = 0.f; r[j] = b; p[j] = invd[j] * b; u0[j] = 0.f; tmp[j] = 0.f;
        if (in) { pbuf[row[j]] = p[j]; A.r[row[j]] = b; }
        a0 += (double)b * (double)b;
        a1 += (double)b * (double)p[j];
    }
    a0 = block_sum<kClusterBlock>(a0, s_buf); if (threadIdx.x == 0) P0[blockIdx.x] = a0;
    a1 = block_sum<kClusterBlock>(a1, s_buf); if (threadIdx.x == 0) P1[blockIdx.x] = a1;
    cluster.sync();
    const float rhsNorm2 = (float)grid_total(P0, nb, s_buf);
    float absNew = (float)grid_total(P1, nb, s_buf);
    const float tol = 1.1920929e-07f;
    float threshold = tol * tol * rhsNorm2;
    if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
    float residualNorm2 = rhsNorm2;
    int it = 0;
    if (rhsNorm2 != 0.f && !(residualNorm2 < threshold)) {
        float gamma = 0.f, beta = 1.f;   // first pass: the published p already is the search direction
        int cur = 0;
        while (it < A.max_iters) {
            const float* pc = pbuf + (size_t)cur * S;
            float* pn = pbuf + (size_t)(cur ^ 1) * S;
            // ---- phase 1: p <- z + beta p (own rows from registers, neighbours recomputed), u_t = A_t * p[stencil(t)]
            int col[RPT][3];
#pragma unroll
            for (int j = 0; j < RPT; ++j)
#pragma unroll
                for (int a = 0; a < 3; ++a) col[j][a] = row[j] < S ? A.nbr_row[a * S + row[j]] : -1;
#pragma unroll
            for (int j = 0; j < RPT; ++j) {
                if (row[j] >= S) continue;
                p[j] = (invd[j] * r[j]) * gamma + beta * p[j];
                pn[row[j]] = p[j];
                float px[4];
                px[0] = p[j];
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const int c = col[j][a];
                    px[a + 1] = c >= 0 ? (A.invd[c] * A.r[c]) * gamma + beta * pc[c] : 0.f;
                }
                float Ab[10];
#pragma unroll
                for (int i = 0; i < 10; ++i) Ab[i] = A.blk[(long long)i * S + row[j]];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float u = 0.f;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float aik = (i >= k) ? Ab[i * (i + 1) / 2 + k] : Ab[k * (k + 1) / 2 + i];
                        if (k == 0 || col[j][k - 1] >= 0) u += aik * px[k];
                    }
                    const bool valid = (i == 0) || col[j][i - 1] >= 0;
                    if (i == 0) u0[j] = u; else A.U[(long long)i * S + row[j]] = valid ? u : 0.f;
                }
            }
            cluster.sync();
            // ---- phase 2: tmp_s = u_s[0] + incident u + damping * diag * p_s ; p . tmp
            double pAp = 0.0;
#pragma unroll
            for (int j = 0; j < RPT; ++j) {
                if (row[j] >= S) continue;
                int tn[6];
#pragma unroll
                for (int h = 0; h < 6; ++h) tn[h] = A.inc[h * S + row[j]];
                float y = u0[j];
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        const int t = tn[h * 3 + a];
                        if (t >= 0) y += A.U[(long long)(a + 1) * S + t];
                    }
                y += dd[j] * p[j];
                tmp[j] = y;
                pAp += (double)p[j] * (double)y;
            }
            pAp = block_sum<kClusterBlock>(pAp, s_buf); if (threadIdx.x == 0) P0[blockIdx.x] = pAp;
            cluster.sync();
            const float alpha = absNew / (float)grid_total(P0, nb, s_buf);
            // ---- phase 3 (registers only): x += alpha p ; residual -= alpha tmp ; norms ; publish the residual
            double rn = 0.0, rz = 0.0;
#pragma unroll
            for (int j = 0; j < RPT; ++j) {
                if (row[j] >= S) continue;
                x[j] += alpha * p[j];
                r[j] = r[j] - alpha * tmp[j];
                A.r[row[j]] = r[j];
                rn += (double)r[j] * (double)r[j];
                rz += (double)r[j] * (double)(invd[j] * r[j]);
            }
            rn = block_sum<kClusterBlock>(rn, s_buf); if (threadIdx.x == 0) P1[blockIdx.x] = rn;
            rz = block_sum<kClusterBlock>(rz, s_buf); if (threadIdx.x == 0) P2[blockIdx.x] = rz;
            cluster.sync();
            residualNorm2 = (float)grid_total(P1, nb, s_buf);
            if (residualNorm2 < threshold) break;
            const float absOld = absNew;
            absNew = (float)grid_total(P2, nb, s_buf);
            beta = absNew / absOld;
            gamma = 1.f;
            cur ^= 1;
            ++it;
        }
    } else if (rhsNorm2 == 0.f) {
        residualNorm2 = 0.f;
    }
#pragma unroll
    for (int j = 0; j < RPT; ++j) if (row[j] < S) A.x[row[j]] = x[j];
    if (gtid == 0) {
        A.result[0] = (float)it;
        A.result[1] = residualNorm2;
        A.result[2] = rhsNorm2;
        A.result[3] = threshold;
    }
}

// ---- the distance block as a dataflow kernel: one persistent launch per rank, no barrier inside a CG iteration but its reduction
// Rank g owns rows [lo, hi) of the surface list (a z-slab; all rows on one GPU).  Its kernel runs k_dist_assemble, the CG,
// Optimizer::updateDist, updateGrad and the regulariser energies on those rows.
// Every vector entry that ANOTHER thread reads -- u = r / diag at a row's finite-difference neighbours, the stencil products
// U1..U3 at the stencils that contain a row -- is stored as ONE 8-byte packet (tag << 32 | float bits).  An 8-byte store is
// indivisible, so a reader needs neither fence nor barrier: it polls the packet until it carries the tag of the phase that
// produced the value (tags count the producing phases of the job, launches included: every rank counts the same, tags never
// repeat).  Own rows' packets live in the rank's planes; rows in a peer's halo (other ranks' rows within one z-plane of its
// range) are ALSO stored, through the peer's mapped pointer (NVLink), into the peer's halo buffer.  Threads therefore run
// ahead as far as their inputs allow; the only barrier of an iteration is its dot-product reduction, which also protects the
// packets: a packet is overwritten only after a reduction that its readers contributed to after reading it.
// The CG runs in the Chronopoulos-Gear arrangement (u = M^-1 r, w = A u, p = u + beta p, s = w + beta s): ONE fused reduction per
// iteration yields (r, u), (w, u) and |r|^2.  Reduction: blocks publish partial sums and arrive on the rank's counter; the rank's
// leader block adds them and sends the rank's sums to every rank (its own included) as six tagged packets (XState::ll,
// double-buffered by reduction parity); EVERY block polls its rank's mailbox until the packets of all ranks carry the
// reduction's tag and adds them in rank order -- identical bits in every block of every rank, hence identical branches.
// No system-scope fence anywhere.  Every wait gives up after `timeout_ns` and raises a sticky abort flag (here and at the peers)
// instead of hanging.  On a single device the ranks of a job (threads of one process, comm.cu) are played by ONE cooperative
// launch: block b belongs to rank b / nb.
enum HaloKind { HK_B2 = 0, HK_B5, HK_B9, HK_B11, HK_B12, HK_B13, HK_Z, HK_U1, HK_U2, HK_U3, HK_D };
static_assert(HK_D + 1 == kHaloKinds, "halo kinds");

struct MgSeg { int peer, b, e, hbase, nh; };   // own rows [b, e) are halo rows hbase + row of rank `peer`, whose halo has nh rows
struct MgRank {
    int rank, lo, hi;
    int halo_lo, halo_hi, nh;       // halo rows [halo_lo, lo) and [hi, halo_hi); nh of them
    int nseg;
    MgSeg seg[2 * (kMaxRanks - 1)];
    int inner_lo, inner_hi;         // rows in [inner_lo, inner_hi) are in nobody's halo
    float* vecs; float* pbuf; XState* xs; unsigned long long* halo;
    unsigned long long* xp;         // 4 planes of S packets: u, U1, U2, U3 of the own rows
    const int* nbr_row; const int* inc; const int* lin;
    VolView vol; SurfView surf;
    double* partials;               // [2][3][nb]
    float* result;                  // iterations, residualNorm2, rhsNorm2, threshold, abort flag
    double* regslot;                // {normal, laplacian, albedo energy sums, accepted updates}: partial sums over the own rows
};
struct MgArgs {
    int S, world, nb, nplayed;      // nb blocks per rank; this launch plays `nplayed` ranks (1, or all on a single device)
    PeerView peer[kMaxRanks];       // every rank's halo buffer and mailbox as addressable from this launch
    MgRank loc[kMaxRanks];          // the ranks played, in block order
    float damping, vs_inv;
    double thr;
    int max_iters, only_on_success, want_reg;
    int flags;
    unsigned long long timeout_ns;
};

constexpr int kMgBlock = 512;
constexpr int kMgRows = 2;      // rows per thread and batch in the gather phases

__device__ __forceinline__ unsigned long long global_ns() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ unsigned long long ld_vol64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_vol64(unsigned long long* p, unsigned long long v) { asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
// packets that only this GPU's threads exchange: gpu scope (read from / written to L2, never through an L1) is enough and cheaper
__device__ __forceinline__ unsigned long long ld_gpu64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_gpu64(unsigned long long* p, unsigned long long v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }

// spin until pred(value) holds; false after the time limit or once the abort flag is up
template <typename Load, typename Pred>
__device__ __forceinline__ bool spin_until(Load load, Pred pred, XState* xs, unsigned long long limit_ns, unsigned long long* out) {
    unsigned spins = 0;
    unsigned long long t0 = 0;
    for (;;) {
        const unsigned long long v = load();
        if (pred(v)) { *out = v; return true; }
        if ((++spins & 255u) == 0) {
            if (ld_vol64(&xs->abort)) { *out = v; return false; }
            const unsigned long long now = global_ns();
            if (!t0) t0 = now;
            else if (now - t0 > limit_ns) { st_vol64(&xs->abort, 1ull); *out = v; return false; }
        }
    }
}

struct MgCtx {
    const MgArgs* A; const MgRank* R;
    int lb, nb;                      // block index within the rank, blocks per rank
    unsigned long long base;         // barriers completed by earlier launches
    unsigned long long base_r;       // ... of which reductions
    unsigned k, kr;                  // barriers / reductions completed by this launch
    unsigned tg;                     // last value tag handed out (a producing phase takes the next one)
    bool ok;
};

// push one value of an own row to every rank that has the row in its halo
__device__ __forceinline__ void halo_push(const MgCtx& c, int kind, int s, float v, unsigned tag) {
    const MgRank& R = *c.R;
    if (s >= R.inner_lo && s < R.inner_hi) return;
    const unsigned long long pkt = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
    for (int i = 0; i < R.nseg; ++i)
        if (s >= R.seg[i].b && s < R.seg[i].e)
            st_vol64(c.A->peer[R.seg[i].peer].halo + (size_t)kind * R.seg[i].nh + (R.seg[i].hbase + s), pkt);
}
// poll a packet until it carries `tag`
__device__ __noinline__ unsigned long long packet_spin(MgCtx& c, const unsigned long long* p, unsigned tag) {
    unsigned long long v;
    const bool local = p >= c.R->xp && p < c.R->xp + 4 * (size_t)c.A->S;   // own planes: gpu scope; halo buffer (written by peers): system scope
    if (!spin_until([&] { return local ? ld_gpu64(p) : ld_vol64(p); }, [&](unsigned long long x) { return (unsigned)(x >> 32) == tag; }, c.R->xs, c.A->timeout_ns, &v)) c.ok = false;
    return v;
}
__device__ __forceinline__ unsigned long long packet_load(const MgCtx& c, int kind, int row) {
    const MgRank& R = *c.R;
    if (row >= R.lo && row < R.hi) return ld_gpu64(R.xp + (size_t)kind * (size_t)c.A->S + row);
    const int h = row < R.lo ? row - R.halo_lo : (R.lo - R.halo_lo) + (row - R.hi);
    return ld_vol64(R.halo + (size_t)(HK_Z + kind) * R.nh + h);
}
// where the packet of `kind` (0 = u, 1..3 = U1..U3) of row `row` lives for this rank: its own planes, or its halo buffer
__device__ __forceinline__ const unsigned long long* packet_addr(const MgCtx& c, int kind, int row) {
    const MgRank& R = *c.R;
    if (row >= R.lo && row < R.hi) return R.xp + (size_t)kind * (size_t)c.A->S + row;
    const int h = row < R.lo ? row - R.halo_lo : (R.lo - R.halo_lo) + (row - R.hi);
    return R.halo + (size_t)(HK_Z + kind) * R.nh + h;
}
// every 256th unsuccessful poll: give up (false) once the abort flag is up or the time limit has passed
__device__ __noinline__ bool wait_may_continue(MgCtx& c, unsigned long long& t0) {
    XState* xs = c.R->xs;
    if (ld_vol64(&xs->abort)) { c.ok = false; return false; }
    const unsigned long long now = global_ns();
    if (!t0) { t0 = now; return true; }
    if (now - t0 > c.A->timeout_ns) { st_vol64(&xs->abort, 1ull); c.ok = false; return false; }
    return true;
}
__device__ __forceinline__ void packet_put(const MgCtx& c, int kind, int s, float v, unsigned tag) {
    st_gpu64(c.R->xp + (size_t)kind * (size_t)c.A->S + s, ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v));
    halo_push(c, HK_Z + kind, s, v, tag);
}
// a halo row's value of `kind` carrying `tag` (prologue / epilogue quantities, which have no planes of their own)
__device__ __forceinline__ float halo_get(MgCtx& c, int kind, int row, unsigned tag) {
    const MgRank& R = *c.R;
    const int h = row < R.lo ? row - R.halo_lo : (R.lo - R.halo_lo) + (row - R.hi);
    const unsigned long long* p = R.halo + (size_t)kind * R.nh + h;
    unsigned long long v = ld_vol64(p);
    if ((unsigned)(v >> 32) != tag) v = packet_spin(c, p, tag);
    return __uint_as_float((unsigned)v);
}

// Barrier of the rank's blocks (REDUCE = false): every block arrives on the rank's counter and polls it, as grid_barrier does.
// REDUCE = true adds the two-scalar all-reduce over all ranks: blocks publish their partial sums and arrive; the rank's leader
// block waits for the arrivals, adds the partials and sends the rank's sums to every rank (its own included) as four tagged
// packets; EVERY block then polls its rank's mailbox until the packets of all ranks carry this barrier's tag and adds them in
// rank order -- identical bits in every block of every rank.  The own rank's packet doubles as the local release: it is sent
// only after all local blocks have arrived.
template <bool REDUCE>
__device__ __forceinline__ double3 mg_barrier(MgCtx& c, double a, double b, double d, double* s_buf /* 3 * NW + 4 */) {
    constexpr int NW = kMgBlock / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const MgArgs& A = *c.A; const MgRank& R = *c.R;
    XState* xs = R.xs;
    // partial sums double-buffered by reduction parity: a block may publish for the next reduction while a slower one still adds up this one
    double* P0 = R.partials + (size_t)((c.kr + 1) & 1u) * 3 * c.nb; double* P1 = P0 + c.nb; double* P2 = P0 + 2 * c.nb;
    if (REDUCE) {
        a = warp_sum(a); b = warp_sum(b); d = warp_sum(d);
        if (lane == 0) { s_buf[warp] = a; s_buf[NW + warp] = b; s_buf[2 * NW + warp] = d; }
    }
    __syncthreads();
    ++c.k;
    const unsigned long long target = (unsigned long long)c.k * (unsigned long long)c.nb;
    // relaxed polls (an acquire per poll would invalidate the L1 every time); one fence after the wait orders what follows
    auto load_arrive = [&] { return ld_gpu64(&xs->arrive); };
    if (!REDUCE) {
        if (threadIdx.x == 0) {
            asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(&xs->arrive) : "memory");
            unsigned long long v;
            if (!spin_until(load_arrive, [&](unsigned long long x) { return x >= target; }, xs, A.timeout_ns, &v)) c.ok = false;
            __threadfence();
            s_buf[3 * NW + 3] = c.ok ? 0.0 : 1.0;
        }
        __syncthreads();
        if (s_buf[3 * NW + 3] != 0.0) c.ok = false;
        __syncthreads();
        return make_double3(0.0, 0.0, 0.0);
    }
    // mailbox parity and packet tag count REDUCTIONS: a peer can be at most one reduction ahead, so two buffers suffice
    ++c.kr;
    const unsigned long long er = c.base_r + c.kr;
    const int par = (int)(er & 1ull);
    const unsigned tag = (unsigned)er;
    if (warp == 0 && A.world == 1) {
        // one rank: every block waits for the arrivals and adds the partials itself, in block order (no mailbox round trip)
        double ta = lane < NW ? s_buf[lane] : 0.0, tb = lane < NW ? s_buf[NW + lane] : 0.0, td = lane < NW ? s_buf[2 * NW + lane] : 0.0;
        ta = warp_sum(ta); tb = warp_sum(tb); td = warp_sum(td);
        bool ok = true;
        if (lane == 0) {
            P0[c.lb] = ta; P1[c.lb] = tb; P2[c.lb] = td;
            asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(&xs->arrive) : "memory");
            unsigned long long v;
            ok = spin_until(load_arrive, [&](unsigned long long x) { return x >= target; }, xs, A.timeout_ns, &v);
            __threadfence();
        }
        ok = __shfl_sync(0xffffffffu, ok ? 1 : 0, 0) != 0;
        double xa = 0.0, xb = 0.0, xd = 0.0;
        for (int i = lane; i < c.nb; i += 32) { xa += __ldcg(P0 + i); xb += __ldcg(P1 + i); xd += __ldcg(P2 + i); }
        xa = warp_sum(xa); xb = warp_sum(xb); xd = warp_sum(xd);
        if (lane == 0) { s_buf[3 * NW] = xa; s_buf[3 * NW + 1] = xb; s_buf[3 * NW + 2] = xd; s_buf[3 * NW + 3] = ok ? 0.0 : 1.0; }
    } else if (warp == 0) {
        double ta = lane < NW ? s_buf[lane] : 0.0, tb = lane < NW ? s_buf[NW + lane] : 0.0, td = lane < NW ? s_buf[2 * NW + lane] : 0.0;
        ta = warp_sum(ta); tb = warp_sum(tb); td = warp_sum(td);
        bool ok = true;
        if (lane == 0) {
            P0[c.lb] = ta; P1[c.lb] = tb; P2[c.lb] = td;
            asm volatile("red.release.gpu.global.add.u64 [%0], 1;" ::"l"(&xs->arrive) : "memory");
        }
        if (c.lb == 0) {
            if (lane == 0) {
                unsigned long long v;
                ok = spin_until(load_arrive, [&](unsigned long long x) { return x >= target; }, xs, A.timeout_ns, &v);
                __threadfence();
            }
            ok = __shfl_sync(0xffffffffu, ok ? 1 : 0, 0) != 0;
            double xa = 0.0, xb = 0.0, xd = 0.0;
            for (int i = lane; i < c.nb; i += 32) { xa += __ldcg(P0 + i); xb += __ldcg(P1 + i); xd += __ldcg(P2 + i); }
            xa = warp_sum(xa); xb = warp_sum(xb); xd = warp_sum(xd);
            __threadfence();   // what the arrivals made visible to this block, before the packet that releases the others
            if (lane < A.world) {
                unsigned long long* dst = &A.peer[lane].xs->ll[par][R.rank][0];
                const unsigned long long th = (unsigned long long)tag << 32;
                if (!ok) st_vol64(&A.peer[lane].xs->abort, 1ull);
                const double v3[3] = { xa, xb, xd };
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    const unsigned long long u = (unsigned long long)__double_as_longlong(v3[i]);
                    st_vol64(dst + 2 * i, th | (u & 0xffffffffull)); st_vol64(dst + 2 * i + 1, th | (u >> 32));
                }
            }
        }
        // every block: the sums of all ranks from the own mailbox
        double xa = 0.0, xb = 0.0, xd = 0.0;
        if (lane < A.world) {
            const unsigned long long* src = &xs->ll[par][lane][0];
            unsigned long long w[6];
            unsigned spins = 0;
            unsigned long long t0 = 0;
            for (;;) {   // the packets of a sender are polled together: one round trip when they are there
                bool all = true;
#pragma unroll
                for (int i = 0; i < 6; ++i) w[i] = ld_vol64(src + i);
#pragma unroll
                for (int i = 0; i < 6; ++i) all = all && (unsigned)(w[i] >> 32) == tag;
                if (all) break;
                if ((++spins & 255u) == 0) {
                    if (ld_vol64(&xs->abort)) { ok = false; break; }
                    const unsigned long long now = global_ns();
                    if (!t0) t0 = now;
                    else if (now - t0 > A.timeout_ns) { st_vol64(&xs->abort, 1ull); ok = false; break; }
                }
            }
            xa = __longlong_as_double((long long)((w[0] & 0xffffffffull) | (w[1] << 32)));
            xb = __longlong_as_double((long long)((w[2] & 0xffffffffull) | (w[3] << 32)));
            xd = __longlong_as_double((long long)((w[4] & 0xffffffffull) | (w[5] << 32)));
        }
        ok = __all_sync(0xffffffffu, ok);
        double sa = 0.0, sb = 0.0, sd = 0.0;
        for (int q = 0; q < A.world; ++q) {   // rank order
            sa += __shfl_sync(0xffffffffu, xa, q); sb += __shfl_sync(0xffffffffu, xb, q); sd += __shfl_sync(0xffffffffu, xd, q);
        }
        __threadfence();   // the packet was a relaxed load: order the other blocks' data behind it
        if (lane == 0) {
            s_buf[3 * NW] = sa; s_buf[3 * NW + 1] = sb; s_buf[3 * NW + 2] = sd;
            if (!ok) st_vol64(&xs->abort, 1ull);
            s_buf[3 * NW + 3] = ok ? 0.0 : 1.0;
        }
    }
    __syncthreads();
    if (s_buf[3 * NW + 3] != 0.0) c.ok = false;
    const double3 out = make_double3(s_buf[3 * NW], s_buf[3 * NW + 1], s_buf[3 * NW + 2]);
    __syncthreads();   // s_buf is reused by the next call
    return out;
}

// ONE_ROW: the rank has at most one row per thread; the row's whole CG state then lives in registers
template <bool ONE_ROW>
__global__ void __launch_bounds__(kMgBlock, ONE_ROW ? 1 : 2) k_dist_block_mg(const MgArgs* __restrict__ Ap) {
    __shared__ double s_buf[3 * (kMgBlock / 32) + 4];
    __shared__ double s_reg[kMgBlock / 32][4];
    const MgArgs& A = *Ap;
    const int nb = A.nb;
    const MgRank& R = A.loc[blockIdx.x / nb];
    MgCtx c;
    c.A = &A; c.R = &R; c.lb = blockIdx.x % nb; c.nb = nb; c.k = 0; c.kr = 0; c.ok = true;
    c.base = __ldcg(&R.xs->epoch); c.base_r = __ldcg(&R.xs->epoch_r); c.tg = (unsigned)__ldcg(&R.xs->epoch_t);
    const int S = A.S, lo = R.lo, hi = R.hi;
    const int T = nb * kMgBlock, ltid = c.lb * kMgBlock + threadIdx.x;
    const size_t PS = (size_t)S;
    float* V = R.vecs;
    float* Pp = R.pbuf;              // search direction p
    float* Sp = R.pbuf + PS;         // s = A p
    const float tol = 1.1920929e-07f;
    auto own = [&](int row) { return row >= lo && row < hi; };
    // vecs planes: 0..13 stencil blocks | 14 rhs | 15 1/diag | 16 damping*diag | 17 x | 18 u = r/diag | 19 r | 20 w = A u | 21 U0

    // ---- the stencil planes the neighbours' assembly reads: rhs of slots 1..3 and the diagonal entries (a+1, a+1)
    const unsigned t_stencil = ++c.tg;
    for (int s = lo + ltid; s < hi; s += T) {
        if (s >= R.inner_lo && s < R.inner_hi) continue;
        const int planes[6] = { 2, 5, 9, 11, 12, 13 };
#pragma unroll
        for (int i = 0; i < 6; ++i) halo_push(c, HK_B2 + i, s, V[planes[i] * PS + s], t_stencil);
    }
    unsigned t_u = 0;
    float rhsNorm2 = 0.f, threshold = 0.f, residualNorm2 = 0.f;
    int it = 0;
    if constexpr (ONE_ROW) {
        // ---- one row per thread: assembly, then the whole solve with the row's state in registers
        const int s = lo + ltid;
        const bool in = s < hi;
        int col[3] = { -1, -1, -1 }, tn[6] = { -1, -1, -1, -1, -1, -1 };
        const unsigned long long* pu[3] = { nullptr, nullptr, nullptr };     // u packets of the finite-difference neighbours
        const unsigned long long* pU[6] = { nullptr, nullptr, nullptr, nullptr, nullptr, nullptr };   // U packets of the incident stencils
        float Ab[10], invd = 0.f, dd = 0.f, x = 0.f, r = 0.f, u = 0.f, p = 0.f, sv = 0.f, w = 0.f;
#pragma unroll
        for (int i = 0; i < 10; ++i) Ab[i] = 0.f;
        t_u = ++c.tg;
        if (in) {
#pragma unroll
            for (int a = 0; a < 3; ++a) { col[a] = __ldg(R.nbr_row + (size_t)a * S + s); if (col[a] >= 0) pu[a] = packet_addr(c, 0, col[a]); }
#pragma unroll
            for (int h = 0; h < 6; ++h) { tn[h] = __ldg(R.inc + (size_t)h * S + s); if (tn[h] >= 0) pU[h] = packet_addr(c, 1 + h % 3, tn[h]); }
#pragma unroll
            for (int i = 0; i < 10; ++i) Ab[i] = __ldg(V + (size_t)i * PS + s);
            float rhs = __ldg(V + 10 * PS + s), diag = Ab[0];
#pragma unroll
            for (int h = 0; h < 6; ++h) {
                const int t = tn[h], a = h % 3;
                if (t < 0) continue;
                if (own(t)) { rhs += __ldg(V + (size_t)(10 + a + 1) * PS + t); diag += __ldg(V + (size_t)tri(a + 1, a + 1) * PS + t); }
                else { rhs += halo_get(c, HK_B11 + a, t, t_stencil); diag += halo_get(c, HK_B2 + a, t, t_stencil); }
            }
            dd = (A.damping != 0.f) ? A.damping * diag : 0.f;
            const float full = diag + dd;
            invd = (full != 0.f) ? 1.f / full : 1.f;
            r = rhs; u = invd * rhs;
            V[14 * PS + s] = rhs;          // diagnostics (psg_opt_last_vectors)
            packet_put(c, 0, s, u, t_u);
        }
        auto matrix_pass1 = [&](double& g_ru, double& d_wu, double& n_rr) {
            const unsigned tU = ++c.tg;
            if (in) {
                // all packets of a phase are polled TOGETHER: one round trip per attempt, however many of them are late
                unsigned long long pk[3];
                {
                    unsigned spins = 0; unsigned long long t0 = 0;
                    for (;;) {
                        bool all = true;
#pragma unroll
                        for (int a = 0; a < 3; ++a) pk[a] = pu[a] ? (own(col[a]) ? ld_gpu64(pu[a]) : ld_vol64(pu[a])) : 0ull;
#pragma unroll
                        for (int a = 0; a < 3; ++a) all = all && (!pu[a] || (unsigned)(pk[a] >> 32) == t_u);
                        if (all) break;
                        if ((++spins & 255u) == 0 && !wait_may_continue(c, t0)) break;
                    }
                }
                float px[4];
                px[0] = u;
#pragma unroll
                for (int a = 0; a < 3; ++a) px[a + 1] = pu[a] ? __uint_as_float((unsigned)pk[a]) : 0.f;
                float u0 = 0.f;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float uu = 0.f;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float aik = (i >= k) ? Ab[i * (i + 1) / 2 + k] : Ab[k * (k + 1) / 2 + i];
                        if (k == 0 || col[k - 1] >= 0) uu += aik * px[k];
                    }
                    const bool valid = (i == 0) || col[i - 1] >= 0;
                    if (i == 0) u0 = uu; else packet_put(c, i, s, valid ? uu : 0.f, tU);
                }
                unsigned long long qk[6];
                {
                    unsigned spins = 0; unsigned long long t0 = 0;
                    for (;;) {
                        bool all = true;
#pragma unroll
                        for (int h = 0; h < 6; ++h) qk[h] = pU[h] ? (own(tn[h]) ? ld_gpu64(pU[h]) : ld_vol64(pU[h])) : 0ull;
#pragma unroll
                        for (int h = 0; h < 6; ++h) all = all && (!pU[h] || (unsigned)(qk[h] >> 32) == tU);
                        if (all) break;
                        if ((++spins & 255u) == 0 && !wait_may_continue(c, t0)) break;
                    }
                }
                w = u0;
#pragma unroll
                for (int h = 0; h < 6; ++h)
                    if (pU[h]) w += __uint_as_float((unsigned)qk[h]);
                w += dd * u;
                g_ru += (double)r * (double)u; d_wu += (double)w * (double)u; n_rr += (double)r * (double)r;
            }
        };
        double g0 = 0.0, d0 = 0.0, n0 = 0.0;
        matrix_pass1(g0, d0, n0);
        const double3 t0 = mg_barrier<true>(c, g0, d0, n0, s_buf);
        rhsNorm2 = (float)t0.z;
        float gamma = (float)t0.x;
        float alpha = gamma / (float)t0.y, beta = 0.f;
        threshold = tol * tol * rhsNorm2;
        if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
        residualNorm2 = rhsNorm2;
        if (c.ok && rhsNorm2 != 0.f && !(residualNorm2 < threshold)) {
            while (it < A.max_iters) {
                t_u = ++c.tg;
                if (in) {
                    p = u + beta * p; sv = w + beta * sv;
                    x += alpha * p; r = r - alpha * sv;
                    u = invd * r;
                    packet_put(c, 0, s, u, t_u);
                }
                double g1 = 0.0, d1 = 0.0, n1 = 0.0;
                matrix_pass1(g1, d1, n1);
                const double3 t1 = mg_barrier<true>(c, g1, d1, n1, s_buf);
                if (!c.ok) break;
                residualNorm2 = (float)t1.z;
                if (residualNorm2 < threshold) break;
                const float gnew = (float)t1.x;
                beta = gnew / gamma;
                alpha = gnew / ((float)t1.y - beta * gnew / alpha);
                gamma = gnew;
                ++it;
            }
        } else if (rhsNorm2 == 0.f) {
            residualNorm2 = 0.f;
        }
        if (in) V[17 * PS + s] = x;
    } else {
        // ---- k_dist_assemble on the own rows; x = 0, r = b, u = b / diag, p = s = 0
        t_u = ++c.tg;
        for (int s = lo + ltid; s < hi; s += T) {
            float rhs = V[10 * PS + s], diag = V[s];
    #pragma unroll
            for (int h = 0; h < 2; ++h)
    #pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const int t = __ldg(R.inc + (size_t)(h * 3 + a) * S + s);
                    if (t < 0) continue;
                    if (own(t)) { rhs += __ldg(V + (size_t)(10 + a + 1) * PS + t); diag += __ldg(V + (size_t)tri(a + 1, a + 1) * PS + t); }
                    else { rhs += halo_get(c, HK_B11 + a, t, t_stencil); diag += halo_get(c, HK_B2 + a, t, t_stencil); }
                }
            const float dd = (A.damping != 0.f) ? A.damping * diag : 0.f;
            const float full = diag + dd;
            const float invd = (full != 0.f) ? 1.f / full : 1.f;
            const float u = invd * rhs;
            V[14 * PS + s] = rhs; V[15 * PS + s] = invd; V[16 * PS + s] = dd; V[17 * PS + s] = 0.f; V[19 * PS + s] = rhs;
            V[18 * PS + s] = u; Pp[s] = 0.f; Sp[s] = 0.f;
            packet_put(c, 0, s, u, t_u);
        }
        // One "matrix pass" = stencil products from u (phase B), gather into w with the three dot products (phase C); no barrier
        // in between: the gathers wait for the packets they need.
        auto matrix_pass = [&](double& g_ru, double& d_wu, double& n_rr) {
            const unsigned tU = ++c.tg;
            // ---- phase B: u_t = A_t * u[stencil(t)], rows in batches (all loads of a batch are issued before any is used)
            for (int base = lo + ltid; base < hi; base += T * kMgRows) {
                int col[kMgRows][3];
                unsigned long long pk[kMgRows][3];
                float u0[kMgRows];
    #pragma unroll
                for (int j = 0; j < kMgRows; ++j) {
                    const int s = base + j * T;
    #pragma unroll
                    for (int a = 0; a < 3; ++a) col[j][a] = s < hi ? __ldg(R.nbr_row + (size_t)a * S + s) : -1;
                }
    #pragma unroll
                for (int j = 0; j < kMgRows; ++j) {
                    const int s = base + j * T;
                    u0[j] = V[18 * PS + (s < hi ? s : lo)];
    #pragma unroll
                    for (int a = 0; a < 3; ++a) pk[j][a] = col[j][a] >= 0 ? packet_load(c, 0, col[j][a]) : 0ull;
                }
                {   // late packets: poll the whole batch together
                    unsigned spins = 0; unsigned long long t0 = 0;
                    for (;;) {
                        bool all = true;
#pragma unroll
                        for (int j = 0; j < kMgRows; ++j)
#pragma unroll
                            for (int a = 0; a < 3; ++a) all = all && (col[j][a] < 0 || (unsigned)(pk[j][a] >> 32) == t_u);
                        if (all) break;
                        if ((++spins & 255u) == 0 && !wait_may_continue(c, t0)) break;
#pragma unroll
                        for (int j = 0; j < kMgRows; ++j)
#pragma unroll
                            for (int a = 0; a < 3; ++a) if (col[j][a] >= 0 && (unsigned)(pk[j][a] >> 32) != t_u) pk[j][a] = packet_load(c, 0, col[j][a]);
                    }
                }
    #pragma unroll
                for (int j = 0; j < kMgRows; ++j) {
                    const int s = base + j * T;
                    if (s >= hi) continue;
                    float px[4];
                    px[0] = u0[j];
    #pragma unroll
                    for (int a = 0; a < 3; ++a) px[a + 1] = col[j][a] >= 0 ? __uint_as_float((unsigned)pk[j][a]) : 0.f;
                    float Ab[10];
    #pragma unroll
                    for (int i = 0; i < 10; ++i) Ab[i] = __ldg(V + (size_t)i * PS + s);
    #pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float u = 0.f;
    #pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float aik = (i >= k) ? Ab[i * (i + 1) / 2 + k] : Ab[k * (k + 1) / 2 + i];
                            if (k == 0 || col[j][k - 1] >= 0) u += aik * px[k];
                        }
                        const bool valid = (i == 0) || col[j][i - 1] >= 0;
                        const float uv = valid ? u : 0.f;
                        if (i == 0) V[21 * PS + s] = uv; else packet_put(c, i, s, uv, tU);
                    }
                }
            }
            // ---- phase C: w = u_s[0] + incident u + damping * diag * u_s ; (r, u), (w, u), (r, r)
            for (int base = lo + ltid; base < hi; base += T * kMgRows) {
                int tn[kMgRows][6];
    #pragma unroll
                for (int j = 0; j < kMgRows; ++j)
    #pragma unroll
                    for (int h = 0; h < 6; ++h) tn[j][h] = (base + j * T < hi) ? __ldg(R.inc + (size_t)h * S + base + j * T) : -1;
                unsigned long long pk[kMgRows][6];
                float uu[kMgRows], us[kMgRows], dg[kMgRows], rs[kMgRows];
    #pragma unroll
                for (int j = 0; j < kMgRows; ++j) {
                    const int s = base + j * T;
                    const int s0 = s < hi ? s : lo;
                    uu[j] = V[21 * PS + s0]; us[j] = V[18 * PS + s0]; dg[j] = V[16 * PS + s0]; rs[j] = V[19 * PS + s0];
    #pragma unroll
                    for (int h = 0; h < 6; ++h) pk[j][h] = tn[j][h] >= 0 ? packet_load(c, 1 + h % 3, tn[j][h]) : 0ull;
                }
                {
                    unsigned spins = 0; unsigned long long t0 = 0;
                    for (;;) {
                        bool all = true;
#pragma unroll
                        for (int j = 0; j < kMgRows; ++j)
#pragma unroll
                            for (int h = 0; h < 6; ++h) all = all && (tn[j][h] < 0 || (unsigned)(pk[j][h] >> 32) == tU);
                        if (all) break;
                        if ((++spins & 255u) == 0 && !wait_may_continue(c, t0)) break;
#pragma unroll
                        for (int j = 0; j < kMgRows; ++j)
#pragma unroll
                            for (int h = 0; h < 6; ++h) if (tn[j][h] >= 0 && (unsigned)(pk[j][h] >> 32) != tU) pk[j][h] = packet_load(c, 1 + h % 3, tn[j][h]);
                    }
                }
    #pragma unroll
                for (int j = 0; j < kMgRows; ++j) {
                    const int s = base + j * T;
                    if (s >= hi) continue;
                    float w = uu[j];
    #pragma unroll
                    for (int h = 0; h < 6; ++h)
                        if (tn[j][h] >= 0) w += __uint_as_float((unsigned)pk[j][h]);
                    w += dg[j] * us[j];
                    V[20 * PS + s] = w;
                    g_ru += (double)rs[j] * (double)us[j];
                    d_wu += (double)w * (double)us[j];
                    n_rr += (double)rs[j] * (double)rs[j];
                }
            }
        };
        double g0 = 0.0, d0 = 0.0, n0 = 0.0;
        matrix_pass(g0, d0, n0);
        const double3 t0 = mg_barrier<true>(c, g0, d0, n0, s_buf);
        rhsNorm2 = (float)t0.z;
        float gamma = (float)t0.x;
        float alpha = gamma / (float)t0.y, beta = 0.f;
        threshold = tol * tol * rhsNorm2;
        if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
        residualNorm2 = rhsNorm2;
        if (c.ok && rhsNorm2 != 0.f && !(residualNorm2 < threshold)) {
            while (it < A.max_iters) {
                // ---- phase A (own rows): p = u + beta p ; s = w + beta s ; x += alpha p ; r -= alpha s ; u = r / diag, published
                t_u = ++c.tg;
                for (int s = lo + ltid; s < hi; s += T) {
                    const float pnew = V[18 * PS + s] + beta * Pp[s];
                    const float snew = V[20 * PS + s] + beta * Sp[s];
                    Pp[s] = pnew; Sp[s] = snew;
                    V[17 * PS + s] = V[17 * PS + s] + alpha * pnew;
                    const float r2 = V[19 * PS + s] - alpha * snew;
                    V[19 * PS + s] = r2;
                    const float u = V[15 * PS + s] * r2;
                    V[18 * PS + s] = u;
                    packet_put(c, 0, s, u, t_u);
                }
                double g1 = 0.0, d1 = 0.0, n1 = 0.0;
                matrix_pass(g1, d1, n1);
                const double3 t1 = mg_barrier<true>(c, g1, d1, n1, s_buf);
                if (!c.ok) break;
                residualNorm2 = (float)t1.z;
                if (residualNorm2 < threshold) break;
                const float gnew = (float)t1.x;
                beta = gnew / gamma;
                alpha = gnew / ((float)t1.y - beta * gnew / alpha);
                gamma = gnew;
                ++it;
            }
        } else if (rhsNorm2 == 0.f) {
            residualNorm2 = 0.f;
        }
    }
    if (ltid == 0) {
        R.result[0] = (float)it; R.result[1] = residualNorm2; R.result[2] = rhsNorm2; R.result[3] = threshold;
    }
    // ---- Optimizer::updateDist (OptimizerAux.cpp:162-188) on the own rows; the new distances of the boundary rows go to the peers
    const float err = rhsNorm2 != 0.f ? sqrtf(residualNorm2 / rhsNorm2) : 0.f;
    const bool apply = c.ok && (!A.only_on_success || err <= tol);
    double acc = 0.0;
    const unsigned t_dist = ++c.tg;
    for (int s = lo + ltid; s < hi; s += T) {
        const int lin = R.lin[s];
        float d = R.vol.dist[lin];
        if (apply) {
            const float x = V[17 * PS + s];
            if (fabs((double)x) < A.thr) { d -= x; R.vol.dist[lin] = d; acc += 1.0; }
        }
        halo_push(c, HK_D, s, d, t_dist);
    }
    // the halo rows' distances into the own distance plane (the finite differences and the Laplacian read it by voxel index)
    {
        const unsigned td = t_dist;
        for (int h = ltid; h < R.nh; h += T) {
            const int row = h < lo - R.halo_lo ? R.halo_lo + h : hi + (h - (lo - R.halo_lo));
            R.vol.dist[R.lin[row]] = halo_get(c, HK_D, row, td);
        }
    }
    mg_barrier<false>(c, 0.0, 0.0, 0.0, s_buf);
    // ---- Optimizer::updateGrad (OptimizerAux.cpp:152-160) + the slot records of the own slots
    if (apply) {
        for (int t = ltid; t < R.surf.nslots; t += T) {
            const int s = __float_as_int(R.surf.rec2[t].w);
            const int lin = __float_as_int(R.surf.rec3[t].w);
            const float dist = __ldcg(R.vol.dist + lin);
            const int pb = __float_as_int(R.surf.rec1[t].w);
            float fd[3], nfd[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int nl = R.surf.nbr_lin[(size_t)a * S + s];
                const float pp = ((pb >> a) & 1) ? 1.f : -1.f;
                const float dn = nl >= 0 ? __ldcg(R.vol.dist + nl) : dist;
                fd[a] = (pp * (dn - dist)) * A.vs_inv;
            }
            R.vol.gx[lin] = fd[0]; R.vol.gy[lin] = fd[1]; R.vol.gz[lin] = fd[2];
            normalized3(fd, nfd);
            R.surf.rec0[t] = make_float4(dist, nfd[0], nfd[1], nfd[2]);
            R.surf.rec2[t] = make_float4(fd[0], fd[1], fd[2], __int_as_float(s));
            R.surf.rec4[t] = make_float4(nfd[0], nfd[1], nfd[2], 0.f);
        }
    }
    // ---- getNormalEnergy / getLaplacianEnergy (Optimizer.cpp:86-119) over the own rows, and the accepted-update count
    double en = 0.0, el = 0.0;
    if (A.want_reg) {
        for (int s = lo + ltid; s < hi; s += T) {
            const int lin = R.lin[s];
            const float dist = __ldcg(R.vol.dist + lin);
            const unsigned char pb = R.surf.pos[s];
            const long long stride[3] = { 1, R.vol.dims[0], (long long)R.vol.dims[0] * R.vol.dims[1] };
            float fd[3], d2[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int nl = R.surf.nbr_lin[(size_t)a * S + s];
                const float pp = ((pb >> a) & 1) ? 1.f : -1.f;
                const float dn = nl >= 0 ? __ldcg(R.vol.dist + nl) : dist;
                fd[a] = (pp * (dn - dist)) * A.vs_inv;
                const long long l1 = (long long)lin + stride[a], l0 = (long long)lin - stride[a];
                const float d1 = (l1 < R.vol.N) ? __ldcg(R.vol.dist + l1) : dist;
                const float d0 = (l0 >= 0) ? __ldcg(R.vol.dist + l0) : dist;
                d2[a] = d1 + d0 - 2 * dist;
            }
            const float e = norm3(fd) - 1;
            en += (double)(e * e);
            const float lap = (d2[0] + d2[1] + d2[2]) * A.vs_inv * A.vs_inv;
            el += (double)(lap * lap);
        }
    }
    {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        en = warp_sum(en); el = warp_sum(el); acc = warp_sum(acc);
        if (lane == 0) { s_reg[warp][0] = en; s_reg[warp][1] = el; s_reg[warp][3] = acc; }
        __syncthreads();
        if (threadIdx.x < 4 && threadIdx.x != 2) {
            double v = 0.0;
            for (int i = 0; i < kMgBlock / 32; ++i) v += s_reg[i][threadIdx.x];
            if (v != 0.0) atomicAdd(&R.regslot[threadIdx.x], v);
        }
    }
    // ---- bookkeeping for the next launch (every block has left its last barrier: nothing reads these any more)
    if (c.lb == 0 && threadIdx.x == 0) {
        R.result[4] = ld_vol64(&R.xs->abort) ? 1.f : 0.f;
        R.xs->arrive = 0ull;
        R.xs->epoch = c.base + c.k;
        R.xs->epoch_r = c.base_r + c.kr;
        R.xs->epoch_t = (unsigned long long)c.tg;
    }
}

// ---- regularised albedo block (reg albedo != 0) ------------------------------------------------------------------------
// {Ps,Led}Optimizer::optimizeAlbedoAll with albedo_reg (PsOptimizer.cpp:85-121, LedOptimizer.cpp:162-196):
//   H = J^T W J + (reg * Jr^T) * Jr,  b = J^T W r + (reg * Jr^T) * r_r,  H.diagonal() += damping * H.diagonal(),
//   Jacobi-preconditioned CG on the 3S x 3S system, Optimizer::updateAlbedo (OptimizerAux.cpp:120-150).
// Jr (Optimizer::albedoRegJacobian, Optimizer.cpp:221-245, 593-647) has one row per surface voxel t and channel c with <= 4
// entries: a0 = (grad rho_c . (-pos / vs)) / |grad rho_c| in the voxel's own column and a_k = grad rho_c[k] * pos[k] / vs / |grad
// rho_c| in the column of its finite-difference neighbour along axis k (only if that neighbour is a surface voxel); the
// residual is |grad rho_c| (Optimizer::computeAlbedoGrad, Optimizer.cpp:396-460).  The reference writes the BLUE row's own
// entry into the GREEN column (3 * row + 1, Optimizer.cpp:618): reproduced, so the blue channel of a voxel is regularised
// only through its neighbours' rows and the green unknown is coupled to the blue rows.
// The matrix is never formed: u = Jr p per row (phase 1), y = D p + reg * Jr^T u gathered per unknown through the same
// incidence lists the distance solve uses (phase 2).  One persistent cooperative launch, grid barriers between phases.
struct AregArgs {
    int S;
    VolView vol;
    const int* lin; const int* nbr_lin; const int* nbr_row; const int* inc; const unsigned char* pos;
    const float* bd; const float* Hd;     // [3 s + c]: J^T W r and the diagonal of J^T W J from the albedo sweep
    float* w;                             // 39 planes of S floats (layout below)
    float* b_out; float* delta_out;       // [3 s + c] diagnostics (psg_opt_last_vectors), may alias bd / Hd
    float reg, damping, vs_inv;
    int max_iters, only_on_success;
    double* partials;                     // [3][gridDim.x]
    unsigned long long* barrier;
    float* result;                        // iterations, residualNorm2, rhsNorm2, threshold
    unsigned long long* accepted;
};
// planes: 0..11 coefficient a_k of channel c at 3 k + c | 12..14 rhs | 15..17 1 / damped diagonal | 18..20 damping * diagonal |
//         21..23 x | 24..26 r | 27..29 p | 30..32 y = H p | 33..35 u = Jr p

template <int BS>
__global__ void __launch_bounds__(BS) k_albedo_reg_solve(AregArgs A) {
    __shared__ double s_buf[2 * (BS / 32)];
    const int S = A.S, nb = gridDim.x, T = nb * BS, gtid = blockIdx.x * BS + threadIdx.x;
    double* P0 = A.partials; double* P1 = A.partials + nb; double* P2 = A.partials + 2 * nb;
    unsigned long long bar_target = 0;
    float* W = A.w;
    auto pl = [&](int plane, int s) -> float* { return W + (size_t)plane * S + s; };

    // ---- set-up A: albedo gradient, its row norms and the Jacobian coefficients of every row
    for (int s = gtid; s < S; s += T) {
        const int lin = A.lin[s];
        const float rho[3] = { A.vol.r[lin], A.vol.g[lin], A.vol.b[lin] };
        const unsigned char pb = A.pos[s];
        float G[9], posf[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const int nl = A.nbr_lin[a * S + s];
            posf[a] = ((pb >> a) & 1) ? 1.f : -1.f;
            const float nr = nl >= 0 ? A.vol.r[nl] : rho[0], ng = nl >= 0 ? A.vol.g[nl] : rho[1], nbb = nl >= 0 ? A.vol.b[nl] : rho[2];
            G[a] = (posf[a] * (nr - rho[0])) * A.vs_inv; G[3 + a] = (posf[a] * (ng - rho[1])) * A.vs_inv; G[6 + a] = (posf[a] * (nbb - rho[2])) * A.vs_inv;
        }
        float rd[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) rd[a] = -A.vs_inv * posf[a];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float gn = norm3(G + 3 * c);
            float a0 = dot3(G + 3 * c, rd);
            float ak[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) ak[a] = G[3 * c + a] * (0.f + A.vs_inv * posf[a]);
            if (gn != 0.f) { a0 /= gn; ak[0] /= gn; ak[1] /= gn; ak[2] /= gn; }
            *pl(c, s) = a0; *pl(3 + c, s) = ak[0]; *pl(6 + c, s) = ak[1]; *pl(9 + c, s) = ak[2];
            *pl(33 + c, s) = gn;   // residual of the row, parked in the u planes until set-up B has used it
        }
    }
    grid_barrier(A.barrier, bar_target, nb);
    // ---- set-up B: right-hand side, diagonal (preconditioner, damping), start vectors
    double a0s = 0.0, a1s = 0.0;
    for (int s = gtid; s < S; s += T) {
        int tn[6];
#pragma unroll
        for (int h = 0; h < 6; ++h) tn[h] = __ldg(A.inc + h * S + s);
        float rhs[3], diag[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) { rhs[c] = A.bd[3 * (long long)s + c]; diag[c] = A.Hd[3 * (long long)s + c]; }
        // own rows: red -> column 0, green -> column 1, blue -> column 1 (the reference's index)
        {
            const float ar = __ldcg(pl(0, s)), ag = __ldcg(pl(1, s)), ab = __ldcg(pl(2, s));
            const float sr = A.reg * ar, sg = A.reg * ag, sb = A.reg * ab;
            rhs[0] += sr * __ldcg(pl(33, s)); diag[0] += sr * ar;
            rhs[1] += sg * __ldcg(pl(34, s)); diag[1] += sg * ag;
            rhs[1] += sb * __ldcg(pl(35, s)); diag[1] += sb * ab;
        }
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                const int t = tn[h * 3 + a];
                if (t < 0) continue;
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float ak = __ldcg(pl(3 * (a + 1) + c, t));
                    const float sk = A.reg * ak;
                    rhs[c] += sk * __ldcg(pl(33 + c, t)); diag[c] += sk * ak;
                }
            }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float dd = (A.damping != 0.f) ? A.damping * diag[c] : 0.f;
            const float full = diag[c] + dd;
            const float invd = (full != 0.f) ? 1.f / full : 1.f;
            const float pv = invd * rhs[c];
            *pl(12 + c, s) = rhs[c]; *pl(15 + c, s) = invd; *pl(18 + c, s) = dd;
            *pl(21 + c, s) = 0.f; *pl(24 + c, s) = rhs[c]; *pl(27 + c, s) = pv;
            a0s += (double)rhs[c] * (double)rhs[c]; a1s += (double)rhs[c] * (double)pv;
            if (A.b_out) A.b_out[3 * (long long)s + c] = rhs[c];
        }
    }
    {
        const double2 t = block_sum2<BS>(a0s, a1s, s_buf);
        if (threadIdx.x == 0) { P0[blockIdx.x] = t.x; P1[blockIdx.x] = t.y; }
    }
    grid_barrier(A.barrier, bar_target, nb);
    const double2 t01 = grid_total2_cg(P0, P1, nb, s_buf);
    const float rhsNorm2 = (float)t01.x;
    float absNew = (float)t01.y;
    const float tol = 1.1920929e-07f;
    float threshold = tol * tol * rhsNorm2;
    if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
    float residualNorm2 = rhsNorm2;
    int it = 0;
    if (rhsNorm2 != 0.f && !(residualNorm2 < threshold)) {
        while (it < A.max_iters) {
            // ---- phase 1: u = Jr p, one value per row (voxel, channel)
            for (int s = gtid; s < S; s += T) {
                int col[3];
#pragma unroll
                for (int a = 0; a < 3; ++a) col[a] = __ldg(A.nbr_row + a * S + s);
                const float pg = __ldcg(pl(28, s));
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    float u = __ldg(pl(c, s)) * (c == 2 ? pg : __ldcg(pl(27 + c, s)));
#pragma unroll
                    for (int a = 0; a < 3; ++a)
                        if (col[a] >= 0) u += __ldg(pl(3 * (a + 1) + c, s)) * __ldcg(pl(27 + c, col[a]));
                    *pl(33 + c, s) = u;
                }
            }
            grid_barrier(A.barrier, bar_target, nb);
            // ---- phase 2: y = D p + reg * Jr^T u + damping * diag * p ; p . y
            double pAp = 0.0;
            for (int s = gtid; s < S; s += T) {
                int tn[6];
#pragma unroll
                for (int h = 0; h < 6; ++h) tn[h] = __ldg(A.inc + h * S + s);
                float acc[3];
                acc[0] = (A.reg * __ldg(pl(0, s))) * __ldcg(pl(33, s));
                acc[1] = (A.reg * __ldg(pl(1, s))) * __ldcg(pl(34, s)) + (A.reg * __ldg(pl(2, s))) * __ldcg(pl(35, s));
                acc[2] = 0.f;
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        const int t = tn[h * 3 + a];
                        if (t < 0) continue;
#pragma unroll
                        for (int c = 0; c < 3; ++c) acc[c] += (A.reg * __ldg(pl(3 * (a + 1) + c, t))) * __ldcg(pl(33 + c, t));
                    }
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float ps = __ldcg(pl(27 + c, s));
                    const float y = A.Hd[3 * (long long)s + c] * ps + acc[c] + __ldg(pl(18 + c, s)) * ps;
                    *pl(30 + c, s) = y;
                    pAp += (double)ps * (double)y;
                }
            }
            pAp = block_sum<BS>(pAp, s_buf); if (threadIdx.x == 0) P0[blockIdx.x] = pAp;
            grid_barrier(A.barrier, bar_target, nb);
            const float alpha = absNew / (float)grid_total_cg(P0, nb, s_buf);
            // ---- phase 3 (own unknowns): x += alpha p ; r -= alpha y ; norms
            double rn = 0.0, rz = 0.0;
            for (int s = gtid; s < S; s += T) {
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const float ps = __ldcg(pl(27 + c, s));
                    *pl(21 + c, s) = __ldcg(pl(21 + c, s)) + alpha * ps;
                    const float r2 = __ldcg(pl(24 + c, s)) - alpha * __ldcg(pl(30 + c, s));
                    *pl(24 + c, s) = r2;
                    rn += (double)r2 * (double)r2;
                    rz += (double)r2 * (double)(__ldg(pl(15 + c, s)) * r2);
                }
            }
            {
                const double2 t = block_sum2<BS>(rn, rz, s_buf);
                if (threadIdx.x == 0) { P1[blockIdx.x] = t.x; P2[blockIdx.x] = t.y; }
            }
            grid_barrier(A.barrier, bar_target, nb);
            const double2 t12 = grid_total2_cg(P1, P2, nb, s_buf);
            residualNorm2 = (float)t12.x;
            if (residualNorm2 < threshold) break;
            const float absOld = absNew;
            absNew = (float)t12.y;
            const float beta = absNew / absOld;
            // ---- phase 4 (own unknowns): p <- z + beta p
            for (int s = gtid; s < S; s += T) {
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    *pl(27 + c, s) = __ldg(pl(15 + c, s)) * __ldcg(pl(24 + c, s)) + beta * __ldcg(pl(27 + c, s));
            }
            grid_barrier(A.barrier, bar_target, nb);
            ++it;
        }
    } else if (rhsNorm2 == 0.f) {
        residualNorm2 = 0.f;
    }
    if (gtid == 0) {
        A.result[0] = (float)it; A.result[1] = residualNorm2; A.result[2] = rhsNorm2; A.result[3] = threshold;
    }
    // ---- Optimizer::updateAlbedo (OptimizerAux.cpp:120-150): a channel is accepted only if it stays strictly inside (0, 1)
    const float err = rhsNorm2 != 0.f ? sqrtf(residualNorm2 / rhsNorm2) : 0.f;
    const bool apply = !A.only_on_success || err <= tol;
    int acc = 0;
    for (int s = gtid; s < S; s += T) {
        const int lin = A.lin[s];
        float* chan[3] = { A.vol.r, A.vol.g, A.vol.b };
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float d = __ldcg(pl(21 + c, s));
            if (A.delta_out) A.delta_out[3 * (long long)s + c] = d;
            if (!apply) continue;
            const float nv = chan[c][lin] - d;
            if ((double)nv > 0.0 && (double)nv < 1.0) { chan[c][lin] = nv; ++acc; }
        }
    }
    {
        // one atomic per block
        __shared__ int s_acc;
        if (threadIdx.x == 0) s_acc = 0;
        __syncthreads();
        if (acc) atomicAdd(&s_acc, acc);
        __syncthreads();
        if (threadIdx.x == 0 && s_acc) atomicAdd(A.accepted, (unsigned long long)s_acc);
    }
}

// Optimizer::updateDist (OptimizerAux.cpp:162-188): d -= delta only where |delta| < sqrt(3) * voxel_size (double compare)
// IterativeSolverBase: info = error <= tolerance ? Success : NoConvergence, error = sqrt(residualNorm2 / rhsNorm2)
__device__ __forceinline__ bool pcg_success(const float* __restrict__ verdict) {
    const float err = verdict[2] != 0.f ? sqrtf(verdict[1] / verdict[2]) : 0.f;
    return err <= 1.1920929e-07f;
}

__global__ void __launch_bounds__(256) k_dist_update(VolView vol, const int* __restrict__ lin, int S, const float* __restrict__ delta,
                                                    double thr, unsigned long long* accepted, const float* __restrict__ verdict,
                                                    int only_on_success) {
    if (only_on_success && !pcg_success(verdict)) return;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    int acc = 0;
    if (s < S) {
        const float d = delta[s];
        if (fabs((double)d) < thr) { vol.dist[lin[s]] -= d; acc = 1; }
    }
    const int n = __syncthreads_count(acc);       // one atomic per block: tens of thousands on one address serialise in L2
    if (threadIdx.x == 0 && n) atomicAdd(accepted, (unsigned long long)n);
}

// Optimizer::updateGrad (OptimizerAux.cpp:152-160): grad <- finite-difference gradient (not normalised); also refreshes
// the compact per-row records (own distance, new gradient, neighbour distances)
__global__ void __launch_bounds__(256) k_grad_update(VolView vol, SurfView surf, float vs_inv, const float* __restrict__ verdict,
                                                    int only_on_success) {
    if (only_on_success && !pcg_success(verdict)) return;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int S = surf.S;
    if (t >= S) return;
    const int s = __float_as_int(surf.rec2[t].w);
    const int lin = __float_as_int(surf.rec3[t].w);
    const float dist = vol.dist[lin];
    const int pb = __float_as_int(surf.rec1[t].w);
    float fd[3], nfd[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int nl = surf.nbr_lin[a * S + s];
        const float p = ((pb >> a) & 1) ? 1.f : -1.f;
        const float dn = nl >= 0 ? vol.dist[nl] : dist;
        fd[a] = (p * (dn - dist)) * vs_inv;
    }
    vol.gx[lin] = fd[0]; vol.gy[lin] = fd[1]; vol.gz[lin] = fd[2];
    normalized3(fd, nfd);   // the stored gradient now IS the finite-difference gradient, so g_hat = normalised fd
    surf.rec0[t] = make_float4(dist, nfd[0], nfd[1], nfd[2]);
    surf.rec2[t] = make_float4(fd[0], fd[1], fd[2], __int_as_float(s));
    surf.rec4[t] = make_float4(nfd[0], nfd[1], nfd[2], 0.f);
}

// launch geometry of the cooperative kernels, queried once per device (the queries cost microseconds each, the solve runs
// every outer iteration)
struct PcgGeometry { int sms = 0, per_sm_small = 0, per_sm_small512 = 0, per_sm_small1024 = 0, per_sm_4_1024 = 0, per_sm_1 = 0, per_sm_4 = 0,
                     per_sm_cg = 0, per_sm_cg512 = 0, per_sm_cg1024 = 0; };
static const PcgGeometry& pcg_geometry() {
    static PcgGeometry g[64];
    int dev = 0;
    PSG_CUDA(cudaGetDevice(&dev));
    PcgGeometry& e = g[dev & 63];
    if (e.sms == 0) {
        PSG_CUDA(cudaDeviceGetAttribute(&e.sms, cudaDevAttrMultiProcessorCount, dev));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&e.per_sm_small, k_dist_solve_small<kPcgBlock>, kPcgBlock, 0));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&e.per_sm_small512, k_dist_solve_small<512>, 512, 0));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&e.per_sm_small1024, k_dist_solve_small<1024>, 1024, 0));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&e.per_sm_cg, k_dist_solve_small_cg<kPcgBlock>, kPcgBlock, 0));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&e.per_sm_cg512, k_dist_solve_small_cg<512>, 512, 0));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&e.per_sm_cg1024, k_dist_solve_small_cg<1024>, 1024, 0));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&e.per_sm_1, k_dist_pcg<kPcgBlock, 1>, kPcgBlock, 0));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&e.per_sm_4, k_dist_pcg<kPcgBlock, 4>, kPcgBlock, 0));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&e.per_sm_4_1024, k_dist_pcg<1024, 4>, 1024, 0));
    }
    return e;
}

// vecs layout (floats, planes of S): 0..13 stencil blocks | 14 rhs | 15 invd | 16 ddiag | 17 x | 18 r | 19 p | 20 tmp | 21..24 U
// queues the assembly and the solve; the verdict lands in mail slot kSlotPcg
void launch_dist_solve(psg_optimizer* o, float damping) {
    const int S = o->S;
    o->red.alloc((size_t)kRedDoubles * kRedCopies + kMailDoubles);
    PSG_CUDA(cudaMemsetAsync(mail_dev(o) + 8 * kSlotPcg, 0, 16 * sizeof(double), o->stream));   // verdict + accepted counter
    if (S == 0) return;
    o->vecs.alloc((size_t)S * 25 + 8);
    float* base = o->vecs.p;
    PcgArgs A;
    A.S = S; A.blk = base; A.nbr_row = o->nbr_row.p; A.inc = o->inc.p;
    A.rhs = base + (size_t)14 * S; A.invd = base + (size_t)15 * S; A.ddiag = base + (size_t)16 * S;
    A.x = base + (size_t)17 * S; A.r = base + (size_t)18 * S; A.p = base + (size_t)19 * S; A.tmp = base + (size_t)20 * S;
    A.U = base + (size_t)21 * S;
    A.result = reinterpret_cast<float*>(mail_dev(o) + 8 * kSlotPcg);
    A.damping = damping;
    A.max_iters = 2 * S;

    constexpr int kClusterMax = 8, kClusterRows = 1 << 16;
    const bool use_cluster = o->pcg_mode == 1 && S <= kClusterRows;
    o->last_pcg_kernel = use_cluster ? 1 : 0;
    const PcgGeometry& geo = pcg_geometry();
    const int sms = geo.sms;
    // grid kernel geometry: one row per thread while that still leaves SMs idle, four rows per thread beyond
    const int rpt = (S <= sms * kPcgBlock * 2) ? 1 : 4;
    int grid = 0;
    // large surfaces: 1024-thread blocks, a quarter of the blocks at every grid barrier (mode 5 keeps 256 for comparison)
    const int bsz = (rpt == 4 && o->pcg_mode != 5 && geo.per_sm_4_1024 >= 1) ? 1024 : kPcgBlock;
    if (!use_cluster) {
        int per_sm = bsz == 1024 ? geo.per_sm_4_1024 : (rpt == 1 ? geo.per_sm_1 : geo.per_sm_4);
        if (per_sm < 1) { set_error("PCG kernel does not fit on an SM"); throw CudaError{-1}; }
        if (per_sm > 4) per_sm = 4;
        grid = sms * per_sm;
        const int need = (S + bsz * rpt - 1) / (bsz * rpt);
        if (grid > need) grid = need;
    }
    o->pcg_partials.alloc((size_t)3 * (use_cluster ? kClusterMax : grid) + 1);
    A.partials = o->pcg_partials.p + 1;
    A.barrier = reinterpret_cast<unsigned long long*>(o->pcg_partials.p);
    o->pcg_pbuf.alloc((size_t)2 * S);
    float* pbuf = o->pcg_pbuf.p;

    { Prof pf(o->stream, KID_PCG_ASSEMBLE); k_dist_assemble<<<(S + kPcgBlock - 1) / kPcgBlock, kPcgBlock, 0, o->stream>>>(A); }
    PSG_LAUNCH_CHECK();

    if (use_cluster) {
        // one thread-block cluster: as many CTAs as there is work for, at most the portable 8
        int nc = (S + kClusterBlock - 1) / kClusterBlock;
        if (nc > kClusterMax) nc = kClusterMax;
        const int threads = nc * kClusterBlock;
        const int crpt = (S + threads - 1) / threads;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(nc); cfg.blockDim = dim3(kClusterBlock); cfg.dynamicSmemBytes = 0; cfg.stream = o->stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = nc; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        Prof pf(o->stream, KID_PCG_SOLVE);
        if (crpt <= 1) PSG_CUDA(cudaLaunchKernelEx(&cfg, k_dist_pcg_cluster<1>, A, pbuf));
        else if (crpt <= 2) PSG_CUDA(cudaLaunchKernelEx(&cfg, k_dist_pcg_cluster<2>, A, pbuf));
        else if (crpt <= 4) PSG_CUDA(cudaLaunchKernelEx(&cfg, k_dist_pcg_cluster<4>, A, pbuf));
        else PSG_CUDA(cudaLaunchKernelEx(&cfg, k_dist_pcg_cluster<8>, A, pbuf));
    } else {
        // cooperative launch = the driver guarantees all blocks are co-resident, which the in-kernel barrier relies on
        void* args[] = { &A, &pbuf };
        Prof pf(o->stream, KID_PCG_SOLVE);
        if (rpt == 1) PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_pcg<kPcgBlock, 1>, dim3(grid), dim3(kPcgBlock), args, 0, o->stream));
        else if (bsz == 1024) PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_pcg<1024, 4>, dim3(grid), dim3(1024), args, 0, o->stream));
        else PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_pcg<kPcgBlock, 4>, dim3(grid), dim3(kPcgBlock), args, 0, o->stream));
    }
}

// Fused path for surfaces with at most one row per resident thread (128^3, 256^3): returns false if it does not apply.
// can a 16-CTA cluster of the DSMEM kernel be scheduled on this device? (queried once)
template <int RPT>
static bool cluster16_available() {
    static int state = -1;
    if (state < 0) {
        state = 0;
        const size_t smem = sizeof(ClSmem<RPT>);
        if (cudaFuncSetAttribute(k_dist_solve_cluster16<RPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) == cudaSuccess &&
            cudaFuncSetAttribute(k_dist_solve_cluster16<RPT>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(kClNC); cfg.blockDim = dim3(kClBS); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = kClNC; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, k_dist_solve_cluster16<RPT>, &cfg) == cudaSuccess && n >= 1) state = 1;
        }
        cudaGetLastError();   // a refusal is an answer, not an error of the caller
    }
    return state == 1;
}

template <int RPT>
static void launch_cluster16(psg_optimizer* o, SolveSmallArgs& G) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kClNC); cfg.blockDim = dim3(kClBS); cfg.dynamicSmemBytes = sizeof(ClSmem<RPT>); cfg.stream = o->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kClNC; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    Prof pf(o->stream, KID_PCG_SOLVE);
    PSG_CUDA(cudaLaunchKernelEx(&cfg, k_dist_solve_cluster16<RPT>, G));
}

bool launch_dist_block_small(psg_optimizer* o, float damping, bool only_on_success, int reg_slot) {
    const int S = o->S;
    if (S == 0 || (o->pcg_mode != 0 && o->pcg_mode != 3 && o->pcg_mode != 4 && o->pcg_mode != 6)) return false;
    const PcgGeometry& geo = pcg_geometry();
    // mode 6: up to 32 768 rows one 16-CTA cluster with the vectors in distributed shared memory
    const bool want_cluster = o->pcg_mode == 6 && S <= 2 * kClNC * kClBS;
    const bool cg = o->pcg_mode == 0;   // default: the fused kernel in the Chronopoulos-Gear arrangement (two barriers per iteration)
    const int cl_rpt = S <= kClNC * kClBS ? 1 : 2;
    const bool use_cluster = want_cluster && (cl_rpt == 1 ? cluster16_available<1>() : cluster16_available<2>());
    if (o->pcg_mode == 6 && !use_cluster) return false;
    // Block size: the solve is a chain of grid barriers (three per iteration), and a barrier costs more the more blocks take
    // part, while a row's arithmetic is negligible.  Measured (profiles/r1_summary.md section 8c): 128^3 (28 k rows) 125 / 114 /
    // 120 us with 256 / 512 / 1024 threads per block, 256^3 (113 k rows) 216 / 164 / 152 us.  So: the smallest block that
    // keeps the grid at or below 64 blocks, else the largest that fits.  Modes 3 / 4 force 512 / 1024.
    int bsz = kPcgBlock;
    if (o->pcg_mode == 3) bsz = 512;
    else if (o->pcg_mode == 4) bsz = 1024;
    else {
        while (bsz < 1024 && (S + bsz - 1) / bsz > 64) bsz *= 2;
    }
    const int sms = geo.sms;
    auto per_of = [&](int b) {
        if (cg) return b == 1024 ? geo.per_sm_cg1024 : b == 512 ? geo.per_sm_cg512 : geo.per_sm_cg;
        return b == 1024 ? geo.per_sm_small1024 : b == 512 ? geo.per_sm_small512 : geo.per_sm_small;
    };
    auto fits = [&](int b) { const int per = per_of(b); return per >= 1 && (S + b - 1) / b <= sms * per; };
    while (bsz > kPcgBlock && !fits(bsz)) bsz /= 2;
    const int per_sm = per_of(bsz);
    const int grid = (S + bsz - 1) / bsz;
    if (per_sm < 1 || grid > sms * per_sm) return false;

    o->red.alloc((size_t)kRedDoubles * kRedCopies + kMailDoubles);
    static_assert(kSlotReg + 1 == kSlotPcg && kSlotPcg + 1 == kSlotCount, "the three slots are cleared with one memset");
    PSG_CUDA(cudaMemsetAsync(mail_dev(o) + 8 * kSlotReg, 0, 24 * sizeof(double), o->stream));   // reg energies, verdict, counter + barrier
    o->vecs.alloc((size_t)S * 25 + 8);
    float* base = o->vecs.p;
    SolveSmallArgs G;
    PcgArgs& A = G.A;
    A.S = S; A.blk = base; A.nbr_row = o->nbr_row.p; A.inc = o->inc.p;
    A.rhs = base + (size_t)14 * S; A.invd = base + (size_t)15 * S; A.ddiag = base + (size_t)16 * S;
    A.x = base + (size_t)17 * S; A.r = base + (size_t)18 * S; A.p = base + (size_t)19 * S; A.tmp = base + (size_t)20 * S;
    A.U = base + (size_t)21 * S;
    A.result = reinterpret_cast<float*>(mail_dev(o) + 8 * kSlotPcg);
    A.damping = damping;
    A.max_iters = 2 * S;
    o->pcg_partials.alloc((size_t)4 * grid + 1);
    A.partials = o->pcg_partials.p + 1;
    A.barrier = reinterpret_cast<unsigned long long*>(mail_dev(o) + 8 * kSlotCount + 1);
    o->pcg_pbuf.alloc((size_t)2 * S);
    float* pbuf = o->pcg_pbuf.p;
    G.vol = o->vol->view();
    G.surf = surf_view(o);
    G.lin = o->lin.p;
    G.thr = sqrt(3.0) * (double)o->vs;
    G.vs_inv = o->vs_inv;
    G.accepted = reinterpret_cast<unsigned long long*>(mail_dev(o) + 8 * kSlotCount);
    G.only_on_success = only_on_success ? 1 : 0;
    G.reg_out = reg_slot >= 0 ? mail_dev(o) + 8 * reg_slot : nullptr;
    G.zbuf = A.r;   // plane 18 is free on this path (the residual lives in registers)
    if (use_cluster) {
        if (cl_rpt == 1) launch_cluster16<1>(o, G); else launch_cluster16<2>(o, G);
        o->last_pcg_kernel = 3;
        return true;
    }
    o->last_pcg_kernel = 2;
    void* args[] = { &G, &pbuf };
    Prof pf(o->stream, KID_PCG_SOLVE);
    if (cg) {
        o->last_pcg_kernel = 5;
        if (bsz == 1024) PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_solve_small_cg<1024>, dim3(grid), dim3(1024), args, 0, o->stream));
        else if (bsz == 512) PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_solve_small_cg<512>, dim3(grid), dim3(512), args, 0, o->stream));
        else PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_solve_small_cg<kPcgBlock>, dim3(grid), dim3(kPcgBlock), args, 0, o->stream));
        return true;
    }
    if (bsz == 1024) PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_solve_small<1024>, dim3(grid), dim3(1024), args, 0, o->stream));
    else if (bsz == 512) PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_solve_small<512>, dim3(grid), dim3(512), args, 0, o->stream));
    else PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_solve_small<kPcgBlock>, dim3(grid), dim3(kPcgBlock), args, 0, o->stream));
    return true;
}

void launch_dist_update(psg_optimizer* o, bool only_on_success) {
    const int S = o->S;
    if (S == 0) return;
    float* base = o->vecs.p;
    const float* delta = base + (size_t)17 * S;
    const float* verdict = reinterpret_cast<const float*>(mail_dev(o) + 8 * kSlotPcg);
    unsigned long long* cnt = reinterpret_cast<unsigned long long*>(mail_dev(o) + 8 * kSlotCount);
    const VolView vol = o->vol->view();
    const SurfView sv = surf_view(o);
    const double thr = sqrt(3.0) * (double)o->vs;
    const int gate = only_on_success ? 1 : 0;
    { Prof pf(o->stream, KID_DIST_UPDATE); k_dist_update<<<(S + 255) / 256, 256, 0, o->stream>>>(vol, o->lin.p, S, delta, thr, cnt, verdict, gate); }
    PSG_LAUNCH_CHECK();
    { Prof pf(o->stream, KID_GRAD_UPDATE); k_grad_update<<<(S + 255) / 256, 256, 0, o->stream>>>(vol, sv, o->vs_inv, verdict, gate); }
    PSG_LAUNCH_CHECK();
}

// ---- sharded job: the distance block after the sweep -------------------------------------------------------------------------
static void fill_mg_rank(psg_optimizer* o, MgRank* R) {
    memset(R, 0, sizeof(*R));
    const int world = o->sharded ? o->world : 1, rank = o->sharded ? o->rank : 0;
    R->rank = rank; R->lo = o->row_lo; R->hi = o->row_hi;
    R->halo_lo = o->halo_lo[rank]; R->halo_hi = o->halo_hi[rank];
    if (!o->sharded) { R->halo_lo = R->lo; R->halo_hi = R->hi; }
    R->nh = (R->lo - R->halo_lo) + (R->halo_hi - R->hi);
    int inner_lo = R->lo, inner_hi = R->hi, n = 0;
    for (int q = 0; q < world; ++q) {
        if (q == rank) continue;
        const int qlo = o->row_bounds[q], qhi = o->row_bounds[q + 1];
        const int qnh = (qlo - o->halo_lo[q]) + (o->halo_hi[q] - qhi);
        // rows of mine below rank q's range that q needs (q's lower halo), and rows of mine above it (q's upper halo)
        const int b0 = std::max(R->lo, o->halo_lo[q]), e0 = std::min(R->hi, qlo);
        if (e0 > b0) {
            R->seg[n] = MgSeg{q, b0, e0, -o->halo_lo[q], qnh}; ++n;
            if (q > rank) inner_hi = std::min(inner_hi, b0); else inner_lo = std::max(inner_lo, e0);
        }
        const int b1 = std::max(R->lo, qhi), e1 = std::min(R->hi, o->halo_hi[q]);
        if (e1 > b1) {
            R->seg[n] = MgSeg{q, b1, e1, (qlo - o->halo_lo[q]) - qhi, qnh}; ++n;
            if (q < rank) inner_lo = std::max(inner_lo, e1); else inner_hi = std::min(inner_hi, b1);
        }
    }
    R->nseg = n;
    if (inner_lo > inner_hi) { inner_lo = R->lo; inner_hi = R->lo; }
    R->inner_lo = inner_lo; R->inner_hi = inner_hi;
    R->vecs = o->vecs.p; R->pbuf = o->pcg_pbuf.p; R->xs = reinterpret_cast<XState*>(o->xstate.p); R->halo = o->halo.p;
    R->xp = o->xplanes.p;
    R->nbr_row = o->nbr_row.p; R->inc = o->inc.p; R->lin = o->lin.p;
    R->vol = o->vol->view(); R->surf = surf_view(o);
    R->partials = o->pcg_partials.p + 1;
    R->result = reinterpret_cast<float*>(mail_dev(o) + 8 * kSlotPcg);
    R->regslot = mail_dev(o) + 8 * kSlotReg;
}

struct MgLaunch { float damping; int only_on_success, want_reg; };

static int mg_blocks_cap() {
    static int cap = 0;
    if (!cap) {
        int per_sm = 0;
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_dist_block_mg<false>, kMgBlock, 0));
        if (per_sm < 2) { set_error("distributed distance solve does not fit on an SM"); throw CudaError{-1}; }
        cap = pcg_geometry().sms;   // one block per SM: the barriers cost more with more blocks than the rows gain
    }
    return cap;
}

static void launch_mg(psg_optimizer* on, const MgRank* ranks, int nplayed, const MgLaunch& L) {
    MgArgs A;
    memset(&A, 0, sizeof(A));
    A.S = on->S; A.world = on->sharded ? on->world : 1; A.nplayed = nplayed;
    for (int q = 0; q < A.world; ++q) A.peer[q] = on->peer[q];
    if (!on->sharded) { A.peer[0].halo = on->halo.p; A.peer[0].xs = reinterpret_cast<XState*>(on->xstate.p); }
    int need = 1;
    for (int i = 0; i < nplayed; ++i) { A.loc[i] = ranks[i]; need = std::max(need, (ranks[i].hi - ranks[i].lo + kMgBlock - 1) / kMgBlock); }
    // one row per thread if the rows fit one block per SM (state in registers); else the general variant, two blocks per SM
    const int cap1 = std::max(1, mg_blocks_cap() / nplayed);
    const bool one_row = need <= cap1;
    A.nb = one_row ? need : std::min(need, std::max(1, 2 * mg_blocks_cap() / nplayed));
    if (A.nb > 4 * kPcgMaxBlocks / 6) A.nb = 4 * kPcgMaxBlocks / 6;   // [2][3][nb] partial sums in 4 * kPcgMaxBlocks doubles
    A.damping = L.damping; A.vs_inv = on->vs_inv; A.thr = sqrt(3.0) * (double)on->vs;
    A.max_iters = 2 * on->S; A.only_on_success = L.only_on_success; A.want_reg = L.want_reg;
    A.flags = on->pcg_mode == 7 ? 1 : 0;
    A.timeout_ns = 10ull * 1000ull * 1000ull * 1000ull;
    trace("[rank %d] distributed solve: %d rank(s) played, %d blocks per rank, rows %d..%d, halo %d..%d (%d), %d segments", on->rank, nplayed, A.nb,
          ranks[0].lo, ranks[0].hi, ranks[0].halo_lo, ranks[0].halo_hi, ranks[0].nh, ranks[0].nseg);
    on->mg_args.alloc(sizeof(MgArgs));
    PSG_CUDA(cudaMemcpyAsync(on->mg_args.p, &A, sizeof(A), cudaMemcpyHostToDevice, on->stream));
    const MgArgs* dev = reinterpret_cast<const MgArgs*>(on->mg_args.p);
    void* args[] = { (void*)&dev };
    Prof pf(on->stream, KID_PCG_SOLVE);
    if (one_row) PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_block_mg<true>, dim3(A.nb * nplayed), dim3(kMgBlock), args, 0, on->stream));
    else PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_block_mg<false>, dim3(A.nb * nplayed), dim3(kMgBlock), args, 0, on->stream));
}

static void mg_leader(psg_optimizer* on, void* const* args, int world, void* user) {
    MgRank ranks[kMaxRanks];
    for (int q = 0; q < world; ++q) ranks[q] = *static_cast<const MgRank*>(args[q]);
    launch_mg(on, ranks, world, *static_cast<const MgLaunch*>(user));
}

void launch_dist_block_mg(psg_optimizer* o, float damping, bool only_on_success, bool want_reg) {
    // verdict (identical on every rank), then the slot whose entries are partial sums over the rank's rows
    PSG_CUDA(cudaMemsetAsync(mail_dev(o) + 8 * kSlotReg, 0, 24 * sizeof(double), o->stream));
    if (o->sharded) o->reduce_reg = true;
    if (o->S == 0) return;
    if (!o->sharded && !o->halo.p) {   // one GPU: no halo, but the kernel's argument block wants valid pointers
        o->halo.alloc(kHaloKinds);
        PSG_CUDA(cudaMemsetAsync(o->halo.p, 0, kHaloKinds * sizeof(unsigned long long), o->stream));
    }
    MgRank R;
    fill_mg_rank(o, &R);
    MgLaunch L; L.damping = damping; L.only_on_success = only_on_success ? 1 : 0; L.want_reg = want_reg ? 1 : 0;
    if (o->group && o->sharded) group_run(o, &R, mg_leader, &L);
    else launch_mg(o, &R, 1, L);
    o->last_pcg_kernel = 4;
    o->vol_unsynced = true;
}

// regularised albedo block: the sweep left b = J^T W r in vecs[0 .. 3S) and the diagonal of J^T W J in vecs[3S .. 6S)
void launch_albedo_reg_solve(psg_optimizer* o, float damping, bool only_on_success) {
    const int S = o->S;
    PSG_CUDA(cudaMemsetAsync(mail_dev(o) + 8 * kSlotAreg, 0, 8 * sizeof(double), o->stream));   // verdict, accepted count, barrier
    if (S == 0) return;
    o->areg.alloc((size_t)S * 39 + 8);
    o->vecs.alloc((size_t)S * 25 + 8);
    const PcgGeometry& geo = pcg_geometry();
    static int per_sm = 0;
    if (!per_sm) PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_albedo_reg_solve<kPcgBlock>, kPcgBlock, 0));
    if (per_sm < 1) { set_error("regularised albedo solve does not fit on an SM"); throw CudaError{-1}; }
    int grid = geo.sms * (per_sm > 2 ? 2 : per_sm);
    const int need = (S + kPcgBlock - 1) / kPcgBlock;
    if (grid > need) grid = need;
    if (grid > kPcgMaxBlocks) grid = kPcgMaxBlocks;
    o->pcg_partials.alloc((size_t)4 * kPcgMaxBlocks + 1);
    AregArgs A;
    A.S = S; A.vol = o->vol->view();
    A.lin = o->lin.p; A.nbr_lin = o->nbr_lin.p; A.nbr_row = o->nbr_row.p; A.inc = o->inc.p; A.pos = o->pos.p;
    A.bd = o->vecs.p; A.Hd = o->vecs.p + (size_t)3 * S;
    A.w = o->areg.p;
    A.b_out = o->vecs.p + (size_t)6 * S; A.delta_out = o->vecs.p + (size_t)9 * S;
    A.reg = o->set.reg_weight_rho; A.damping = damping; A.vs_inv = o->vs_inv;
    A.max_iters = 2 * 3 * S; A.only_on_success = only_on_success ? 1 : 0;
    A.partials = o->pcg_partials.p + 1;
    double* slot = mail_dev(o) + 8 * kSlotAreg;
    A.result = reinterpret_cast<float*>(slot);
    A.accepted = reinterpret_cast<unsigned long long*>(slot + 5);
    A.barrier = reinterpret_cast<unsigned long long*>(slot + 6);
    void* args[] = { &A };
    { Prof pf(o->stream, KID_PCG_SOLVE);
      PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_albedo_reg_solve<kPcgBlock>, dim3(grid), dim3(kPcgBlock), args, 0, o->stream)); }
    o->vol->generation++;   // the albedo planes changed under the slot records: re-gathered before the next sweep
}

void read_albedo_reg_verdict(const psg_optimizer* o, int* success, int* iters, float* err, long long* accepted) {
    *success = 1; *iters = 0; *err = 0.f; *accepted = 0;
    if (o->S == 0) return;
    const double* slot = mail_slot(o, kSlotAreg);
    const float* res = reinterpret_cast<const float*>(slot);
    *iters = (int)res[0];
    *err = res[2] != 0.f ? sqrtf(res[1] / res[2]) : 0.f;
    *success = (*err <= 1.1920929e-07f) ? 1 : 0;
    *accepted = (long long)*reinterpret_cast<const unsigned long long*>(slot + 5);
}

void read_dist_verdict(const psg_optimizer* o, int* success, int* iters, float* err, long long* accepted) {
    *success = 1; *iters = 0; *err = 0.f; *accepted = 0;
    if (o->S == 0) return;
    const float* res = reinterpret_cast<const float*>(mail_slot(o, kSlotPcg));
    *iters = (int)res[0];
    const float tol = 1.1920929e-07f;
    *err = res[2] != 0.f ? sqrtf(res[1] / res[2]) : 0.f;
    *success = (*err <= tol) ? 1 : 0;   // IterativeSolverBase: info = error <= tolerance ? Success : NoConvergence
    if (o->last_pcg_kernel == 4) {
        if (res[4] != 0.f) { set_error("the distance solve gave up waiting for a peer (10 s limit)"); throw CudaError{-1}; }
        *accepted = (long long)(mail_slot(o, kSlotReg)[3] + 0.5);   // summed over the ranks by the flush
    } else {
        *accepted = (long long)*reinterpret_cast<const unsigned long long*>(mail_slot(o, kSlotCount));
    }
}

}  // namespace psg
