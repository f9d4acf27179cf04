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
    float* result;           // [0] iterations, [1] residualNorm2, [2] rhsNorm2, [3] threshold
    float damping;
    int max_iters;
};

// slot pair (i >= k) -> index into the lower triangle stored in planes 0..9
__device__ __forceinline__ int tri(int i, int k) { return i * (i + 1) / 2 + k; }

__global__ void __launch_bounds__(kPcgBlock) k_dist_assemble(PcgArgs A) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    const int S = A.S;
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

// Large surfaces: cooperative launch over the whole GPU, grid.sync() between phases, vectors in global memory.
// Same three-barrier iteration as the cluster kernel below (the search-direction update is folded into the stencil
// phase: neighbours' new p is recomputed from the published r, 1/diag and previous p).  A thread walks its rows in
// groups of RPT and issues the loads of a whole group before using any of them, so a phase costs two dependent memory
// round trips per group instead of two per row.
template <int BS, int RPT>
__global__ void __launch_bounds__(BS) k_dist_pcg(PcgArgs A, float* __restrict__ pbuf /* 2 x S */) {
    cg::grid_group grid = cg::this_grid();
    __shared__ double s_buf[BS / 32];
    const int S = A.S;
    const int nb = gridDim.x;
    const int T = nb * BS;
    const int gtid = blockIdx.x * BS + threadIdx.x;
    double* P0 = A.partials;
    double* P1 = A.partials + nb;
    double* P2 = A.partials + 2 * nb;

    double a0 = 0.0, a1 = 0.0;
    for (int s = gtid; s < S; s += T) {
        const float b = A.rhs[s];
        const float pv = A.invd[s] * b;
        A.x[s] = 0.f; A.r[s] = b; pbuf[s] = pv;
        a0 += (double)b * (double)b;
        a1 += (double)b * (double)pv;
    }
    a0 = block_sum<BS>(a0, s_buf); if (threadIdx.x == 0) P0[blockIdx.x] = a0;
    a1 = block_sum<BS>(a1, s_buf); if (threadIdx.x == 0) P1[blockIdx.x] = a1;
    grid.sync();
    const float rhsNorm2 = (float)grid_total(P0, nb, s_buf);
    float absNew = (float)grid_total(P1, nb, s_buf);
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
                    for (int a = 0; a < 3; ++a) col[j][a] = in ? A.nbr_row[a * S + row[j]] : -1;
                    pj[j] = in ? (A.invd[row[j]] * A.r[row[j]]) * gamma + beta * pc[row[j]] : 0.f;
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
                        A.U[(long long)i * S + row[j]] = valid ? u : 0.f;
                    }
                }
            }
            grid.sync();
            // ---- phase 2: tmp_s = u_s[0] + incident u + damping * diag * p_s ; p . tmp
            double pAp = 0.0;
            for (int base = gtid; base < S; base += T * RPT) {
                int tn[RPT][6];
#pragma unroll
                for (int j = 0; j < RPT; ++j)
#pragma unroll
                    for (int h = 0; h < 6; ++h) tn[j][h] = (base + j * T < S) ? A.inc[h * S + base + j * T] : -1;
#pragma unroll
                for (int j = 0; j < RPT; ++j) {
                    const int s = base + j * T;
                    if (s >= S) continue;
                    float y = A.U[s];
#pragma unroll
                    for (int h = 0; h < 2; ++h)
#pragma unroll
                        for (int a = 0; a < 3; ++a) {
                            const int t = tn[j][h * 3 + a];
                            if (t >= 0) y += A.U[(long long)(a + 1) * S + t];
                        }
                    const float ps = pn[s];
                    y += A.ddiag[s] * ps;
                    A.tmp[s] = y;
                    pAp += (double)ps * (double)y;
                }
            }
            pAp = block_sum<BS>(pAp, s_buf); if (threadIdx.x == 0) P0[blockIdx.x] = pAp;
            grid.sync();
            const float alpha = absNew / (float)grid_total(P0, nb, s_buf);
            // ---- phase 3: x += alpha p ; residual -= alpha tmp ; norms
            double rn = 0.0, rz = 0.0;
            for (int base = gtid; base < S; base += T * RPT) {
                float xv[RPT], pv[RPT], rv[RPT], tv[RPT], iv[RPT];
#pragma unroll
                for (int j = 0; j < RPT; ++j) {
                    const int s = base + j * T;
                    const bool in = s < S;
                    xv[j] = in ? A.x[s] : 0.f; pv[j] = in ? pn[s] : 0.f; rv[j] = in ? A.r[s] : 0.f;
                    tv[j] = in ? A.tmp[s] : 0.f; iv[j] = in ? A.invd[s] : 0.f;
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
            rn = block_sum<BS>(rn, s_buf); if (threadIdx.x == 0) P1[blockIdx.x] = rn;
            rz = block_sum<BS>(rz, s_buf); if (threadIdx.x == 0) P2[blockIdx.x] = rz;
            grid.sync();
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
    if (gtid == 0) {
        A.result[0] = (float)it;
        A.result[1] = residualNorm2;
        A.result[2] = rhsNorm2;
        A.result[3] = threshold;
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

// Optimizer::updateDist (OptimizerAux.cpp:162-188): d -= delta only where |delta| < sqrt(3) * voxel_size (double compare)
__global__ void __launch_bounds__(256) k_dist_update(VolView vol, const int* __restrict__ lin, int S, const float* __restrict__ delta,
                                                    double thr, unsigned long long* accepted) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    int acc = 0;
    if (s < S) {
        const float d = delta[s];
        if (fabs((double)d) < thr) { vol.dist[lin[s]] -= d; acc = 1; }
    }
    const unsigned m = __ballot_sync(0xffffffffu, acc);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(accepted, (unsigned long long)__popc(m));
}

// Optimizer::updateGrad (OptimizerAux.cpp:152-160): grad <- finite-difference gradient (not normalised); also refreshes
// the compact per-row records (own distance, new gradient, neighbour distances)
__global__ void __launch_bounds__(256) k_grad_update(VolView vol, SurfView surf, float vs_inv) {
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

// vecs layout (floats, planes of S): 0..13 stencil blocks | 14 rhs | 15 invd | 16 ddiag | 17 x | 18 r | 19 p | 20 tmp | 21..24 U
void solve_dist(psg_optimizer* o, float damping, int* success, int* iters, float* err) {
    const int S = o->S;
    *success = 1; *iters = 0; *err = 0.f;
    if (S == 0) return;
    o->vecs.alloc((size_t)S * 25 + 8);
    float* base = o->vecs.p;
    PcgArgs A;
    A.S = S; A.blk = base; A.nbr_row = o->nbr_row.p; A.inc = o->inc.p;
    A.rhs = base + (size_t)14 * S; A.invd = base + (size_t)15 * S; A.ddiag = base + (size_t)16 * S;
    A.x = base + (size_t)17 * S; A.r = base + (size_t)18 * S; A.p = base + (size_t)19 * S; A.tmp = base + (size_t)20 * S;
    A.U = base + (size_t)21 * S;
    A.result = base + (size_t)25 * S;
    A.damping = damping;
    A.max_iters = 2 * S;

    { Prof pf(o->stream, KID_PCG_ASSEMBLE); k_dist_assemble<<<(S + kPcgBlock - 1) / kPcgBlock, kPcgBlock, 0, o->stream>>>(A); }
    PSG_LAUNCH_CHECK();

    constexpr int kClusterMax = 8, kClusterRows = 1 << 16;
    if (S <= kClusterRows && o->pcg_mode != 2) {
        // one thread-block cluster: as many CTAs as there is work for, at most the portable 8
        int nc = (S + kClusterBlock - 1) / kClusterBlock;
        if (nc > kClusterMax) nc = kClusterMax;
        const int threads = nc * kClusterBlock;
        const int rpt = (S + threads - 1) / threads;
        o->pcg_partials.alloc((size_t)3 * nc);
        A.partials = o->pcg_partials.p;
        o->pcg_pbuf.alloc((size_t)2 * S);
        float* pbuf = o->pcg_pbuf.p;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(nc); cfg.blockDim = dim3(kClusterBlock); cfg.dynamicSmemBytes = 0; cfg.stream = o->stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = nc; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        Prof pf(o->stream, KID_PCG_SOLVE);
        if (rpt <= 1) PSG_CUDA(cudaLaunchKernelEx(&cfg, k_dist_pcg_cluster<1>, A, pbuf));
        else if (rpt <= 2) PSG_CUDA(cudaLaunchKernelEx(&cfg, k_dist_pcg_cluster<2>, A, pbuf));
        else if (rpt <= 4) PSG_CUDA(cudaLaunchKernelEx(&cfg, k_dist_pcg_cluster<4>, A, pbuf));
        else PSG_CUDA(cudaLaunchKernelEx(&cfg, k_dist_pcg_cluster<8>, A, pbuf));
    } else {
        int dev = 0, sms = 0, per_sm = 0;
        PSG_CUDA(cudaGetDevice(&dev));
        PSG_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_dist_pcg<kPcgBlock, 4>, kPcgBlock, 0));
        if (per_sm < 1) { set_error("PCG kernel does not fit on an SM"); throw CudaError{-1}; }
        if (per_sm > 4) per_sm = 4;
        int grid = sms * per_sm;
        const int need = (S + kPcgBlock * 4 - 1) / (kPcgBlock * 4);
        if (grid > need) grid = need;
        o->pcg_partials.alloc((size_t)3 * grid);
        A.partials = o->pcg_partials.p;
        o->pcg_pbuf.alloc((size_t)2 * S);
        float* pbuf = o->pcg_pbuf.p;
        void* args[] = { &A, &pbuf };
        { Prof pf(o->stream, KID_PCG_SOLVE);
          PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_pcg<kPcgBlock, 4>, dim3(grid), dim3(kPcgBlock), args, 0, o->stream)); }
    }
    o->ensure_host_staging();
    float* res = reinterpret_cast<float*>(o->host_staging);   // pinned: a pageable destination makes the copy synchronous and slow
    PSG_CUDA(cudaMemcpyAsync(res, A.result, 4 * sizeof(float), cudaMemcpyDeviceToHost, o->stream));
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    *iters = (int)res[0];
    const float tol = 1.1920929e-07f;
    *err = res[2] != 0.f ? sqrtf(res[1] / res[2]) : 0.f;
    *success = (*err <= tol) ? 1 : 0;   // IterativeSolverBase: info = error <= tolerance ? Success : NoConvergence
}

void apply_dist_update(psg_optimizer* o, long long* accepted) {
    const int S = o->S;
    *accepted = 0;
    if (S == 0) return;
    float* base = o->vecs.p;
    const float* delta = base + (size_t)17 * S;
    o->red.alloc(kMaxFrames * 64 + 8);
    unsigned long long* cnt = reinterpret_cast<unsigned long long*>(o->red.p);
    PSG_CUDA(cudaMemsetAsync(cnt, 0, sizeof(unsigned long long), o->stream));
    const VolView vol = o->vol->view();
    const SurfView sv = surf_view(o);
    const double thr = sqrt(3.0) * (double)o->vs;
    { Prof pf(o->stream, KID_DIST_UPDATE); k_dist_update<<<(S + 255) / 256, 256, 0, o->stream>>>(vol, o->lin.p, S, delta, thr, cnt); }
    PSG_LAUNCH_CHECK();
    { Prof pf(o->stream, KID_GRAD_UPDATE); k_grad_update<<<(S + 255) / 256, 256, 0, o->stream>>>(vol, sv, o->vs_inv); }
    PSG_LAUNCH_CHECK();
    o->ensure_host_staging();
    unsigned long long* h = reinterpret_cast<unsigned long long*>(o->host_staging);
    PSG_CUDA(cudaMemcpyAsync(h, cnt, sizeof(*h), cudaMemcpyDeviceToHost, o->stream));
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    *accepted = (long long)*h;
}

}  // namespace psg
