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

__device__ __forceinline__ double block_sum(double v, double* s_buf) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) s_buf[warp] = v;
    __syncthreads();
    double t = 0.0;
    if (warp == 0) {
        t = lane < (kPcgBlock / 32) ? s_buf[lane] : 0.0;
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

__global__ void __launch_bounds__(kPcgBlock) k_dist_pcg(PcgArgs A) {
    cg::grid_group grid = cg::this_grid();
    __shared__ double s_buf[kPcgBlock / 32];
    const int S = A.S;
    const int nb = gridDim.x;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
    const int gstride = nb * blockDim.x;
    double* P0 = A.partials;
    double* P1 = A.partials + nb;
    double* P2 = A.partials + 2 * nb;

    // ---- init: x = 0, residual = rhs, p = precond.solve(residual)
    double a0 = 0.0, a1 = 0.0;
    for (int s = gtid; s < S; s += gstride) {
        const float b = A.rhs[s];
        const float pv = A.invd[s] * b;
        A.x[s] = 0.f; A.r[s] = b; A.p[s] = pv;
        a0 += (double)b * (double)b;
        a1 += (double)b * (double)pv;
    }
    a0 = block_sum(a0, s_buf); if (threadIdx.x == 0) P0[blockIdx.x] = a0;
    a1 = block_sum(a1, s_buf); if (threadIdx.x == 0) P1[blockIdx.x] = a1;
    grid.sync();
    const float rhsNorm2 = (float)grid_total(P0, nb, s_buf);
    float absNew = (float)grid_total(P1, nb, s_buf);
    const float tol = 1.1920929e-07f;   // NumTraits<float>::epsilon(), the default tolerance
    float threshold = tol * tol * rhsNorm2;
    if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
    float residualNorm2 = rhsNorm2;
    int it = 0;
    if (rhsNorm2 != 0.f && !(residualNorm2 < threshold)) {
        while (it < A.max_iters) {
            // ---- phase 1: u_t = A_t * p[stencil(t)]
            for (int t = gtid; t < S; t += gstride) {
                int col[4]; float px[4];
                col[0] = t; px[0] = A.p[t];
#pragma unroll
                for (int a = 0; a < 3; ++a) { col[a + 1] = A.nbr_row[a * S + t]; px[a + 1] = col[a + 1] >= 0 ? A.p[col[a + 1]] : 0.f; }
                float Ab[10];
#pragma unroll
                for (int i = 0; i < 10; ++i) Ab[i] = A.blk[(long long)i * S + t];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float u = 0.f;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float aik = (i >= k) ? Ab[i * (i + 1) / 2 + k] : Ab[k * (k + 1) / 2 + i];
                        if (col[k] >= 0) u += aik * px[k];
                    }
                    A.U[(long long)i * S + t] = (col[i] >= 0) ? u : 0.f;
                }
            }
            grid.sync();
            // ---- phase 2: tmp_s = sum of incident u + damping * diag * p_s ; p . tmp
            double pAp = 0.0;
            for (int s = gtid; s < S; s += gstride) {
                float y = A.U[s];
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int a = 0; a < 3; ++a) {
                        const int t = A.inc[(h * 3 + a) * S + s];
                        if (t >= 0) y += A.U[(long long)(a + 1) * S + t];
                    }
                const float ps = A.p[s];
                y += A.ddiag[s] * ps;
                A.tmp[s] = y;
                pAp += (double)ps * (double)y;
            }
            pAp = block_sum(pAp, s_buf); if (threadIdx.x == 0) P0[blockIdx.x] = pAp;
            grid.sync();
            const float alpha = absNew / (float)grid_total(P0, nb, s_buf);
            // ---- phase 3: x += alpha p ; residual -= alpha tmp ; norms
            double rn = 0.0, rz = 0.0;
            for (int s = gtid; s < S; s += gstride) {
                A.x[s] += alpha * A.p[s];
                const float rv = A.r[s] - alpha * A.tmp[s];
                A.r[s] = rv;
                rn += (double)rv * (double)rv;
                rz += (double)rv * (double)(A.invd[s] * rv);
            }
            rn = block_sum(rn, s_buf); if (threadIdx.x == 0) P1[blockIdx.x] = rn;
            rz = block_sum(rz, s_buf); if (threadIdx.x == 0) P2[blockIdx.x] = rz;
            grid.sync();
            residualNorm2 = (float)grid_total(P1, nb, s_buf);
            if (residualNorm2 < threshold) break;
            const float absOld = absNew;
            absNew = (float)grid_total(P2, nb, s_buf);
            const float beta = absNew / absOld;
            // ---- phase 4: p = z + beta p
            for (int s = gtid; s < S; s += gstride) A.p[s] = A.invd[s] * A.r[s] + beta * A.p[s];
            ++it;
            grid.sync();
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

// Optimizer::updateGrad (OptimizerAux.cpp:152-160): grad <- finite-difference gradient (not normalised)
__global__ void __launch_bounds__(256) k_grad_update(VolView vol, SurfView surf, float vs_inv) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    const int S = surf.S;
    if (s >= S) return;
    const int lin = surf.lin[s];
    const float dist = vol.dist[lin];
    const unsigned char pb = surf.pos[s];
    float fd[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int nl = surf.nbr_lin[a * S + s];
        const float p = ((pb >> a) & 1) ? 1.f : -1.f;
        const float dn = nl >= 0 ? vol.dist[nl] : dist;
        fd[a] = (p * (dn - dist)) * vs_inv;
    }
    vol.gx[lin] = fd[0]; vol.gy[lin] = fd[1]; vol.gz[lin] = fd[2];
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

    int dev = 0, sms = 0, per_sm = 0;
    PSG_CUDA(cudaGetDevice(&dev));
    PSG_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    PSG_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_dist_pcg, kPcgBlock, 0));
    if (per_sm < 1) { set_error("PCG kernel does not fit on an SM"); throw CudaError{-1}; }
    if (per_sm > 2) per_sm = 2;   // enough to hide latency; fewer blocks = cheaper grid.sync
    int grid = sms * per_sm;
    const int need = (S + kPcgBlock - 1) / kPcgBlock;
    if (grid > need) grid = need;
    o->pcg_partials.alloc((size_t)3 * grid);
    A.partials = o->pcg_partials.p;
    void* args[] = { &A };
    { Prof pf(o->stream, KID_PCG_SOLVE);
      PSG_CUDA(cudaLaunchCooperativeKernel((void*)k_dist_pcg, dim3(grid), dim3(kPcgBlock), args, 0, o->stream)); }
    float res[4];
    PSG_CUDA(cudaMemcpyAsync(res, A.result, sizeof(res), cudaMemcpyDeviceToHost, o->stream));
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
    SurfView sv;
    sv.S = S; sv.lin = o->lin.p; sv.vis = o->svis.p; sv.nbr_lin = o->nbr_lin.p; sv.nbr_row = o->nbr_row.p; sv.pos = o->pos.p;
    const double thr = sqrt(3.0) * (double)o->vs;
    { Prof pf(o->stream, KID_DIST_UPDATE); k_dist_update<<<(S + 255) / 256, 256, 0, o->stream>>>(vol, o->lin.p, S, delta, thr, cnt); }
    PSG_LAUNCH_CHECK();
    { Prof pf(o->stream, KID_GRAD_UPDATE); k_grad_update<<<(S + 255) / 256, 256, 0, o->stream>>>(vol, sv, o->vs_inv); }
    PSG_LAUNCH_CHECK();
    unsigned long long h = 0;
    PSG_CUDA(cudaMemcpyAsync(&h, cnt, sizeof(h), cudaMemcpyDeviceToHost, o->stream));
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    *accepted = (long long)h;
}

}  // namespace psg
