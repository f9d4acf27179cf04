// Host-side small solves of the light and pose blocks.  Plain C++ (compiled by the host compiler, not through nvcc's front
// end, which drops the unroll hints the 6 x 6 and 9 x 9 block products need to stay in registers).
#include <cmath>
#include <cstddef>
#include <vector>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include "host_solve.h"

namespace psg {

// sum_i a[i] * b[i] in the order Eigen's vectorised reduction adds them on the reference's x86-64 build (SSE2, 4 floats per
// packet, no FMA): two packet accumulators over strides of 8, their sum, one more packet if the length allows, the horizontal sum
// (l0 + l2) + (l1 + l3), then the tail element by element (third/eigen/Eigen/src/Core/Redux.h:236-281, arch/SSE/PacketMath.h
// predux).  Besides matching the reference's rounding it keeps eight independent accumulation chains in flight.
#if defined(__SSE2__)
// the same packet operations Eigen issues (pmul, padd, predux of arch/SSE/PacketMath.h)
static inline float eigen_dot(const float* a, const float* b, int n) {
    if (n < 4) {
        float r = a[0] * b[0];
        for (int i = 1; i < n; ++i) r += a[i] * b[i];
        return r;
    }
    const int aligned = n & ~3, aligned2 = n & ~7;
    __m128 p0 = _mm_mul_ps(_mm_loadu_ps(a), _mm_loadu_ps(b));
    if (aligned > 4) {
        __m128 p1 = _mm_mul_ps(_mm_loadu_ps(a + 4), _mm_loadu_ps(b + 4));
        for (int i = 8; i < aligned2; i += 8) {
            p0 = _mm_add_ps(p0, _mm_mul_ps(_mm_loadu_ps(a + i), _mm_loadu_ps(b + i)));
            p1 = _mm_add_ps(p1, _mm_mul_ps(_mm_loadu_ps(a + i + 4), _mm_loadu_ps(b + i + 4)));
        }
        p0 = _mm_add_ps(p0, p1);
        if (aligned > aligned2) p0 = _mm_add_ps(p0, _mm_mul_ps(_mm_loadu_ps(a + aligned2), _mm_loadu_ps(b + aligned2)));
    }
    const __m128 t = _mm_add_ps(p0, _mm_movehl_ps(p0, p0));
    float r = _mm_cvtss_f32(_mm_add_ss(t, _mm_shuffle_ps(t, t, 1)));
    for (int i = aligned; i < n; ++i) r += a[i] * b[i];
    return r;
}
#else
static inline float eigen_dot(const float* a, const float* b, int n) {
    if (n < 4) {
        float r = a[0] * b[0];
        for (int i = 1; i < n; ++i) r += a[i] * b[i];
        return r;
    }
    const int aligned = n & ~3, aligned2 = n & ~7;
    float p0[4], p1[4];
    for (int l = 0; l < 4; ++l) p0[l] = a[l] * b[l];
    if (aligned > 4) {
        for (int l = 0; l < 4; ++l) p1[l] = a[4 + l] * b[4 + l];
        for (int i = 8; i < aligned2; i += 8) {
            for (int l = 0; l < 4; ++l) p0[l] += a[i + l] * b[i + l];
            for (int l = 0; l < 4; ++l) p1[l] += a[i + 4 + l] * b[i + 4 + l];
        }
        for (int l = 0; l < 4; ++l) p0[l] += p1[l];
        if (aligned > aligned2) for (int l = 0; l < 4; ++l) p0[l] += a[aligned2 + l] * b[aligned2 + l];
    }
    float r = (p0[0] + p0[2]) + (p0[1] + p0[3]);
    for (int i = aligned; i < n; ++i) r += a[i] * b[i];
    return r;
}
#endif

// tmp = H.selfadjointView<Lower>() * p for block-diagonal H: Eigen walks column j, adds the diagonal term to res[j], scatters
// l * p[j] into res[i] for the rows i > j below it while gathering res_j += l * p[i], and adds res_j to res[j] last
// (SparseSelfAdjointView.h:299-331, column-major lower view).  BS > 0 keeps a block's accumulators in registers.
template <int BS>
static inline void selfadjoint_lower_times(int nblocks, const float* __restrict__ H, const float* __restrict__ p, float* __restrict__ out,
                                           int bs_runtime = BS) {
    const int bs = BS > 0 ? BS : bs_runtime;
    for (int blk = 0; blk < nblocks; ++blk) {
        const float* __restrict__ Hb = H + (size_t)blk * bs * bs;
        const float* __restrict__ pp = p + blk * bs;
        float t[BS > 0 ? BS : 16];
        float* tt = BS > 0 ? t : out + blk * bs;
#pragma GCC unroll 16
        for (int i = 0; i < bs; ++i) tt[i] = 0.f;
#pragma GCC unroll 16
        for (int j = 0; j < bs; ++j) {
            tt[j] += Hb[j * bs + j] * pp[j];
            const float rhs_j = pp[j];
            float res_j = 0.f;
#pragma GCC unroll 16
            for (int i = j + 1; i < bs; ++i) {
                const float l = Hb[i * bs + j];
                res_j += l * pp[i];
                tt[i] += l * rhs_j;
            }
            tt[j] += res_j;
        }
        if (BS > 0) {
#pragma GCC unroll 16
            for (int i = 0; i < bs; ++i) out[blk * bs + i] = t[i];
        }
    }
}

// Eigen::ConjugateGradient<SparseMatrix<float>, Lower, DiagonalPreconditioner<float>> on a small dense symmetric system
// (the light and pose blocks: F independent 4x4 / 9x9 / 6x6 blocks solved jointly, exactly as the reference does,
// PsOptimizer.cpp:184-191, 217-227), statement by statement (third/eigen/Eigen/src/IterativeLinearSolvers/
// ConjugateGradient.h:29-98): the product with the lower selfadjoint view walks the columns and scatters
// (SparseCore/SparseSelfAdjointView.h:299-331), the inner products follow eigen_dot.  Only the lower triangle is read.
// Returns Eigen's verdict.
CgVerdict eigen_like_cg(int nblocks, int bs, const float* H /* nblocks * bs * bs */, const float* b, float* x) {
    const int n = nblocks * bs;
    thread_local static std::vector<float> scratch;
    if (scratch.size() < (size_t)5 * n) scratch.resize((size_t)5 * n);
    float* res = scratch.data(); float* p = res + n; float* z = p + n; float* tmp = z + n; float* invd = tmp + n;
    for (int blk = 0, i = 0; blk < nblocks; ++blk)
        for (int k = 0; k < bs; ++k, ++i) {
            x[i] = 0.f;
            res[i] = b[i];
            const float d = H[(size_t)blk * bs * bs + k * bs + k];
            invd[i] = d != 0.f ? 1.f / d : 1.f;       // DiagonalPreconditioner::factorize (BasicPreconditioners.h:67-80)
        }
    const float tol = 1.1920929e-07f;
    const int maxIters = 2 * n;
    const float rhsNorm2 = eigen_dot(b, b, n);
    if (rhsNorm2 == 0.f) return {1, 0, 0.f};
    float threshold = tol * tol * rhsNorm2;
    if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
    float residualNorm2 = eigen_dot(res, res, n);
    if (residualNorm2 < threshold) return {1, 0, std::sqrt(residualNorm2 / rhsNorm2)};
    for (int i = 0; i < n; ++i) p[i] = invd[i] * res[i];
    float absNew = eigen_dot(res, p, n);
    int it = 0;
    while (it < maxIters) {
        // tmp = H.selfadjointView<Lower>() * p, column by column
        switch (bs) {
            case 3: selfadjoint_lower_times<3>(nblocks, H, p, tmp); break;
            case 4: selfadjoint_lower_times<4>(nblocks, H, p, tmp); break;
            case 6: selfadjoint_lower_times<6>(nblocks, H, p, tmp); break;
            case 9: selfadjoint_lower_times<9>(nblocks, H, p, tmp); break;
            default: selfadjoint_lower_times<0>(nblocks, H, p, tmp, bs); break;
        }
        const float alpha = absNew / eigen_dot(p, tmp, n);
        for (int i = 0; i < n; ++i) { x[i] += alpha * p[i]; res[i] -= alpha * tmp[i]; }
        residualNorm2 = eigen_dot(res, res, n);
        if (residualNorm2 < threshold) break;
        for (int i = 0; i < n; ++i) z[i] = invd[i] * res[i];
        const float absOld = absNew;
        absNew = eigen_dot(res, z, n);
        const float beta = absNew / absOld;
        for (int i = 0; i < n; ++i) p[i] = z[i] + beta * p[i];
        ++it;
    }
    const float err = std::sqrt(residualNorm2 / rhsNorm2);
    return {err <= tol ? 1 : 0, it, err};
}

}  // namespace psg
