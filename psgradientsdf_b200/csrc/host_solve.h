// Host-side small solves (host_solve.cpp).
#pragma once

namespace psg {

struct CgVerdict { int success; int iters; float err; };

// Eigen::ConjugateGradient<SparseMatrix<float>, Lower, DiagonalPreconditioner<float>> on nblocks dense bs x bs blocks (row major,
// lower triangle read), operation by operation as the reference's build executes it: same iterates, iteration count and verdict.
CgVerdict eigen_like_cg(int nblocks, int bs, const float* H, const float* b, float* x);

}  // namespace psg
