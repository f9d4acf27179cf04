// TEST INFRASTRUCTURE ONLY -- oracle build stand-in for cpp/include/normals/NormalEstimator.h.
//
// The reference's estimator is OpenCV box-filter arithmetic (un-vendored, version unpinned), which
// is outside the hot path.  The parity contract is stated on identical float
// inputs, so this stand-in simply hands the fusion loop the normal maps the test supplies.
#ifndef PSG_ORACLE_NORMAL_ESTIMATOR_SHIM_H
#define PSG_ORACLE_NORMAL_ESTIMATOR_SHIM_H
#include <opencv2/core/core.hpp>
namespace cv {
template <typename T>
class NormalEstimator {
    Mat nx_, ny_, nz_, n_sq_inv_;
public:
    NormalEstimator() {}
    void set_maps(const Mat& nx, const Mat& ny, const Mat& nz, const Mat& n_sq_inv) {
        nx_ = nx; ny_ = ny; nz_ = nz; n_sq_inv_ = n_sq_inv;
    }
    void compute(const Mat&, Mat& nx, Mat& ny, Mat& nz) const { nx = nx_; ny = ny_; nz = nz_; }
    Mat* n_sq_inv_ptr() { return &n_sq_inv_; }
};
}  // namespace cv
#endif
