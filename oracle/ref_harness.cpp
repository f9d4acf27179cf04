// TEST INFRASTRUCTURE ONLY.  Never linked into, loaded by, or called from the product library.
//
// C-callable harness around the UNMODIFIED reference sources (compiled where they lie under
// /root/reference/cpp by oracle/Makefile into oracle/_ref/libpsgref.so).  It exposes the state the
// reference keeps private (volume, visibility vectors, surface list, residuals, per-block normal
// equations, CG results) so that tests can compare the CUDA path and the C restatement
// (oracle/psg_oracle.c) against the reference itself.
//
// What is the reference's own code here and what is harness glue:
//   * fusion                 -> VolumetricGradSdf::update            (VolumetricGradSdf.cpp:51-138), unmodified
//   * select_vis / surface   -> Optimizer::select_vis/getSurfaceVoxel (Optimizer.cpp:30, OptimizerAux.cpp:237)
//   * residual / Jacobians   -> {Ps,Led}Optimizer::computeResidual(), *Jacobian()
//   * block solves           -> ref_opt_block_reference() calls optimize*All() directly;
//                               ref_opt_block() re-plays the same dozen Eigen statements
//                               (PsOptimizer.cpp:85-234, LedOptimizer.cpp:134-275) so that H, b and delta
//                               can be handed out; a CPU test checks both leave bit-identical state.
//   * outer loop             -> {Ps,Led}Optimizer::alternatingOptimize, unmodified
// Normals are injected (oracle/shim/normals/NormalEstimator.h): the parity contract is on identical
// float inputs.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include <map>

// Pull in every third-party header first so that the keyword overrides below only touch the
// reference's own class definitions (their include guards make the later includes no-ops).
#include "mat.h"   // first: it sets EIGEN_DONT_ALIGN_STATICALLY before Eigen is seen (mat.h:4-6)
#include <Eigen/Dense>
#include <Eigen/Sparse>
#include <Eigen/StdVector>
#include <Eigen/IterativeLinearSolvers>
#include <sophus/se3.hpp>
#include <nlohmann/json.hpp>
#include <opencv2/core/core.hpp>
#include "normals/NormalEstimator.h"

#define private public
#define protected public
#define class struct
#include "sdf_tracker/VolumetricGradSdf.h"
#include "ps_optimizer/Optimizer.h"
#include "ps_optimizer/PsOptimizer.h"
#include "ps_optimizer/LedOptimizer.h"
#include "sdf_tracker/RigidPointOptimizer.h"
#include "mesh/MarchingCubes.h"
#undef class
#undef private
#undef protected
#include "ps_optimizer/Auxilary.h"   // interpolateImage, computeImageGradient (static free functions)

namespace {

struct RefVol {
    VolumetricGradSdf* sdf = nullptr;
    cv::NormalEstimator<float> nest;
};

struct RefOpt {
    RefVol* vol = nullptr;
    OptimizerSettings* set = nullptr;
    Optimizer* opt = nullptr;
    PsOptimizer* ps = nullptr;
    LedOptimizer* led = nullptr;
    std::vector<std::shared_ptr<cv::Mat>> images;
    // last block solve
    Eigen::SparseMatrix<float> H;
    Eigen::VectorXf b, delta;
    int info = -1, iters = 0;
    float err = 0.f;
    std::streambuf* saved_cout = nullptr;
};

std::ostringstream g_sink;
bool g_quiet = true;
struct Quiet {
    std::streambuf* old = nullptr;
    Quiet() { if (g_quiet) { g_sink.str(""); old = std::cout.rdbuf(g_sink.rdbuf()); } }
    ~Quiet() { if (old) std::cout.rdbuf(old); }
};

Mat3f K_from(const float* K9) {
    Mat3f K;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) K(r, c) = K9[3 * r + c];
    return K;
}
Mat4f M4_from(const float* p16) {
    Mat4f M;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) M(r, c) = p16[4 * r + c];
    return M;
}

}  // namespace

extern "C" {

void ref_set_quiet(int q) { g_quiet = q != 0; }

// ---------------------------------------------------------------- volume
void* ref_vol_create(const int* dims, float voxel_size, const float* shift, float T, float zmin, float zmax) {
    Quiet q;
    RefVol* v = new RefVol;
    v->sdf = new VolumetricGradSdf(Vec3i(dims[0], dims[1], dims[2]), voxel_size,
                                   Vec3f(shift[0], shift[1], shift[2]), T);
    v->sdf->set_zmin(zmin);
    v->sdf->set_zmax(zmax);
    return v;
}

void ref_vol_destroy(void* h) {
    RefVol* v = static_cast<RefVol*>(h);
    delete v->sdf;
    delete v;
}

// What main_ps.cpp's SE3(poses[i]) hands to update(): rotation goes matrix -> unit quaternion -> matrix.
void ref_se3_from_matrix(const float* pose16, float* R9, float* t3) {
    SE3 pose(M4_from(pose16));
    const Mat3f R = pose.rotationMatrix();
    const Vec3f t = pose.translation();
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) R9[3 * r + c] = R(r, c);
    for (int r = 0; r < 3; ++r) t3[r] = t[r];
}

// main_ps.cpp:207,218-219: (increase_counter();) update(color, depth, K, SE3(pose), NEst)
void ref_vol_integrate(void* h, const float* color_bgr, const float* depth, const float* nx, const float* ny,
                       const float* nz, const float* n_sq_inv, int H, int W, const float* K9,
                       const float* pose16, int increase_counter) {
    Quiet q;
    RefVol* v = static_cast<RefVol*>(h);
    cv::Mat color = cv::Mat::wrap(H, W, CV_32FC3, const_cast<float*>(color_bgr));
    cv::Mat dep = cv::Mat::wrap(H, W, CV_32FC1, const_cast<float*>(depth));
    v->nest.set_maps(cv::Mat::wrap(H, W, CV_32FC1, const_cast<float*>(nx)),
                     cv::Mat::wrap(H, W, CV_32FC1, const_cast<float*>(ny)),
                     cv::Mat::wrap(H, W, CV_32FC1, const_cast<float*>(nz)),
                     cv::Mat::wrap(H, W, CV_32FC1, const_cast<float*>(n_sq_inv)));
    if (increase_counter) v->sdf->increase_counter();
    v->sdf->update(color, dep, K_from(K9), SE3(M4_from(pose16)), &v->nest);
}

long long ref_vol_num_voxels(void* h) { return static_cast<long long>(static_cast<RefVol*>(h)->sdf->num_voxels_); }
int ref_vol_counter(void* h) { return static_cast<int>(static_cast<RefVol*>(h)->sdf->counter_); }
void ref_vol_dims(void* h, int* dims, float* voxel_size, float* origin) {
    VolumetricGradSdf* s = static_cast<RefVol*>(h)->sdf;
    for (int i = 0; i < 3; ++i) { dims[i] = s->grid_dim_[i]; origin[i] = s->origin_[i]; }
    *voxel_size = s->voxel_size_;
}

// AoS rows of 8 floats: dist gx gy gz weight r g b  (Sdfvoxel.h:6-13)
void ref_vol_download(void* h, float* aos) {
    VolumetricGradSdf* s = static_cast<RefVol*>(h)->sdf;
    for (size_t i = 0; i < s->num_voxels_; ++i) {
        const SdfVoxel& v = s->tsdf_[i];
        float* o = aos + 8 * i;
        o[0] = v.dist; o[1] = v.grad[0]; o[2] = v.grad[1]; o[3] = v.grad[2];
        o[4] = v.weight; o[5] = v.r; o[6] = v.g; o[7] = v.b;
    }
}
void ref_vol_upload(void* h, const float* aos) {
    VolumetricGradSdf* s = static_cast<RefVol*>(h)->sdf;
    for (size_t i = 0; i < s->num_voxels_; ++i) {
        SdfVoxel& v = s->tsdf_[i];
        const float* o = aos + 8 * i;
        v.dist = o[0]; v.grad = Vec3f(o[1], o[2], o[3]);
        v.weight = o[4]; v.r = o[5]; v.g = o[6]; v.b = o[7];
    }
}
// vis_[lin] as little-endian bit words, word-major: words[w * N + lin]; bit k of word w = vis[64 w + k].
// Also returns each vector's length (the reference's select_vis depends on it).
void ref_vol_vis(void* h, uint64_t* words, int nwords, int* lengths) {
    VolumetricGradSdf* s = static_cast<RefVol*>(h)->sdf;
    const size_t N = s->num_voxels_;
    std::memset(words, 0, sizeof(uint64_t) * N * nwords);
    for (size_t i = 0; i < N; ++i) {
        const std::vector<bool>& vis = s->vis_[i];
        if (lengths) lengths[i] = static_cast<int>(vis.size());
        for (size_t k = 0; k < vis.size() && k < static_cast<size_t>(64 * nwords); ++k)
            if (vis[k]) words[(k >> 6) * N + i] |= (uint64_t(1) << (k & 63));
    }
}
void ref_vol_set_vis(void* h, const uint64_t* words, int nwords, const int* lengths) {
    VolumetricGradSdf* s = static_cast<RefVol*>(h)->sdf;
    const size_t N = s->num_voxels_;
    for (size_t i = 0; i < N; ++i) {
        std::vector<bool>& vis = s->vis_[i];
        vis.assign(lengths[i], false);
        for (int k = 0; k < lengths[i] && k < 64 * nwords; ++k)
            vis[k] = (words[(k >> 6) * N + i] >> (k & 63)) & 1;
    }
}

// ---------------------------------------------------------------- optimizer
// model: 0 SH1, 1 SH2, 2 LED.  loss: LossFunction enum value (OptimizerSettings.h:7-14).
void* ref_opt_create(void* vol, int model, float voxel_size, const float* K9, const char* save_path, int loss,
                     float lambda, float damping, float reg_rho, float reg_n, float reg_l, int max_it,
                     float conv_threshold, int upsample) {
    Quiet q;
    RefOpt* o = new RefOpt;
    o->vol = static_cast<RefVol*>(vol);
    o->set = new OptimizerSettings();
    o->set->model = model == 0 ? ModelType::SH1 : (model == 1 ? ModelType::SH2 : ModelType::LED);
    o->set->order = model == 1 ? 2 : 1;
    o->set->loss = static_cast<LossFunction>(loss);
    o->set->lambda = lambda;
    o->set->lambda_sq = lambda * lambda;
    o->set->damping = damping;
    o->set->reg_weight_rho = reg_rho;
    o->set->reg_weight_n = reg_n;
    o->set->reg_weight_l = reg_l;
    o->set->max_it = max_it;
    o->set->conv_threshold = conv_threshold;
    o->set->upsample = upsample != 0;
    if (model == 2) {
        o->led = new LedOptimizer(o->vol->sdf, voxel_size, K_from(K9), save_path, o->set);
        o->opt = o->led;
    } else {
        o->ps = new PsOptimizer(o->vol->sdf, voxel_size, K_from(K9), save_path, o->set);
        o->opt = o->ps;
    }
    return o;
}
void ref_opt_destroy(void* h) {
    RefOpt* o = static_cast<RefOpt*>(h);
    delete o->opt;
    delete o->set;
    delete o;
}

// main_ps.cpp:323-326
void ref_opt_set_keyframes(void* h, int F, const int* frame_idx, const float* poses16, const float* images_bgr,
                           int H, int W) {
    RefOpt* o = static_cast<RefOpt*>(h);
    std::vector<int> kf(frame_idx, frame_idx + F);
    std::vector<std::string> stamps;
    std::vector<Eigen::Matrix4f, Eigen::aligned_allocator<Eigen::Matrix4f>> poses;
    o->images.clear();
    for (int f = 0; f < F; ++f) {
        stamps.push_back(std::to_string(frame_idx[f]));
        poses.push_back(M4_from(poses16 + 16 * f));
        cv::Mat m(H, W, CV_32FC3);
        std::memcpy(m.data, images_bgr + static_cast<size_t>(f) * H * W * 3, sizeof(float) * H * W * 3);
        o->images.push_back(std::make_shared<cv::Mat>(m));
    }
    o->opt->setImages(o->images);
    o->opt->setKeyframes(kf);
    o->opt->setKeytimestamps(stamps);
    o->opt->setPoses(poses);
}
// main_ps.cpp:325 with caller-chosen strings (ref_opt_set_keyframes installs the frame ordinals)
void ref_opt_set_keytimestamps(void* h, int F, const char* const* stamps) {
    std::vector<std::string> st(stamps, stamps + F);
    static_cast<RefOpt*>(h)->opt->setKeytimestamps(st);
}
// Auxilary.h:41-123 on keyframe f at n (row, col) positions: interpolateImage(row, col) and computeImageGradient(row, col, img, 0 | 1),
// the way getIntensity / poseJacobian / distJacobian call them.  The corner cell (last row AND last column) reads past the image in the
// reference and must not be probed.
void ref_opt_image_probe(void* h, int f, int n, const float* rows, const float* cols, float* rgb3, float* g0, float* g1) {
    RefOpt* o = static_cast<RefOpt*>(h);
    const cv::Mat& img = *o->images[f];
    for (int i = 0; i < n; ++i) {
        if (rgb3) { const Vec3f c = interpolateImage(rows[i], cols[i], img); rgb3[3 * i] = c[0]; rgb3[3 * i + 1] = c[1]; rgb3[3 * i + 2] = c[2]; }
        const Vec3f a = computeImageGradient(rows[i], cols[i], img, 0), b = computeImageGradient(rows[i], cols[i], img, 1);
        for (int c = 0; c < 3; ++c) { g0[3 * i + c] = a[c]; g1[3 * i + c] = b[c]; }
    }
}
void ref_opt_init(void* h) { Quiet q; static_cast<RefOpt*>(h)->opt->init(); }

int ref_opt_num_surface(void* h) { return static_cast<int>(static_cast<RefOpt*>(h)->opt->surface_points_.size()); }
void ref_opt_surface(void* h, int* out) {
    const std::vector<int>& s = static_cast<RefOpt*>(h)->opt->surface_points_;
    std::copy(s.begin(), s.end(), out);
}
int ref_opt_num_frames(void* h) { return static_cast<int>(static_cast<RefOpt*>(h)->opt->num_frames_); }
float ref_opt_voxel_size(void* h) { return static_cast<RefOpt*>(h)->opt->voxel_size_; }

// light: SH -> F x basis floats; LED -> 3 floats
int ref_opt_light_len(void* h) {
    RefOpt* o = static_cast<RefOpt*>(h);
    if (o->led) return 3;
    return static_cast<int>(o->ps->light_.size()) * (o->set->order == 2 ? 9 : 4);
}
void ref_opt_get_light(void* h, float* out) {
    RefOpt* o = static_cast<RefOpt*>(h);
    if (o->led) { for (int i = 0; i < 3; ++i) out[i] = o->led->light_[i]; return; }
    const int nb = o->set->order == 2 ? 9 : 4;
    for (size_t f = 0; f < o->ps->light_.size(); ++f)
        for (int i = 0; i < nb; ++i) out[f * nb + i] = o->ps->light_[f][i];
}
void ref_opt_set_light(void* h, const float* in) {
    RefOpt* o = static_cast<RefOpt*>(h);
    if (o->led) { for (int i = 0; i < 3; ++i) o->led->light_[i] = in[i]; return; }
    const int nb = o->set->order == 2 ? 9 : 4;
    for (size_t f = 0; f < o->ps->light_.size(); ++f)
        for (int i = 0; i < nb; ++i) o->ps->light_[f][i] = in[f * nb + i];
}
void ref_opt_get_poses(void* h, float* out16) {
    RefOpt* o = static_cast<RefOpt*>(h);
    for (size_t f = 0; f < o->opt->poses_.size(); ++f)
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out16[16 * f + 4 * r + c] = o->opt->poses_[f](r, c);
}
void ref_opt_set_poses(void* h, const float* in16) {
    RefOpt* o = static_cast<RefOpt*>(h);
    for (size_t f = 0; f < o->opt->poses_.size(); ++f) o->opt->poses_[f] = M4_from(in16 + 16 * f);
}
void ref_opt_get_reg(void* h, float* out3) {
    RefOpt* o = static_cast<RefOpt*>(h);
    out3[0] = o->set->reg_weight_rho; out3[1] = o->set->reg_weight_n; out3[2] = o->set->reg_weight_l;
}
void ref_opt_set_reg(void* h, const float* in3) {
    RefOpt* o = static_cast<RefOpt*>(h);
    o->set->reg_weight_rho = in3[0]; o->set->reg_weight_n = in3[1]; o->set->reg_weight_l = in3[2];
}

void ref_opt_init_albedo(void* h) { static_cast<RefOpt*>(h)->opt->initAlbedo(); }

// getPSEnergy() as the reference computes it (single float, serial), and the same per-pair float
// losses re-summed in double, plus the active pair count.
float ref_opt_energy(void* h, double* E64, long long* n_active) {
    RefOpt* o = static_cast<RefOpt*>(h);
    Optimizer* p = o->opt;
    const float E = p->getPSEnergy();
    double acc = 0.0;
    long long cnt = 0;
    VolumetricGradSdf* s = o->vol->sdf;
    for (int lin : p->surface_points_) {
        const SdfVoxel& v = s->tsdf_[lin];
        const std::vector<bool>& vis = s->vis_[lin];
        const Vec3i idx = s->line2idx(lin);
        for (size_t f = 0; f < p->num_frames_; ++f) {
            if (!vis[f]) continue;
            Vec3f intensity;
            if (!p->getIntensity(idx, v, p->getRotation(f), p->getTranslation(f), *p->images_[f], intensity)) continue;
            Vec3f rend = o->ps ? o->ps->renderedIntensity(v, idx, static_cast<int>(f))
                               : o->led->renderedIntensity(v, idx, static_cast<int>(f));
            acc += static_cast<double>(p->computeLoss(intensity - rend));
            ++cnt;
        }
    }
    if (E64) *E64 = acc / static_cast<double>(p->surface_points_.size());
    if (n_active) *n_active = cnt;
    return E;
}
// getPSEnergy() alone (timing leg of bench.py)
float ref_opt_ps_energy(void* h) { return static_cast<RefOpt*>(h)->opt->getPSEnergy(); }
float ref_opt_normal_energy(void* h) { return static_cast<RefOpt*>(h)->opt->getNormalEnergy(); }
float ref_opt_laplacian_energy(void* h) { return static_cast<RefOpt*>(h)->opt->getLaplacianEnergy(); }
float ref_opt_albedo_reg_energy(void* h) { return static_cast<RefOpt*>(h)->opt->getAlbedoRegEnergy(); }

// Per-pair probe: projected pixel (col m, row n) and in-bounds flag of getIntensity (OptimizerAux.cpp:207-231),
// sampled intensity and rendered intensity.  Output arrays are S x F (row-major s, f).
void ref_opt_probe_pairs(void* h, float* m_out, float* n_out, unsigned char* ok_out, float* I_out, float* rend_out) {
    RefOpt* o = static_cast<RefOpt*>(h);
    Optimizer* p = o->opt;
    VolumetricGradSdf* s = o->vol->sdf;
    const float fx = p->K_(0, 0), fy = p->K_(1, 1), cx = p->K_(0, 2), cy = p->K_(1, 2);
    const size_t F = p->num_frames_;
    size_t row = 0;
    for (int lin : p->surface_points_) {
        const SdfVoxel& v = s->tsdf_[lin];
        const std::vector<bool>& vis = s->vis_[lin];
        const Vec3i idx = s->line2idx(lin);
        for (size_t f = 0; f < F; ++f) {
            const size_t k = row * F + f;
            m_out[k] = n_out[k] = 0.f; ok_out[k] = 0;
            for (int c = 0; c < 3; ++c) { I_out[3 * k + c] = 0.f; rend_out[3 * k + c] = 0.f; }
            if (!vis[f]) continue;
            const Mat3f R = p->getRotation(f);
            const Vec3f t = p->getTranslation(f);
            // same statements as getIntensity()
            Mat3f Rt = R.transpose();
            Vec3f point = Rt * (s->voxel2world(idx) - v.dist * v.grad.normalized() - t);
            const float z_inv = 1. / point[2];
            m_out[k] = fx * point[0] * z_inv + cx;
            n_out[k] = fy * point[1] * z_inv + cy;
            Vec3f intensity;
            if (!p->getIntensity(idx, v, R, t, *p->images_[f], intensity)) continue;
            ok_out[k] = 1;
            Vec3f rend = o->ps ? o->ps->renderedIntensity(v, idx, static_cast<int>(f))
                               : o->led->renderedIntensity(v, idx, static_cast<int>(f));
            for (int c = 0; c < 3; ++c) { I_out[3 * k + c] = intensity[c]; rend_out[3 * k + c] = rend[c]; }
        }
        ++row;
    }
}

// computeResidual(): r (3*F*S, index 3*(f*S+s)+c) and the diagonal of W.
void ref_opt_residual(void* h, float* r_out, float* w_out) {
    RefOpt* o = static_cast<RefOpt*>(h);
    std::pair<Eigen::VectorXf, Eigen::SparseMatrix<float>> pr =
        o->ps ? o->ps->computeResidual() : o->led->computeResidual();
    const Eigen::Index n = pr.first.size();
    for (Eigen::Index i = 0; i < n; ++i) { r_out[i] = pr.first[i]; w_out[i] = 0.f; }
    for (int k = 0; k < pr.second.outerSize(); ++k)
        for (Eigen::SparseMatrix<float>::InnerIterator it(pr.second, k); it; ++it)
            if (it.row() == it.col()) w_out[it.row()] = it.value();
}

// FD gradient and direction per surface voxel: computeDistGrad (Optimizer.cpp:287-364). out: S x 6.
void ref_opt_dist_grad(void* h, float* out) {
    RefOpt* o = static_cast<RefOpt*>(h);
    VolumetricGradSdf* s = o->vol->sdf;
    size_t row = 0;
    for (int lin : o->opt->surface_points_) {
        std::pair<Vec3f, Vec3f> g = o->opt->computeDistGrad(s->tsdf_[lin], s->line2idx(lin));
        for (int i = 0; i < 3; ++i) { out[6 * row + i] = g.first[i]; out[6 * row + 3 + i] = g.second[i]; }
        ++row;
    }
}

// Jacobian of one block as CSC (query with null outputs to get nnz/cols).
// which: 0 albedo, 1 light, 2 dist, 3 pose, 4 distReg (rows S), 5 distLaplacian, 6 albedoReg
static Eigen::SparseMatrix<float> block_jacobian(RefOpt* o, int which, Eigen::VectorXf* reg_res) {
    switch (which) {
        case 0: return o->ps ? o->ps->albedoJacobian() : o->led->albedoJacobian();
        case 1: return o->ps ? o->ps->lightJacobian() : o->led->lightJacobian();
        case 2: return o->ps ? o->ps->distJacobian() : o->led->distJacobian();
        case 3: return o->ps ? o->ps->poseJacobian() : o->led->poseJacobian();
        case 4: { auto pr = o->opt->distRegJacobian(); if (reg_res) *reg_res = pr.second; return pr.first; }
        case 5: { auto pr = o->opt->distLaplacianJacobian(); if (reg_res) *reg_res = pr.second; return pr.first; }
        default: { auto pr = o->opt->albedoRegJacobian(); if (reg_res) *reg_res = pr.second; return pr.first; }
    }
}

static void csc_out(Eigen::SparseMatrix<float>& M, long long* nnz, int* rows, int* cols, long long* colptr,
                    int* rowidx, float* vals) {
    M.makeCompressed();
    *nnz = M.nonZeros(); *rows = static_cast<int>(M.rows()); *cols = static_cast<int>(M.cols());
    if (colptr) for (int c = 0; c <= M.cols(); ++c) colptr[c] = M.outerIndexPtr()[c];
    if (rowidx) for (long long k = 0; k < *nnz; ++k) rowidx[k] = M.innerIndexPtr()[k];
    if (vals) for (long long k = 0; k < *nnz; ++k) vals[k] = M.valuePtr()[k];
}

static Eigen::SparseMatrix<float> g_lastJ;
void ref_opt_jacobian(void* h, int which, long long* nnz, int* rows, int* cols) {
    Quiet q;
    g_lastJ = block_jacobian(static_cast<RefOpt*>(h), which, nullptr);
    csc_out(g_lastJ, nnz, rows, cols, nullptr, nullptr, nullptr);
}
void ref_last_jacobian_fetch(long long* colptr, int* rowidx, float* vals) {
    long long nnz; int r, c;
    csc_out(g_lastJ, &nnz, &r, &c, colptr, rowidx, vals);
}

// One block step, re-playing optimize*All (PsOptimizer.cpp:85-234, LedOptimizer.cpp:134-275) so that H, b,
// delta and the CG verdict can be inspected.  which: 0 albedo, 1 light, 2 dist, 3 pose.  If apply != 0 the
// reference's update rule for that block/model is executed (including its Success gating).
// Returns the system size n.
int ref_opt_block(void* h, int which, int apply, float damping, int albedo_reg, int normal_reg, int laplacian_reg) {
    Quiet q;
    RefOpt* o = static_cast<RefOpt*>(h);
    OptimizerSettings* st = o->set;
    Eigen::SparseMatrix<float> J = block_jacobian(o, which, nullptr);
    std::pair<Eigen::VectorXf, Eigen::SparseMatrix<float>> pr =
        o->ps ? o->ps->computeResidual() : o->led->computeResidual();
    Eigen::VectorXf r = pr.first;
    Eigen::SparseMatrix<float> w = pr.second;
    Eigen::SparseMatrix<float> H = J.transpose() * w * J;
    Eigen::VectorXf b = J.transpose() * w * r;
    if (which == 0 && albedo_reg) {
        std::pair<Eigen::SparseMatrix<float>, Eigen::VectorXf> tmp = o->opt->albedoRegJacobian();
        Eigen::SparseMatrix<float> Jr_r = tmp.first;
        Eigen::VectorXf r_r = tmp.second;
        H += st->reg_weight_rho * Jr_r.transpose() * Jr_r;
        b += st->reg_weight_rho * Jr_r.transpose() * r_r;
    }
    if (which == 2 && normal_reg) {
        std::pair<Eigen::SparseMatrix<float>, Eigen::VectorXf> pair = o->opt->distRegJacobian();
        Eigen::SparseMatrix<float> Jr_d = pair.first;
        Eigen::VectorXf r_n = pair.second;
        H += st->reg_weight_n * (Jr_d.transpose() * Jr_d);
        b += st->reg_weight_n * (Jr_d.transpose() * r_n);
    }
    if (which == 2 && laplacian_reg) {
        std::pair<Eigen::SparseMatrix<float>, Eigen::VectorXf> pair = o->opt->distLaplacianJacobian();
        Eigen::SparseMatrix<float> Jl_d = pair.first;
        Eigen::VectorXf r_n = pair.second;
        H += st->reg_weight_l * (Jl_d.transpose() * Jl_d);
        b += st->reg_weight_l * (Jl_d.transpose() * r_n);
    }
    // SH light block has no damping (PsOptimizer.cpp:175-203); every other block has.
    const bool damp = !(which == 1 && o->ps);
    if (damp && damping != 0.0) H.diagonal() += damping * H.diagonal();

    Eigen::ConjugateGradient<Eigen::SparseMatrix<float>> solver;
    solver.compute(H);
    Eigen::VectorXf delta = solver.solve(b);
    o->H = H; o->b = b; o->delta = delta;
    o->info = static_cast<int>(solver.info());
    o->iters = static_cast<int>(solver.iterations());
    o->err = static_cast<float>(solver.error());

    if (apply) {
        const bool ok = solver.info() == Eigen::Success;
        if (which == 0) { if (o->led || ok) o->opt->updateAlbedo(delta); }
        else if (which == 1) {
            if (o->led) o->led->light_ -= delta;
            else {
                const size_t basis = st->order == 2 ? 9 : 4;
                for (size_t f = 0; f < o->opt->num_frames_; ++f) o->ps->light_[f] -= delta.segment(f * basis, basis);
            }
        }
        else if (which == 2) { if (o->led || ok) o->opt->updateDist(delta, true); }
        else { if (o->ps || ok) o->opt->updatePose(delta); }
    }
    return static_cast<int>(b.size());
}
void ref_opt_block_result(void* h, float* b, float* delta, int* info, int* iters, float* err) {
    RefOpt* o = static_cast<RefOpt*>(h);
    if (b) for (Eigen::Index i = 0; i < o->b.size(); ++i) b[i] = o->b[i];
    if (delta) for (Eigen::Index i = 0; i < o->delta.size(); ++i) delta[i] = o->delta[i];
    if (info) *info = o->info;
    if (iters) *iters = o->iters;
    if (err) *err = o->err;
}
void ref_opt_block_H(void* h, long long* nnz, int* rows, int* cols, long long* colptr, int* rowidx, float* vals) {
    RefOpt* o = static_cast<RefOpt*>(h);
    csc_out(o->H, nnz, rows, cols, colptr, rowidx, vals);
}

// The reference's own block step, untouched.
void ref_opt_block_reference(void* h, int which, float damping, int albedo_reg, int normal_reg, int laplacian_reg) {
    Quiet q;
    RefOpt* o = static_cast<RefOpt*>(h);
    if (o->ps) {
        if (which == 0) o->ps->optimizeAlbedoAll(albedo_reg, damping);
        else if (which == 1) o->ps->optimizeLightAll();
        else if (which == 2) o->ps->optimizeDistAll(normal_reg, laplacian_reg, damping);
        else o->ps->optimizePosesAll(damping);
    } else {
        if (which == 0) o->led->optimizeAlbedoAll(albedo_reg, damping);
        else if (which == 1) o->led->optimizeLightAll(damping);
        else if (which == 2) o->led->optimizeDistAll(normal_reg, laplacian_reg, damping);
        else o->led->optimizePosesAll(damping);
    }
}

// The reference's outer loop, untouched; files go to save_path (must exist).  Returns its bool, and the
// wall time in seconds.  Captured stdout (energies, timers) can be fetched with ref_last_stdout().
int ref_opt_alternating(void* h, int light, int albedo, int distance, int pose, double* seconds) {
    Quiet q;
    RefOpt* o = static_cast<RefOpt*>(h);
    const auto t0 = std::chrono::steady_clock::now();
    const bool r = o->opt->alternatingOptimize(light, albedo, distance, pose);
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    return r ? 1 : 0;
}
static std::string g_last_out;
const char* ref_last_stdout() { g_last_out = g_sink.str(); return g_last_out.c_str(); }

void ref_opt_subsampling(void* h) { Quiet q; static_cast<RefOpt*>(h)->opt->subsampling(); }


// ---- rigid depth-to-SDF tracker (sdf_tracker/RigidPointOptimizer.cpp:11-78), SURVEY.md 8f next-2 ---------------------
// One evaluation of optimize_sampled's inner loop (:38-62) at a given pose, re-played statement by statement against the
// reference's own Sdf::weights / Sdf::tsdf (VolumetricGradSdf.h:76-93): cost, gradient, approximate Hessian, pixel count.
void ref_tracker_system(void* h, const float* depth, int H, int W, const float* K9, const float* pose16, int sampling,
                        float* E_out, float* g6, float* H36, long long* count) {
    RefVol* v = static_cast<RefVol*>(h);
    Sdf* tSDF_ = v->sdf;
    SE3 pose_(M4_from(pose16));
    const float z_min = tSDF_->z_min_, z_max = tSDF_->z_max_;
    const float fx = K9[0], fy = K9[4], cx = K9[2], cy = K9[5];
    const float fx_inv = 1.f / fx;
    const float fy_inv = 1.f / fy;
    Mat3f R = pose_.rotationMatrix();
    Vec3f t = pose_.translation();
    float E = 0;
    Vec6f g = Vec6f::Zero();
    Mat6f Hm = Mat6f::Zero();
    size_t counter = 0;
    for (int y = 0; y < H; y += sampling) for (int x = 0; x < W; x += sampling) {
        const float z = depth[y * W + x];
        if (z <= z_min || z >= z_max) continue;
        const float x0 = (float(x) - cx) * fx_inv;
        const float y0 = (float(y) - cy) * fy_inv;
        Vec3f point(x0 * z, y0 * z, z);
        point = R * point + t;
        float w0 = tSDF_->weights(point);
        if (w0 > 0) {
            Vec3f grad_curr;
            float phi0 = tSDF_->tsdf(point, &grad_curr);
            E += phi0 * phi0;
            Vec6f grad_xi;
            grad_xi << grad_curr, point.cross(grad_curr);
            g += phi0 * grad_xi;
            Hm += grad_xi * grad_xi.transpose();
            ++counter;
        }
    }
    *E_out = E;
    for (int i = 0; i < 6; ++i) g6[i] = g[i];
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) H36[6 * i + j] = Hm(i, j);
    *count = (long long)counter;
}

// xi = damping * H.llt().solve(g); pose <- SE3::exp(-xi) * pose (RigidPointOptimizer.cpp:67-75); returns |xi|^2
float ref_tracker_update(const float* g6, const float* H36, float damping, const float* pose_in16, float* pose_out16, float* xi6) {
    Vec6f g; Mat6f Hm;
    for (int i = 0; i < 6; ++i) g[i] = g6[i];
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) Hm(i, j) = H36[6 * i + j];
    Vec6f xi = damping * Hm.llt().solve(g);
    SE3 pose_(M4_from(pose_in16));
    pose_ = SE3::exp(-xi) * pose_;
    const Mat4f M = pose_.matrix();
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) pose_out16[4 * r + c] = M(r, c);
    if (xi6) for (int i = 0; i < 6; ++i) xi6[i] = xi[i];
    return xi.squaredNorm();
}

// The reference's own RigidPointOptimizer::optimize (main_ps.cpp:236-243 path) from a given start pose
int ref_tracker_optimize(void* h, const float* depth, int H, int W, const float* K9, const float* pose_in16, int num_iterations,
                         float conv_threshold, float damping, float* pose_out16) {
    Quiet q;
    RefVol* v = static_cast<RefVol*>(h);
    RigidPointOptimizer opt(num_iterations, conv_threshold, damping, v->sdf);
    opt.set_pose(SE3(M4_from(pose_in16)));
    cv::Mat dep = cv::Mat::wrap(H, W, CV_32FC1, const_cast<float*>(depth));
    Mat3f K;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) K(r, c) = K9[3 * r + c];
    const bool conv = opt.optimize(dep, K);
    const Mat4f M = opt.pose().matrix();
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) pose_out16[4 * r + c] = M(r, c);
    return conv ? 1 : 0;
}


// compute_centroid (voxel_ps/src/main_ps.cpp:346-375) lives in the driver translation unit, which cannot be built here
// (imread etc.); its statements are re-played with the same Eigen types.
void ref_compute_centroid(const float* K9, const float* depth_ptr, int h, int w, const float* pose16, float* out3) {
    Mat3f K;
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) K(r, c) = K9[3 * r + c];
    const Mat4f Trans = M4_from(pose16);
    Vec3f centroid(0., 0., 0.);
    int counter = 0;
    Mat3f R = Trans.topLeftCorner(3, 3);
    Vec3f t = Trans.topRightCorner(3, 1);
    float fx = K(0, 0);
    float fy = K(1, 1);
    float cx = K(0, 2);
    float cy = K(1, 2);
    float fx_inv = 1.f / fx;
    float fy_inv = 1.f / fy;
    for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) {
        float z = depth_ptr[y * w + x];
        if (z > 0.0) {
            float x0 = (float(x) - cx) * fx_inv;
            float y0 = (float(y) - cy) * fy_inv;
            centroid += R * Vec3f(x0 * z, y0 * z, z) + t;
            ++counter;
        }
    }
    const Vec3f c = centroid / float(counter);
    for (int a = 0; a < 3; ++a) out3[a] = c[a];
}

// ---- surface extraction ------------------------------------------------------------------------------
// third/mesh/MarchingCubes.cpp:314-699, unmodified, on caller-supplied arrays.  Query with null outputs for the counts.
// vertices: 3 floats each, colors: 3 bytes each, faces: 3 ints each.
int ref_marching_cubes(const int* dims, const float* size3, const float* origin3, const float* tsdf, const float* weights,
                       const unsigned char* red, const unsigned char* green, const unsigned char* blue, float iso,
                       long long* n_vertices, long long* n_faces, float* vertices, unsigned char* colors, int* faces) {
    MarchingCubes mc(Vec3i(dims[0], dims[1], dims[2]), Vec3f(size3[0], size3[1], size3[2]), Vec3f(origin3[0], origin3[1], origin3[2]));
    const bool ok = mc.computeIsoSurface(tsdf, weights, red, green, blue, iso);
    if (n_vertices) *n_vertices = (long long)mc.vertices_.size();
    if (n_faces) *n_faces = (long long)mc.faces_.size();
    if (vertices) for (size_t i = 0; i < mc.vertices_.size(); ++i) for (int a = 0; a < 3; ++a) vertices[3 * i + a] = mc.vertices_[i][a];
    if (colors) for (size_t i = 0; i < mc.colors_.size(); ++i) for (int a = 0; a < 3; ++a) colors[3 * i + a] = mc.colors_[i][a];
    if (faces) for (size_t i = 0; i < mc.faces_.size(); ++i) for (int a = 0; a < 3; ++a) faces[3 * i + a] = mc.faces_[i][a];
    return ok ? 1 : 0;
}
// the reference's own writers, unmodified: the files they produce are what the tests compare byte by byte
int ref_vol_extract_mesh(void* h, const char* filename) { Quiet q; return static_cast<RefVol*>(h)->sdf->extract_mesh(filename) ? 1 : 0; }
int ref_vol_extract_pc(void* h, const char* filename) { Quiet q; return static_cast<RefVol*>(h)->sdf->extract_pc(filename) ? 1 : 0; }
int ref_vol_save_sdf(void* h, const char* filename) { Quiet q; return static_cast<RefVol*>(h)->sdf->saveSDF(filename) ? 1 : 0; }
// these write <save_path><name>_mesh.ply / _obj.ply / _pointcloud.ply / <save_path><name>
int ref_opt_extract_mesh(void* h, const char* name) { Quiet q; return static_cast<RefOpt*>(h)->opt->extract_mesh(name) ? 1 : 0; }
int ref_opt_save_white_mesh(void* h, const char* name) { Quiet q; return static_cast<RefOpt*>(h)->opt->save_white_mesh(name) ? 1 : 0; }
int ref_opt_save_pointcloud(void* h, const char* name) { Quiet q; return static_cast<RefOpt*>(h)->opt->save_pointcloud(name) ? 1 : 0; }
int ref_opt_save_sdf(void* h, const char* name) { Quiet q; return static_cast<RefOpt*>(h)->opt->saveSDF(name) ? 1 : 0; }

}  // extern "C"
