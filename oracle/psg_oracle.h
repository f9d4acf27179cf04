/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("oracle") of the PSgradientSDF photometric-optimizer
 * hot path.  Plain C99, single thread, O(1) surface membership (dense row map) instead of the
 * reference's std::find, otherwise the reference's arithmetic in the reference's evaluation order.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library.  The product (psgradientsdf_b200/) never does.
 *
 * Parity pin: checked against oracle/_ref/libpsgref.so (the reference's own sources compiled
 * unmodified + ref_harness.cpp) by tests/test_oracle_vs_ref.py, and against the committed fixtures in
 * tests/golden/ (generated from the reference by tests/golden/make_golden.py).  The reference itself
 * ships no tests or golden vectors.
 */
#ifndef PSG_ORACLE_H
#define PSG_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct or_volume or_volume;
typedef struct or_opt or_opt;

enum { OR_SH1 = 0, OR_SH2 = 1, OR_LED = 2 };
enum { OR_L2 = 0, OR_CAUCHY = 1, OR_HUBER = 2, OR_TUKEY = 3, OR_TRUNC_L2 = 4 }; /* OptimizerSettings.h:7-14 */
enum { OR_ALBEDO = 0, OR_LIGHT = 1, OR_DIST = 2, OR_POSE = 3 };

/* ---- volume (VolumetricGradSdf.h:65-70, VolumetricGradSdf.cpp:14-38) ---- */
or_volume* or_volume_create(const int dims[3], float voxel_size, const float shift[3], float T, float zmin,
                            float zmax, int max_frames);
void or_volume_destroy(or_volume*);
long long or_volume_num_voxels(const or_volume*);
void or_volume_geometry(const or_volume*, int dims[3], float* voxel_size, float origin[3]);
/* VolumetricGradSdf::update (VolumetricGradSdf.cpp:51-138); R row-major 3x3 camera-to-world, t translation */
void or_volume_integrate(or_volume*, const float* color_bgr, const float* depth, const float* nx, const float* ny,
                         const float* nz, const float* n_sq_inv, int H, int W, const float K[9], const float R[9],
                         const float t[3], int increase_counter);
void or_volume_download(const or_volume*, float* aos8);
void or_volume_upload(or_volume*, const float* aos8);
int or_volume_vis_words(const or_volume*);
void or_volume_vis(const or_volume*, uint64_t* words /* nwords x N */, int* lengths /* N */);
void or_volume_set_vis(or_volume*, const uint64_t* words, const int* lengths);

/* ---- optimizer (Optimizer.h:24-182) ---- */
or_opt* or_opt_create(or_volume*, int model, float voxel_size, const float K[9], int loss, float lambda,
                      float reg_rho, float reg_n, float reg_l);
void or_opt_destroy(or_opt*);
/* setImages/setKeyframes/setPoses (main_ps.cpp:323-326); images F x H x W x 3 float BGR; poses F x 16 row-major */
void or_opt_set_keyframes(or_opt*, int F, const int* frame_idx, const float* poses16, const float* images_bgr, int H,
                          int W);
/* {Ps,Led}Optimizer::init: initial light + select_vis + getSurfaceVoxel (+ LED computeLightIntensive) */
void or_opt_init(or_opt*);
int or_opt_num_surface(const or_opt*);
void or_opt_surface(const or_opt*, int* lin_out);
void or_opt_vis_keyframes(const or_opt*, uint64_t* vis_out /* S */);
int or_opt_light_len(const or_opt*);
void or_opt_get_light(const or_opt*, float*);
void or_opt_set_light(or_opt*, const float*);
void or_opt_get_poses(const or_opt*, float* poses16);
void or_opt_set_poses(or_opt*, const float* poses16);
void or_opt_get_reg(const or_opt*, float out3[3]);
void or_opt_set_reg(or_opt*, const float in3[3]);
float or_opt_voxel_size(const or_opt*);

void or_opt_init_albedo(or_opt*);                                   /* Optimizer.cpp:50-81 */
float or_opt_energy(or_opt*, double* E64, long long* n_active);     /* PsOptimizer.cpp:47-78 */
float or_opt_normal_energy(or_opt*, double* E64);                   /* Optimizer.cpp:86-103 */
float or_opt_laplacian_energy(or_opt*, double* E64);                /* Optimizer.cpp:106-119 */
float or_opt_albedo_reg_energy(or_opt*, double* E64);               /* Optimizer.cpp:122-136 */
/* per pair: getIntensity's pixel (col m, row n), in-bounds flag, sampled and rendered intensity; S x F */
void or_opt_probe_pairs(or_opt*, float* m, float* n, unsigned char* ok, float* I3, float* rend3);
void or_opt_residual(or_opt*, float* r3 /* S x F x 3 */, float* w3);
void or_opt_dist_grad(or_opt*, float* out6 /* S x 6: grad, pos */);
/* interpolateImage + computeImageGradient (Auxilary.h:41-123) of keyframe f at n (row, col) positions; rgb3 may be NULL */
void or_opt_image_probe(const or_opt*, int f, int n, const float* rows, const float* cols, float* rgb3, float* g0, float* g1);

/* One block step (PsOptimizer.cpp:85-234 / LedOptimizer.cpp:134-275).  Returns system size n.
 * Afterwards: or_opt_block_result() gives b, delta, CG verdict; or_opt_block_H() the matrix as COO triplets
 * (full symmetric pattern). */
/* Diagnostic switch (not reference behaviour): sum the per-keyframe light / pose normal equations in double instead of
 * the reference's serial float, to separate summation-order noise from real differences. */
void or_opt_set_accum64(or_opt*, int on);
int or_opt_block(or_opt*, int which, int apply, float damping, int albedo_reg, int normal_reg, int laplacian_reg);
void or_opt_block_result(const or_opt*, float* b, float* delta, int* success, int* iters, float* err);
long long or_opt_block_H(const or_opt*, int* rows, int* cols, float* vals); /* returns nnz; pass NULLs to query */

void or_opt_subsampling(or_opt*);                                   /* OptimizerAux.cpp:622-684 */

#ifdef __cplusplus
}
#endif
/* RigidPointOptimizer::optimize_sampled's inner loop (sdf_tracker/RigidPointOptimizer.cpp:38-62) at pose (R, t) */
void or_tracker_system(const or_volume*, const float* depth, int H, int W, const float K[9], const float R[9],
                       const float t[3], int sampling, float* E, float g6[6], float H36[36], long long* count);

#endif
