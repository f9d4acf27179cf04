/* psg.h -- C ABI of the B200-native PSgradientSDF photometric-optimizer hot path.
 *
 * Drop-in boundary: the reference reaches this path through two C++ classes,
 *   Sdf / VolumetricGradSdf   (cpp/include/sdf_tracker/Sdf.h:24-141, VolumetricGradSdf.h:65-122)
 *   Optimizer / PsOptimizer / LedOptimizer (cpp/include/ps_optimizer/Optimizer.h:24-182)
 * called from cpp/voxel_ps/src/main_ps.cpp:183-330.  Every entry point below names the reference
 * interface it replaces.  Plain pointers and sizes only; all pointers are HOST pointers unless a
 * parameter says otherwise; the library owns all device memory.  The host keeps ownership of every
 * buffer it passes (images are copied at psg_opt_set_keyframes).
 *
 * Conventions: matrices are row-major; poses are camera-to-world 4x4 (main_ps.cpp:219);
 * colour images are float32 BGR (ImageLoader.h:167-188); depth is float32 metres; voxels are ordered
 * lin = i + W*j + W*H*k (VoxelGrid.h:79-81); a voxel record is 8 floats
 * {dist, grad.x, grad.y, grad.z, weight, r, g, b} (Sdfvoxel.h:6-13).
 *
 * Errors: every function returns 0 on success and a non-zero psg_status otherwise;
 * psg_last_error() returns a description of the last failure on the calling thread.  There is no CPU
 * fallback: without a usable sm_100 device every call that needs one fails with PSG_ERR_CUDA.
 * Not thread-safe per handle (the reference is single-threaded).
 */
#ifndef PSG_H
#define PSG_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct psg_volume psg_volume;
typedef struct psg_optimizer psg_optimizer;

typedef enum {
    PSG_OK = 0,
    PSG_ERR_INVALID = 1,     /* bad argument */
    PSG_ERR_CUDA = 2,        /* CUDA runtime / device error */
    PSG_ERR_STATE = 3,       /* call order violated (e.g. step before init) */
    PSG_ERR_UNSUPPORTED = 4  /* feature of the reference not provided on this path */
} psg_status;

/* ModelType / LossFunction (cpp/include/ps_optimizer/OptimizerSettings.h:7-21) */
typedef enum { PSG_MODEL_SH1 = 0, PSG_MODEL_SH2 = 1, PSG_MODEL_LED = 2 } psg_model;
typedef enum { PSG_LOSS_L2 = 0, PSG_LOSS_CAUCHY = 1, PSG_LOSS_HUBER = 2, PSG_LOSS_TUKEY = 3, PSG_LOSS_TRUNC_L2 = 4 } psg_loss;
typedef enum { PSG_BLOCK_ALBEDO = 0, PSG_BLOCK_LIGHT = 1, PSG_BLOCK_DIST = 2, PSG_BLOCK_POSE = 3 } psg_block;

/* struct OptimizerSettings (OptimizerSettings.h:23-51), the fields the path reads. */
typedef struct {
    int model;           /* psg_model */
    int loss;            /* psg_loss */
    float lambda;        /* robust-loss scale */
    float reg_weight_rho;
    float reg_weight_n;
    float reg_weight_l;
} psg_settings;

/* What one optimize*All call did (PsOptimizer.cpp:85-234): size of the system, whether the CG reported
 * Eigen::Success, its iteration count and error estimate, and whether the update was applied (the SH path
 * applies albedo/distance only on Success, the LED path applies pose only on Success). */
typedef struct {
    int n;
    int cg_success;
    int cg_iterations;
    float cg_error;
    int applied;
    long long n_accepted;   /* albedo channels / distances whose gated update was accepted (OptimizerAux.cpp:149,183) */
    /* getPSEnergy() of the state the block started from: the block's sweep forms every residual anyway, so the value the
     * reference logs after the PREVIOUS block comes for free (same definition as psg_opt_energy). */
    double energy_before;
    long long n_active_before;
} psg_step_info;

const char* psg_last_error(void);
/* Number of kernels this library has launched in the calling process so far (for bench.py's gpu_launches). */
long long psg_kernel_launches(void);
int psg_device_count(int* count);

/* Built-in launch timer (no counterpart in the reference, which only has wall-clock Timer::tic/toc, Timer.h:29-49).
 * When enabled, every kernel launch is bracketed by CUDA events on the stream it is launched on; psg_profile_get
 * returns the launch count and summed device time of kernel `id` (0 <= id < psg_profile_kernels()). */
int psg_profile_enable(int on);
int psg_profile_reset(void);
int psg_profile_kernels(void);
const char* psg_profile_name(int id);
int psg_profile_get(int id, long long* launches, double* total_ms);

/* ------------------------------------------------------------------ volume (Sdf / VolumetricGradSdf) */

/* VolumetricGradSdf(grid_dim, voxel_size, shift, T) + set_zmin/set_zmax (VolumetricGradSdf.h:65-70,
 * Sdf.h:105-111, main_ps.cpp:183-186).  max_frames bounds the per-voxel visibility record (one bit per
 * input frame, VolumetricGradSdf.h:42); device selects the CUDA device. */
int psg_volume_create(psg_volume** out, const int dims[3], float voxel_size, const float shift[3], float T,
                      float zmin, float zmax, int max_frames, int device);
int psg_volume_destroy(psg_volume* vol);
/* grid_dim_, voxel_size_, origin_ (VoxelGrid.h:30-34) -- these change after psg_opt_upsample. */
int psg_volume_geometry(const psg_volume* vol, int dims[3], float* voxel_size, float origin[3]);

/* Sdf::increase_counter() (optional) followed by VolumetricGradSdf::update(color, depth, K, pose, NEst)
 * (VolumetricGradSdf.cpp:51-138).  nx, ny, nz, n_sq_inv are what NormalEstimator::compute / n_sq_inv_ptr()
 * hand to update() (NormalEstimator.h:150-176,183).  R, t = pose.rotationMatrix(), pose.translation(). */
int psg_volume_integrate(psg_volume* vol, const float* color_bgr, const float* depth, const float* nx,
                         const float* ny, const float* nz, const float* n_sq_inv, int H, int W,
                         const float K[9], const float R[9], const float t[3], int increase_counter);

/* tsdf_ as an array of SdfVoxel records, N x 8 floats (replaces the optimizer's `friend` access,
 * VolumetricGradSdf.h:26-30; consumed by the PLY / marching-cubes / .sdf writers). */
int psg_volume_download(psg_volume* vol, float* aos8);
int psg_volume_upload(psg_volume* vol, const float* aos8);
/* vis_ (VolumetricGradSdf.h:42) as bit words: words[w * N + lin], bit k of word w = vis[64 w + k];
 * nwords = ceil(max_frames / 64).  The vector's length is (index of highest set bit) + 1. */
int psg_volume_vis_words(const psg_volume* vol, int* nwords);
int psg_volume_download_vis(psg_volume* vol, uint64_t* words);
int psg_volume_upload_vis(psg_volume* vol, const uint64_t* words, int frame_counter);

/* ------------------------------------------------------------------ optimizer (Optimizer / PsOptimizer / LedOptimizer) */

/* PsOptimizer / LedOptimizer(tSDF, voxel_size, K, save_path, settings) (main_ps.cpp:193-202).  The optimizer
 * keeps a non-owning pointer to vol, like the reference.  At most 10 optimizers can be alive per device at a time (each holds a
 * copy of its per-keyframe pose / light table in the device's constant memory); psg_opt_destroy gives the place back. */
int psg_opt_create(psg_optimizer** out, psg_volume* vol, float voxel_size, const float K[9],
                   const psg_settings* settings);
int psg_opt_destroy(psg_optimizer* opt);

/* setImages + setKeyframes + setPoses (Optimizer.h:137-154, main_ps.cpp:323-326): F keyframes (F <= 64),
 * frame_idx[k] = input-frame ordinal of keyframe k, poses F x 16, images F x H x W x 3 float BGR. */
int psg_opt_set_keyframes(psg_optimizer* opt, int F, const int* frame_idx, const float* poses16,
                          const float* images_bgr, int H, int W);
/* setKeytimestamps (Optimizer.h:151-154, called at main_ps.cpp:325): one NUL-terminated string per keyframe, copied.
 * Optimizer::savePoses (OptimizerAux.cpp:580-599) prints key_stamps_[i] as the first column of
 * after_poses_opt_<it>.txt.  F must equal the keyframe count of the last psg_opt_set_keyframes call; the strings survive
 * later psg_opt_set_keyframes calls with the same F and are dropped when F changes.  Where the reference would index an
 * unset key_stamps_ (undefined behaviour), this path prints the keyframe's frame_idx. */
int psg_opt_set_keytimestamps(psg_optimizer* opt, int F, const char* const* stamps);

/* {Ps,Led}Optimizer::init() (PsOptimizer.cpp:25-42, LedOptimizer.cpp:25-36): initial light, select_vis()
 * (Optimizer.cpp:30-47), getSurfaceVoxel() (OptimizerAux.cpp:237-257), LED computeLightIntensive(). */
int psg_opt_init(psg_optimizer* opt);
int psg_opt_num_surface(const psg_optimizer* opt, int* S);
/* surface_points_ (Optimizer.h:44), ascending lin indices, and the keyframe-indexed visibility of each. */
int psg_opt_get_surface(psg_optimizer* opt, int* lin, uint64_t* vis_keyframes /* may be NULL */);
/* The SdfVoxel records (8 floats: dist, grad[3], weight, r, g, b -- Sdfvoxel.h:6-13) of the surface voxels, in the order
 * of psg_opt_get_surface: what Optimizer::save_pointcloud / extract_pc read through tsdf_[surface_points_[i]]
 * (OptimizerAux.cpp:456-511) without downloading the whole volume. */
int psg_opt_get_surface_voxels(psg_optimizer* opt, float* aos8 /* S x 8 */);

int psg_opt_light_len(const psg_optimizer* opt, int* n);  /* SH: F*4 or F*9 ; LED: 3 */
int psg_opt_get_light(psg_optimizer* opt, float* light);
int psg_opt_set_light(psg_optimizer* opt, const float* light);
int psg_opt_get_poses(psg_optimizer* opt, float* poses16);
int psg_opt_set_poses(psg_optimizer* opt, const float* poses16);
/* settings_->reg_weight_{rho,n,l}: the reference's outer loop rescales them (PsOptimizer.cpp:277,283,403) */
int psg_opt_get_reg(const psg_optimizer* opt, float reg3[3]);
int psg_opt_set_reg(psg_optimizer* opt, const float reg3[3]);

/* Overlapped keyframe upload (no counterpart in the reference, whose setImages deep-copies on the host): with on != 0,
 * psg_opt_set_keyframes(_u8) only queues the copy + conversion on a second stream and returns; psg_opt_init builds the work
 * list meanwhile and waits for the images before it returns.  The image buffer handed to set_keyframes must stay valid and
 * unmodified until psg_opt_init has returned (use pinned memory, otherwise the driver stages the copy synchronously). */
int psg_opt_set_upload_overlap(psg_optimizer* opt, int on);
/* Tuning knob without a counterpart in the reference: how many warps share the keyframes of one group of 32 surface
 * voxels in the sweep kernels.  0 (default) = chosen per launch from the surface size; 1 = one warp walks all keyframes
 * in ascending order, which reproduces the reference's per-voxel float summation order bit for bit; 2, 4, 8 = split. */
int psg_opt_set_sweep_split(psg_optimizer* opt, int warps);
/* Distance solve: 0 (default) = up to one row per resident thread (128^3, 256^3) the fused cooperative-grid kernel in the
 * Chronopoulos-Gear arrangement (one fused reduction and two grid barriers per CG iteration; block size chosen from the row count:
 * fewer, larger blocks make the grid barriers cheaper); beyond that the general persistent cooperative-grid kernel.  2 = always the
 * general kernel; 3 / 4 = the fused grid kernel in Eigen's arrangement (three barriers per iteration) with 512 / 1024 threads per
 * block; 6 = up to 32 768 rows one 16-CTA thread-block cluster with the CG vectors in distributed shared memory (hardware cluster
 * barrier, no global-memory traffic inside the solve), else the general kernel; 1 = one portable cluster of at most 8 CTAs with the
 * vectors in global memory; 5 = the general kernel with 256-thread blocks on large surfaces (default there: 1024); 8 = the dataflow
 * kernel of the sharded path on one GPU; 7 = measurement switch of that kernel.  All run the same iteration (in exact arithmetic)
 * and return the same iterates up to rounding, with the same iteration counts on every system tried. */
int psg_opt_set_pcg_mode(psg_optimizer* opt, int mode);

/* Optimizer::initAlbedo() (Optimizer.cpp:50-81) */
int psg_opt_init_albedo(psg_optimizer* opt);

/* getPSEnergy() (PsOptimizer.cpp:47-78, LedOptimizer.cpp:40-71): sum of robust losses over active
 * (surface voxel, keyframe) pairs divided by S, accumulated in double; n_active = number of such pairs. */
int psg_opt_energy(psg_optimizer* opt, double* E, long long* n_active);
/* Same result as psg_opt_energy, for hosts that must not trigger device allocations once they are running (all GPU
 * memory already committed to another owner, or a loop that cannot absorb the stall of an allocation).  It has no
 * counterpart in the reference.  The kernel's scratch (results, per-block partial sums, arrival counters) lives in one
 * ~4.2 MB array inside the loaded module, present on every GPU that loads the library until exit; the per-keyframe
 * poses and light are read from the optimizer's table in constant memory, which (like its copy in device memory) was
 * set aside when the optimizer was created and is refreshed by a 6 KB copy when poses or light changed.  Nothing is
 * allocated or freed.  Limits and rules:
 *   - at most 2 097 152 surface voxels and device ordinals below 64; larger inputs are rejected with PSG_ERR_INVALID;
 *   - the first call on a device looks up the array's device address, and the driver may load the kernel's code then
 *     (a driver-side, one-time cost); make that first call before any stream capture and before the phase in which
 *     allocations must not happen;
 *   - all calls on a device share the array (it keeps the previous call's contents, the parts that must be zero are
 *     cleared at the start of every call), so two calls that overlap on the GPU race -- serialise them. */
int psg_opt_energy_noalloc(psg_optimizer* opt, double* E, long long* n_active);

/* getNormalEnergy / getLaplacianEnergy / getAlbedoRegEnergy (Optimizer.cpp:86-136); any pointer may be NULL. */
int psg_opt_reg_energies(psg_optimizer* opt, double* E_normal, double* E_laplacian, double* E_albedo);

/* computeResidual() (PsOptimizerJa.cpp:567-626): residual and IRLS weight per (surface row, keyframe, channel),
 * S x F x 3 each, zeros for inactive pairs.  Diagnostic / parity probe; not used by the step functions. */
int psg_opt_residual(psg_optimizer* opt, float* r3, float* w3);
/* Per pair: getIntensity's projected pixel (col m, row n) and in-bounds flag (OptimizerAux.cpp:207-231),
 * sampled RGB and rendered RGB.  S x F (x 3).  Diagnostic / parity probe. */
int psg_opt_probe_pairs(psg_optimizer* opt, float* m, float* n, unsigned char* ok, float* I3, float* rend3);
/* Diagnostic: interpolateImage(row, col) and computeImageGradient(row, col, img, 0 | 1) (Auxilary.h:41-123) of keyframe `frame`
 * at n positions inside the image, evaluated by the device functions the sweeps use: rgb3, g0 (along columns), g1 (along
 * rows), 3 floats each per position.  In the bottom-right corner cell the reference reads one element past the image; this
 * path uses the centre texel there. */
int psg_debug_image_probe(psg_optimizer* opt, int frame, int n, const float* rows, const float* cols, float* rgb3, float* g0,
                          float* g1);

/* optimizeAlbedoAll / optimizeLightAll / optimizeDistAll / optimizePosesAll
 * (PsOptimizer.cpp:85-234, LedOptimizer.cpp:134-275).  flags: PSG_REG_* bits select the regularisers
 * the reference passes as booleans. */
#define PSG_REG_ALBEDO 1
#define PSG_REG_NORMAL 2
#define PSG_REG_LAPLACIAN 4
int psg_opt_step(psg_optimizer* opt, int block, float damping, int flags, psg_step_info* info /* may be NULL */);
/* ---- surface extraction ---------------------------------------------------------------------------------
 * Marching cubes on the device with the reference's exact output: the crop box of voxels with |dist| <= sqrt(3) * voxel_size
 * (OptimizerAux.cpp:278-308), MarchingCubes::computeIsoSurface over the (box - 2)^3 cells of it with size = voxel_size * box and
 * origin = -voxel_size * box_min (OptimizerAux.cpp:344-345, third/mesh/MarchingCubes.cpp:314-541): three new vertices per
 * triangle, cells in x, y, z order, triangles in table order, degenerate triangles dropped; vertices and 8-bit colours are
 * bit-identical to the reference's vertices_ / colors_ and face t is (3t, 3t+1, 3t+2).  voxel_size <= 0 means the volume's own. */
int psg_volume_mesh_box(psg_volume* vol, float voxel_size, int lo[3], int hi[3], int* any /* 0: no voxel near the surface */);
/* white != 0: save_white_mesh (OptimizerAux.cpp:365-453), every colour 255.  The triangles stay on the device until the next
 * extraction; *n_triangles = 0 also when the crop box is empty. */
int psg_volume_mesh_extract(psg_volume* vol, float voxel_size, int white, long long* n_triangles);
int psg_volume_mesh_download(psg_volume* vol, float* vertices /* n_triangles x 9, may be NULL */,
                             unsigned char* colors /* n_triangles x 9 (r, g, b per vertex), may be NULL */);
/* VolumetricGradSdf::extract_mesh (VolumetricGradSdf.cpp:234-318) + MarchingCubes::savePly (MarchingCubes.cpp:653-699): the
 * ASCII PLY the reference writes, byte for byte.  No file is written when there is no triangle (savePly returns false then);
 * *n_triangles (may be NULL) tells.  PSG_ERR_INVALID when the file cannot be opened. */
int psg_volume_extract_mesh(psg_volume* vol, const char* filename, int white, long long* n_triangles);
/* VolumetricGradSdf::extract_pc (VolumetricGradSdf.cpp:320-376): voxels within sqrt(3) * voxel_size of the surface seen by at
 * least one frame, in lin order: grid-relative position, unit gradient, colour. */
int psg_volume_extract_pc(psg_volume* vol, const char* filename, long long* n_points /* may be NULL */);
/* VolumetricGradSdf::saveSDF (VolumetricGradSdf.cpp:379-460): crop box dimensions, lower corner, voxel size, -dist values. */
int psg_volume_save_sdf(psg_volume* vol, const char* filename);
/* Optimizer::extract_mesh / save_white_mesh / saveSDF / save_pointcloud (OptimizerAux.cpp:278-577) with the full file name
 * instead of save_path_ + name + suffix; they read Optimizer::voxel_size_ where the volume's own writers read theirs. */
int psg_opt_extract_mesh(psg_optimizer* opt, const char* filename, int white, long long* n_triangles /* may be NULL */);
int psg_opt_save_sdf(psg_optimizer* opt, const char* filename);
int psg_opt_save_pointcloud(psg_optimizer* opt, const char* filename);

/* One pass of the reference's while-body (PsOptimizer.cpp:303-360, LedOptimizer.cpp:343-403): the enabled blocks in the
 * reference's order (SH: albedo, light, distance, pose; LED: light, albedo, distance, pose), the regulariser energies after
 * the distance block and the closing getPSEnergy().  Same kernels as four psg_opt_step calls + psg_opt_reg_energies +
 * psg_opt_energy, queued back to back: every block's solve and update runs on the device (the light / pose systems are solved
 * there by Eigen's CG in Eigen's operation order, the poses / light updated in the device's keyframe table), so the host waits
 * once, for everything, instead of 8 times.  The distance update's CG gate (applied only on Success, PsOptimizer.cpp:168-170)
 * and the LED pose update's (LedOptimizer.cpp:271-273) are evaluated on the device. */
typedef struct {
    int light, albedo, distance, pose;   /* alternatingOptimize(bool light, bool albedo, bool distance, bool pose) */
    float damping;
    int reg_flags;                       /* PSG_REG_NORMAL | PSG_REG_LAPLACIAN for the distance block, PSG_REG_ALBEDO for the albedo block */
    int closing_energy;                  /* evaluate the energy even if no block is enabled */
} psg_iteration_config;
typedef struct {
    int ran[4];                          /* indexed by PSG_BLOCK_* */
    psg_step_info block[4];
    double energy_after[4];              /* getPSEnergy() after each executed block */
    double reg_normal_after_dist, reg_laplacian_after_dist;   /* getNormalEnergy / getLaplacianEnergy after the distance block */
    double energy_end;                   /* = energy_after of the last executed block */
    long long n_active_end;
    double reg_albedo_after_albedo;      /* getAlbedoRegEnergy after the albedo block (reg_flags has PSG_REG_ALBEDO) */
} psg_iteration_info;
int psg_opt_iteration(psg_optimizer* opt, const psg_iteration_config* cfg, psg_iteration_info* info /* may be NULL */);

/* Normal equations of the last psg_opt_step for the small dense blocks (diagnostic / parity probe):
 * light: F blocks of nb x nb (+ nb rhs) or LED 3 (+3); pose: F blocks of 6 x 6 (+6).  Row-major, after damping. */
int psg_opt_last_system(psg_optimizer* opt, double* H_blocks, double* b, double* delta);
/* Distance / albedo block of the last step: rhs b, solution delta (n floats each). */
int psg_opt_last_vectors(psg_optimizer* opt, float* b, float* delta);

/* Optimizer::subsampling() (OptimizerAux.cpp:622-684): every voxel splits into 8; the volume handle is
 * updated in place (dims x2, voxel size /2, origin shifted, VoxelGrid.h:143-149) and the surface is re-extracted. */
int psg_opt_upsample(psg_optimizer* opt);

/* ------------------------------------------------------------------ frame front end on the device
 * cv::NormalEstimator<float> (cpp/include/normals/NormalEstimator.h:125-176; main_ps.cpp:115 builds it with an 11 x 11
 * window) and the loader's pixel conversions (img_loader/ImageLoader.h:130-146 depth, :166-188 colour), so that
 * VolumetricGradSdf::update (VolumetricGradSdf.cpp:51-138) can be fed the 16-bit depth and 8-bit colour the loader read
 * from disk.  Results are OpenCV's bit for bit (tests/golden/normals.npz, OpenCV 4.13). */
typedef struct psg_normals psg_normals;
/* NormalEstimator(width, height, K, Size(window, window)): builds the per-pixel tables of cache() (:52-123). */
int psg_normals_create(psg_normals** out, int width, int height, const float K[9], int window, int device);
int psg_normals_destroy(psg_normals* ne);
/* The cached tables, 9 planes of height*width floats: x0_n_sq_inv, y0_n_sq_inv, n_sq_inv, Q11, Q12, Q13, Q22, Q23, Q33
 * (plane 2 is what update() receives through NEst->n_sq_inv_ptr(), VolumetricGradSdf.cpp:66). */
int psg_normals_tables(psg_normals* ne, float* planes9);
/* NormalEstimator::compute(depth, nx, ny, nz) (:148-176): host depth in metres (0 = no measurement) -> host unit normals;
 * NaN where the whole window is empty, as OpenCV >= 4 produces. */
int psg_normals_compute(psg_normals* ne, const float* depth, float* nx, float* ny, float* nz);
/* load_depth + load_color + NEst->compute + update() for one frame in the loader's raw format: depth_u16 (height*width,
 * metres = value * depth_unit), color_bgr_u8 (height*width*3).  Same result as converting and estimating on the host and
 * calling psg_volume_integrate; 5 bytes per pixel cross PCIe instead of 32. */
int psg_volume_integrate_raw(psg_volume* vol, psg_normals* ne, const unsigned char* color_bgr_u8, const unsigned short* depth_u16,
                             float depth_unit, const float K[9], const float R[9], const float t[3], int increase_counter);
/* psg_opt_set_keyframes with the keyframe images as the loader read them (8-bit BGR, F*H*W*3 bytes): converted on the
 * device with load_color's float(src) * (1.0f / 255.0f), so the texels equal those of the float entry point. */
int psg_opt_set_keyframes_u8(psg_optimizer* opt, int F, const int* frame_idx, const float* poses4x4, const unsigned char* images_bgr_u8,
                             int H, int W);

/* ------------------------------------------------------------------ keyframe gate and volume placement
 * sharpDetector's measure, modifiedLaplacian(img) (cpp/include/ps_optimizer/SharpDetector.h:12-37): mean over the blue channel
 * of |G_col(M_row(x))| + |M_col(G_row(x))|, M = (-1, 2, -1), G = (0.25, 0.5, 0.25), reflect-101 border -- what
 * cv::sepFilter2D + cv::mean(FM).val[0] return for the CV_32FC3 colour image (bit-equal to OpenCV 4.13 in the fixture).
 * The image has the size the estimator was created with; the caller compares with its sharpness threshold. */
int psg_normals_sharpness(psg_normals* ne, const float* color_bgr, float* measure);
int psg_normals_sharpness_u8(psg_normals* ne, const unsigned char* color_bgr_u8, float* measure);
/* compute_centroid(K, depth, Trans) (voxel_ps/src/main_ps.cpp:346-375): where main() centres the volume.  Host code: the
 * serial float sum is part of the result. */
int psg_compute_centroid(const float K[9], const float* depth, int H, int W, const float pose4x4[16], float centroid[3]);
/* sampleKeyFrame(..., max_num) (main_ps.cpp:392-421, called with 40): positions of the keyframes that are kept (kept must
 * hold max(n, max_num) ints). */
int psg_sample_keyframes(int n, int max_num, int* kept, int* n_kept);

/* ------------------------------------------------------------------ rigid depth-to-SDF tracking
 * The per-pixel part of RigidPointOptimizer::optimize_sampled (cpp/include/sdf_tracker/RigidPointOptimizer.cpp:38-62) with
 * Sdf::weights / Sdf::tsdf of the gradient-SDF volume (VolumetricGradSdf.h:76-93).  The host keeps the reference's 6 x 6
 * solve and pose update (:67-75, Eigen LLT + Sophus SE3::exp) and calls psg_tracker_system once per Gauss-Newton iteration. */
/* the frame to track: depth in metres (H*W floats, host), or as the loader read it (16-bit, metres = value * unit) */
int psg_tracker_set_depth(psg_volume* vol, const float* depth, int H, int W);
int psg_tracker_set_depth_u16(psg_volume* vol, const unsigned short* depth_u16, float depth_unit, int H, int W);
/* at pose (R, t) (camera to world, SE3::rotationMatrix / translation): E = sum phi^2 (the reference divides by count
 * afterwards), g = sum phi J, H = sum J J^T (row-major 6 x 6), count = pixels that hit a fused voxel; zmin / zmax are the
 * volume's.  Sums are formed in double; which pixels count is decided bit-identically to the reference. */
int psg_tracker_system(psg_volume* vol, const float K[9], const float R[9], const float t[3], int sampling, double* E_sum,
                       double g6[6], double H36[36], long long* count);

/* RigidPointOptimizer(num_iterations, conv_threshold, damping, tSDF)::optimize_sampled(depth, K, sampling) from pose_in
 * (row-major 4 x 4, camera to world) on the frame given to psg_tracker_set_depth: Gauss-Newton with the device sums above,
 * xi = damping * H.llt().solve(g), converged when |xi|^2 < conv_threshold^2, else pose <- SE3::exp(-xi) * pose
 * (RigidPointOptimizer.cpp:26-77).  The 6 x 6 solve and the SE(3) update are float host code written after Eigen's LLT and
 * Sophus' SE3::exp; against the reference they agree to rounding, not bit for bit. */
typedef struct { int num_iterations; float conv_threshold; float damping; int sampling; } psg_tracker_config;
int psg_tracker_optimize(psg_volume* vol, const float K[9], const float pose_in[16], const psg_tracker_config* cfg,
                         float pose_out[16], int* converged, int* iterations);

/* ---- several GPUs (no counterpart in the reference, which is one CPU process; BASELINE.json north_star) --------------------
 * One rank per GPU.  Every rank creates its volume, fuses the same frames and hands over the same keyframes; after
 * psg_opt_comm_init the calls psg_opt_init / init_albedo / energy / step / iteration / alternating_optimize / upsample /
 * reg_energies / sync_volume / get_surface_voxels / set_shard_policy / comm_destroy are COLLECTIVE: all ranks call them in the
 * same order with the same arguments.
 *
 * Sharded mode: the surface list is in ascending voxel order, so a contiguous ROW RANGE of it is a z-slab
 * of the volume; rank g owns one range (equal numbers of visible pairs per rank) and does all per-row work on it -- sweeps,
 * distance stencil blocks, the CG of the distance block, the gated updates, the regulariser energies.  Across ranks go: the
 * energies and the per-keyframe light / pose normal equations (three small all-reduces per outer iteration, stream ordered, as
 * tagged packets written into the peers' memory), and, inside the persistent distance-solve kernel, the one-voxel halo of the CG
 * vectors plus three scalars per reduction, also written straight into the peers' memory over NVLink (CUDA IPC mappings) as
 * tagged packets.  NCCL carries the set-up (unique id, handle exchange, barriers) and psg_opt_sync_volume's all-gather.  Energies, light, poses and CG verdicts are identical on
 * all ranks; of the VOLUME a rank keeps only its own rows (and the distance of its halo) current until psg_opt_sync_volume.
 * Diagnostics that return per-row data (psg_opt_probe_pairs, psg_opt_residual, psg_opt_last_vectors) cover the rank's own rows.
 * reg albedo != 0 is not provided in this mode.
 * Replicated mode: every rank sweeps everything and rank 0's reduced sums are broadcast before the small solves, so that all
 * ranks stay in the same state bit for bit; nothing is gained, nothing lost -- the fallback for work lists too small to split. */
int psg_comm_unique_id(unsigned char id[128]);                                   /* rank 0; send the bytes to the other ranks */
int psg_opt_comm_init(psg_optimizer* opt, const unsigned char id[128], int rank, int world);   /* one process per GPU, NCCL */
/* The ranks of a job as THREADS of one process whose optimizers share a device: opts[q] becomes rank q.  Same code path as
 * above with the peers' pointers taken directly and the collectives done by whichever thread arrives last; every collective
 * call must then be made by `world` threads concurrently, one per optimizer.  For testing the multi-rank logic on one GPU:
 * the distributed distance solve runs as one cooperative launch whose blocks play all ranks. */
int psg_opt_comm_init_local(psg_optimizer* const* opts, int world);
int psg_opt_comm_destroy(psg_optimizer* opt);
/* Sharding policy.  0 (default): shard when the work list has at least three million visible (voxel, keyframe) pairs (below
 * that the cross-GPU latency of the distance solve costs more than the split sweeps save: 128^3 and 256^3 x 30 keyframes stay
 * replicated, 512^3 x 30 and 1024^3 x 64 are sharded), 1: always shard, 2: never.  Switching re-partitions at the next call
 * that touches the work list. */
int psg_opt_set_shard_policy(psg_optimizer* opt, int policy);
int psg_opt_is_sharded(const psg_optimizer* opt, int* sharded, long long* visible_pairs);
/* The rows [row_lo, row_hi) of the surface list this rank owns and the rows [halo_lo, row_lo) + [row_hi, halo_hi) it receives
 * from its neighbours (all rows / no halo unless sharded); any pointer may be NULL. */
int psg_opt_row_range(psg_optimizer* opt, int* row_lo, int* row_hi, int* halo_lo, int* halo_hi);
/* Sharded mode: every rank receives the other ranks' rows of the volume planes (distance, gradient, albedo), so that
 * psg_volume_download, the mesh / point-cloud / .sdf writers and psg_opt_upsample see the whole reconstruction.  Called by
 * psg_opt_alternating_optimize before every dump and at the end, by psg_opt_upsample, psg_opt_get_surface_voxels and
 * psg_opt_comm_destroy.  A no-op when nothing changed or the job is not sharded. */
int psg_opt_sync_volume(psg_optimizer* opt);
/* The partition rule on host arrays (no device involved): cut points bounds[world + 1] over a surface list lin[S] (ascending)
 * with keyframe visibility vis[S], equal visible-pair counts per range, and per rank the first / one-past-last row of its halo
 * for a z-plane of `plane` = W * H voxels. */
int psg_partition_rows(int S, const int* lin, const uint64_t* vis, long long plane, int world, int* bounds, int* halo_lo, int* halo_hi);
/* Which 32-slot blocks of a slot list `rank` of `world` takes in a round-robin split (blocks first, first + stride, ...):
 * used by psg_opt_energy_partial to emulate ranks on one list. */
int psg_shard_blocks(int S, int rank, int world, int* first, int* stride, int* count);
/* Diagnostic: the partial energy sum and pair count over every world-th block of this rank's slot list, computed locally
 * without any communication (the partials of all `world` emulated ranks add up to the list's energy sum). */
int psg_opt_energy_partial(psg_optimizer* opt, int rank, int world, double* loss_sum, long long* n_active);

/* {Ps,Led}Optimizer::alternatingOptimize(light, albedo, distance, pose) (PsOptimizer.cpp:239-428,
 * LedOptimizer.cpp:279-478) with the settings it reads from OptimizerSettings (max_it, conv_threshold, damping, upsample,
 * OptimizerSettings.h:24-33).  Runs initAlbedo, the regulariser-weight normalisation and the outer loop; writes
 * <save_path>optimizer_doc.txt, after_poses_opt_<it>.txt, after_iter_<it>_pointcloud.ply, upsample_after_5_pointcloud.ply,
 * final_refined_pointcloud.ply (LED: refined_sdf.sdf) when write_artifacts != 0 (save_path must exist and end in '/',
 * as in the reference), and beside every point cloud the mesh the reference extracts there (extract_mesh, PsOptimizer.cpp:373,380,
 * 398,422): after_iter_<it>_mesh.ply, upsample_after_5_mesh.ply, final_refined_mesh.ply. */
typedef struct {
    int light, albedo, distance, pose;   /* the --light/--albedo/--distance/--pose switches (main_ps.cpp:79-82) */
    int max_it;
    float conv_threshold;
    float damping;
    int upsample;
    int write_artifacts;
    int verbose;                         /* echo the reference's stdout lines */
} psg_loop_config;
typedef struct {
    int iterations;                      /* outer iterations executed */
    int converged;                       /* rel_diff < conv_threshold (the reference returns true) */
    int diverged;                        /* total energy went up (the reference returns false) */
    float last_rel_diff;
    double E, E_normal, E_laplacian, E_albedo, E_total;   /* after the last block */
    double ms_per_iteration;             /* host wall clock over the loop body, file dumps excluded */
    float iteration_ms[64];
} psg_loop_result;
/* Diagnostic: microseconds the calling thread spent inside psg_opt_iteration since the last reset, split into
 * us[0] queuing kernels, us[1] waiting for the device's results, us[2] bookkeeping after the wait (the step records). */
int psg_debug_host_times(double us[3], long long* iterations, int reset);
/* Diagnostic: CUDA events on the optimizer's own stream (events of another library see only their own stream).  begin
 * records an event; end records a second one, waits for it and returns the device time between the two in milliseconds:
 * the device-side duration of whatever was called in between (idle gaps while the host works included). */
/* Diagnostic: which distance-solve kernel the last distance block used: 0 general cooperative grid, 1 portable 8-CTA cluster,
 * 2 fused grid kernel, 3 16-CTA cluster over distributed shared memory, 4 dataflow kernel (sharded job, or mode 8), 5 fused grid
 * kernel in the Chronopoulos-Gear arrangement; -1 none yet. */
int psg_debug_last_pcg_kernel(const psg_optimizer* opt, int* kernel);
int psg_debug_timer_begin(psg_optimizer* opt);
int psg_debug_timer_end(psg_optimizer* opt, float* ms);
/* Diagnostic, host only: the Jacobi-preconditioned CG the light and pose blocks are solved with (Eigen::ConjugateGradient<
 * SparseMatrix<float>, Lower, DiagonalPreconditioner>, PsOptimizer.cpp:184-191,217-227) on nblocks dense bs x bs blocks (row
 * major, lower triangle read). */
int psg_debug_small_cg(int nblocks, int bs, const float* H, const float* b, float* x, int* success, int* iterations, float* error);
int psg_opt_alternating_optimize(psg_optimizer* opt, const psg_loop_config* cfg, const char* save_path,
                                 psg_loop_result* result /* may be NULL */);

#ifdef __cplusplus
}
#endif
#endif /* PSG_H */
