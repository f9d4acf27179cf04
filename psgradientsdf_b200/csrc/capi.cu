// C ABI (include/psg.h) and the host-side control of the optimizer blocks: small dense solves, pose / light updates.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <chrono>
#include <cstdlib>
#include <cstring>

#include <atomic>
#include <mutex>

#include "host_solve.h"
#include "psg_host.h"

namespace psg {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void count_launch(int n) { g_launches += n; }
void trace(const char* fmt, ...) {
    static const int on = getenv("PSG_TRACE") ? 1 : 0;
    if (!on) return;
    va_list ap;
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    fputc('\n', stderr);
    fflush(stderr);
}

// ---- launch timer: CUDA events on the launching stream around every kernel, summed per kernel id on request
static const char* const kKernelNames[KID_COUNT] = {
    "k_fill_volume", "k_pack_frame", "k_integrate", "k_aos_soa", "k_surface_mask", "k_scan_counts", "k_surface_scatter",
    "k_neighbors", "k_upsample", "k_sweep<ENERGY>", "k_sweep<INIT_ALBEDO>", "k_sweep<LED_INIT>", "k_sweep<PROBE>",
    "k_sweep<ALBEDO>", "k_sweep<LIGHT>", "k_sweep<POSE>", "k_sweep<DIST>", "k_reg_energy", "k_dist_assemble", "k_dist_pcg",
    "k_dist_update", "k_grad_update", "k_pad_images", "k_normals", "k_comm_scatter", "allreduce_small_sums", "nccl_allgather", "k_track", "k_mesh_bbox", "k_mesh_cells<count>", "k_mesh_scan", "k_mesh_cells<emit>", "k_mesh_misc", "k_small_solve" };
static bool g_prof_on = false;
struct ProfRec { int id; cudaEvent_t e0, e1; };
static std::vector<ProfRec> g_prof_recs;
static long long g_prof_count[KID_COUNT] = {0};
static double g_prof_ms[KID_COUNT] = {0};

static std::mutex g_prof_mutex;                 // ranks as threads of one process (comm.cu) launch concurrently
static std::vector<cudaEvent_t> g_event_pool;   // events are recycled: creating two per launch costs more than the launch
static cudaEvent_t pool_get() {
    if (!g_event_pool.empty()) { cudaEvent_t e = g_event_pool.back(); g_event_pool.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}
Prof::Prof(cudaStream_t stream, int kernel_id) : st(stream), id(kernel_id) {
    if (kernel_id != KID_COMM_ALLREDUCE && kernel_id != KID_COMM_ALLGATHER) g_launches += 1;   // NCCL's kernels are not ours
    if (g_prof_on) {
        { std::lock_guard<std::mutex> lk(g_prof_mutex); e0 = pool_get(); e1 = pool_get(); }
        cudaEventRecord(e0, st);
    }
}
Prof::~Prof() {
    if (e0) { cudaEventRecord(e1, st); std::lock_guard<std::mutex> lk(g_prof_mutex); g_prof_recs.push_back({id, e0, e1}); }
}
static void prof_collect() {
    std::lock_guard<std::mutex> lk(g_prof_mutex);
    for (const ProfRec& r : g_prof_recs) {
        float ms = 0.f;
        if (cudaEventSynchronize(r.e1) == cudaSuccess && cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
            g_prof_count[r.id] += 1; g_prof_ms[r.id] += (double)ms;
        }
        g_event_pool.push_back(r.e0); g_event_pool.push_back(r.e1);
    }
    g_prof_recs.clear();
}

static void sh_eval_host(const float* n, int nb, float* sh) {
    sh[0] = 1.f; sh[1] = n[0]; sh[2] = n[1]; sh[3] = n[2];
    if (nb == 9) {
        sh[4] = n[0] * n[1]; sh[5] = n[0] * n[2]; sh[6] = n[1] * n[2];
        sh[7] = n[0] * n[0] - n[1] * n[1]; sh[8] = n[0] * n[0] - n[2] * n[2];
    }
}

// All transfers run on the handle's (non-blocking) stream: a plain cudaMemcpy would not order against it.
static void copy_sync(void* dst, const void* src, size_t bytes, cudaMemcpyKind kind, cudaStream_t st) {
    PSG_CUDA(cudaMemcpyAsync(dst, src, bytes, kind, st));
    PSG_CUDA(cudaStreamSynchronize(st));
}

// float4-padded BGR texels
__global__ void __launch_bounds__(256) k_pad_images(const float* __restrict__ src, float4* __restrict__ dst, long long npix) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    dst[i] = make_float4(src[3 * i], src[3 * i + 1], src[3 * i + 2], 0.f);
}

}  // namespace psg

using namespace psg;

static void ensure_upload_stream(psg_optimizer* o) {
    if (!o->upload_stream) PSG_CUDA(cudaStreamCreateWithFlags(&o->upload_stream, cudaStreamNonBlocking));
    if (!o->upload_done) PSG_CUDA(cudaEventCreateWithFlags(&o->upload_done, cudaEventDisableTiming));
}
// the keyframe texels are complete on the device and the caller's buffer is free again
static void finish_upload(psg_optimizer* o) {
    if (!o->upload_pending) return;
    PSG_CUDA(cudaStreamWaitEvent(o->stream, o->upload_done, 0));
    PSG_CUDA(cudaEventSynchronize(o->upload_done));
    o->upload_pending = false;
}

void psg_optimizer::ensure_host_staging() {
    if (host_staging) return;
    void* p = nullptr;
    const size_t bytes = (kRedDoubles + kMailDoubles) * sizeof(double);
    PSG_CUDA(cudaMallocHost(&p, bytes));
    host_staging = static_cast<double*>(p);
    PSG_CUDA(cudaMallocHost(&p, sizeof(psg::FrameTable)));
    host_frames = static_cast<psg::FrameTable*>(p);
}

#define PSG_TRY try {
#define PSG_CATCH                                                      \
    }                                                                  \
    catch (const psg::CudaError&) { return PSG_ERR_CUDA; }             \
    catch (const std::exception& e) { psg::set_error("%s", e.what()); return PSG_ERR_INVALID; }

extern "C" {

const char* psg_last_error(void) { return g_err; }
long long psg_kernel_launches(void) { return g_launches.load(); }

int psg_profile_enable(int on) {
    prof_collect();
    g_prof_on = on != 0;
    return PSG_OK;
}
int psg_profile_reset(void) {
    prof_collect();
    for (int i = 0; i < KID_COUNT; ++i) { g_prof_count[i] = 0; g_prof_ms[i] = 0.0; }
    return PSG_OK;
}
int psg_profile_kernels(void) { return KID_COUNT; }
const char* psg_profile_name(int id) { return (id >= 0 && id < KID_COUNT) ? kKernelNames[id] : ""; }
int psg_profile_get(int id, long long* launches, double* total_ms) {
    if (id < 0 || id >= KID_COUNT) { set_error("psg_profile_get: bad kernel id %d", id); return PSG_ERR_INVALID; }
    prof_collect();
    if (launches) *launches = g_prof_count[id];
    if (total_ms) *total_ms = g_prof_ms[id];
    return PSG_OK;
}

int psg_device_count(int* count) {
    PSG_TRY
    PSG_CUDA(cudaGetDeviceCount(count));
    return PSG_OK;
    PSG_CATCH
}

// ------------------------------------------------------------------ volume
int psg_volume_create(psg_volume** out, const int dims[3], float voxel_size, const float shift[3], float T, float zmin,
                      float zmax, int max_frames, int device) {
    if (!out || !dims || !shift) { set_error("psg_volume_create: null argument"); return PSG_ERR_INVALID; }
    if (dims[0] <= 0 || dims[1] <= 0 || dims[2] <= 0 || (long long)dims[0] * dims[1] * dims[2] > (1ll << 31) - 64 ||
        dims[1] > 65535 || dims[2] > 65535) {   // H and D are CUDA grid dimensions of the fusion launch
        set_error("psg_volume_create: grid %d x %d x %d not supported (needs 0 < W*H*D < 2^31, H and D <= 65535)", dims[0], dims[1], dims[2]);
        return PSG_ERR_INVALID;
    }
    if (max_frames < 1) { set_error("psg_volume_create: max_frames must be >= 1"); return PSG_ERR_INVALID; }
    psg_volume* v = nullptr;
    PSG_TRY
    int ndev = 0;
    PSG_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { set_error("psg_volume_create: device %d not available (%d devices)", device, ndev); return PSG_ERR_CUDA; }
    PSG_CUDA(cudaSetDevice(device));
    v = new psg_volume;
    v->device = device;
    for (int a = 0; a < 3; ++a) { v->dims[a] = dims[a]; v->shift[a] = shift[a]; }
    v->N = (long long)dims[0] * dims[1] * dims[2];
    v->vs = voxel_size;
    // VoxelGrid(grid_dim, voxel_size, shift): origin = shift - 0.5 * voxel_size * dim (VoxelGrid.h:127-133)
    const float half_vs = (float)(0.5 * (double)voxel_size);
    for (int a = 0; a < 3; ++a) v->origin[a] = shift[a] - half_vs * (float)dims[a];
    v->T = T; v->invT = (float)(1.0 / (double)T);   // Sdf(T): inv_T_(1./T_) (Sdf.h:83-87)
    v->zmin = zmin; v->zmax = zmax;
    v->max_frames = max_frames; v->nwords = (max_frames + 63) / 64;
    PSG_CUDA(cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking));
    for (int a = 0; a < 8; ++a) v->f[a].alloc((size_t)v->N);
    v->vis.alloc((size_t)v->N * v->nwords);
    launch_fill_volume(v);
    PSG_CUDA(cudaStreamSynchronize(v->stream));
    *out = v;
    return PSG_OK;
    }
    catch (const psg::CudaError&) { if (v) psg_volume_destroy(v); return PSG_ERR_CUDA; }
}

int psg_volume_destroy(psg_volume* v) {
    if (!v) return PSG_OK;
    cudaSetDevice(v->device);
    for (int a = 0; a < 8; ++a) v->f[a].release();
    v->vis.release(); v->frame.release(); v->staging.release();
    v->brick_list.release(); v->tiles.release(); v->track_depth.release(); v->track_partials.release();
    v->mesh_counts.release(); v->mesh_crossed.release(); v->mesh_box.release(); v->mesh_vertices.release(); v->mesh_colors.release();
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
    return PSG_OK;
}

int psg_volume_geometry(const psg_volume* v, int dims[3], float* voxel_size, float origin[3]) {
    if (!v) { set_error("null volume"); return PSG_ERR_INVALID; }
    for (int a = 0; a < 3; ++a) { if (dims) dims[a] = v->dims[a]; if (origin) origin[a] = v->origin[a]; }
    if (voxel_size) *voxel_size = v->vs;
    return PSG_OK;
}

int psg_volume_integrate(psg_volume* v, const float* color_bgr, const float* depth, const float* nx, const float* ny,
                         const float* nz, const float* n_sq_inv, int H, int W, const float K[9], const float R[9],
                         const float t[3], int increase_counter) {
    if (!v || !color_bgr || !depth || !nx || !ny || !nz || !n_sq_inv || !K || !R || !t || H <= 0 || W <= 0) {
        set_error("psg_volume_integrate: null or empty argument");
        return PSG_ERR_INVALID;
    }
    const int counter = v->counter + (increase_counter ? 1 : 0);
    if (counter >= 64 * v->nwords) {
        set_error("psg_volume_integrate: frame %d exceeds max_frames %d given at creation", counter, v->max_frames);
        return PSG_ERR_INVALID;
    }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(v->device));
    const size_t HW = (size_t)H * W;
    v->staging.alloc(HW * 8);
    float* st = v->staging.p;
    PSG_CUDA(cudaMemcpyAsync(st, color_bgr, HW * 3 * sizeof(float), cudaMemcpyHostToDevice, v->stream));
    PSG_CUDA(cudaMemcpyAsync(st + 3 * HW, depth, HW * sizeof(float), cudaMemcpyHostToDevice, v->stream));
    PSG_CUDA(cudaMemcpyAsync(st + 4 * HW, nx, HW * sizeof(float), cudaMemcpyHostToDevice, v->stream));
    PSG_CUDA(cudaMemcpyAsync(st + 5 * HW, ny, HW * sizeof(float), cudaMemcpyHostToDevice, v->stream));
    PSG_CUDA(cudaMemcpyAsync(st + 6 * HW, nz, HW * sizeof(float), cudaMemcpyHostToDevice, v->stream));
    PSG_CUDA(cudaMemcpyAsync(st + 7 * HW, n_sq_inv, HW * sizeof(float), cudaMemcpyHostToDevice, v->stream));
    launch_pack_frame(v, st, H, W);
    v->counter = counter;
    v->generation++;
    launch_integrate(v, H, W, K, R, t, counter);
    PSG_CUDA(cudaStreamSynchronize(v->stream));
    return PSG_OK;
    PSG_CATCH
}

static const long long kChunk = 1ll << 24;   // voxels per staging chunk (512 MiB of AoS)

int psg_volume_download(psg_volume* v, float* aos8) {
    if (!v || !aos8) { set_error("psg_volume_download: null argument"); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(v->device));
    const long long chunk = v->N < kChunk ? v->N : kChunk;
    v->staging.alloc((size_t)chunk * 8);
    for (long long b = 0; b < v->N; b += chunk) {
        const long long c = (v->N - b) < chunk ? (v->N - b) : chunk;
        launch_soa_to_aos(v, v->staging.p, b, c);
        PSG_CUDA(cudaMemcpyAsync(aos8 + 8 * b, v->staging.p, (size_t)c * 8 * sizeof(float), cudaMemcpyDeviceToHost, v->stream));
        PSG_CUDA(cudaStreamSynchronize(v->stream));
    }
    return PSG_OK;
    PSG_CATCH
}

int psg_volume_upload(psg_volume* v, const float* aos8) {
    if (!v || !aos8) { set_error("psg_volume_upload: null argument"); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(v->device));
    const long long chunk = v->N < kChunk ? v->N : kChunk;
    v->staging.alloc((size_t)chunk * 8);
    for (long long b = 0; b < v->N; b += chunk) {
        const long long c = (v->N - b) < chunk ? (v->N - b) : chunk;
        PSG_CUDA(cudaMemcpyAsync(v->staging.p, aos8 + 8 * b, (size_t)c * 8 * sizeof(float), cudaMemcpyHostToDevice, v->stream));
        launch_aos_to_soa(v, v->staging.p, b, c);
        PSG_CUDA(cudaStreamSynchronize(v->stream));
    }
    v->generation++;
    return PSG_OK;
    PSG_CATCH
}

int psg_volume_vis_words(const psg_volume* v, int* nwords) {
    if (!v || !nwords) { set_error("null argument"); return PSG_ERR_INVALID; }
    *nwords = v->nwords;
    return PSG_OK;
}
int psg_volume_download_vis(psg_volume* v, uint64_t* words) {
    if (!v || !words) { set_error("null argument"); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(v->device));
    copy_sync(words, v->vis.p, (size_t)v->N * v->nwords * sizeof(uint64_t), cudaMemcpyDeviceToHost, v->stream);
    return PSG_OK;
    PSG_CATCH
}
int psg_volume_upload_vis(psg_volume* v, const uint64_t* words, int frame_counter) {
    if (!v || !words) { set_error("null argument"); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(v->device));
    copy_sync(v->vis.p, words, (size_t)v->N * v->nwords * sizeof(uint64_t), cudaMemcpyHostToDevice, v->stream);
    v->counter = frame_counter;
    return PSG_OK;
    PSG_CATCH
}

// ------------------------------------------------------------------ optimizer
int psg_opt_create(psg_optimizer** out, psg_volume* vol, float voxel_size, const float K[9], const psg_settings* set) {
    if (!out || !vol || !K || !set) { set_error("psg_opt_create: null argument"); return PSG_ERR_INVALID; }
    if (set->model < 0 || set->model > 2) { set_error("psg_opt_create: unknown model %d", set->model); return PSG_ERR_INVALID; }
    if (set->loss < 0 || set->loss > 4) { set_error("psg_opt_create: unknown loss %d", set->loss); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(vol->device));
    const int cslot = psg::acquire_const_slot(vol->device);   // fails when kConstSlots optimizers are alive on this device
    psg_optimizer* o = new psg_optimizer;
    o->cslot = cslot;
    o->vol = vol; o->set = *set;
    o->nb = set->model == PSG_MODEL_SH2 ? 9 : 4;
    o->vs = voxel_size; o->vs_inv = 1.f / voxel_size;   // Optimizer.cpp:22-23
    memcpy(o->K, K, sizeof(o->K));
    o->stream = vol->stream;
    o->frames_dev.alloc(psg::kMaxFrames + 1);
    o->small_sys.alloc((size_t)psg::kMaxFrames * 81 + 2 * psg::kMaxFrames * 9);
    *out = o;
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_destroy(psg_optimizer* o) {
    if (!o) return PSG_OK;
    cudaSetDevice(o->vol->device);
    psg::comm_destroy(o, /*collective=*/false);
    o->xstate.release(); o->comm_buf.release();
    if (o->upload_stream) { cudaStreamSynchronize(o->upload_stream); cudaStreamDestroy(o->upload_stream); }
    if (o->upload_done) cudaEventDestroy(o->upload_done);
    if (o->timer0) { cudaEventDestroy(o->timer0); cudaEventDestroy(o->timer1); }
    o->upload_staging.release();
    o->images.release(); o->lin.release(); o->nbr_lin.release(); o->nbr_row.release(); o->inc.release();
    o->svis.release(); o->pos.release(); o->rec0.release(); o->rec1.release(); o->rec2.release(); o->rec3.release(); o->rec4.release();
    o->pvis.release(); o->perm.release(); o->sort_hist.release(); o->perm2.release(); o->block_pairs.release(); o->mask.release(); o->prefix.release(); o->block_counts.release();
    if (o->host_staging) cudaFreeHost(o->host_staging);
    if (o->host_frames) cudaFreeHost(o->host_frames);
    o->frames_dev.release(); o->small_sys.release();
    psg::release_const_slot(o->vol->device, o->cslot);
    o->areg.release(); o->mg_args.release(); o->xplanes.release(); o->halo.release();
    o->red.release(); o->sweep_partials.release(); o->sweep_counter.release(); o->vecs.release(); o->pcg_partials.release(); o->pcg_pbuf.release();
    delete o;
    return PSG_OK;
}

int psg_opt_set_keyframes(psg_optimizer* o, int F, const int* frame_idx, const float* poses16, const float* images,
                          int H, int W) {
    if (!o || (F > 0 && (!frame_idx || !poses16 || !images)) || F < 0 || H <= 0 || W <= 0) {
        set_error("psg_opt_set_keyframes: null or empty argument");
        return PSG_ERR_INVALID;
    }
    if (F > kMaxFrames) { set_error("psg_opt_set_keyframes: %d keyframes given, at most %d supported", F, kMaxFrames); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    if (F != o->F) o->key_stamps.clear();
    o->F = F; o->H = H; o->W = W;
    o->frame_idx.assign(frame_idx, frame_idx + F);
    o->poses.assign(poses16, poses16 + 16 * (size_t)F);
    const long long npix = (long long)F * H * W;
    o->images.alloc((size_t)(npix > 0 ? npix : 1));
    if (o->upload_overlap && npix > 0) {
        // all keyframes in one copy + one conversion on the upload stream; psg_opt_init waits for it after it has built the
        // work list (which does not need the images)
        ensure_upload_stream(o);
        o->upload_staging.alloc((size_t)npix * 3);
        PSG_CUDA(cudaMemcpyAsync(o->upload_staging.p, images, (size_t)npix * 3 * sizeof(float), cudaMemcpyHostToDevice, o->upload_stream));
        { Prof pf(o->upload_stream, KID_PAD_IMAGES);
          k_pad_images<<<(unsigned)((npix + 255) / 256), 256, 0, o->upload_stream>>>(o->upload_staging.p, o->images.p, npix); }
        PSG_LAUNCH_CHECK();
        PSG_CUDA(cudaEventRecord(o->upload_done, o->upload_stream));
        o->upload_pending = true;
        o->frames_dirty = true;
        o->inited = false;
        return PSG_OK;
    }
    // stage one keyframe at a time through the volume's staging buffer
    const long long per = (long long)H * W;
    o->vol->staging.alloc((size_t)per * 3);
    for (int f = 0; f < F; ++f) {
        PSG_CUDA(cudaMemcpyAsync(o->vol->staging.p, images + (size_t)f * per * 3, (size_t)per * 3 * sizeof(float),
                                 cudaMemcpyHostToDevice, o->stream));
        { Prof pf(o->stream, KID_PAD_IMAGES);
          k_pad_images<<<(unsigned)((per + 255) / 256), 256, 0, o->stream>>>(o->vol->staging.p, o->images.p + (size_t)f * per, per); }
        PSG_LAUNCH_CHECK();
    }
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    o->frames_dirty = true;
    o->inited = false;
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_set_keyframes_u8(psg_optimizer* o, int F, const int* frame_idx, const float* poses16, const unsigned char* images_bgr_u8,
                             int H, int W) {
    if (!o || (F > 0 && (!frame_idx || !poses16 || !images_bgr_u8)) || F < 0 || H <= 0 || W <= 0) {
        set_error("psg_opt_set_keyframes_u8: null or empty argument");
        return PSG_ERR_INVALID;
    }
    if (F > kMaxFrames) { set_error("psg_opt_set_keyframes_u8: %d keyframes given, at most %d supported", F, kMaxFrames); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    if (F != o->F) o->key_stamps.clear();
    o->F = F; o->H = H; o->W = W;
    o->frame_idx.assign(frame_idx, frame_idx + F);
    o->poses.assign(poses16, poses16 + 16 * (size_t)F);
    const long long npix = (long long)F * H * W;
    o->images.alloc((size_t)(npix > 0 ? npix : 1));
    if (o->upload_overlap && npix > 0) {
        ensure_upload_stream(o);
        o->upload_staging.alloc((size_t)(npix * 3 + 3) / 4);
        unsigned char* raw = reinterpret_cast<unsigned char*>(o->upload_staging.p);
        PSG_CUDA(cudaMemcpyAsync(raw, images_bgr_u8, (size_t)npix * 3, cudaMemcpyHostToDevice, o->upload_stream));
        launch_texels_u8(o->upload_stream, raw, o->images.p, npix);
        PSG_CUDA(cudaEventRecord(o->upload_done, o->upload_stream));
        o->upload_pending = true;
        o->frames_dirty = true;
        o->inited = false;
        return PSG_OK;
    }
    if (npix > 0) {
        // the 8-bit pixels of all keyframes in one copy (3 B / pixel instead of 12), converted on the device
        o->vol->staging.alloc((size_t)(npix * 3 + 3) / 4);
        unsigned char* raw = reinterpret_cast<unsigned char*>(o->vol->staging.p);
        PSG_CUDA(cudaMemcpyAsync(raw, images_bgr_u8, (size_t)npix * 3, cudaMemcpyHostToDevice, o->stream));
        launch_texels_u8(o->stream, raw, o->images.p, npix);
    }
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    o->frames_dirty = true;
    o->inited = false;
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_set_keytimestamps(psg_optimizer* o, int F, const char* const* stamps) {
    if (!o || F < 0 || (F > 0 && !stamps)) { set_error("psg_opt_set_keytimestamps: null or empty argument"); return PSG_ERR_INVALID; }
    if (F != o->F) { set_error("psg_opt_set_keytimestamps: %d stamps for %d keyframes", F, o->F); return PSG_ERR_INVALID; }
    for (int i = 0; i < F; ++i)
        if (!stamps[i]) { set_error("psg_opt_set_keytimestamps: stamp %d is null", i); return PSG_ERR_INVALID; }
    o->key_stamps.assign(stamps, stamps + F);
    return PSG_OK;
}

int psg_opt_init(psg_optimizer* o) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    const bool led = o->set.model == PSG_MODEL_LED;
    if (led) {
        o->light.assign(3, 1.f);
    } else {
        // PsOptimizer::init (PsOptimizer.cpp:30-37): l_f = SH(R_f * (0,0,-1)), l_f[0] = 0.02
        o->light.assign((size_t)o->F * o->nb, 0.f);
        for (int f = 0; f < o->F; ++f) {
            const float* P = o->poses.data() + 16 * f;
            const float s[3] = {0.f, 0.f, -1.f};
            float d[3];
            for (int a = 0; a < 3; ++a) d[a] = P[4 * a] * s[0] + (P[4 * a + 1] * s[1] + P[4 * a + 2] * s[2]);
            sh_eval_host(d, o->nb, o->light.data() + (size_t)f * o->nb);
            o->light[(size_t)f * o->nb] = 0.02f;
        }
    }
    o->frames_dirty = true;
    if (o->sharded && o->vol_unsynced) sync_volume(o);   // the new work list is made from the WHOLE volume: every rank needs the others' rows first
    build_surface(o, true);
    finish_upload(o);   // overlapped keyframe upload: the images are needed from here on
    o->inited = true;
    if (led && o->S > 0) {
        // LedOptimizer::computeLightIntensive (LedOptimizer.cpp:76-112)
        double sums[6];
        sweep_led_light_init(o, sums);
        for (int c = 0; c < 3; ++c) o->light[c] = (float)sums[c] / (float)sums[3 + c];
        o->frames_dirty = true;
    }
    return PSG_OK;
    PSG_CATCH
}

#define PSG_NEED_INIT(o)                                                                   \
    if (!(o)) { set_error("null optimizer"); return PSG_ERR_INVALID; }                     \
    if (!(o)->inited) { set_error("psg_opt_init must be called first"); return PSG_ERR_STATE; }

int psg_opt_num_surface(const psg_optimizer* o, int* S) {
    PSG_NEED_INIT(o)
    if (!S) { set_error("psg_opt_num_surface: null argument"); return PSG_ERR_INVALID; }
    *S = o->S;
    return PSG_OK;
}

int psg_opt_get_surface(psg_optimizer* o, int* lin, uint64_t* vis) {
    PSG_NEED_INIT(o)
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    if (o->S > 0) {
        if (lin) copy_sync(lin, o->lin.p, (size_t)o->S * sizeof(int), cudaMemcpyDeviceToHost, o->stream);
        if (vis) copy_sync(vis, o->svis.p, (size_t)o->S * sizeof(uint64_t), cudaMemcpyDeviceToHost, o->stream);
    }
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_get_surface_voxels(psg_optimizer* o, float* aos8) {
    PSG_NEED_INIT(o)
    if (!aos8) { set_error("psg_opt_get_surface_voxels: null argument"); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    sync_volume(o);   // sharded: collective
    if (o->S > 0) {
        o->vol->staging.alloc((size_t)o->S * 8);
        launch_gather_voxels(o, o->vol->staging.p);
        copy_sync(aos8, o->vol->staging.p, (size_t)o->S * 8 * sizeof(float), cudaMemcpyDeviceToHost, o->stream);
    }
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_light_len(const psg_optimizer* o, int* n) {
    if (!o || !n) { set_error("null argument"); return PSG_ERR_INVALID; }
    *n = o->set.model == PSG_MODEL_LED ? 3 : o->F * o->nb;
    return PSG_OK;
}
int psg_opt_get_light(psg_optimizer* o, float* light) {
    PSG_NEED_INIT(o)
    if (!light) { set_error("psg_opt_get_light: null argument"); return PSG_ERR_INVALID; }
    memcpy(light, o->light.data(), o->light.size() * sizeof(float));
    return PSG_OK;
}
int psg_opt_set_light(psg_optimizer* o, const float* light) {
    PSG_NEED_INIT(o)
    if (!light) { set_error("psg_opt_set_light: null argument"); return PSG_ERR_INVALID; }
    memcpy(o->light.data(), light, o->light.size() * sizeof(float));
    o->frames_dirty = true;
    return PSG_OK;
}
int psg_opt_get_poses(psg_optimizer* o, float* poses16) {
    if (!o || !poses16) { set_error("null argument"); return PSG_ERR_INVALID; }
    memcpy(poses16, o->poses.data(), o->poses.size() * sizeof(float));
    return PSG_OK;
}
int psg_opt_set_poses(psg_optimizer* o, const float* poses16) {
    if (!o || !poses16) { set_error("null argument"); return PSG_ERR_INVALID; }
    memcpy(o->poses.data(), poses16, o->poses.size() * sizeof(float));
    o->frames_dirty = true;
    return PSG_OK;
}
int psg_opt_get_reg(const psg_optimizer* o, float r[3]) {
    if (!o || !r) { set_error("null argument"); return PSG_ERR_INVALID; }
    r[0] = o->set.reg_weight_rho; r[1] = o->set.reg_weight_n; r[2] = o->set.reg_weight_l;
    return PSG_OK;
}
int psg_opt_set_reg(psg_optimizer* o, const float r[3]) {
    if (!o || !r) { set_error("null argument"); return PSG_ERR_INVALID; }
    o->set.reg_weight_rho = r[0]; o->set.reg_weight_n = r[1]; o->set.reg_weight_l = r[2];
    return PSG_OK;
}

int psg_opt_set_sweep_split(psg_optimizer* o, int warps) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (warps != 0 && warps != 1 && warps != 2 && warps != 4 && warps != 8) {
        set_error("psg_opt_set_sweep_split: %d is not one of 0, 1, 2, 4, 8", warps);
        return PSG_ERR_INVALID;
    }
    o->sweep_split = warps;
    return PSG_OK;
}

int psg_opt_set_upload_overlap(psg_optimizer* o, int on) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    o->upload_overlap = on != 0;
    return PSG_OK;
}

int psg_opt_set_pcg_mode(psg_optimizer* o, int mode) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (mode < 0 || mode > 8) { set_error("psg_opt_set_pcg_mode: %d is not 0 .. 8", mode); return PSG_ERR_INVALID; }
    o->pcg_mode = mode;
    return PSG_OK;
}

int psg_opt_init_albedo(psg_optimizer* o) {
    PSG_NEED_INIT(o)
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    sweep_init_albedo(o);
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_energy(psg_optimizer* o, double* E, long long* n_active) {
    PSG_NEED_INIT(o)
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    double sum = 0.0; long long na = 0;
    sweep_energy(o, &sum, &na);
    if (E) *E = sum / (double)o->S;   // E / surface_points_.size() (PsOptimizer.cpp:77)
    if (n_active) *n_active = na;
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_reg_energies(psg_optimizer* o, double* En, double* El, double* Er) {
    PSG_NEED_INIT(o)
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    double a = 0, b = 0, c = 0;
    reg_energies(o, &a, &b, &c);
    if (En) *En = a / (double)o->S;
    if (El) *El = b / (double)o->S;
    if (Er) *Er = c / (double)o->S;
    return PSG_OK;
    PSG_CATCH
}

static int probe_common(psg_optimizer* o, float* m, float* n, unsigned char* ok, float* I3, float* rend3, float* r3, float* w3) {
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    const size_t SF = (size_t)o->S * o->F;
    if (SF == 0) return PSG_OK;
    DevBuf<float> buf; DevBuf<unsigned char> okb;
    buf.alloc(SF * 14); okb.alloc(SF);
    PSG_CUDA(cudaMemsetAsync(buf.p, 0, SF * 14 * sizeof(float), o->stream));
    PSG_CUDA(cudaMemsetAsync(okb.p, 0, SF, o->stream));
    float* dm = buf.p; float* dn = dm + SF; float* dI = dn + SF; float* dr = dI + 3 * SF; float* dres = dr + 3 * SF; float* dw = dres + 3 * SF;
    sweep_probe(o, dm, dn, okb.p, dI, dr, dres, dw);
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    if (m) copy_sync(m, dm, SF * 4, cudaMemcpyDeviceToHost, o->stream);
    if (n) copy_sync(n, dn, SF * 4, cudaMemcpyDeviceToHost, o->stream);
    if (ok) copy_sync(ok, okb.p, SF, cudaMemcpyDeviceToHost, o->stream);
    if (I3) copy_sync(I3, dI, SF * 12, cudaMemcpyDeviceToHost, o->stream);
    if (rend3) copy_sync(rend3, dr, SF * 12, cudaMemcpyDeviceToHost, o->stream);
    if (r3) copy_sync(r3, dres, SF * 12, cudaMemcpyDeviceToHost, o->stream);
    if (w3) copy_sync(w3, dw, SF * 12, cudaMemcpyDeviceToHost, o->stream);
    buf.release(); okb.release();
    return PSG_OK;
    PSG_CATCH
}

int psg_debug_image_probe(psg_optimizer* o, int frame, int n, const float* rows, const float* cols, float* rgb3, float* g0, float* g1) {
    PSG_NEED_INIT(o)
    if (frame < 0 || frame >= o->F || n < 0 || !rows || !cols || !rgb3 || !g0 || !g1) { set_error("psg_debug_image_probe: bad argument"); return PSG_ERR_INVALID; }
    for (int i = 0; i < n; ++i)
        if (!(rows[i] >= 0.f && rows[i] < (float)o->H && cols[i] >= 0.f && cols[i] < (float)o->W)) {
            set_error("psg_debug_image_probe: position %d (%g, %g) is outside the %d x %d image", i, rows[i], cols[i], o->H, o->W);
            return PSG_ERR_INVALID;
        }
    if (n == 0) return PSG_OK;
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    DevBuf<float> buf;
    buf.alloc((size_t)n * 11);
    float* d = buf.p;
    PSG_CUDA(cudaMemcpyAsync(d, rows, (size_t)n * 4, cudaMemcpyHostToDevice, o->stream));
    PSG_CUDA(cudaMemcpyAsync(d + n, cols, (size_t)n * 4, cudaMemcpyHostToDevice, o->stream));
    image_probe(o, frame, n, d, d + n, d + 2 * (size_t)n, d + 5 * (size_t)n, d + 8 * (size_t)n);
    copy_sync(rgb3, d + 2 * (size_t)n, (size_t)n * 12, cudaMemcpyDeviceToHost, o->stream);
    copy_sync(g0, d + 5 * (size_t)n, (size_t)n * 12, cudaMemcpyDeviceToHost, o->stream);
    copy_sync(g1, d + 8 * (size_t)n, (size_t)n * 12, cudaMemcpyDeviceToHost, o->stream);
    buf.release();
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_residual(psg_optimizer* o, float* r3, float* w3) {
    PSG_NEED_INIT(o)
    return probe_common(o, nullptr, nullptr, nullptr, nullptr, nullptr, r3, w3);
}
int psg_opt_probe_pairs(psg_optimizer* o, float* m, float* n, unsigned char* ok, float* I3, float* rend3) {
    PSG_NEED_INIT(o)
    return probe_common(o, m, n, ok, I3, rend3, nullptr, nullptr);
}

// ---- one block = queue (no host wait) + finish (after fetch_mail) ------------------------------------------------
namespace {

// where the host thread's time goes inside psg_opt_iteration (psg_debug_host_times): launching, waiting for the device, solving
double g_host_us[3] = {0.0, 0.0, 0.0};
long long g_host_iterations = 0;
struct HostClock {
    int slot; std::chrono::steady_clock::time_point t0;
    explicit HostClock(int s) : slot(s), t0(std::chrono::steady_clock::now()) {}
    ~HostClock() { g_host_us[slot] += std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count(); }
};

struct Pending { int block = -1; int slot = 0; size_t nred = 0; float damping = 0.f; bool albedo_reg = false; };

// queues the kernels of one block (none of them waits for the host)
Pending queue_block(psg_optimizer* o, int block, float damping, int flags, int slot, bool reg_energies_after = false,
                    bool keep_vectors = true) {
    Pending p; p.block = block; p.slot = slot; p.damping = damping;
    const bool led = o->set.model == PSG_MODEL_LED;
    o->last_block = block;
    if (block == PSG_BLOCK_ALBEDO && (flags & PSG_REG_ALBEDO)) {
        // reg albedo != 0: the sweep hands over J^T W J / J^T W r, the regularised 3S x 3S system is solved by its own PCG;
        // SH applies the update only on Success (PsOptimizer.cpp:118), LED unconditionally (LedOptimizer.cpp:195)
        const size_t n = (size_t)3 * o->S;
        o->vecs.alloc((size_t)(o->S > 0 ? o->S : 1) * 25 + 8);
        launch_albedo(o, damping, o->vecs.p, o->vecs.p + n, slot, /*regularised=*/true);
        launch_albedo_reg_solve(o, damping, /*only_on_success=*/!led);
        if (reg_energies_after) launch_reg_energies(o, kSlotReg);   // getAlbedoRegEnergy after the update
        p.albedo_reg = true;
    } else if (block == PSG_BLOCK_ALBEDO) {
        const size_t n = (size_t)3 * o->S;
        o->vecs.alloc((size_t)(o->S > 0 ? o->S : 1) * 25 + 8);
        // rhs / delta per voxel are only kept for psg_opt_last_vectors (scattered 12-byte writes): the outer loop skips them
        launch_albedo(o, damping, keep_vectors ? o->vecs.p : nullptr, keep_vectors ? o->vecs.p + n : nullptr, slot);
        if (!keep_vectors) o->last_block = -1;
    } else if (block == PSG_BLOCK_LIGHT || block == PSG_BLOCK_POSE) {
        // the sweep leaves the per-keyframe sums on the device; one thread block solves them there and updates the keyframe
        // table (small_solve.cu), so the next block can be queued right behind
        launch_small_block(o, block, slot);
        reduce_pending(o);   // several ranks: the solve needs the job's sums
        launch_small_solve(o, block, slot, damping);
    } else {
        launch_dist_sweep(o, flags, slot);
        // SH: updateDist only on Success (PsOptimizer.cpp:168-170); LED: unconditionally (LedOptimizer.cpp:242).  The
        // verdict stays on the device: the update kernels read it, the host learns it with the next round trip.
        if (o->sharded || o->pcg_mode == 8) {   // 8: the dataflow kernel on one GPU
            launch_dist_block_mg(o, damping, /*only_on_success=*/!led, reg_energies_after);
        } else if (!launch_dist_block_small(o, damping, /*only_on_success=*/!led, reg_energies_after ? kSlotReg : -1)) {
            launch_dist_solve(o, damping);
            launch_dist_update(o, /*only_on_success=*/!led);
            if (reg_energies_after) launch_reg_energies(o, kSlotReg);   // getNormalEnergy / getLaplacianEnergy after the update
        }
    }
    return p;
}

// host part of a block once fetch_mail has brought its results: the step record
void finish_block(psg_optimizer* o, const Pending& p, psg_step_info* out) {
    const bool led = o->set.model == PSG_MODEL_LED;
    const int block = p.block;
    psg_step_info inf;
    memset(&inf, 0, sizeof(inf));
    if (block == PSG_BLOCK_ALBEDO && p.albedo_reg) {
        int ok = 1, iters = 0; float err = 0.f; long long acc = 0;
        read_albedo_reg_verdict(o, &ok, &iters, &err, &acc);
        inf.n = 3 * o->S; inf.cg_success = ok; inf.cg_iterations = iters; inf.cg_error = err;
        inf.applied = (led || ok) ? 1 : 0; inf.n_accepted = acc;
        o->last_n = 3 * o->S;
    } else if (block == PSG_BLOCK_ALBEDO) {
        const size_t n = (size_t)3 * o->S;
        inf.n = (int)n; inf.cg_success = 1; inf.cg_iterations = 0; inf.cg_error = 0.f; inf.applied = 1;
        inf.n_accepted = (long long)(mail_slot(o, p.slot)[5] + 0.5);
        o->last_n = (int)n;
    } else if (block == PSG_BLOCK_LIGHT || block == PSG_BLOCK_POSE) {
        // solved and applied on the device; fetch_mail has already brought the updated light / poses into the host copies
        const int bs = block == PSG_BLOCK_POSE ? 6 : (led ? 3 : o->nb);
        const int nblk = (block == PSG_BLOCK_LIGHT && led) ? 1 : o->F;
        int ok = 1, iters = 0, applied = 1; float err = 0.f;
        read_small_verdict(o, block, &ok, &iters, &err, &applied);
        inf.n = bs * nblk; inf.cg_success = ok; inf.cg_iterations = iters; inf.cg_error = err; inf.applied = applied;
        o->last_n = bs * nblk;
    } else {
        int ok = 1, iters = 0; float err = 0.f; long long acc = 0;
        read_dist_verdict(o, &ok, &iters, &err, &acc);
        inf.n = o->S; inf.cg_success = ok; inf.cg_iterations = iters; inf.cg_error = err;
        inf.applied = (led || ok) ? 1 : 0; inf.n_accepted = acc;
        o->last_n = o->S;
    }
    const double* sc = mail_slot(o, p.slot);
    o->last_energy_sum = sc[6];
    o->last_active = (long long)(sc[7] + 0.5);
    inf.energy_before = o->S > 0 ? sc[6] / (double)o->S : 0.0;
    inf.n_active_before = o->last_active;
    if (out) *out = inf;
}

}  // namespace

int psg_opt_step(psg_optimizer* o, int block, float damping, int flags, psg_step_info* info) {
    PSG_NEED_INIT(o)
    if (block < 0 || block > 3) { set_error("psg_opt_step: unknown block %d", block); return PSG_ERR_INVALID; }
    if (block == PSG_BLOCK_ALBEDO && (flags & PSG_REG_ALBEDO) && o->sharded) {
        set_error("psg_opt_step: the albedo-gradient regulariser (reg albedo != 0) is not provided on a work list sharded over "
                  "several GPUs; use psg_opt_set_shard_policy(opt, 2)");
        return PSG_ERR_UNSUPPORTED;
    }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    const Pending p = queue_block(o, block, damping, flags, 0);
    o->last_albedo_reg = p.albedo_reg;
    fetch_mail(o, p.nred);
    finish_block(o, p, info);
    return PSG_OK;
    PSG_CATCH
}

// One pass of the reference's while-body as one queue of kernels and one host wait: every block's solve and update runs on the
// device (albedo in its sweep, light and pose in k_small_solve, distance in the PCG kernels), each reporting into its own mail
// slot; the closing energy rides behind them and a single round trip brings all of it back.
int psg_opt_iteration(psg_optimizer* o, const psg_iteration_config* cfg, psg_iteration_info* out) {
    PSG_NEED_INIT(o)
    if (!cfg) { set_error("psg_opt_iteration: null configuration"); return PSG_ERR_INVALID; }
    if ((cfg->reg_flags & PSG_REG_ALBEDO) && cfg->albedo && o->sharded) {
        set_error("psg_opt_iteration: the albedo-gradient regulariser (reg albedo != 0) is not provided on a work list sharded "
                  "over several GPUs; use psg_opt_set_shard_policy(opt, 2)");
        return PSG_ERR_UNSUPPORTED;
    }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    const bool led = o->set.model == PSG_MODEL_LED;
    psg_iteration_info R;
    memset(&R, 0, sizeof(R));
    int order[4];
    if (led) { order[0] = PSG_BLOCK_LIGHT; order[1] = PSG_BLOCK_ALBEDO; }
    else { order[0] = PSG_BLOCK_ALBEDO; order[1] = PSG_BLOCK_LIGHT; }
    order[2] = PSG_BLOCK_DIST; order[3] = PSG_BLOCK_POSE;
    const int enabled[4] = { cfg->albedo, cfg->light, cfg->distance, cfg->pose };   // indexed by PSG_BLOCK_*
    static_assert(PSG_BLOCK_ALBEDO == 0 && PSG_BLOCK_LIGHT == 1 && PSG_BLOCK_DIST == 2 && PSG_BLOCK_POSE == 3, "block ids");

    Pending queued[4];
    int nq = 0, slot = 0;
    bool reg_queued = false;
    int last_run = -1;
    for (int k = 0; k < 4; ++k) {
        const int block = order[k];
        if (!enabled[block]) continue;
        const bool areg = (cfg->reg_flags & PSG_REG_ALBEDO) != 0;
        const int flags = block == PSG_BLOCK_DIST ? (cfg->reg_flags & (PSG_REG_NORMAL | PSG_REG_LAPLACIAN))
                                                  : (block == PSG_BLOCK_ALBEDO ? (cfg->reg_flags & PSG_REG_ALBEDO) : 0);
        // regulariser energies after the block that changes them (the albedo term does not depend on the distance, so a
        // later distance block in the same flush recomputes the same value into the same slot)
        const bool want_reg = (block == PSG_BLOCK_DIST && ((cfg->reg_flags & (PSG_REG_NORMAL | PSG_REG_LAPLACIAN)) || areg)) ||
                              (block == PSG_BLOCK_ALBEDO && areg);
        { HostClock hc(0); queued[nq] = queue_block(o, block, cfg->damping, flags, slot, want_reg, /*keep_vectors=*/false); }
        ++nq; ++slot;
        last_run = block;
        if (want_reg) reg_queued = true;
    }
    // closing energy (the convergence test needs it); blocks still queued ride on the same round trip
    if (last_run >= 0 || cfg->closing_energy) {
        { HostClock hc(0); launch_energy(o, slot); }
        static_assert(kSlotAreg > 4, "four blocks and the closing energy use mail slots 0..4");
        { HostClock hc(1); fetch_mail(o, 0); }
        HostClock hc(2);
        const double* sc = mail_slot(o, slot);
        R.energy_end = o->S > 0 ? sc[6] / (double)o->S : 0.0;
        R.n_active_end = (long long)(sc[7] + 0.5);
        for (int i = 0; i < nq; ++i) {
            finish_block(o, queued[i], &R.block[queued[i].block]);
            R.ran[queued[i].block] = 1;
        }
        if (reg_queued) {
            const double* r = mail_slot(o, kSlotReg);
            R.reg_normal_after_dist = o->S > 0 ? r[0] / (double)o->S : 0.0;
            R.reg_laplacian_after_dist = o->S > 0 ? r[1] / (double)o->S : 0.0;
            R.reg_albedo_after_albedo = o->S > 0 ? r[2] / (double)o->S : 0.0;
        }
        o->last_energy_sum = sc[6];
        o->last_active = R.n_active_end;
    }
    // energy after block k = energy the next executed block started from (or the closing energy)
    {
        double next = R.energy_end;
        for (int k = 3; k >= 0; --k) {
            const int block = order[k];
            if (!R.ran[block]) continue;
            R.energy_after[block] = next;
            next = R.block[block].energy_before;
        }
    }
    ++g_host_iterations;
    if (out) *out = R;
    return PSG_OK;
    PSG_CATCH
}

// The host's small solve on caller-supplied blocks (no device involved): what finish_block runs for the light and pose blocks.
int psg_debug_small_cg(int nblocks, int bs, const float* H, const float* b, float* x, int* success, int* iterations, float* error) {
    if (nblocks < 1 || bs < 1 || !H || !b || !x) { set_error("psg_debug_small_cg: bad argument"); return PSG_ERR_INVALID; }
    const CgVerdict v = eigen_like_cg(nblocks, bs, H, b, x);
    if (success) *success = v.success;
    if (iterations) *iterations = v.iters;
    if (error) *error = v.err;
    return PSG_OK;
}

int psg_debug_host_times(double us[3], long long* iterations, int reset) {
    if (us) for (int i = 0; i < 3; ++i) us[i] = g_host_us[i];
    if (iterations) *iterations = g_host_iterations;
    if (reset) { g_host_us[0] = g_host_us[1] = g_host_us[2] = 0.0; g_host_iterations = 0; }
    return PSG_OK;
}

int psg_debug_last_pcg_kernel(const psg_optimizer* o, int* kernel) {
    if (!o || !kernel) { set_error("null argument"); return PSG_ERR_INVALID; }
    *kernel = o->last_pcg_kernel;
    return PSG_OK;
}

int psg_debug_timer_begin(psg_optimizer* o) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    if (!o->timer0) { PSG_CUDA(cudaEventCreate(&o->timer0)); PSG_CUDA(cudaEventCreate(&o->timer1)); }
    PSG_CUDA(cudaEventRecord(o->timer0, o->stream));
    return PSG_OK;
    PSG_CATCH
}
int psg_debug_timer_end(psg_optimizer* o, float* ms) {
    if (!o || !ms) { set_error("null argument"); return PSG_ERR_INVALID; }
    if (!o->timer0) { set_error("psg_debug_timer_end without psg_debug_timer_begin"); return PSG_ERR_STATE; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    PSG_CUDA(cudaEventRecord(o->timer1, o->stream));
    PSG_CUDA(cudaEventSynchronize(o->timer1));
    PSG_CUDA(cudaEventElapsedTime(ms, o->timer0, o->timer1));
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_last_system(psg_optimizer* o, double* H, double* b, double* delta) {
    PSG_NEED_INIT(o)
    if (o->last_block != PSG_BLOCK_LIGHT && o->last_block != PSG_BLOCK_POSE) { set_error("last step was not a light or pose block"); return PSG_ERR_STATE; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    // the solve kernel keeps H | b | x of the last light / pose block on the device
    const bool led = o->set.model == PSG_MODEL_LED;
    const int bs = o->last_block == PSG_BLOCK_POSE ? 6 : (led ? 3 : o->nb);
    const int nblk = (o->last_block == PSG_BLOCK_LIGHT && led) ? 1 : o->F;
    const size_t nH = (size_t)nblk * bs * bs, n = (size_t)nblk * bs;
    std::vector<float> sys(nH + 2 * n);
    copy_sync(sys.data(), o->small_sys.p, sys.size() * sizeof(float), cudaMemcpyDeviceToHost, o->stream);
    if (H) for (size_t i = 0; i < nH; ++i) H[i] = sys[i];
    if (b) for (size_t i = 0; i < n; ++i) b[i] = sys[nH + i];
    if (delta) for (size_t i = 0; i < n; ++i) delta[i] = sys[nH + n + i];
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_last_vectors(psg_optimizer* o, float* b, float* delta) {
    PSG_NEED_INIT(o)
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    const size_t S = (size_t)o->S;
    if (o->last_block == PSG_BLOCK_ALBEDO) {
        const size_t off = o->last_albedo_reg ? 6 * S : 0;   // the regularised solve keeps its rhs / solution behind the sweep's
        if (b) copy_sync(b, o->vecs.p + off, 3 * S * 4, cudaMemcpyDeviceToHost, o->stream);
        if (delta) copy_sync(delta, o->vecs.p + off + 3 * S, 3 * S * 4, cudaMemcpyDeviceToHost, o->stream);
    } else if (o->last_block == PSG_BLOCK_DIST) {
        if (b) copy_sync(b, o->vecs.p + 14 * S, S * 4, cudaMemcpyDeviceToHost, o->stream);
        if (delta) copy_sync(delta, o->vecs.p + 17 * S, S * 4, cudaMemcpyDeviceToHost, o->stream);
    } else { set_error("last step was not an albedo or distance block"); return PSG_ERR_STATE; }
    return PSG_OK;
    PSG_CATCH
}

int psg_opt_upsample(psg_optimizer* o) {
    PSG_NEED_INIT(o)
    if (8 * o->vol->N > (1ll << 31) - 64) { set_error("psg_opt_upsample: upsampled grid exceeds 2^31 voxels"); return PSG_ERR_INVALID; }
    PSG_TRY
    PSG_CUDA(cudaSetDevice(o->vol->device));
    if (o->sharded) { sync_volume(o); release_peers(o); }   // children are made from the whole volume; its planes are reallocated
    launch_upsample(o);
    build_surface(o, false);
    return PSG_OK;
    PSG_CATCH
}

}  // extern "C"
