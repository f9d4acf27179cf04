// Host-side state behind the opaque C handles (include/psg.h) and the launcher prototypes each .cu file exports.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/psg.h"
#include "psg_device.cuh"

namespace psg {

void set_error(const char* fmt, ...);
void trace(const char* fmt, ...);   // stderr, only when PSG_TRACE is set in the environment (debugging aid)
void count_launch(int n = 1);

// Kernel ids for the built-in launch timer (psg_profile_* in include/psg.h).
enum KernelId {
    KID_FILL = 0, KID_PACK_FRAME, KID_FUSE, KID_AOS, KID_SURF_MASK, KID_SURF_SCAN, KID_SURF_SCATTER, KID_SURF_NEIGHBORS,
    KID_UPSAMPLE, KID_SWEEP_ENERGY, KID_SWEEP_INIT_ALBEDO, KID_SWEEP_LED_INIT, KID_SWEEP_PROBE, KID_SWEEP_ALBEDO,
    KID_SWEEP_LIGHT, KID_SWEEP_POSE, KID_SWEEP_DIST, KID_REG_ENERGY, KID_PCG_ASSEMBLE, KID_PCG_SOLVE, KID_DIST_UPDATE,
    KID_GRAD_UPDATE, KID_PAD_IMAGES, KID_NORMALS, KID_COMM_SCATTER, KID_COMM_ALLREDUCE, KID_COMM_ALLGATHER, KID_TRACK, KID_MESH_BOX, KID_MESH_COUNT, KID_MESH_SCAN, KID_MESH_EMIT, KID_MESH_MISC, KID_SMALL_SOLVE, KID_COUNT
};
// Counts one launch and, when profiling is enabled, brackets it with CUDA events on the launching stream.
struct Prof {
    cudaStream_t st; int id; cudaEvent_t e0 = nullptr, e1 = nullptr;
    Prof(cudaStream_t stream, int kernel_id);
    ~Prof();
};

struct CudaError { int code; };
#define PSG_CUDA(call)                                                                          \
    do {                                                                                        \
        cudaError_t e__ = (call);                                                               \
        if (e__ != cudaSuccess) {                                                               \
            psg::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e__)); \
            throw psg::CudaError{(int)e__};                                                     \
        }                                                                                       \
    } while (0)
#define PSG_LAUNCH_CHECK() PSG_CUDA(cudaGetLastError())

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    void alloc(size_t count) {
        if (count <= n && p) return;
        release();
        if (count == 0) count = 1;
        PSG_CUDA(cudaMalloc((void**)&p, count * sizeof(T)));
        n = count;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr; n = 0;
    }
};

}  // namespace psg

namespace psg {
constexpr int kMaxRanks = 8;     // ranks of one sharded job (one 8-GPU NVSwitch box)

// Per-rank synchronisation block of the distributed distance solve, in memory the peers can write (CUDA IPC mapping, or the
// same address space when the ranks of a job live in one process).  Zeroed once when it is allocated; every counter and tag is
// monotonic across launches, so nothing is reset while a peer may still be writing.
struct XState {
    unsigned long long epoch;              // barriers completed by earlier launches (written by the owner at the end of a launch)
    unsigned long long epoch_r;            // ... of which barriers with a cross-rank reduction (mailbox parity and packet tags)
    unsigned long long epoch_t;            // value tags handed out by earlier launches (every vector entry that another thread reads carries one)
    unsigned long long arrive;             // arrivals of the owner's blocks at the current barrier (reset by the owner at the end of a launch)
    unsigned long long abort;              // sticky: a wait ran into its time limit
    unsigned long long ll[2][kMaxRanks][6];   // [reduction parity][sender][word]: three doubles as six (tag << 32 | 32 data bits) packets, written by the sender
};

// what a rank exposes to its peers
struct PeerView {
    unsigned long long* halo = nullptr;   // kHaloKinds planes of nh (tag << 32 | float bits) packets: the owner's halo rows, written by their owners
    XState* xs = nullptr;
};
// All-reduce of the small sums of a sharded job (comm.cu: k_packet_allreduce): behind a rank's XState, in the same peer-mapped
// allocation, an inbox of (epoch << 32 | 32 data bits) packets per reduction parity and sender, two packets per double.
constexpr int kArMaxDoubles = 64 * 64 + 64;        // per-keyframe sums (kRedDoubles) + mail slots in flight
constexpr size_t kArInboxOffset = (sizeof(XState) + 255) / 256 * 256;
constexpr size_t kArInboxBytes = (size_t)2 * kMaxRanks * 2 * kArMaxDoubles * sizeof(unsigned long long);
constexpr int kHaloKinds = 11;   // stencil planes 2, 5, 9, 11, 12, 13 | u = r / diag | U1..U3 | distance
}  // namespace psg

struct psg_volume {
    int device = 0;
    int dims[3] = {0, 0, 0};
    long long N = 0;
    float vs = 0.f, shift[3] = {0, 0, 0}, origin[3] = {0, 0, 0}, T = 0.f, invT = 0.f, zmin = 0.f, zmax = 0.f;
    int max_frames = 64, nwords = 1, counter = 0;
    unsigned long long generation = 1;       // bumped whenever the host or fusion rewrites voxel records
    psg::DevBuf<float> f[8];                 // dist gx gy gz w r g b
    psg::DevBuf<unsigned long long> vis;     // nwords * N
    psg::DevBuf<float> frame;                // packed per-pixel fusion inputs (8 floats / pixel)
    psg::DevBuf<float> staging;              // H2D staging for planar inputs / AoS transfers
    psg::DevBuf<int> brick_list;             // fusion: bricks of the current frame that survive the cull, then their count
    psg::DevBuf<float> tiles;                // fusion: largest valid depth per 8 x 8 pixel tile of the current frame
    psg::DevBuf<float> track_depth;          // rigid tracker: current frame's depth in metres
    psg::DevBuf<double> track_partials;
    int track_H = 0, track_W = 0;
    // surface extraction (mesh.cu): per-block counts / offsets, crop box, and the triangles of the last extraction
    psg::DevBuf<unsigned> mesh_counts, mesh_crossed;
    psg::DevBuf<int> mesh_box;
    psg::DevBuf<float> mesh_vertices;        // 9 floats per triangle
    psg::DevBuf<unsigned char> mesh_colors;  // 9 bytes per triangle
    long long mesh_triangles = -1;           // -1: nothing extracted yet
    cudaStream_t stream = nullptr;
    psg::VolView view() const;
};

struct psg_optimizer {
    psg_volume* vol = nullptr;
    psg_settings set{};
    int nb = 4;                              // SH basis size
    float vs = 0.f, vs_inv = 0.f;            // Optimizer::voxel_size_, voxel_size_inv_
    float K[9] = {0};
    int F = 0, H = 0, W = 0;
    std::vector<int> frame_idx;
    std::vector<std::string> key_stamps;     // Optimizer::key_stamps_ (first column of savePoses); empty = print frame_idx
    std::vector<float> poses;                // F x 16 (host copy, authoritative)
    std::vector<float> light;                // F x nb or 3
    psg::DevBuf<float4> images;              // F x H x W texels (b, g, r, 0)
    psg::FrameTable frame_table;             // per-keyframe R, t, light: host staging of the device table
    psg::DevBuf<psg::FrameParams> frames_dev;  // the table the sweeps read and the small solves update (kMaxFrames + 1 entries)
    bool frames_dirty = true;                // the host copy (poses / light) changed: upload before the next sweep
    bool const_dirty = true;                 // frames_dev changed: refresh this optimizer's slot of the constant-memory copy
    int cslot = -1;                          // that slot (sweep.cu: c_frames)
    // light / pose blocks solved on the device (small_solve.cu): the dense blocks, rhs and solution of the last one, and what the
    // next fetch_mail has to bring back into the host copies
    psg::DevBuf<float> small_sys;            // H (nblk * bs * bs) | b (n) | x (n)
    int small_pending = 0;                   // bit PSG_BLOCK_LIGHT / PSG_BLOCK_POSE: the device table holds newer values than the host
    psg::FrameTable* host_frames = nullptr;  // pinned landing area of the table
    // overlapped keyframe upload (psg_opt_set_upload_overlap): copy + conversion on their own stream while psg_opt_init
    // builds the work list; the caller's image buffer must then stay valid until psg_opt_init returns
    bool upload_overlap = false, upload_pending = false;
    cudaStream_t upload_stream = nullptr;
    cudaEvent_t upload_done = nullptr;
    psg::DevBuf<float> upload_staging;
    bool inited = false;
    // surface work list
    int S = 0;
    psg::DevBuf<int> lin, nbr_lin, nbr_row, inc;
    psg::DevBuf<unsigned long long> svis;
    psg::DevBuf<unsigned char> pos;
    psg::DevBuf<float4> rec0, rec1, rec2, rec3, rec4;   // slot-ordered voxel records (SurfView)
    psg::DevBuf<unsigned long long> pvis;
    psg::DevBuf<int> perm, perm2;
    psg::DevBuf<unsigned> sort_hist;         // radix sort of the slot order: per-chunk bucket offsets of the pass in flight
    unsigned long long vol_gen_seen = 0;     // volume generation the records were gathered from
    psg::DevBuf<unsigned int> mask, prefix;  // membership bitmask over the volume + exclusive popcount prefix
    psg::DevBuf<unsigned int> block_counts, block_pairs;
    // reductions / systems
    psg::DevBuf<double> red;                 // scratch for reductions (energy, JtJ blocks)
    psg::DevBuf<double> sweep_partials;      // per-block scalar partials of the sweeps
    psg::DevBuf<unsigned> sweep_counter;     // arrival counter of the last-block reduction
    void* comm = nullptr;                    // ncclComm_t: one process per GPU
    void* group = nullptr;                   // psg::LocalGroup*: the ranks of the job live in this process (threads), see comm.cu
    psg::DevBuf<float> comm_buf;             // staging of psg_opt_sync_volume: world chunks of packed rows
    int rank = 0, world = 1;
    bool sharded = false;                    // every rank owns a contiguous row range of the surface list (a z-slab); else replicated
    // row-range sharding (comm.cu): own rows [row_lo, row_hi) = the slots of this rank's sweeps; halo = rows of other ranks within
    // one z-plane (W * H voxels) of the range, whose vector entries the peers push into this rank's copies
    int row_lo = 0, row_hi = 0, n_slots = 0;
    std::vector<int> row_bounds;             // world + 1 cut points (identical on every rank)
    std::vector<int> halo_lo, halo_hi;       // per rank q: its halo rows are [halo_lo[q], row_bounds[q]) and [row_bounds[q+1], halo_hi[q])
    bool layout_dirty = true;                // slot list / peer tables must be rebuilt (mode switch, new surface)
    bool vol_unsynced = false;               // sharded: other ranks' rows of the volume planes are stale (psg_opt_sync_volume)
    psg::DevBuf<unsigned char> xstate;       // psg::XState of this rank
    psg::DevBuf<unsigned long long> halo;    // this rank's halo packets (PeerView::halo), kHaloKinds x (rows of its halo)
    psg::DevBuf<unsigned long long> xplanes; // u and U1..U3 of all rows as (tag, value) packets: what other threads of this rank read (pcg.cu)
    psg::PeerView peer[psg::kMaxRanks];      // peer[q]: rank q's buffers as addressable from this process (peer[rank] = own)
    bool peers_ready = false;                // every rank has mapped every rank's buffers (exchange_peers .. release_peers)
    unsigned ar_epoch = 0;                   // packet all-reduces done so far (the same on every rank: collectives run in lockstep)
    std::vector<void*> ipc_opened;           // base pointers of the imported allocations (closed before anything is freed)
    int shard_policy = 0;                    // 0 = by work per rank, 1 = always, 2 = never
    long long vis_pairs = 0;                 // visible (surface voxel, keyframe) pairs of the current work list
    int reduce_lo = 1 << 30, reduce_hi = -1; // mail slots whose values are still per-rank partial sums
    size_t reduce_red = 0;                   // ... and how many doubles of o->red
    bool reduce_reg = false;                 // ... and the regulariser / accepted-count slot (kSlotReg)
    int last_pcg_kernel = -1;                // which distance-solve kernel ran last: 0 general grid, 1 portable cluster, 2 fused grid, 3 16-CTA cluster (DSMEM), 4 distributed
    int pcg_mode = 0;                        // 0 / 2 = cooperative-grid kernel, 1 = cluster kernel for small surfaces
    int sweep_split = 0;                     // 0 = choose the keyframe split per launch; else force W (1, 2, 4, 8)
    psg::DevBuf<float> vecs;                 // per-voxel normal equations + PCG vectors
    psg::DevBuf<double> pcg_partials;
    psg::DevBuf<float> pcg_pbuf;             // double-buffered search direction of the cluster PCG
    psg::DevBuf<unsigned char> mg_args;      // argument block of the distributed distance solve (pcg.cu)
    psg::DevBuf<float> areg;                 // workspace of the regularised albedo solve (reg albedo != 0), 39 planes of S
    std::vector<double> lastH, lastb, lastdelta;   // small dense systems of the last step
    int last_n = 0, last_block = -1;
    bool last_albedo_reg = false;            // the last albedo step went through the regularised solve
    // pinned host staging: the mailbox (kMailDoubles) followed by kRedDoubles of per-keyframe sums
    double* host_staging = nullptr;
    void ensure_host_staging();
    double last_energy_sum = 0.0;            // sum of robust losses seen by the last sweep (state before its update)
    long long last_active = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t timer0 = nullptr, timer1 = nullptr;   // psg_debug_timer_begin / _end
};

namespace psg {

// o->red: the mailbox, then kRedCopies x kRedDoubles of per-keyframe sums (copy 0 holds the result); the mailbox (10 slots x 8 doubles) every kernel of an outer iteration reports
// into: slots 0..4 sweep scalars (one per sweep in flight), then the three below
constexpr int kRedDoubles = kMaxFrames * 64;
constexpr int kRedCopies = 8;     // private copies of the per-keyframe sums on large grids (same-address fp64 atomics serialise in L2)
constexpr int kMailDoubles = 80;
constexpr int kPcgMaxBlocks = 1024;   // most blocks a cooperative distance-solve launch uses (4 per SM, <= 256 SMs)
constexpr int kSlotAreg = 5;     // regularised albedo solve: 4 floats verdict (as kSlotPcg), [5] accepted updates, [6] grid-barrier counter
constexpr int kSlotReg = 6;      // regulariser energies {normal, laplacian, albedo}
constexpr int kSlotPcg = 7;      // 4 floats: iterations, residualNorm2, rhsNorm2, threshold
constexpr int kSlotCount = 8;    // [0] unsigned long long: accepted distance updates, [1] arrival counter of the PCG's grid barrier
constexpr int kSlotSmall = 9;    // small solves: 4 floats {success, iterations, error, applied} of the light block, 4 of the pose block; last double: comm.cu

// ---- small_solve.cu: the light / pose blocks' Eigen-ordered CG and update on the device
void launch_small_solve(psg_optimizer* o, int block, int slot, float damping);
void read_small_verdict(const psg_optimizer* o, int block, int* success, int* iterations, float* error, int* applied);
void adopt_device_frames(psg_optimizer* o);

// ---- volume.cu
void launch_fill_volume(psg_volume* v);
void launch_integrate(psg_volume* v, int H, int W, const float* K, const float* R, const float* t, int counter);
void launch_pack_frame(psg_volume* v, const float* planar_dev, int H, int W);
void launch_aos_to_soa(psg_volume* v, const float* aos_dev, long long begin, long long count);
void launch_soa_to_aos(psg_volume* v, float* aos_dev, long long begin, long long count);

// ---- normals.cu
void launch_texels_u8(cudaStream_t st, const unsigned char* bgr8_dev, float4* dst, long long n);

// ---- surface.cu
void build_surface(psg_optimizer* o, bool reindex_vis);
void refresh_records(psg_optimizer* o);      // re-gather rec0..2 from the volume (no allocation)
SurfView surf_view(psg_optimizer* o);        // refreshes the records first if the volume changed under them
void launch_upsample(psg_optimizer* o);
void launch_gather_voxels(psg_optimizer* o, float* out_dev);   // S x 8 floats (SdfVoxel layout) in surface-list order

// ---- sweep.cu
void reduce_pending(psg_optimizer* o);      // stream-ordered reduction over the ranks of the sums queued so far
void refresh_frame_table(psg_optimizer* o);
int acquire_const_slot(int device);
void release_const_slot(int device, int slot);
void shard_blocks(int S, int rank, int world, int* first, int* stride, int* count);
void comm_allreduce_doubles(psg_optimizer* o, double* const* ptr, const size_t* count, int n);   // in place, n <= 3 arrays in one step
void comm_broadcast_doubles(psg_optimizer* o, double* a, size_t n);   // rank 0's values to every rank
void update_shard_policy(psg_optimizer* o);
int comm_destroy(psg_optimizer* o, bool collective);
// ranks as threads of one process: rendezvous of the members; `leader` runs once with every member's argument, on the last
// arriver's stream, ordered after everything the members had queued; the members' streams continue after it
void group_run(psg_optimizer* o, void* my_arg, void (*leader)(psg_optimizer* on, void* const* args, int world, void* user), void* user);
void partition_rows(psg_optimizer* o);       // row_lo / row_hi / n_slots and the halo ranges of every rank
void exchange_peers(psg_optimizer* o);       // sharded: fill o->peer[] (CUDA IPC between processes, plain pointers inside one); collective
void build_layout(psg_optimizer* o);          // row ranges, slot list, records and (sharded) the peer tables; collective when sharded
void release_peers(psg_optimizer* o);         // close the imported peer buffers before anything they map is freed; collective
void sync_volume(psg_optimizer* o);           // sharded: every rank receives the other ranks' rows of the volume planes; collective
void sweep_energy_partial(psg_optimizer* o, int rank, int world, double* sum, long long* n_active);
void sweep_energy(psg_optimizer* o, double* E_sum, long long* n_active);
void sweep_init_albedo(psg_optimizer* o);
void sweep_led_light_init(psg_optimizer* o, double sums[6]);
void image_probe(psg_optimizer* o, int f, int n, const float* rows_dev, const float* cols_dev, float* rgb3, float* g0, float* g1);
void sweep_probe(psg_optimizer* o, float* m, float* n, unsigned char* ok, float* I3, float* rend3, float* r3, float* w3);
// queue-only pieces (no host wait) and the one round trip that collects their results
double* mail_dev(psg_optimizer* o);
double* red_dev(psg_optimizer* o);
const double* red_host(const psg_optimizer* o);
const double* fetch_mail(psg_optimizer* o, size_t n_red_doubles);
const double* mail_slot(const psg_optimizer* o, int slot);
void launch_albedo(psg_optimizer* o, float damping, float* b_dev, float* delta_dev, int slot, bool regularised = false);
size_t launch_small_block(psg_optimizer* o, int block, int slot);
void launch_dist_sweep(psg_optimizer* o, int flags, int slot);
void launch_energy(psg_optimizer* o, int slot);
void launch_reg_energies(psg_optimizer* o, int slot);
void reg_energies(psg_optimizer* o, double* En, double* El, double* Er);

// ---- mesh.cu
bool write_mesh_file(psg_volume* v, float vs, const char* path, int white, long long* n_triangles);   // false: file not opened
bool write_sdf_file(psg_volume* v, float vs, const char* path, bool* any);

// ---- pcg.cu
void launch_dist_block_mg(psg_optimizer* o, float damping, bool only_on_success, bool want_reg);   // sharded: the whole distance block after the sweep, one kernel
bool launch_dist_block_small(psg_optimizer* o, float damping, bool only_on_success, int reg_slot);   // fused assemble + PCG + update (+ reg energies)
void launch_dist_solve(psg_optimizer* o, float damping);              // assemble + PCG; verdict -> mail slot kSlotPcg
void launch_dist_update(psg_optimizer* o, bool only_on_success);     // gated on the device by the PCG verdict; count -> kSlotCount
void read_dist_verdict(const psg_optimizer* o, int* success, int* iters, float* err, long long* accepted);   // after fetch_mail
void launch_albedo_reg_solve(psg_optimizer* o, float damping, bool only_on_success);   // after launch_albedo(..., regularised = true)
void read_albedo_reg_verdict(const psg_optimizer* o, int* success, int* iters, float* err, long long* accepted);

}  // namespace psg
