// Multi-GPU plumbing: one process per GPU, NCCL over NVLink / NVSwitch.
//
// The reference is a single process; this is the B200 part of the design.  Every rank keeps the whole
// volume and work list; the sweeps are split by contiguous slot ranges of the visibility-sorted work list
// (psg_shard_range).  What has to cross ranks per outer iteration is small:
//   * energy / counters and the per-keyframe light (4x4 / 9x9) and pose (6x6) normal equations: all-reduce of <= 4104 doubles;
//   * albedo after initAlbedo / the albedo block: all-gather of the {r, g, b} records (16 B per surface voxel);
//   * distance: all-reduce of the 14 stencil-block planes (56 B per surface voxel), after which every rank runs the same
//     deterministic distance solve and applies the same update, so no further exchange is needed.
// All collectives are enqueued on the optimizer's stream, so they order with the kernels without host synchronisation.
#include <dlfcn.h>
#include <nccl.h>

#include "psg_host.h"

namespace psg {

// NCCL is bound at first use, not at link time: a process that also imports PyTorch must end up with ONE libnccl.so.2
// (PyTorch bundles a newer one than the system's, and whichever is loaded first wins the soname).  Order of preference:
// a copy that is already loaded (RTLD_NOLOAD), then $PSG_NCCL_LIB, then the system library.
struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};
static NcclApi g_nccl;

static bool nccl_load() {
    if (g_nccl.ok) return true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) { const char* env = getenv("PSG_NCCL_LIB"); if (env && *env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL); }
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { set_error("NCCL not available: %s", dlerror()); return false; }
#define PSG_SYM(field, name)                                                                        \
    g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name));                        \
    if (!g_nccl.field) { set_error("NCCL symbol %s not found", name); return false; }
    PSG_SYM(GetUniqueId, "ncclGetUniqueId")
    PSG_SYM(Broadcast, "ncclBroadcast")
    PSG_SYM(CommInitRank, "ncclCommInitRank")
    PSG_SYM(CommDestroy, "ncclCommDestroy")
    PSG_SYM(AllReduce, "ncclAllReduce")
    PSG_SYM(AllGather, "ncclAllGather")
    PSG_SYM(GroupStart, "ncclGroupStart")
    PSG_SYM(GroupEnd, "ncclGroupEnd")
    PSG_SYM(GetErrorString, "ncclGetErrorString")
#undef PSG_SYM
    g_nccl.ok = true;
    return true;
}

#define PSG_NCCL(call)                                                                                   \
    do {                                                                                                 \
        ncclResult_t r__ = (call);                                                                       \
        if (r__ != ncclSuccess) {                                                                        \
            psg::set_error("%s failed at %s:%d: %s", #call, __FILE__, __LINE__, g_nccl.GetErrorString(r__)); \
            throw psg::CudaError{(int)r__};                                                              \
        }                                                                                                \
    } while (0)

// The work list is cut into blocks of 32 slots; rank r sweeps blocks r, r + world, r + 2 world, ...  The list is sorted by
// visibility (most keyframes first), so contiguous ranges would give the first ranks the expensive end; round-robin blocks
// give every rank the same mix.
void shard_blocks(int S, int rank, int world, int* first, int* stride, int* count) {
    const int nblocks = (S + 31) / 32;
    *first = rank; *stride = world;
    *count = rank < nblocks ? (nblocks - rank + world - 1) / world : 0;
}

// slots per rank in the packed all-gather layout (the same for every rank: padded to the largest share)
static size_t chunk_slots(const psg_optimizer* o) {
    const int nblocks = (o->S + 31) / 32;
    return (size_t)((nblocks + o->world - 1) / o->world) * 32;
}

void comm_allreduce_doubles(psg_optimizer* o, double* a, size_t na, double* b, size_t nb) {
    ncclComm_t comm = static_cast<ncclComm_t>(o->comm);
    Prof pf(o->stream, KID_COMM_ALLREDUCE);
    PSG_NCCL(g_nccl.GroupStart());
    if (na) PSG_NCCL(g_nccl.AllReduce(a, a, na, ncclDouble, ncclSum, comm, o->stream));
    if (nb) PSG_NCCL(g_nccl.AllReduce(b, b, nb, ncclDouble, ncclSum, comm, o->stream));
    PSG_NCCL(g_nccl.GroupEnd());
}

void comm_broadcast_doubles(psg_optimizer* o, double* a, size_t n) {
    if (!n) return;
    Prof pf(o->stream, KID_COMM_ALLREDUCE);
    PSG_NCCL(g_nccl.Broadcast(a, a, n, ncclDouble, 0, static_cast<ncclComm_t>(o->comm), o->stream));
}

// Sharding pays only when a rank's share of the sweeps outweighs the collectives it adds (~7 per outer iteration, 20-400 us
// each): measured on B200 x 2..8, the break-even is around two million visible pairs per rank (512^3 x 30 keyframes on two
// GPUs).  Below it every rank sweeps the whole list and only rank 0's reduced sums are broadcast, so that all ranks still
// solve identical small systems and stay in the same state bit for bit.
constexpr long long kShardMinPairsPerRank = 2000000;
void update_shard_policy(psg_optimizer* o) {
    if (o->world <= 1 || !o->comm) { o->sharded = false; return; }
    if (o->shard_policy == 1) o->sharded = true;
    else if (o->shard_policy == 2) o->sharded = false;
    else o->sharded = o->vis_pairs / o->world >= kShardMinPairsPerRank;
}

// packed index (rank r, local j) -> work-list slot
__device__ __forceinline__ int packed_slot(int r, int j, int world) { return ((j >> 5) * world + r) * 32 + (j & 31); }

// albedo of every slot -> rec1 + dense volume planes (own slots were written by the sweep itself)
__global__ void __launch_bounds__(256) k_scatter_rho(VolView vol, SurfView surf, const float4* __restrict__ buf, int chunk, int world, int own) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= chunk * world) return;
    const int r = t / chunk, j = t - r * chunk;
    if (r == own) return;
    const int slot = packed_slot(r, j, world);
    if (slot >= surf.S) return;
    const float4 r1 = buf[t];
    surf.rec1[slot] = r1;
    const int lin = __float_as_int(surf.rec3[slot].w);
    vol.r[lin] = r1.x; vol.g[lin] = r1.y; vol.b[lin] = r1.z;
}

// distance stencil blocks of every slot -> 14 planes indexed by surface row
__global__ void __launch_bounds__(256) k_scatter_stencil(SurfView surf, const float* __restrict__ buf, float* __restrict__ blocks, int chunk, int world) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= chunk * world) return;
    const int r = t / chunk, j = t - r * chunk;
    const int slot = packed_slot(r, j, world);
    if (slot >= surf.S) return;
    const int s = __float_as_int(surf.rec2[slot].w);
    const float* src = buf + (size_t)r * 14 * chunk + j;
#pragma unroll
    for (int i = 0; i < 14; ++i) blocks[(size_t)i * surf.S + s] = src[(size_t)i * chunk];
}

float4* comm_pack_rho(psg_optimizer* o) {
    const size_t chunk = chunk_slots(o);
    o->comm_buf.alloc(chunk * o->world * 14);   // sized for the larger user (stencil); 16-byte aligned by cudaMalloc
    return reinterpret_cast<float4*>(o->comm_buf.p) + (size_t)o->rank * chunk;
}

void comm_allgather_rho(psg_optimizer* o) {
    if (o->S == 0) return;
    const size_t chunk = chunk_slots(o);
    float4* base = reinterpret_cast<float4*>(o->comm_buf.p);
    { Prof pf(o->stream, KID_COMM_ALLGATHER);
      PSG_NCCL(g_nccl.AllGather(base + (size_t)o->rank * chunk, base, chunk * 4, ncclFloat, static_cast<ncclComm_t>(o->comm), o->stream)); }
    const SurfView sv = surf_view(o);
    const int n = (int)(chunk * o->world);
    { Prof pf(o->stream, KID_COMM_SCATTER); k_scatter_rho<<<(n + 255) / 256, 256, 0, o->stream>>>(o->vol->view(), sv, base, (int)chunk, o->world, o->rank); }
    PSG_LAUNCH_CHECK();
}

float* comm_pack_stencil(psg_optimizer* o, int* stride) {
    const size_t chunk = chunk_slots(o);
    o->comm_buf.alloc(chunk * o->world * 14);
    *stride = (int)chunk;
    return o->comm_buf.p + (size_t)o->rank * 14 * chunk;
}

void comm_allgather_stencil(psg_optimizer* o, float* blocks) {
    if (o->S == 0) return;
    const size_t chunk = chunk_slots(o);
    float* base = o->comm_buf.p;
    { Prof pf(o->stream, KID_COMM_ALLGATHER);
      PSG_NCCL(g_nccl.AllGather(base + (size_t)o->rank * 14 * chunk, base, 14 * chunk, ncclFloat, static_cast<ncclComm_t>(o->comm), o->stream)); }
    const SurfView sv = surf_view(o);
    const int n = (int)(chunk * o->world);
    { Prof pf(o->stream, KID_COMM_SCATTER); k_scatter_stencil<<<(n + 255) / 256, 256, 0, o->stream>>>(sv, base, blocks, (int)chunk, o->world); }
    PSG_LAUNCH_CHECK();
}

}  // namespace psg

using namespace psg;

extern "C" {

int psg_comm_unique_id(unsigned char id[128]) {
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is expected to be 128 bytes");
    if (!id) { set_error("psg_comm_unique_id: null argument"); return PSG_ERR_INVALID; }
    if (!nccl_load()) return PSG_ERR_CUDA;
    ncclUniqueId u;
    const ncclResult_t r = g_nccl.GetUniqueId(&u);
    if (r != ncclSuccess) { set_error("ncclGetUniqueId failed: %s", g_nccl.GetErrorString(r)); return PSG_ERR_CUDA; }
    memcpy(id, &u, 128);
    return PSG_OK;
}

int psg_opt_comm_init(psg_optimizer* o, const unsigned char id[128], int rank, int world) {
    if (!o || !id) { set_error("psg_opt_comm_init: null argument"); return PSG_ERR_INVALID; }
    if (world < 1 || world > 64 || rank < 0 || rank >= world) { set_error("psg_opt_comm_init: rank %d of %d is not valid (1 <= world <= 64)", rank, world); return PSG_ERR_INVALID; }
    if (o->comm) { set_error("psg_opt_comm_init: a communicator already exists"); return PSG_ERR_STATE; }
    if (!nccl_load()) return PSG_ERR_CUDA;
    try {
        PSG_CUDA(cudaSetDevice(o->vol->device));
        ncclUniqueId u;
        memcpy(&u, id, 128);
        ncclComm_t comm = nullptr;
        PSG_NCCL(g_nccl.CommInitRank(&comm, world, u, rank));
        o->comm = comm; o->rank = rank; o->world = world;
        update_shard_policy(o);
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

int psg_opt_comm_destroy(psg_optimizer* o) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (o->comm) {
        cudaSetDevice(o->vol->device);
        cudaStreamSynchronize(o->stream);
        g_nccl.CommDestroy(static_cast<ncclComm_t>(o->comm));
        o->comm = nullptr;
    }
    o->rank = 0; o->world = 1; o->sharded = false;
    return PSG_OK;
}

int psg_opt_set_shard_policy(psg_optimizer* o, int policy) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (policy < 0 || policy > 2) { set_error("psg_opt_set_shard_policy: %d is not 0 (by work), 1 (always) or 2 (never)", policy); return PSG_ERR_INVALID; }
    o->shard_policy = policy;
    update_shard_policy(o);
    return PSG_OK;
}

int psg_opt_is_sharded(const psg_optimizer* o, int* sharded, long long* visible_pairs) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (sharded) *sharded = o->sharded ? 1 : 0;
    if (visible_pairs) *visible_pairs = o->vis_pairs;
    return PSG_OK;
}

int psg_shard_blocks(int S, int rank, int world, int* first, int* stride, int* count) {
    if (S < 0 || world < 1 || rank < 0 || rank >= world || !first || !stride || !count) { set_error("psg_shard_blocks: bad argument"); return PSG_ERR_INVALID; }
    shard_blocks(S, rank, world, first, stride, count);
    return PSG_OK;
}

int psg_opt_energy_partial(psg_optimizer* o, int rank, int world, double* loss_sum, long long* n_active) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (!o->inited) { set_error("psg_opt_init must be called first"); return PSG_ERR_STATE; }
    if (world < 1 || rank < 0 || rank >= world) { set_error("psg_opt_energy_partial: rank %d of %d is not valid", rank, world); return PSG_ERR_INVALID; }
    try {
        PSG_CUDA(cudaSetDevice(o->vol->device));
        double s = 0.0; long long n = 0;
        sweep_energy_partial(o, rank, world, &s, &n);
        if (loss_sum) *loss_sum = s;
        if (n_active) *n_active = n;
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

}  // extern "C"
