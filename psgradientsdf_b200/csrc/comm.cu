// Multi-GPU plumbing: one rank per GPU, NCCL over NVLink / NVSwitch for the small reductions, peer-mapped memory for the
// distributed distance solve.
//
// The reference is a single process; this is the B200 part of the design (BASELINE.json north_star: "the volume
// is partitioned across the 8 x B200 box in z-slabs with one-voxel halos exchanged over NVLink for the stencils, and the small
// light / pose JtJ blocks are NCCL-allreduced each iteration").
//
// Partition.  The surface list is in ascending lin = i + W j + W H k order, so a contiguous ROW RANGE of it is a z-slab of the
// volume (with ragged first / last planes).  Rank g owns rows [row_bounds[g], row_bounds[g + 1]), cut so that every rank gets
// the same number of visible (voxel, keyframe) pairs.  Everything per row is produced and consumed by its owner: the sweeps
// walk the owner's rows only, the distance stencil blocks stay where the sweep wrote them, the PCG, the gated update, the
// gradient refresh and the regulariser energies run on the owner's rows.  What crosses ranks:
//   * energy / counters and the per-keyframe light (4x4 / 9x9) and pose (6x6) normal equations: one NCCL all-reduce of <= 4104
//     doubles per host round trip (fetch_mail);
//   * the one-voxel halo: entries of z, p and the stencil products of the rows within one z-plane (W H voxels) of a cut, pushed by
//     their owner into the neighbour's halo buffer over NVLink inside the persistent solve kernel (pcg.cu, k_dist_block_mg) as
//     8-byte (value, tag) packets -- the reader spins on the tag, so no fence and no barrier is needed for them -- and the
//     updated distances of those rows after the solve;
//   * two scalars per CG reduction, exchanged as tagged packets through the peers' XState mailboxes by the same kernel (no NCCL
//     call inside the solve).
// Every rank still holds the whole (fused) volume; only the rows it owns stay current between psg_opt_sync_volume calls, plus
// the distance plane of its halo.
//
// Ranks are either processes (one per GPU; NCCL communicator + CUDA IPC mappings of the peers' buffers) or threads of one
// process that share a device (psg_opt_comm_init_local: the same code path with the peers' pointers taken directly and the
// collectives done by a leader thread -- used to test the multi-rank logic on a single GPU, where the distributed solve runs
// as ONE cooperative launch whose blocks play all ranks).
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <condition_variable>
#include <cstring>
#include <mutex>

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

// ---- ranks as threads of one process ---------------------------------------------------------------------------------
// A collective is a rendezvous of the member threads: every member records an event on its stream and publishes what the
// operation needs; the last one to arrive (the leader) makes its stream wait for all events, runs the operation for everybody
// on its own stream, records a completion event and wakes the others, whose streams then wait for that event.
struct LocalGroup {
    int world = 0;
    psg_optimizer* member[kMaxRanks] = { nullptr };
    cudaEvent_t ev[kMaxRanks] = { nullptr };
    cudaEvent_t done[2] = { nullptr, nullptr };
    std::mutex m;
    std::condition_variable cv;
    int arrived = 0;
    long long generation = 0;
    void* arg[kMaxRanks] = { nullptr };   // per-member argument of the collective in flight
    PeerView views[kMaxRanks];            // filled by the leader of exchange_peers
    int failed = 0;
};

template <typename F>
static void group_collective(psg_optimizer* o, void* my_arg, F&& leader_fn) {
    LocalGroup* g = static_cast<LocalGroup*>(o->group);
    PSG_CUDA(cudaEventRecord(g->ev[o->rank], o->stream));
    std::unique_lock<std::mutex> lk(g->m);
    g->arg[o->rank] = my_arg;
    const long long gen = g->generation;
    cudaEvent_t done = g->done[gen & 1];
    if (++g->arrived == g->world) {
        int fail = 0;
        try {
            for (int q = 0; q < g->world; ++q) PSG_CUDA(cudaStreamWaitEvent(o->stream, g->ev[q], 0));
            leader_fn(g);
            PSG_CUDA(cudaEventRecord(done, o->stream));
        } catch (const CudaError&) { fail = 1; }
        g->failed = fail;
        g->arrived = 0;
        ++g->generation;
        g->cv.notify_all();
    } else {
        g->cv.wait(lk, [&] { return g->generation != gen; });
    }
    const int fail = g->failed;
    lk.unlock();
    if (fail) { set_error("a collective of the local rank group failed"); throw CudaError{-1}; }
    PSG_CUDA(cudaStreamWaitEvent(o->stream, done, 0));
}

// the same rendezvous for code outside this file: `leader` runs once, on the last member's thread and stream, with every
// member's argument
void group_run(psg_optimizer* o, void* my_arg, void (*leader)(psg_optimizer* on, void* const* args, int world, void* user), void* user) {
    group_collective(o, my_arg, [&](LocalGroup* g) { leader(o, g->arg, g->world, user); });
}

// out[q][i] = sum over members of in[q'][i], same association (member order) for everybody
__global__ void __launch_bounds__(256) k_group_allreduce(double* const* bufs /* world pointers */, int world, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double t = 0.0;
    for (int q = 0; q < world; ++q) t += bufs[q][i];
    for (int q = 0; q < world; ++q) bufs[q][i] = t;
}
__global__ void __launch_bounds__(256) k_group_broadcast(double* const* bufs, int world, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double t = bufs[0][i];
    for (int q = 1; q < world; ++q) bufs[q][i] = t;
}

struct PtrTable { DevBuf<double*> dev; };
static PtrTable& ptr_table() { static thread_local PtrTable t; return t; }

struct ReduceArg { double* ptr[3]; size_t cnt[3]; int n; };

// ---- the small sums of a sharded job over peer memory ---------------------------------------------------------------------------
// One thread per double.  It stores its value into every peer's inbox as two 8-byte packets (epoch << 32 | half of the bits:
// value and tag travel in one aligned store, so no fence and no flag), then reads the packets the peers stored into its own
// rank's inbox, polling until they carry this reduction's epoch, and adds the ranks' values in rank order -- the same
// association on every rank, identical bits everywhere.  One NVLink hop; NCCL's all-reduce of the same ~1000 doubles takes
// 17 us on two and 42 us on eight B200.  Inboxes alternate by epoch parity: a peer can be at most one reduction ahead (it
// cannot finish reduction e + 1 without this rank's packets of e + 1, which are stored after this rank has read e).
struct PacketReduceArgs {
    double* ptr[3]; unsigned cnt[3]; int n;
    unsigned long long* inbox[kMaxRanks];   // inbox[q]: rank q's inbox as addressable from here (inbox[rank] = own)
    XState* own;
    int rank, world;
    unsigned epoch;                         // >= 1
};
__device__ __forceinline__ void ar_store(unsigned long long* p, unsigned long long v) { asm volatile("st.volatile.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ unsigned long long ar_load(const unsigned long long* p) { unsigned long long v; asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
__global__ void __launch_bounds__(256) k_packet_allreduce(const PacketReduceArgs A) {
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned flat = i;
    int k = 0;
    while (k < A.n && i >= A.cnt[k]) { i -= A.cnt[k]; ++k; }
    if (k >= A.n) return;
    double* mine = A.ptr[k] + i;
    const double v = *mine;
    const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
    const unsigned long long tag = (unsigned long long)A.epoch << 32;
    const size_t lane = ((size_t)(A.epoch & 1u) * kMaxRanks + (size_t)A.rank) * (2 * (size_t)kArMaxDoubles) + 2 * (size_t)flat;
    for (int q = 0; q < A.world; ++q) {
        if (q == A.rank) continue;
        ar_store(A.inbox[q] + lane, tag | (bits & 0xffffffffull));
        ar_store(A.inbox[q] + lane + 1, tag | (bits >> 32));
    }
    double t = 0.0;
    const long long t0 = clock64();
    for (int q = 0; q < A.world; ++q) {
        if (q == A.rank) { t += v; continue; }
        const unsigned long long* src = A.inbox[A.rank] + ((size_t)(A.epoch & 1u) * kMaxRanks + (size_t)q) * (2 * (size_t)kArMaxDoubles) + 2 * (size_t)flat;
        unsigned long long lo, hi;
        for (;;) {
            lo = ar_load(src); hi = ar_load(src + 1);
            if ((lo >> 32) == A.epoch && (hi >> 32) == A.epoch) break;
            if (clock64() - t0 > 8000000000ll) {   // ~4 s: a peer is gone; make the failure loud instead of waiting forever
                A.own->abort = 1ull;
                __trap();
            }
        }
        t += __longlong_as_double((long long)((hi << 32) | (lo & 0xffffffffull)));
    }
    *mine = t;
}

void comm_allreduce_doubles(psg_optimizer* o, double* const* ptr, const size_t* count, int n) {
    if (n <= 0) return;
    if (o->group) {
        ReduceArg mine; mine.n = n;
        for (int k = 0; k < n; ++k) { mine.ptr[k] = ptr[k]; mine.cnt[k] = count[k]; }
        group_collective(o, &mine, [&](LocalGroup* g) {
            PtrTable& t = ptr_table();
            t.dev.alloc((size_t)kMaxRanks * 3);
            double* host[kMaxRanks * 3];
            for (int k = 0; k < n; ++k)
                for (int q = 0; q < g->world; ++q) host[k * kMaxRanks + q] = static_cast<ReduceArg*>(g->arg[q])->ptr[k];
            PSG_CUDA(cudaMemcpyAsync(t.dev.p, host, sizeof(host), cudaMemcpyHostToDevice, o->stream));
            for (int k = 0; k < n; ++k) {
                k_group_allreduce<<<(unsigned)((count[k] + 255) / 256), 256, 0, o->stream>>>(t.dev.p + k * kMaxRanks, g->world, count[k]);
                PSG_LAUNCH_CHECK();
            }
            PSG_CUDA(cudaStreamSynchronize(o->stream));   // `host` lives on this frame
        });
        return;
    }
    size_t total = 0;
    for (int k = 0; k < n; ++k) total += count[k];
    if (o->peers_ready && o->sharded && n <= 3 && total <= (size_t)kArMaxDoubles) {
        // peers' inboxes are mapped: one kernel, one NVLink hop
        PacketReduceArgs A;
        memset(&A, 0, sizeof(A));
        A.n = n;
        for (int k = 0; k < n; ++k) { A.ptr[k] = ptr[k]; A.cnt[k] = (unsigned)count[k]; }
        for (int q = 0; q < o->world; ++q)
            A.inbox[q] = reinterpret_cast<unsigned long long*>(reinterpret_cast<unsigned char*>(o->peer[q].xs) + kArInboxOffset);
        A.own = reinterpret_cast<XState*>(o->xstate.p);
        A.rank = o->rank; A.world = o->world;
        A.epoch = ++o->ar_epoch;
        if (total == 0) return;
        Prof pf(o->stream, KID_COMM_ALLREDUCE);
        k_packet_allreduce<<<(unsigned)((total + 255) / 256), 256, 0, o->stream>>>(A);
        count_launch(1);
        PSG_LAUNCH_CHECK();
        return;
    }
    ncclComm_t comm = static_cast<ncclComm_t>(o->comm);
    Prof pf(o->stream, KID_COMM_ALLREDUCE);
    PSG_NCCL(g_nccl.GroupStart());
    for (int k = 0; k < n; ++k)
        if (count[k]) PSG_NCCL(g_nccl.AllReduce(ptr[k], ptr[k], count[k], ncclDouble, ncclSum, comm, o->stream));
    PSG_NCCL(g_nccl.GroupEnd());
}

void comm_broadcast_doubles(psg_optimizer* o, double* a, size_t n) {
    if (!n) return;
    if (o->group) {
        ReduceArg mine; mine.n = 1; mine.ptr[0] = a; mine.cnt[0] = n;
        group_collective(o, &mine, [&](LocalGroup* g) {
            PtrTable& t = ptr_table();
            t.dev.alloc((size_t)kMaxRanks * 3);
            double* host[kMaxRanks * 3] = { nullptr };
            for (int q = 0; q < g->world; ++q) host[q] = static_cast<ReduceArg*>(g->arg[q])->ptr[0];
            PSG_CUDA(cudaMemcpyAsync(t.dev.p, host, sizeof(host), cudaMemcpyHostToDevice, o->stream));
            k_group_broadcast<<<(unsigned)((n + 255) / 256), 256, 0, o->stream>>>(t.dev.p, g->world, n);
            PSG_LAUNCH_CHECK();
            PSG_CUDA(cudaStreamSynchronize(o->stream));
        });
        return;
    }
    Prof pf(o->stream, KID_COMM_ALLREDUCE);
    PSG_NCCL(g_nccl.Broadcast(a, a, n, ncclDouble, 0, static_cast<ncclComm_t>(o->comm), o->stream));
}

// host-side barrier of the ranks (with the streams drained): a one-double all-reduce
static void comm_barrier(psg_optimizer* o) {
    if (o->world <= 1) return;
    o->red.alloc((size_t)kRedDoubles * kRedCopies + kMailDoubles);
    double* p[1] = { o->red.p + kMailDoubles - 1 };   // last double of the mailbox (slot kSlotSmall holds floats in its first half only): never summed otherwise
    const size_t c[1] = { 1 };
    PSG_CUDA(cudaMemsetAsync(p[0], 0, sizeof(double), o->stream));
    comm_allreduce_doubles(o, p, c, 1);
    PSG_CUDA(cudaStreamSynchronize(o->stream));
}

// The 32-slot blocks of a slot list that `rank` of `world` would take in a round-robin split: only used to emulate ranks on one
// rank's list (psg_opt_energy_partial); a sharded job splits ROWS (partition_rows).
void shard_blocks(int S, int rank, int world, int* first, int* stride, int* count) {
    const int nblocks = (S + 31) / 32;
    *first = rank; *stride = world;
    *count = rank < nblocks ? (nblocks - rank + world - 1) / world : 0;
}

// Sharding divides the sweeps by the number of ranks but makes every CG iteration of the distance block wait for three
// dependent NVLink hops (halo of u, halo of the stencil products, the reduction): measured on 8 x B200 that costs 150-250 us per
// outer iteration whatever the size (profiles/r2_summary.md), so it pays once the sweeps of an iteration take more than about
// twice that -- around three million visible pairs (512^3 x 30 keyframes has 4.2 million).  Below it every rank sweeps the whole
// list and only rank 0's reduced sums are broadcast, so that all ranks still solve identical small systems and stay in the same
// state bit for bit.
constexpr long long kShardMinPairs = 3000000;
void update_shard_policy(psg_optimizer* o) {
    bool want = false;
    if (o->world > 1 && (o->comm || o->group)) {
        if (o->shard_policy == 1) want = true;
        else if (o->shard_policy == 2) want = false;
        else want = o->vis_pairs >= kShardMinPairs;
    }
    if (want != o->sharded) { o->sharded = want; o->layout_dirty = true; }
}

// ---- row partition ----------------------------------------------------------------------------------------------------
// Pure host arithmetic on the (replicated) surface list, exported for the CPU tests as psg_partition_rows.
// bounds[world + 1]: cut points with equal visible-pair counts per range (every row weighs popcount(vis) + 1, so rows without
// pairs still spread); halo_lo / halo_hi[world]: rank q needs the rows [halo_lo[q], bounds[q]) below and [bounds[q + 1],
// halo_hi[q]) above its range -- all rows of other ranks whose voxel lies within `plane` (= W * H) linear indices of the
// range's first / last voxel: that covers the finite-difference neighbours (+-1, +-W, +-W H), the stencils that contain a
// row, and the 6-point Laplacian.
static void partition_rows_host(int S, const int* lin, const unsigned long long* vis, long long plane, int world, int* bounds,
                                int* halo_lo, int* halo_hi) {
    std::vector<long long> pre((size_t)S + 1, 0);
    for (int s = 0; s < S; ++s) pre[s + 1] = pre[s] + __builtin_popcountll(vis[s]) + 1;
    bounds[0] = 0;
    for (int g = 1; g < world; ++g) {
        const long long target = pre[S] * g / world;
        int b = (int)(std::lower_bound(pre.begin(), pre.end(), target) - pre.begin());
        if (b < bounds[g - 1]) b = bounds[g - 1];
        if (b > S) b = S;
        bounds[g] = b;
    }
    bounds[world] = S;
    for (int q = 0; q < world; ++q) {
        const int lo = bounds[q], hi = bounds[q + 1];
        if (hi <= lo) { halo_lo[q] = lo; halo_hi[q] = hi; continue; }   // owns nothing, needs nothing
        const long long first = lin[lo], last = lin[hi - 1];
        halo_lo[q] = (int)(std::lower_bound(lin, lin + lo, (int)std::max<long long>(first - plane, 0)) - lin);
        const long long top = std::min<long long>(last + plane, 0x7fffffffll);
        halo_hi[q] = (int)(std::upper_bound(lin + hi, lin + S, (int)top) - lin);
    }
}

void partition_rows(psg_optimizer* o) {
    const int S = o->S, world = o->sharded ? o->world : 1;
    o->row_bounds.assign(world + 1, 0); o->halo_lo.assign(world, 0); o->halo_hi.assign(world, 0);
    if (!o->sharded) {
        o->row_bounds[1] = S; o->halo_hi[0] = S;
        o->row_lo = 0; o->row_hi = S; o->n_slots = S;
        return;
    }
    std::vector<int> lin((size_t)(S > 0 ? S : 1));
    std::vector<unsigned long long> vis((size_t)(S > 0 ? S : 1));
    if (S > 0) {
        PSG_CUDA(cudaMemcpyAsync(lin.data(), o->lin.p, (size_t)S * sizeof(int), cudaMemcpyDeviceToHost, o->stream));
        PSG_CUDA(cudaMemcpyAsync(vis.data(), o->svis.p, (size_t)S * sizeof(unsigned long long), cudaMemcpyDeviceToHost, o->stream));
        PSG_CUDA(cudaStreamSynchronize(o->stream));
    }
    const long long plane = (long long)o->vol->dims[0] * o->vol->dims[1];
    partition_rows_host(S, lin.data(), vis.data(), plane, world, o->row_bounds.data(), o->halo_lo.data(), o->halo_hi.data());
    o->row_lo = o->row_bounds[o->rank]; o->row_hi = o->row_bounds[o->rank + 1];
    o->n_slots = o->row_hi - o->row_lo;
}

// ---- peers' views of each other's buffers ----------------------------------------------------------------------------------
struct IpcRecord { cudaIpcMemHandle_t h; unsigned long long offset; };
struct IpcPacket { IpcRecord halo, xs; };

typedef int (*cuMemGetAddressRange_t)(unsigned long long*, size_t*, unsigned long long);
static cuMemGetAddressRange_t address_range_fn() {
    static cuMemGetAddressRange_t fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qr;
        PSG_CUDA(cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &qr));
        if (!p || qr != cudaDriverEntryPointSuccess) { set_error("cuMemGetAddressRange not available"); throw CudaError{-1}; }
        fn = reinterpret_cast<cuMemGetAddressRange_t>(p);
    }
    return fn;
}

// an IPC handle always names a whole allocation; small cudaMalloc blocks are carved out of larger ones, so the offset travels too
static IpcRecord ipc_export(const void* ptr) {
    IpcRecord r;
    memset(&r, 0, sizeof(r));
    unsigned long long base = 0; size_t size = 0;
    const int st = address_range_fn()(&base, &size, (unsigned long long)(uintptr_t)ptr);
    if (st != 0) { set_error("cuMemGetAddressRange failed (%d)", st); throw CudaError{st}; }
    PSG_CUDA(cudaIpcGetMemHandle(&r.h, reinterpret_cast<void*>((uintptr_t)base)));
    r.offset = (unsigned long long)(uintptr_t)ptr - base;
    return r;
}

struct IpcCache { std::vector<std::pair<cudaIpcMemHandle_t, void*>> open; };

static void* ipc_import(psg_optimizer* o, IpcCache& cache, const IpcRecord& r) {
    for (auto& e : cache.open)
        if (memcmp(&e.first, &r.h, sizeof(r.h)) == 0) return static_cast<char*>(e.second) + r.offset;
    void* base = nullptr;
    PSG_CUDA(cudaIpcOpenMemHandle(&base, r.h, cudaIpcMemLazyEnablePeerAccess));
    cache.open.push_back({r.h, base});
    o->ipc_opened.push_back(base);
    return static_cast<char*>(base) + r.offset;
}

void release_peers(psg_optimizer* o) {
    bool had = false;
    o->peers_ready = false;
    for (int q = 0; q < kMaxRanks; ++q) { if (o->peer[q].xs && q != o->rank) had = true; o->peer[q] = PeerView(); }
    if (!o->ipc_opened.empty()) {
        PSG_CUDA(cudaStreamSynchronize(o->stream));
        for (void* b : o->ipc_opened) cudaIpcCloseMemHandle(b);
        o->ipc_opened.clear();
    }
    // nobody may free or reallocate a buffer before every peer has let go of it
    if (had && o->world > 1 && (o->comm || o->group)) comm_barrier(o);
}

void exchange_peers(psg_optimizer* o) {
    o->peers_ready = false;
    for (int q = 0; q < kMaxRanks; ++q) o->peer[q] = PeerView();
    if (!o->sharded) return;
    trace("[rank %d] exchange_peers: rows %d..%d of %d", o->rank, o->row_lo, o->row_hi, o->S);
    if (o->world > kMaxRanks) { set_error("a sharded job has at most %d ranks (%d given)", kMaxRanks, o->world); throw CudaError{-1}; }
    // halo packets of this rank: one per halo row and kind, tags start at zero (= never written)
    const int nh = (o->row_lo - o->halo_lo[o->rank]) + (o->halo_hi[o->rank] - o->row_hi);
    const size_t hcount = (size_t)kHaloKinds * (size_t)(nh > 0 ? nh : 1);
    if (o->halo.n < hcount) o->halo.alloc(hcount);
    PSG_CUDA(cudaMemsetAsync(o->halo.p, 0, o->halo.n * sizeof(unsigned long long), o->stream));
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    PeerView mine;
    mine.halo = o->halo.p; mine.xs = reinterpret_cast<XState*>(o->xstate.p);
    if (o->group) {
        // same address space: take the members' pointers (the rendezvous doubles as the barrier after allocation)
        group_collective(o, &mine, [&](LocalGroup* g) { for (int q = 0; q < g->world; ++q) g->views[q] = *static_cast<PeerView*>(g->arg[q]); });
        LocalGroup* g = static_cast<LocalGroup*>(o->group);
        for (int q = 0; q < o->world; ++q) o->peer[q] = g->views[q];
        group_collective(o, nullptr, [&](LocalGroup*) {});   // everybody has copied before the table is reused
        return;
    }
    // processes: all-gather the IPC handles through NCCL, then map
    IpcPacket pk;
    pk.halo = ipc_export(mine.halo); pk.xs = ipc_export(mine.xs);
    DevBuf<unsigned char> buf;
    buf.alloc(sizeof(IpcPacket) * (size_t)o->world);
    PSG_CUDA(cudaMemcpyAsync(buf.p + sizeof(IpcPacket) * (size_t)o->rank, &pk, sizeof(pk), cudaMemcpyHostToDevice, o->stream));
    PSG_NCCL(g_nccl.AllGather(buf.p + sizeof(IpcPacket) * (size_t)o->rank, buf.p, sizeof(IpcPacket), ncclChar,
                              static_cast<ncclComm_t>(o->comm), o->stream));
    std::vector<IpcPacket> all((size_t)o->world);
    PSG_CUDA(cudaMemcpyAsync(all.data(), buf.p, sizeof(IpcPacket) * (size_t)o->world, cudaMemcpyDeviceToHost, o->stream));
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    buf.release();
    IpcCache cache;
    for (int q = 0; q < o->world; ++q) {
        if (q == o->rank) { o->peer[q] = mine; continue; }
        o->peer[q].halo = static_cast<unsigned long long*>(ipc_import(o, cache, all[q].halo));
        o->peer[q].xs = static_cast<XState*>(ipc_import(o, cache, all[q].xs));
    }
    comm_barrier(o);   // every rank has mapped everything before anyone launches a kernel that writes to a peer
    o->peers_ready = true;   // from here on the small sums travel as packets (k_packet_allreduce)
    trace("[rank %d] exchange_peers done", o->rank);
}

// ---- psg_opt_sync_volume: the rows every rank owns -> everybody ---------------------------------------------------------------
// packed per rank: 7 planes {dist, gx, gy, gz, r, g, b} of `chunk` rows (chunk = the largest range, so that equal-sized
// messages can be all-gathered)
__global__ void __launch_bounds__(256) k_pack_rows(VolView vol, const int* __restrict__ lin, int row0, int n, int chunk, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int l = lin[row0 + i];
    out[i] = vol.dist[l]; out[chunk + i] = vol.gx[l]; out[2 * (size_t)chunk + i] = vol.gy[l]; out[3 * (size_t)chunk + i] = vol.gz[l];
    out[4 * (size_t)chunk + i] = vol.r[l]; out[5 * (size_t)chunk + i] = vol.g[l]; out[6 * (size_t)chunk + i] = vol.b[l];
}
__global__ void __launch_bounds__(256) k_unpack_rows(VolView vol, const int* __restrict__ lin, int row0, int n, int chunk, const float* __restrict__ in) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int l = lin[row0 + i];
    vol.dist[l] = in[i]; vol.gx[l] = in[chunk + i]; vol.gy[l] = in[2 * (size_t)chunk + i]; vol.gz[l] = in[3 * (size_t)chunk + i];
    vol.r[l] = in[4 * (size_t)chunk + i]; vol.g[l] = in[5 * (size_t)chunk + i]; vol.b[l] = in[6 * (size_t)chunk + i];
}

struct GatherArg { float* buf; };

void sync_volume(psg_optimizer* o) {
    if (!o->sharded || !o->vol_unsynced) return;
    o->vol_unsynced = false;
    if (o->S == 0) return;
    const int world = o->world;
    int chunk = 1;
    for (int q = 0; q < world; ++q) chunk = std::max(chunk, o->row_bounds[q + 1] - o->row_bounds[q]);
    o->comm_buf.alloc((size_t)chunk * 7 * world);
    const VolView vol = o->vol->view();
    float* mine = o->comm_buf.p + (size_t)o->rank * 7 * chunk;
    if (o->n_slots > 0) {
        k_pack_rows<<<(o->n_slots + 255) / 256, 256, 0, o->stream>>>(vol, o->lin.p, o->row_lo, o->n_slots, chunk, mine);
        count_launch(); PSG_LAUNCH_CHECK();
    }
    if (o->group) {
        GatherArg ga; ga.buf = o->comm_buf.p;
        group_collective(o, &ga, [&](LocalGroup* g) {
            for (int q = 0; q < world; ++q)
                for (int p = 0; p < world; ++p) {
                    if (p == q) continue;
                    float* src = static_cast<GatherArg*>(g->arg[p])->buf + (size_t)p * 7 * chunk;
                    float* dst = static_cast<GatherArg*>(g->arg[q])->buf + (size_t)p * 7 * chunk;
                    PSG_CUDA(cudaMemcpyAsync(dst, src, (size_t)7 * chunk * sizeof(float), cudaMemcpyDeviceToDevice, o->stream));
                }
        });
    } else {
        Prof pf(o->stream, KID_COMM_ALLGATHER);
        PSG_NCCL(g_nccl.AllGather(mine, o->comm_buf.p, (size_t)7 * chunk, ncclFloat, static_cast<ncclComm_t>(o->comm), o->stream));
    }
    for (int q = 0; q < world; ++q) {
        const int n = o->row_bounds[q + 1] - o->row_bounds[q];
        if (q == o->rank || n <= 0) continue;
        { Prof pf(o->stream, KID_COMM_SCATTER);
          k_unpack_rows<<<(n + 255) / 256, 256, 0, o->stream>>>(vol, o->lin.p, o->row_bounds[q], n, chunk, o->comm_buf.p + (size_t)q * 7 * chunk); }
        PSG_LAUNCH_CHECK();
    }
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    // the slot records cover own rows only and those did not change
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
    if (o->comm || o->group) { set_error("psg_opt_comm_init: a communicator already exists"); return PSG_ERR_STATE; }
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

int psg_opt_comm_init_local(psg_optimizer* const* opts, int world) {
    if (!opts || world < 1 || world > kMaxRanks) { set_error("psg_opt_comm_init_local: 1 <= world <= %d optimizers expected", kMaxRanks); return PSG_ERR_INVALID; }
    for (int q = 0; q < world; ++q) {
        if (!opts[q]) { set_error("psg_opt_comm_init_local: optimizer %d is null", q); return PSG_ERR_INVALID; }
        if (opts[q]->comm || opts[q]->group) { set_error("psg_opt_comm_init_local: optimizer %d already has a communicator", q); return PSG_ERR_STATE; }
        if (opts[q]->vol->device != opts[0]->vol->device) { set_error("psg_opt_comm_init_local: all members must live on one device"); return PSG_ERR_INVALID; }
    }
    try {
        PSG_CUDA(cudaSetDevice(opts[0]->vol->device));
        LocalGroup* g = new LocalGroup;
        g->world = world;
        for (int q = 0; q < world; ++q) { g->member[q] = opts[q]; PSG_CUDA(cudaEventCreateWithFlags(&g->ev[q], cudaEventDisableTiming)); }
        for (int k = 0; k < 2; ++k) PSG_CUDA(cudaEventCreateWithFlags(&g->done[k], cudaEventDisableTiming));
        for (int q = 0; q < world; ++q) {
            opts[q]->group = g; opts[q]->rank = q; opts[q]->world = world;
            update_shard_policy(opts[q]);
        }
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

}  // extern "C"

// collective = every rank is inside this call (psg_opt_comm_destroy); psg_opt_destroy on a handle that still has a communicator
// only lets go of the local resources
int psg::comm_destroy(psg_optimizer* o, bool collective) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    try {
        if (o->comm || o->group) {
            cudaSetDevice(o->vol->device);
            cudaStreamSynchronize(o->stream);
            if (collective) {
                if (o->sharded) sync_volume(o);
                release_peers(o);
            } else {
                for (void* b : o->ipc_opened) cudaIpcCloseMemHandle(b);
                o->ipc_opened.clear();
                o->peers_ready = false;
                for (int q = 0; q < kMaxRanks; ++q) o->peer[q] = PeerView();
            }
        }
        if (o->comm) {
            g_nccl.CommDestroy(static_cast<ncclComm_t>(o->comm));
            o->comm = nullptr;
        }
        if (o->group) {
            LocalGroup* g = static_cast<LocalGroup*>(o->group);
            bool last = false;
            {
                std::lock_guard<std::mutex> lk(g->m);
                g->member[o->rank] = nullptr;
                last = true;
                for (int q = 0; q < g->world; ++q) if (g->member[q]) last = false;
            }
            o->group = nullptr;
            if (last) {
                for (int q = 0; q < g->world; ++q) if (g->ev[q]) cudaEventDestroy(g->ev[q]);
                for (int k = 0; k < 2; ++k) if (g->done[k]) cudaEventDestroy(g->done[k]);
                delete g;
            }
        }
    } catch (const psg::CudaError&) { o->comm = nullptr; o->group = nullptr; o->rank = 0; o->world = 1; o->sharded = false; o->layout_dirty = true; return PSG_ERR_CUDA; }
    o->rank = 0; o->world = 1;
    if (o->sharded) { o->sharded = false; o->layout_dirty = true; }
    return PSG_OK;
}

extern "C" {

int psg_opt_comm_destroy(psg_optimizer* o) { return psg::comm_destroy(o, true); }

int psg_opt_set_shard_policy(psg_optimizer* o, int policy) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (policy < 0 || policy > 2) { set_error("psg_opt_set_shard_policy: %d is not 0 (by work), 1 (always) or 2 (never)", policy); return PSG_ERR_INVALID; }
    try {
        PSG_CUDA(cudaSetDevice(o->vol->device));
        if (o->sharded && policy == 2) sync_volume(o);   // leaving the sharded mode: every rank gets the whole state back
        o->shard_policy = policy;
        update_shard_policy(o);
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

int psg_opt_is_sharded(const psg_optimizer* o, int* sharded, long long* visible_pairs) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (sharded) *sharded = o->sharded ? 1 : 0;
    if (visible_pairs) *visible_pairs = o->vis_pairs;
    return PSG_OK;
}

int psg_opt_row_range(psg_optimizer* o, int* row_lo, int* row_hi, int* halo_lo, int* halo_hi) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (!o->inited) { set_error("psg_opt_init must be called first"); return PSG_ERR_STATE; }
    try {
        PSG_CUDA(cudaSetDevice(o->vol->device));
        if (o->layout_dirty) build_layout(o);
        if (row_lo) *row_lo = o->row_lo;
        if (row_hi) *row_hi = o->row_hi;
        if (halo_lo) *halo_lo = o->halo_lo[o->sharded ? o->rank : 0];
        if (halo_hi) *halo_hi = o->halo_hi[o->sharded ? o->rank : 0];
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

int psg_opt_sync_volume(psg_optimizer* o) {
    if (!o) { set_error("null optimizer"); return PSG_ERR_INVALID; }
    if (!o->inited) { set_error("psg_opt_init must be called first"); return PSG_ERR_STATE; }
    try {
        PSG_CUDA(cudaSetDevice(o->vol->device));
        if (o->layout_dirty) build_layout(o);
        sync_volume(o);
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

int psg_partition_rows(int S, const int* lin, const uint64_t* vis, long long plane, int world, int* bounds, int* halo_lo, int* halo_hi) {
    if (S < 0 || world < 1 || (S > 0 && (!lin || !vis)) || plane < 1 || !bounds || !halo_lo || !halo_hi) { set_error("psg_partition_rows: bad argument"); return PSG_ERR_INVALID; }
    partition_rows_host(S, lin, reinterpret_cast<const unsigned long long*>(vis), plane, world, bounds, halo_lo, halo_hi);
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
