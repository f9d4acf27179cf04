// Keyframe-visibility selection, surface-voxel compaction into a dense ordered work list, finite-difference
// neighbour tables, and 2x upsampling.
//
// Replaces Optimizer::select_vis (cpp/include/ps_optimizer/Optimizer.cpp:30-47), Optimizer::getSurfaceVoxel
// (OptimizerAux.cpp:237-257), the std::find membership tests of Optimizer::ifValidDirection
// (Optimizer.cpp:462-474) and Optimizer::subsampling (OptimizerAux.cpp:622-684).
//
// Membership is a bitmask over the volume (1 bit / voxel) plus an exclusive popcount prefix per 32-voxel word
// (a rank structure, 0.25 B / voxel instead of a 4 B / voxel row map): member(l) is a bit test and
// row(l) = prefix[l/32] + popc(mask[l/32] below bit l%32), which is also what keeps the list in ascending lin order.
#include "psg_host.h"

namespace psg {

struct SelectParams {
    int F;
    int frame_idx[kMaxFrames];
    double thr;   // sqrt(3.0) * voxel_size, evaluated in double (OptimizerAux.cpp:249)
};

// Optimizer::select_vis for one voxel: keyframe slot k takes vis[frame_idx[k]] while frame_idx[k] < vis.size()
// (slots are filled consecutively, exactly like the push_back loop), the rest is padded with false.
__device__ __forceinline__ unsigned long long select_vis(const VolView& v, long long lin, const SelectParams& P) {
    int len = 0;
    for (int w = v.nwords - 1; w >= 0; --w) {
        const unsigned long long word = v.vis[(long long)w * v.N + lin];
        if (word) { len = 64 * w + (64 - __clzll((long long)word)); break; }
    }
    if (len == 0) return 0ull;
    unsigned long long out = 0ull;
    int posn = 0;
    for (int k = 0; k < P.F; ++k) {
        const int fi = P.frame_idx[k];
        if ((unsigned)fi < (unsigned)len) {
            const unsigned long long word = v.vis[(long long)(fi >> 6) * v.N + lin];
            out |= ((word >> (fi & 63)) & 1ull) << posn;
            ++posn;
        }
    }
    return out;
}

// Pass 1: surface predicate per voxel -> membership words + per-block counts.
// Each warp owns consecutive 32-voxel words; a block owns words [blockIdx.x * wpb, (blockIdx.x + 1) * wpb).
__global__ void __launch_bounds__(256) k_surface_mask(VolView v, SelectParams P, unsigned* __restrict__ mask,
                                                     unsigned* __restrict__ block_counts, unsigned* __restrict__ block_pairs,
                                                     long long nwords_total, int wpb) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const long long w0 = (long long)blockIdx.x * wpb;
    unsigned cnt = 0, pairs = 0;
    for (long long w = w0 + warp; w < w0 + wpb && w < nwords_total; w += nwarps) {
        const long long lin = w * 32 + lane;
        unsigned long long kv = 0ull;
        if (lin < v.N) {
            const float d = v.dist[lin];
            if (fabs((double)d) <= P.thr) kv = select_vis(v, lin, P);
        }
        const unsigned m = __ballot_sync(0xffffffffu, kv != 0ull);
        pairs += (unsigned)__popcll(kv);
        if (lane == 0) { mask[w] = m; cnt += __popc(m); }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) pairs += __shfl_xor_sync(0xffffffffu, pairs, o);
    __shared__ unsigned s_cnt[8], s_pairs[8];
    if (lane == 0) { s_cnt[warp] = cnt; s_pairs[warp] = pairs; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned t = 0, tp = 0;
        for (int i = 0; i < nwarps; ++i) { t += s_cnt[i]; tp += s_pairs[i]; }
        block_counts[blockIdx.x] = t;
        block_pairs[blockIdx.x] = tp;      // visible (voxel, keyframe) pairs of the block's surface voxels
    }
}

// Pass 2: exclusive scan of the per-block counts (single block, chunked with a carry); total -> counts[n].
__global__ void __launch_bounds__(1024) k_scan_counts(unsigned* __restrict__ counts, int n, const unsigned* __restrict__ block_pairs,
                                                     unsigned long long* __restrict__ totals /* [0] surface voxels, [1] visible pairs */) {
    __shared__ unsigned s_warp[32];
    __shared__ unsigned s_carry;
    __shared__ unsigned long long s_pairs[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = 0;
    {
        unsigned long long tp = 0ull;
        for (int i = threadIdx.x; i < n; i += 1024) tp += block_pairs[i];
#pragma unroll
        for (int o = 16; o; o >>= 1) tp += __shfl_xor_sync(0xffffffffu, tp, o);
        if (lane == 0) s_pairs[warp] = tp;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long tp = 0ull;
        for (int i = 0; i < 32; ++i) tp += s_pairs[i];
        totals[1] = tp;
    }
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const unsigned x = i < n ? counts[i] : 0u;
        unsigned inc = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            unsigned ws = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, ws, o); if (lane >= o) ws += y; }
            s_warp[lane] = ws;   // inclusive
        }
        __syncthreads();
        const unsigned carry = s_carry;
        const unsigned warp_off = warp ? s_warp[warp - 1] : 0u;
        if (i < n) counts[i] = carry + warp_off + inc - x;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + warp_off + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) { counts[n] = s_carry; totals[0] = s_carry; }
}

// Pass 3: per-word exclusive prefix, then ordered scatter of lin and keyframe visibility.
__global__ void __launch_bounds__(256) k_surface_scatter(VolView v, SelectParams P, const unsigned* __restrict__ mask,
                                                        const unsigned* __restrict__ block_offsets,
                                                        unsigned* __restrict__ prefix, long long nwords_total, int wpb,
                                                        int* __restrict__ lin_out, unsigned long long* __restrict__ vis_out) {
    __shared__ unsigned s_warp[8];
    __shared__ unsigned s_carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const long long w0 = (long long)blockIdx.x * wpb;
    if (threadIdx.x == 0) s_carry = block_offsets[blockIdx.x];
    __syncthreads();
    // phase A: prefix per word, one word per thread per chunk
    for (long long base = w0; base < w0 + wpb && base < nwords_total; base += blockDim.x) {
        const long long w = base + threadIdx.x;
        const bool in = (w < w0 + wpb) && (w < nwords_total);
        const unsigned x = in ? __popc(mask[w]) : 0u;
        unsigned inc = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        unsigned warp_off = 0;
        for (int i = 0; i < warp; ++i) warp_off += s_warp[i];
        const unsigned carry = s_carry;
        if (in) prefix[w] = carry + warp_off + inc - x;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) s_carry = carry + warp_off + inc;
        __syncthreads();
    }
    // phase B: scatter (prefix[] written above by this block; visible after the barrier)
    for (long long w = w0 + warp; w < w0 + wpb && w < nwords_total; w += nwarps) {
        const unsigned m = mask[w];
        if (m == 0u) continue;
        if ((m >> lane) & 1u) {
            const unsigned row = prefix[w] + __popc(m & ((1u << lane) - 1u));
            const long long lin = w * 32 + lane;
            lin_out[row] = (int)lin;
            vis_out[row] = select_vis(v, lin, P);
        }
    }
}

__device__ __forceinline__ bool is_member(const unsigned* mask, long long l) { return (mask[l >> 5] >> (l & 31)) & 1u; }
__device__ __forceinline__ int row_of(const unsigned* mask, const unsigned* prefix, long long l) {
    const unsigned m = mask[l >> 5];
    return (int)(prefix[l >> 5] + __popc(m & ((1u << (l & 31)) - 1u)));
}

// Pass 4: finite-difference neighbour per axis as Optimizer::computeDistGrad chooses it (Optimizer.cpp:287-347):
// forward neighbour if it is a surface voxel, else the backward neighbour unchecked.  Neighbours are addressed by
// linear offset exactly as idx2line does.  A backward neighbour below index 0 (undefined behaviour in the reference)
// is recorded as -1 and treated as "same value as the centre".
// inc[6*S]: surface rows whose 4-voxel stencil contains this row as neighbour along axis a (see pcg.cu).
__global__ void __launch_bounds__(256) k_neighbors(VolView v, const unsigned* __restrict__ mask, const unsigned* __restrict__ prefix,
                                                  const int* __restrict__ lin_in, int S, int* __restrict__ nbr_lin,
                                                  int* __restrict__ nbr_row, unsigned char* __restrict__ pos, int* __restrict__ inc) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const long long lin = lin_in[s];
    const long long stride[3] = { 1, v.dims[0], (long long)v.dims[0] * v.dims[1] };
    unsigned char pb = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const long long fwd = lin + stride[a];
        const bool valid = fwd < v.N && is_member(mask, fwd);   // idx[a] + 1 <= dim[a] always holds inside the grid
        long long nl = valid ? fwd : lin - stride[a];
        if (nl < 0) nl = -1;
        if (valid) pb |= (unsigned char)(1u << a);
        nbr_lin[a * S + s] = (int)nl;
        nbr_row[a * S + s] = (nl >= 0 && is_member(mask, nl)) ? row_of(mask, prefix, nl) : -1;
        // incident stencils
        const long long tm = lin - stride[a];
        inc[a * S + s] = (tm >= 0 && is_member(mask, tm)) ? row_of(mask, prefix, tm) : -1;
        int ip = -1;
        if (valid) {   // t+ = fwd is a member; it points back at us iff its own forward neighbour is not a member
            const long long f2 = fwd + stride[a];
            const bool fwd_has_fwd = f2 < v.N && is_member(mask, f2);
            if (!fwd_has_fwd) ip = row_of(mask, prefix, fwd);
        }
        inc[(3 + a) * S + s] = ip;
    }
    pos[s] = pb;
}

// Optimizer::subsampling / VolumetricGradSdf::subsample (OptimizerAux.cpp:622-684, VolumetricGradSdf.cpp:469-494):
// one thread per CHILD voxel gathers from its parent.
__global__ void __launch_bounds__(256) k_upsample(VolView src, VolView dst, float vs4) {
    const long long lin = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (lin >= dst.N) return;
    const int W0 = dst.dims[0], W1 = dst.dims[1];
    const unsigned ul = (unsigned)lin;               // N < 2^31: 32-bit divisions
    const int i = (int)(ul % (unsigned)W0);
    const unsigned rest = ul / (unsigned)W0;
    const int j = (int)(rest % (unsigned)W1);
    const int k = (int)(rest / (unsigned)W1);
    const long long pl = (long long)(i >> 1) + (long long)src.dims[0] * (j >> 1) + (long long)src.dims[0] * src.dims[1] * (k >> 1);
    const float pd = src.dist[pl];
    if (pd == src.T) {   // never-fused parents are skipped; children keep the fill value (dist = T, rgb = 0.5)
        dst.dist[lin] = src.T; dst.gx[lin] = 0.f; dst.gy[lin] = 0.f; dst.gz[lin] = 0.f;
        dst.w[lin] = 0.f; dst.r[lin] = 0.5f; dst.g[lin] = 0.5f; dst.b[lin] = 0.5f;
        for (int w = 0; w < dst.nwords; ++w) dst.vis[(long long)w * dst.N + lin] = 0ull;
        return;
    }
    const float g0[3] = { src.gx[pl], src.gy[pl], src.gz[pl] };
    float g[3];
    normalized3(g0, g);
    const float a = (i & 1) ? g[0] : -g[0], b = (j & 1) ? g[1] : -g[1], c = (k & 1) ? g[2] : -g[2];
    dst.dist[lin] = pd + vs4 * ((a + b) + c);
    dst.gx[lin] = g0[0]; dst.gy[lin] = g0[1]; dst.gz[lin] = g0[2];
    dst.w[lin] = src.w[pl]; dst.r[lin] = src.r[pl]; dst.g[lin] = src.g[pl]; dst.b[lin] = src.b[pl];
    for (int w = 0; w < dst.nwords; ++w) dst.vis[(long long)w * dst.N + lin] = src.vis[(long long)w * src.N + pl];
}

// Slot-ordered records gathered from the volume planes (slot t -> surface row perm[t]).
__global__ void __launch_bounds__(256) k_gather_records(VolView v, const int* __restrict__ perm, const int* __restrict__ lin_in,
                                                       const unsigned long long* __restrict__ vis_in, const int* __restrict__ nbr_lin,
                                                       const unsigned char* __restrict__ pos, int S, float vs_inv, SurfView out,
                                                       unsigned long long* __restrict__ pvis) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= out.nslots) return;
    const int s = perm[t];
    const int lin = lin_in[s];
    int idx[3];
    line2idx(lin, v.dims, idx);
    const float d = v.dist[lin];
    const float g[3] = { v.gx[lin], v.gy[lin], v.gz[lin] };
    float gh[3], fd[3], nfd[3];
    normalized3(g, gh);
    const unsigned char pb = pos[s];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int nl = nbr_lin[a * S + s];
        const float p = ((pb >> a) & 1) ? 1.f : -1.f;
        const float dn = nl >= 0 ? v.dist[nl] : d;
        fd[a] = (p * (dn - d)) * vs_inv;      // Optimizer::computeDistGrad (Optimizer.cpp:346-351)
    }
    normalized3(fd, nfd);
    out.rec0[t] = make_float4(d, gh[0], gh[1], gh[2]);
    out.rec1[t] = make_float4(v.r[lin], v.g[lin], v.b[lin], __int_as_float((int)pb));
    out.rec2[t] = make_float4(fd[0], fd[1], fd[2], __int_as_float(s));
    out.rec3[t] = make_float4(v.origin[0] + v.vs * (float)idx[0], v.origin[1] + v.vs * (float)idx[1],
                              v.origin[2] + v.vs * (float)idx[2], __int_as_float(lin));   // VoxelGrid::voxel2world
    out.rec4[t] = make_float4(nfd[0], nfd[1], nfd[2], 0.f);
    pvis[t] = vis_in[s];
}

static SurfView raw_view(psg_optimizer* o) {
    SurfView sv;
    sv.S = o->S; sv.nslots = o->n_slots; sv.lin = o->lin.p; sv.vis = o->svis.p; sv.nbr_lin = o->nbr_lin.p; sv.nbr_row = o->nbr_row.p; sv.pos = o->pos.p;
    sv.pvis = o->pvis.p;
    sv.rec0 = o->rec0.p; sv.rec1 = o->rec1.p; sv.rec2 = o->rec2.p; sv.rec3 = o->rec3.p; sv.rec4 = o->rec4.p;
    return sv;
}

void refresh_records(psg_optimizer* o) {
    if (o->n_slots > 0) {
        { Prof pf(o->stream, KID_SURF_NEIGHBORS);
          k_gather_records<<<(o->n_slots + 255) / 256, 256, 0, o->stream>>>(o->vol->view(), o->perm.p, o->lin.p, o->svis.p, o->nbr_lin.p,
                                                                      o->pos.p, o->S, o->vs_inv, raw_view(o), o->pvis.p); }
        PSG_LAUNCH_CHECK();
    }
    o->vol_gen_seen = o->vol->generation;
}

SurfView surf_view(psg_optimizer* o) {
    if (o->layout_dirty) build_layout(o);
    if (o->vol_gen_seen != o->vol->generation) refresh_records(o);
    return raw_view(o);
}

// ---- slot order: LSD radix sort of the rows by (64 - visible keyframes, visibility word) ---------------------------------------
// The sweeps walk 32 slots per warp, all lanes on the same keyframe, over the union of the 32 visibility masks; lanes whose voxel
// does not see the keyframe idle.  Sorting by the whole mask puts equal masks next to each other (measured: sorting by the
// keyframe count alone, rows otherwise in voxel order, makes the sweeps 1.3-1.5x slower), and the count on top puts the rows
// with most work first (no heavy blocks left for the tail of the launch).  ceil(F / 8) stable counting-sort passes over the mask
// bytes, low to high, and a last one over the count; per pass: per-chunk histogram, one scan block, stable scatter (warp match +
// per-warp offsets).  Rows are sorted as indices; the key is fetched through the index.
constexpr int kSortBuckets = 256;
constexpr int kSortChunk = 2048;     // rows per block

__device__ __forceinline__ int sort_digit(unsigned long long m, int pass, int npass) {
    return pass == npass - 1 ? 64 - __popcll(m) : (int)((m >> (8 * pass)) & 255ull);
}

__global__ void __launch_bounds__(256) k_sort_hist(const unsigned long long* __restrict__ vis, const int* __restrict__ in, int row0, int n,
                                                  int pass, int npass, unsigned* __restrict__ hist /* [nchunks][256] */) {
    __shared__ unsigned s_h[kSortBuckets];
    s_h[threadIdx.x] = 0u;
    __syncthreads();
    const int c0 = blockIdx.x * kSortChunk;
    for (int i = c0 + threadIdx.x; i < c0 + kSortChunk && i < n; i += 256) {
        const int row = in ? in[i] : row0 + i;
        atomicAdd(&s_h[sort_digit(vis[row], pass, npass)], 1u);
    }
    __syncthreads();
    hist[(size_t)blockIdx.x * kSortBuckets + threadIdx.x] = s_h[threadIdx.x];
}
// hist[c][b] -> first output position of chunk c's rows of bucket b
__global__ void __launch_bounds__(256) k_sort_scan(unsigned* __restrict__ hist, int nchunks) {
    __shared__ unsigned s_tot[kSortBuckets];
    const int b = threadIdx.x;
    unsigned run = 0;
    for (int c = 0; c < nchunks; ++c) { const unsigned v = hist[(size_t)c * kSortBuckets + b]; hist[(size_t)c * kSortBuckets + b] = run; run += v; }
    s_tot[b] = run;
    __syncthreads();
    unsigned base = 0;
    for (int k = 0; k < b; ++k) base += s_tot[k];
    for (int c = 0; c < nchunks; ++c) hist[(size_t)c * kSortBuckets + b] += base;
}
__global__ void __launch_bounds__(256) k_sort_scatter(const unsigned long long* __restrict__ vis, const int* __restrict__ in, int row0, int n,
                                                     int pass, int npass, const unsigned* __restrict__ hist, int* __restrict__ out) {
    __shared__ unsigned s_off[kSortBuckets];          // next output position per bucket
    __shared__ unsigned s_wcnt[8][kSortBuckets];      // rows of the current tile per warp and bucket
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    s_off[threadIdx.x] = hist[(size_t)blockIdx.x * kSortBuckets + threadIdx.x];
    const int c0 = blockIdx.x * kSortChunk;
    for (int t0 = c0; t0 < c0 + kSortChunk && t0 < n; t0 += 256) {
#pragma unroll
        for (int w = 0; w < 8; ++w) s_wcnt[w][threadIdx.x] = 0u;
        __syncthreads();
        const int i = t0 + threadIdx.x;
        const bool inr = i < n && i < c0 + kSortChunk;
        const int row = inr ? (in ? in[i] : row0 + i) : -1;
        const int b = inr ? sort_digit(vis[row], pass, npass) : -1;
        const unsigned peers = __match_any_sync(0xffffffffu, b);
        const unsigned rank_in_warp = __popc(peers & ((1u << lane) - 1u));
        if (inr && rank_in_warp == 0) s_wcnt[warp][b] = __popc(peers);
        __syncthreads();
        if (inr) {
            unsigned pos = s_off[b] + rank_in_warp;
            for (int w = 0; w < warp; ++w) pos += s_wcnt[w][b];
            out[pos] = row;
        }
        __syncthreads();
        {
            unsigned add = 0;
#pragma unroll
            for (int w = 0; w < 8; ++w) add += s_wcnt[w][threadIdx.x];
            s_off[threadIdx.x] += add;
        }
        __syncthreads();
    }
}

void build_surface(psg_optimizer* o, bool /*reindex_vis*/) {
    psg_volume* vol = o->vol;
    const VolView v = vol->view();
    SelectParams P;
    P.F = o->F;
    for (int k = 0; k < o->F; ++k) P.frame_idx[k] = o->frame_idx[k];
    P.thr = sqrt(3.0) * (double)o->vs;

    const long long nw = (v.N + 31) / 32;
    // words per block: at least 32 (1024 voxels), grown so that the single-block scan sees <= 16384 counts
    long long wpb = 32;
    while ((nw + wpb - 1) / wpb > 16384) wpb *= 2;
    const int nblocks = (int)((nw + wpb - 1) / wpb);
    o->mask.alloc((size_t)nw);
    o->prefix.alloc((size_t)nw);
    o->block_counts.alloc((size_t)nblocks + 1);

    o->block_pairs.alloc((size_t)nblocks + 8);
    unsigned long long* totals_dev = reinterpret_cast<unsigned long long*>(o->block_pairs.p + (((size_t)nblocks + 1) & ~(size_t)1));   // 8-byte aligned
    { Prof pf(o->stream, KID_SURF_MASK); k_surface_mask<<<nblocks, 256, 0, o->stream>>>(v, P, o->mask.p, o->block_counts.p, o->block_pairs.p, nw, (int)wpb); }
    PSG_LAUNCH_CHECK();
    { Prof pf(o->stream, KID_SURF_SCAN); k_scan_counts<<<1, 1024, 0, o->stream>>>(o->block_counts.p, nblocks, o->block_pairs.p, totals_dev); }
    PSG_LAUNCH_CHECK();
    // the one host round trip of the build: the row count sizes everything below, the pair count decides the sharding policy
    unsigned long long totals[2] = {0ull, 0ull};
    PSG_CUDA(cudaMemcpyAsync(totals, totals_dev, sizeof(totals), cudaMemcpyDeviceToHost, o->stream));
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    o->S = (int)totals[0];
    o->vis_pairs = (long long)totals[1];
    const size_t S = (size_t)(o->S > 0 ? o->S : 1);
    o->lin.alloc(S); o->svis.alloc(S); o->nbr_lin.alloc(3 * S); o->nbr_row.alloc(3 * S); o->pos.alloc(S); o->inc.alloc(6 * S);
    o->rec0.alloc(S); o->rec1.alloc(S + 32 * 65);   // rec1 is all-gathered in equal, 32-aligned chunks (comm.cu)
    o->rec0.alloc(S); o->rec2.alloc(S); o->rec3.alloc(S); o->rec4.alloc(S);
    o->pvis.alloc(S); o->perm.alloc(S);
    if (o->S > 0) {
        { Prof pf(o->stream, KID_SURF_SCATTER);
          k_surface_scatter<<<nblocks, 256, 0, o->stream>>>(v, P, o->mask.p, o->block_counts.p, o->prefix.p, nw, (int)wpb,
                                                            o->lin.p, o->svis.p); }
        PSG_LAUNCH_CHECK();
        { Prof pf(o->stream, KID_SURF_NEIGHBORS);
          k_neighbors<<<(o->S + 255) / 256, 256, 0, o->stream>>>(v, o->mask.p, o->prefix.p, o->lin.p, o->S, o->nbr_lin.p,
                                                                 o->nbr_row.p, o->pos.p, o->inc.p); }
        PSG_LAUNCH_CHECK();
    }
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    update_shard_policy(o);
    o->layout_dirty = true;
    build_layout(o);
}

// Everything that depends on how the rows are divided among the ranks: the row range this rank owns (all rows unless the
// work list is sharded), its slot list = its rows in sweep order (radix sort by keyframe visibility), the slot records,
// the buffers an outer iteration uses, and -- sharded -- the peers' views of each other's buffers.  Collective when sharded.
void build_layout(psg_optimizer* o) {
    o->layout_dirty = false;
    release_peers(o);                 // nothing below may reallocate a buffer a peer still maps
    partition_rows(o);                // row_lo / row_hi / n_slots, halo ranges (comm.cu); whole list when not sharded
    if (o->n_slots > 0) {
        // slot order: radix sort by (64 - visible keyframes, visibility word), see above
        const int n = o->n_slots;
        const int nchunks = (n + kSortChunk - 1) / kSortChunk;
        const int npass = (o->F + 7) / 8 + 1;
        o->sort_hist.alloc((size_t)nchunks * kSortBuckets);
        o->perm2.alloc((size_t)(o->S > 0 ? o->S : 1));
        // ping-pong so that the last pass lands in o->perm
        int* bufs[2] = { (npass & 1) ? o->perm.p : o->perm2.p, (npass & 1) ? o->perm2.p : o->perm.p };
        const int* in = nullptr;
        for (int pass = 0; pass < npass; ++pass) {
            int* out = bufs[pass & 1];
            k_sort_hist<<<nchunks, 256, 0, o->stream>>>(o->svis.p, in, o->row_lo, n, pass, npass, o->sort_hist.p);
            PSG_LAUNCH_CHECK();
            k_sort_scan<<<1, 256, 0, o->stream>>>(o->sort_hist.p, nchunks);
            PSG_LAUNCH_CHECK();
            k_sort_scatter<<<nchunks, 256, 0, o->stream>>>(o->svis.p, in, o->row_lo, n, pass, npass, o->sort_hist.p, out);
            PSG_LAUNCH_CHECK();
            in = out;
        }
        count_launch(3 * npass);
    }
    refresh_records(o);
    // everything the blocks of an outer iteration use is sized here, once per work list, so that the iteration path itself
    // never allocates (DevBuf::alloc is a no-op once sized; growing it frees and reallocates, which synchronises the device)
    {
        const size_t Sr = (size_t)(o->S > 0 ? o->S : 1);
        o->red.alloc((size_t)kRedDoubles * kRedCopies + kMailDoubles);
        o->vecs.alloc(Sr * 25 + 8);
        o->pcg_pbuf.alloc(2 * Sr);
        o->pcg_partials.alloc((size_t)4 * kPcgMaxBlocks + 1);
        // dataflow distance solve (pcg.cu): packets of u, U1..U3 per row, tags start at zero (= never written)
        if (o->xplanes.n < 4 * Sr) o->xplanes.alloc(4 * Sr);
        PSG_CUDA(cudaMemsetAsync(o->xplanes.p, 0, o->xplanes.n * sizeof(unsigned long long), o->stream));
        if (!o->xstate.p) {
            o->xstate.alloc(kArInboxOffset + kArInboxBytes);   // the XState, then the all-reduce inbox (comm.cu)
            PSG_CUDA(cudaMemsetAsync(o->xstate.p, 0, kArInboxOffset + kArInboxBytes, o->stream));
        }
        const size_t grid = (Sr + 31) / 32 + 1, ngroups = grid / 128 + 2;
        o->sweep_partials.alloc((grid + ngroups) * 8);
        if (o->sweep_counter.n < ngroups + 1) {
            o->sweep_counter.alloc(ngroups + 1);
            PSG_CUDA(cudaMemsetAsync(o->sweep_counter.p, 0, (ngroups + 1) * sizeof(unsigned), o->stream));
        }
    }
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    exchange_peers(o);                // sharded: map the peers' buffers (CUDA IPC), or take their pointers inside one process
}

void launch_upsample(psg_optimizer* o) {
    psg_volume* vol = o->vol;
    const VolView src = vol->view();
    psg_volume nv_store;   // temporary owner of the new planes
    DevBuf<float> nf[8];
    DevBuf<unsigned long long> nvis;
    const long long N8 = 8 * vol->N;
    for (int a = 0; a < 8; ++a) nf[a].alloc((size_t)N8);
    nvis.alloc((size_t)N8 * vol->nwords);
    VolView dst = src;
    for (int a = 0; a < 3; ++a) dst.dims[a] = 2 * src.dims[a];
    dst.N = N8;
    dst.dist = nf[0].p; dst.gx = nf[1].p; dst.gy = nf[2].p; dst.gz = nf[3].p;
    dst.w = nf[4].p; dst.r = nf[5].p; dst.g = nf[6].p; dst.b = nf[7].p; dst.vis = nvis.p;
    const float vs4 = (float)(0.25 * (double)vol->vs);
    { Prof pf(o->stream, KID_UPSAMPLE); k_upsample<<<(unsigned)((N8 + 255) / 256), 256, 0, o->stream>>>(src, dst, vs4); }
    PSG_LAUNCH_CHECK();
    PSG_CUDA(cudaStreamSynchronize(o->stream));
    for (int a = 0; a < 8; ++a) { vol->f[a].release(); vol->f[a] = nf[a]; nf[a].p = nullptr; }
    vol->vis.release(); vol->vis = nvis; nvis.p = nullptr;
    // VoxelGrid::grid_subsample (VoxelGrid.h:143-149)
    vol->vs = (float)((double)vol->vs * 0.5);
    for (int a = 0; a < 3; ++a) vol->dims[a] *= 2;
    vol->N = N8;
    vol->generation++;
    const float half_vs = (float)(0.5 * (double)vol->vs);
    for (int a = 0; a < 3; ++a) vol->origin[a] = (vol->shift[a] - half_vs * (float)vol->dims[a]) - half_vs * 1.f;
    // Optimizer: voxel_size_ *= 0.5; voxel_size_inv_ = 1.0 / voxel_size_ (OptimizerAux.cpp:679-680)
    o->vs = (float)((double)o->vs * 0.5);
    o->vs_inv = (float)(1.0 / (double)o->vs);
    (void)nv_store;
}

// the eight floats of every surface voxel (SdfVoxel layout: dist, grad, weight, r, g, b) in surface-list order, so that the
// writers (save_pointcloud, extract_pc) need S records from the device instead of the whole volume
__global__ void __launch_bounds__(256) k_gather_voxels(VolView vol, const int* __restrict__ lin, int S, float* __restrict__ out) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const int l = lin[s];
    reinterpret_cast<float4*>(out)[2 * (long long)s] = make_float4(vol.dist[l], vol.gx[l], vol.gy[l], vol.gz[l]);
    reinterpret_cast<float4*>(out)[2 * (long long)s + 1] = make_float4(vol.w[l], vol.r[l], vol.g[l], vol.b[l]);
}

void launch_gather_voxels(psg_optimizer* o, float* out_dev) {
    if (o->S == 0) return;
    { Prof pf(o->stream, KID_AOS); k_gather_voxels<<<(o->S + 255) / 256, 256, 0, o->stream>>>(o->vol->view(), o->lin.p, o->S, out_dev); }
    PSG_LAUNCH_CHECK();
}

}  // namespace psg
