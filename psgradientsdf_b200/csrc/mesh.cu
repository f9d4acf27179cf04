// Surface extraction on the device: crop box, marching cubes, the cropped -dist box of the .sdf dump and the
// volume point cloud; the ASCII writers stay on the host and produce the reference's files byte for byte.
//
// Replaces Optimizer::extract_mesh / save_white_mesh / saveSDF (cpp/include/ps_optimizer/OptimizerAux.cpp:278-453, 513-577),
// VolumetricGradSdf::extract_mesh / extract_pc / saveSDF (cpp/include/sdf_tracker/VolumetricGradSdf.cpp:234-376, 379-460) and
// MarchingCubes::computeIsoSurface / savePly (cpp/third/mesh/MarchingCubes.cpp:314-699).
//
// The reference crops the volume to the box of voxels within sqrt(3) * vs of the surface, copies -dist / weight / int(255 * rgb)
// into dense host arrays and walks the (dim - 2)^3 cells of that box serially, appending three new vertices per triangle.  Here
// nothing is copied: a cell reads its 8 corners from the volume planes in place, and the serial append order (cells in x, y, z
// order, triangles in table order, degenerate ones dropped) is recovered with a count pass, an exclusive scan over per-block
// counts and an emit pass that writes every triangle at its final position.  Only the triangles leave the device (45 B each)
// instead of the volume (32 B per voxel: 4 GiB at 512^3).
//
// Bit-exactness: vertex = v0 + mu * (v1 - v0) with mu a double holding a FLOAT quotient, the sum in double, rounded once
// (MarchingCubes.cpp:589-609); colours go through float / 255.0f, the same interpolation, * 255.0f and a truncating cast
// (:622-640), green and blue read one and two entries further along the CROPPED array than red (green_[idx + 1], blue_[idx + 2]),
// which wraps to the next row of the box at its +x face.  -fmad=false keeps every product and sum separately rounded.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "mc_table.h"
#include "psg_host.h"

namespace psg {

__device__ const unsigned long long d_tri[256] = PSG_MC_TABLE_INIT;

constexpr int kCellsPerBlock = 1024;   // four rounds of 256 threads; one count per block keeps the scan short at 1024^3

struct MeshView {
    const float* dist; const float* w; const float* r; const float* g; const float* b;
    int W, WH;                 // volume strides (every voxel index fits an int: psg_volume_create keeps N below 2^31)
    int lin0;                  // voxel index of the crop box's lower corner
    int lo[3];                 // lower corner of the crop box in the volume
    int bd[3];                 // crop box dimensions
    int cd[3];                 // cells per axis: bd - 2 (the reference's loops stop at dim - 2, MarchingCubes.cpp:333-337)
    long long ncells;
    float vsz[3], org[3];      // MarchingCubes::voxelSize_ = (vs * dim) / dim, origin_ = -vs * lo
    int white;                 // save_white_mesh: all colours 255
};

// ---- crop box ------------------------------------------------------------------------------------------------------------------
// One warp per x-row of the volume; box[0..2] = lower corner (atomicMin), box[3..5] = upper corner (atomicMax).  thr is the
// largest float <= sqrt(3) * vs: for a float x, x > thr decides exactly like the reference's double comparison x > sqrt(3) * vs.
__global__ void __launch_bounds__(256) k_mesh_bbox(const float* __restrict__ dist, int W, int H, int D, float thr, int* __restrict__ box) {
    const int lane = threadIdx.x & 31;
    const long long rows = (long long)H * D;
    int ilo = INT_MAX, ihi = INT_MIN, jlo = INT_MAX, jhi = INT_MIN, klo = INT_MAX, khi = INT_MIN;
    for (long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); row < rows; row += (long long)gridDim.x * 8) {
        const int j = (int)(row % H), k = (int)(row / H);
        const float* p = dist + row * W;
        bool any = false;
        if ((W & 3) == 0) {                      // rows are 16-byte aligned: four voxels per load
            const float4* p4 = reinterpret_cast<const float4*>(p);
#pragma unroll 4
            for (int q = lane; q < (W >> 2); q += 32) {
                const float4 d = p4[q];
                const float dv[4] = { d.x, d.y, d.z, d.w };
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (!(fabsf(dv[c]) > thr)) { any = true; ilo = min(ilo, 4 * q + c); ihi = max(ihi, 4 * q + c); }
            }
        } else {
            for (int i = lane; i < W; i += 32) {
                if (!(fabsf(p[i]) > thr)) { any = true; ilo = min(ilo, i); ihi = max(ihi, i); }
            }
        }
        if (__any_sync(0xffffffffu, any)) { jlo = min(jlo, j); jhi = max(jhi, j); klo = min(klo, k); khi = max(khi, k); }
    }
    ilo = __reduce_min_sync(0xffffffffu, ilo); ihi = __reduce_max_sync(0xffffffffu, ihi);
    if (lane == 0 && ihi >= ilo) {
        // the box only grows: an atomic is needed only where this warp would still move a face (a stale read errs on the safe side)
        const volatile int* cur = box;
        if (ilo < cur[0]) atomicMin(box + 0, ilo);
        if (jlo < cur[1]) atomicMin(box + 1, jlo);
        if (klo < cur[2]) atomicMin(box + 2, klo);
        if (ihi > cur[3]) atomicMax(box + 3, ihi);
        if (jhi > cur[4]) atomicMax(box + 4, jhi);
        if (khi > cur[5]) atomicMax(box + 5, khi);
    }
}

__global__ void k_mesh_box_init(int* box) {
    if (threadIdx.x < 3) box[threadIdx.x] = INT_MAX;
    else if (threadIdx.x < 6) box[threadIdx.x] = INT_MIN;
}

// ---- marching cubes ------------------------------------------------------------------------------------------------------------
// corner b of a cell at (x, y, z): bit 0 (x+1, y+1, z), 1 (x+1, y, z), 2 (x, y, z), 3 (x, y+1, z), 4..7 the same at z+1
// (MarchingCubes::computeLutIndex, :549-585); edge e joins corners kEdgeA[e] -> kEdgeB[e] in the order getVertex receives them
// (:342-506); getColor receives edges 2, 3, 6, 7 the other way round.
__device__ __forceinline__ int corner_dx(int b) { return (0x33 >> b) & 1; }
__device__ __forceinline__ int corner_dy(int b) { return (0x99 >> b) & 1; }
__device__ __forceinline__ int corner_dz(int b) { return b >> 2; }
__device__ __forceinline__ int edge_first(int e) { return e < 8 ? e : e - 8; }
__device__ __forceinline__ int edge_second(int e) { return e < 4 ? ((e + 1) & 3) : e < 8 ? 4 + ((e + 1) & 3) : e - 4; }
constexpr unsigned kColorSwapped = 0x0CC;   // edges 2, 3, 6, 7 (MarchingCubes.cpp:374,386,422,434)

// MarchingCubes::interpolate (:589-609) with isoValue = 0.0f (every caller's default)
__device__ __forceinline__ void mc_interp(float t0, float t1, const float v0[3], const float v1[3], float out[3]) {
    const float iso = 0.0f;
    if ((double)fabsf(iso - t0) < 1e-7) { out[0] = v0[0]; out[1] = v0[1]; out[2] = v0[2]; return; }
    if ((double)fabsf(iso - t1) < 1e-7) { out[0] = v1[0]; out[1] = v1[1]; out[2] = v1[2]; return; }
    if ((double)fabsf(t0 - t1) < 1e-7) { out[0] = v0[0]; out[1] = v0[1]; out[2] = v0[2]; return; }
    double mu = (double)__fdiv_rn(iso - t0, t1 - t0);
    if (mu > 1.0) mu = 1.0; else if (mu < 0) mu = 0.0;
#pragma unroll
    for (int a = 0; a < 3; ++a) out[a] = (float)__dadd_rn((double)v0[a], __dmul_rn(mu, (double)__fsub_rn(v1[a], v0[a])));
}

// the 8-bit colour the crop copy would hold: int(255 * c) stored to an unsigned char (OptimizerAux.cpp:335-337)
__device__ __forceinline__ float crop_color(const float* __restrict__ plane, int lin, int white) {
    return white ? 255.0f : (float)(unsigned char)(int)__fmul_rn(255.0f, plane[lin]);
}

// volume index of entry (idx + k) of the cropped array, idx the crop index of box voxel (x, y, z), k = 0..2
__device__ __forceinline__ int crop_shift(const MeshView& m, int x, int y, int z, int k) {
    x += k;
    if (x >= m.bd[0]) { x -= m.bd[0]; ++y; }   // corners have y <= bd[1] - 2: no second wrap
    return m.lin0 + x + m.W * y + m.WH * z;
}

// offset of corner b from a cell's (x, y, z) voxel
__device__ __forceinline__ int corner_off(const MeshView& m, int b) { return corner_dx(b) + m.W * corner_dy(b) + m.WH * corner_dz(b); }

// MarchingCubes::computeLutIndex (:549-585): 0 when any corner was never fused (weight 0), else bit b = tsdf > 0 at corner b.
// All sixteen loads are issued before the first test (the reference's short-circuit has no side effect to preserve).
__device__ __forceinline__ int cell_case(const MeshView& m, int base) {
    const float* __restrict__ wp = m.w + base;
    const float* __restrict__ dp = m.dist + base;
    float w[8], d[8];
#pragma unroll
    for (int b = 0; b < 8; ++b) { const int o = corner_off(m, b); w[b] = wp[o]; d[b] = dp[o]; }
    int ci = 0;
    bool seen = true;
#pragma unroll
    for (int b = 0; b < 8; ++b) { seen = seen && !(w[b] == 0.0f); if (d[b] < 0.0f) ci |= 1 << b; }   // -d > 0.0f
    return seen ? ci : 0;
}

constexpr int kChunk = 64;   // active cells whose edges sit in shared memory at a time

// cell o (< kCellsPerBlock) after the block's first cell (x0, y0, z0), in x, y, z order; 32-bit arithmetic only
__device__ __forceinline__ void cell_xyz(const MeshView& m, int x0, int y0, int z0, unsigned o, int& x, int& y, int& z) {
    const unsigned xo = (unsigned)x0 + o, cy = xo / (unsigned)m.cd[0];
    x = (int)(xo - cy * (unsigned)m.cd[0]);
    const unsigned yo = (unsigned)y0 + cy, cz = yo / (unsigned)m.cd[1];
    y = (int)(yo - cz * (unsigned)m.cd[1]);
    z = z0 + (int)cz;
}
// the cell 256 further on
__device__ __forceinline__ void cell_advance(const MeshView& m, int& x, int& y, int& z) {
    x += 256;
    if (x < m.cd[0]) return;
    if (m.cd[0] >= 128) { do { x -= m.cd[0]; ++y; } while (x >= m.cd[0]); }
    else { const unsigned q = (unsigned)x / (unsigned)m.cd[0]; x -= (int)q * m.cd[0]; y += (int)q; }
    if (y >= m.cd[1]) { const unsigned q = (unsigned)y / (unsigned)m.cd[1]; y -= (int)q * m.cd[1]; z += (int)q; }
}

// Count pass (EMIT = false): triangles per block of kCellsPerBlock cells -> counts[blockIdx.x].
// Emit pass (EMIT = true): counts[] holds the exclusive scan; every triangle is written at its serial position.
//
// A block first classifies its 1024 cells (one cell per thread, four rounds) and compacts the ones the surface crosses, in cell
// order, into shared memory -- a few dozen out of 1024.  Those are then worked on by the whole block: one thread per (cell, edge)
// interpolates the vertex (and its colour when emitting) into shared memory, one thread per cell drops the degenerate triangles
// (computeTriangles, :613-641) and takes part in the scan that gives the cell its output position, one thread per
// (cell, triangle, vertex) writes.  The serial reference order -- cells in x, y, z order, triangles in table order -- is the order
// of the compaction and of the scans.
template <bool EMIT>
__global__ void __launch_bounds__(256) k_mesh_cells(MeshView m, unsigned* __restrict__ counts, unsigned* __restrict__ crossed,
                                                    float* __restrict__ vertices, unsigned char* __restrict__ colors) {
    __shared__ unsigned long long s_list[kCellsPerBlock]; // x | y << 16 | z << 32 | case << 48 of the crossed cells, in cell order
    __shared__ unsigned s_cnt[4 * 8 + 1];                 // active cells per (round, warp), then their exclusive scan
    __shared__ float s_ep[kChunk][12][3];
    __shared__ unsigned char s_ec[kChunk][12][3];
    __shared__ unsigned s_keep[kChunk], s_off[kChunk], s_wsum[2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long c0 = (long long)blockIdx.x * kCellsPerBlock;
    unsigned running = 0;
    if (EMIT) {
        running = counts[blockIdx.x];
        if (counts[blockIdx.x + 1] == running) return;      // nothing in this block (the common case away from the surface)
    }
    const long long cxy = (long long)m.cd[0] * m.cd[1];
    const int z0 = (int)(c0 / cxy);
    const int y0 = (int)((c0 - (long long)z0 * cxy) / m.cd[0]), x0 = (int)(c0 - (long long)z0 * cxy - (long long)y0 * m.cd[0]);
    // ---- classify and compact
    int my_case[4]; unsigned my_rank[4];
    int cx, cy, cz;
    cell_xyz(m, x0, y0, z0, threadIdx.x, cx, cy, cz);
#pragma unroll
    for (int round = 0; round < 4; ++round) {
        int ci = 0;
        // the count pass leaves one bit per cell ("the surface crosses it"); the emit pass classifies those cells only
        const size_t word = (size_t)blockIdx.x * 32 + round * 8 + warp;
        const bool look = EMIT ? ((crossed[word] >> lane) & 1u) != 0u : c0 + round * 256 + threadIdx.x < m.ncells;
        if (look) {
            ci = cell_case(m, m.lin0 + cx + m.W * cy + m.WH * cz);
            if (ci == 255) ci = 0;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, ci != 0);
        if (!EMIT && lane == 0) crossed[word] = bal;
        my_case[round] = ci;
        my_rank[round] = __popc(bal & ((1u << lane) - 1u));
        if (lane == 0) s_cnt[round * 8 + warp] = __popc(bal);
        if (round < 3) cell_advance(m, cx, cy, cz);
    }
    __syncthreads();
    if (warp == 0) {
        const unsigned x = s_cnt[lane];
        unsigned inc = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
        s_cnt[lane] = inc - x;
        if (lane == 31) s_cnt[32] = inc;
    }
    __syncthreads();
#pragma unroll
    for (int round = 0; round < 4; ++round)
        if (my_case[round]) {
            int x, y, z;
            cell_xyz(m, x0, y0, z0, (unsigned)(round * 256 + threadIdx.x), x, y, z);
            s_list[s_cnt[round * 8 + warp] + my_rank[round]] = (unsigned long long)x | ((unsigned long long)y << 16) | ((unsigned long long)z << 32) |
                                                                ((unsigned long long)my_case[round] << 48);
        }
    const int n_active = (int)s_cnt[32];
    __syncthreads();
    // ---- the active cells, kChunk at a time
    unsigned block_total = 0;
    for (int j0 = 0; j0 < n_active; j0 += kChunk) {
        const int nj = min(kChunk, n_active - j0);
        // one thread per (cell, edge)
        for (int q = threadIdx.x; q < nj * 12; q += 256) {
            const int j = q / 12, e = q - 12 * j;
            const unsigned long long ent = s_list[j0 + j];
            const int ci = (int)(ent >> 48);
            const int a = edge_first(e), b = edge_second(e);
            if (!(((ci >> a) ^ (ci >> b)) & 1)) continue;           // the surface does not cross this edge
            const int x = (int)(ent & 0xFFFF), y = (int)((ent >> 16) & 0xFFFF), z = (int)((ent >> 32) & 0xFFFF);
            const int base = m.lin0 + x + m.W * y + m.WH * z;
            const float ta = -m.dist[base + corner_off(m, a)], tb = -m.dist[base + corner_off(m, b)];
            const int ax = x + corner_dx(a), ay = y + corner_dy(a), az = z + corner_dz(a);
            const int bx = x + corner_dx(b), by = y + corner_dy(b), bz = z + corner_dz(b);
            // MarchingCubes::voxelToWorld (:646-650)
            const float pa[3] = { __fsub_rn(__fmul_rn((float)ax, m.vsz[0]), m.org[0]), __fsub_rn(__fmul_rn((float)ay, m.vsz[1]), m.org[1]),
                                  __fsub_rn(__fmul_rn((float)az, m.vsz[2]), m.org[2]) };
            const float pb[3] = { __fsub_rn(__fmul_rn((float)bx, m.vsz[0]), m.org[0]), __fsub_rn(__fmul_rn((float)by, m.vsz[1]), m.org[1]),
                                  __fsub_rn(__fmul_rn((float)bz, m.vsz[2]), m.org[2]) };
            float p[3];
            mc_interp(ta, tb, pa, pb, p);
            s_ep[j][e][0] = p[0]; s_ep[j][e][1] = p[1]; s_ep[j][e][2] = p[2];
            if (EMIT) {
                // MarchingCubes::getColor (:622-640)
                const bool sw = (kColorSwapped >> e) & 1;
                const int x1 = sw ? bx : ax, y1 = sw ? by : ay, z1 = sw ? bz : az;
                const int x2 = sw ? ax : bx, y2 = sw ? ay : by, z2 = sw ? az : bz;
                const float k1[3] = { __fdiv_rn(crop_color(m.r, crop_shift(m, x1, y1, z1, 0), m.white), 255.0f),
                                      __fdiv_rn(crop_color(m.g, crop_shift(m, x1, y1, z1, 1), m.white), 255.0f),
                                      __fdiv_rn(crop_color(m.b, crop_shift(m, x1, y1, z1, 2), m.white), 255.0f) };
                const float k2[3] = { __fdiv_rn(crop_color(m.r, crop_shift(m, x2, y2, z2, 0), m.white), 255.0f),
                                      __fdiv_rn(crop_color(m.g, crop_shift(m, x2, y2, z2, 1), m.white), 255.0f),
                                      __fdiv_rn(crop_color(m.b, crop_shift(m, x2, y2, z2, 2), m.white), 255.0f) };
                float cv[3];
                mc_interp(sw ? tb : ta, sw ? ta : tb, k1, k2, cv);
#pragma unroll
                for (int a3 = 0; a3 < 3; ++a3) s_ec[j][e][a3] = (unsigned char)(int)__fmul_rn(cv[a3], 255.0f);
            }
        }
        __syncthreads();
        // one thread per cell: which triangles survive (p1 != p2 && p1 != p3 && p2 != p3), and the cell's place in the output
        unsigned keep = 0, n = 0;
        if (threadIdx.x < nj) {
            const unsigned long long tri = d_tri[s_list[j0 + threadIdx.x] >> 48];
            int k = 0;
            for (unsigned long long q = tri; (q & 0xF) != 0xF; q >>= 12, ++k) {
                const float *p1 = s_ep[threadIdx.x][q & 0xF], *p2 = s_ep[threadIdx.x][(q >> 4) & 0xF], *p3 = s_ep[threadIdx.x][(q >> 8) & 0xF];
                const bool ne12 = p1[0] != p2[0] || p1[1] != p2[1] || p1[2] != p2[2];
                const bool ne13 = p1[0] != p3[0] || p1[1] != p3[1] || p1[2] != p3[2];
                const bool ne23 = p2[0] != p3[0] || p2[1] != p3[1] || p2[2] != p3[2];
                if (ne12 && ne13 && ne23) { keep |= 1u << k; ++n; }
            }
        }
        if (threadIdx.x < kChunk) {                                   // warps 0 and 1
            unsigned inc = n;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
            if (lane == 31) s_wsum[warp] = inc;
            s_keep[threadIdx.x] = keep;
            s_off[threadIdx.x] = inc - n;                             // within the warp; the second warp adds s_wsum[0] below
        }
        __syncthreads();
        const unsigned chunk_total = s_wsum[0] + s_wsum[1];
        if (EMIT) {
            // one thread per (cell, triangle slot, vertex)
            for (int q = threadIdx.x; q < nj * 15; q += 256) {
                const int j = q / 15, r = q - 15 * j, k = r / 3, vtx = r - 3 * k;
                const unsigned kp = s_keep[j];
                if (!(kp >> k & 1)) continue;
                const unsigned ord = running + block_total + s_off[j] + (j >= 32 ? s_wsum[0] : 0u) + __popc(kp & ((1u << k) - 1u));
                const int e = (int)((d_tri[s_list[j0 + j] >> 48] >> (4 * r)) & 0xF);
                float* vp = vertices + (size_t)ord * 9 + 3 * vtx;
                unsigned char* cp = colors + (size_t)ord * 9 + 3 * vtx;
                vp[0] = s_ep[j][e][0]; vp[1] = s_ep[j][e][1]; vp[2] = s_ep[j][e][2];
                cp[0] = s_ec[j][e][0]; cp[1] = s_ec[j][e][1]; cp[2] = s_ec[j][e][2];
            }
        }
        block_total += chunk_total;
        __syncthreads();
    }
    if (!EMIT && threadIdx.x == 0) counts[blockIdx.x] = block_total;
}

// Exclusive scan of n per-block counts in place (single block, chunked with a carry); the total lands in counts[n].
__global__ void __launch_bounds__(1024) k_mesh_scan(unsigned* __restrict__ counts, long long n) {
    __shared__ unsigned s_warp[32];
    __shared__ unsigned s_carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    for (long long base = 0; base < n; base += 4096) {
        // four consecutive counts per thread
        const long long i0 = base + 4ll * threadIdx.x;
        unsigned x[4], sum = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) { x[q] = (i0 + q < n) ? counts[i0 + q] : 0u; sum += x[q]; }
        unsigned inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            unsigned ws = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, ws, o); if (lane >= o) ws += y; }
            s_warp[lane] = ws;
        }
        __syncthreads();
        const unsigned carry = s_carry;
        unsigned run = carry + (warp ? s_warp[warp - 1] : 0u) + inc - sum;
#pragma unroll
        for (int q = 0; q < 4; ++q) { if (i0 + q < n) counts[i0 + q] = run; run += x[q]; }
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = run;
        __syncthreads();
    }
    if (threadIdx.x == 0) counts[n] = s_carry;
}

// ---- .sdf dump: -dist of the crop box, x fastest (OptimizerAux.cpp:569-573) ----------------------------------------------------
__global__ void __launch_bounds__(256) k_crop_neg_dist(MeshView m, float* __restrict__ out, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long bxy = (long long)m.bd[0] * m.bd[1];
    const int z = (int)(i / bxy);
    const long long rem = i - (long long)z * bxy;
    const int y = (int)(rem / m.bd[0]), x = (int)(rem - (long long)y * m.bd[0]);
    out[i] = -m.dist[m.lin0 + x + m.W * y + m.WH * z];
}

// ---- volume point cloud (VolumetricGradSdf::extract_pc, VolumetricGradSdf.cpp:320-346) -----------------------------------------
// Voxels with |dist| <= sqrt(3.0) * vs seen by at least one frame, in lin order: {p.xyz, g_hat.xyz, r, g, b}.
template <bool EMIT>
__global__ void __launch_bounds__(256) k_volume_points(VolView v, double thr, unsigned* __restrict__ counts, float* __restrict__ out) {
    __shared__ unsigned s_warp[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long long l0 = (long long)blockIdx.x * kCellsPerBlock;
    unsigned running = 0, mine = 0;
    if (EMIT) {
        running = counts[blockIdx.x];
        if (counts[blockIdx.x + 1] == running) return;
    }
    for (int round = 0; round < kCellsPerBlock / 256; ++round) {
        const long long l = l0 + round * 256 + threadIdx.x;
        bool on = false;
        if (l < v.N && !(fabs((double)v.dist[l]) > thr)) {
            for (int w = 0; w < v.nwords && !on; ++w) on = v.vis[(long long)w * v.N + l] != 0ull;
        }
        if (!EMIT) { mine += on; continue; }
        const unsigned bal = __ballot_sync(0xffffffffu, on);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        unsigned warp_off = 0, total = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) { if (i < warp) warp_off += s_warp[i]; total += s_warp[i]; }
        if (on) {
            const unsigned row = running + warp_off + __popc(bal & ((1u << lane) - 1u));
            const long long WH = (long long)v.dims[0] * v.dims[1];
            const int k = (int)(l / WH);
            const long long rem = l - k * WH;
            const int j = (int)(rem / v.dims[0]), i = (int)(rem - (long long)j * v.dims[0]);
            const float d = v.dist[l];
            float g[3] = { v.gx[l], v.gy[l], v.gz[l] };
            const float z = __fadd_rn(__fmul_rn(g[0], g[0]), __fadd_rn(__fmul_rn(g[1], g[1]), __fmul_rn(g[2], g[2])));
            if (z > 0.f) { const float s = __fsqrt_rn(z); g[0] = __fdiv_rn(g[0], s); g[1] = __fdiv_rn(g[1], s); g[2] = __fdiv_rn(g[2], s); }
            const float idx[3] = { (float)i, (float)j, (float)k };
            float* o = out + (size_t)row * 9;
#pragma unroll
            for (int a = 0; a < 3; ++a) { o[a] = __fsub_rn(__fmul_rn(v.vs, idx[a]), __fmul_rn(d, g[a])); o[3 + a] = g[a]; }
            o[6] = v.r[l]; o[7] = v.g[l]; o[8] = v.b[l];
        }
        running += total;
        __syncthreads();
    }
    if (!EMIT) {
        mine = __reduce_add_sync(0xffffffffu, mine);
        if (lane == 0) s_warp[warp] = mine;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned tsum = 0;
            for (int i = 0; i < 8; ++i) tsum += s_warp[i];
            counts[blockIdx.x] = tsum;
        }
    }
}

}  // namespace psg

using namespace psg;

namespace {

struct Box { int lo[3], hi[3]; bool any; };

// the crop box of extract_mesh / saveSDF: voxels with |dist| <= sqrt(3) * vs, compared in double (std::sqrt(3) is the double
// overload, OptimizerAux.cpp:298)
Box crop_box(psg_volume* v, float vs) {
    PSG_CUDA(cudaSetDevice(v->device));
    v->mesh_box.alloc(8);
    const double thr64 = std::sqrt(3.0) * (double)vs;
    float thr = (float)thr64;
    if ((double)thr > thr64) thr = std::nextafterf(thr, -INFINITY);   // round down: see k_mesh_bbox
    const long long rows = (long long)v->dims[1] * v->dims[2];
    const unsigned nb = (unsigned)std::min<long long>((rows + 7) / 8, 148 * 8);
    { Prof pf(v->stream, KID_MESH_MISC); k_mesh_box_init<<<1, 32, 0, v->stream>>>(v->mesh_box.p); }
    { Prof pf(v->stream, KID_MESH_BOX); k_mesh_bbox<<<nb, 256, 0, v->stream>>>(v->f[0].p, v->dims[0], v->dims[1], v->dims[2], thr, v->mesh_box.p); }
    PSG_LAUNCH_CHECK();
    int h[6];
    PSG_CUDA(cudaMemcpyAsync(h, v->mesh_box.p, sizeof h, cudaMemcpyDeviceToHost, v->stream));
    PSG_CUDA(cudaStreamSynchronize(v->stream));
    Box b;
    for (int a = 0; a < 3; ++a) { b.lo[a] = h[a]; b.hi[a] = h[3 + a]; }
    b.any = h[3] >= h[0];
    return b;
}

MeshView mesh_view(psg_volume* v, float vs, const Box& b, int white) {
    MeshView m{};
    m.dist = v->f[0].p; m.w = v->f[4].p; m.r = v->f[5].p; m.g = v->f[6].p; m.b = v->f[7].p;
    m.W = v->dims[0]; m.WH = v->dims[0] * v->dims[1];
    m.lin0 = b.lo[0] + m.W * b.lo[1] + m.WH * b.lo[2];
    m.ncells = 1;
    for (int a = 0; a < 3; ++a) {
        m.lo[a] = b.lo[a];
        m.bd[a] = b.hi[a] - b.lo[a] + 1;
        m.cd[a] = m.bd[a] - 2 > 0 ? m.bd[a] - 2 : 0;
        m.ncells *= m.cd[a];
        const float size = vs * (float)m.bd[a];        // MarchingCubes(dim, voxel_size_ * dim.cast<float>(), ...)
        m.vsz[a] = size / (float)m.bd[a];              // voxelSize_ = size_.cwiseQuotient(dim_.cast<float>())
        m.org[a] = -vs * (float)b.lo[a];               // origin = -voxel_size_ * Vec3f(xmin, ymin, zmin)
    }
    m.white = white;
    return m;
}

// Marching cubes over the crop box; the triangles stay on the device (v->mesh_vertices / mesh_colors).  -1: empty crop box.
long long run_marching_cubes(psg_volume* v, float vs, int white, Box* box_out) {
    const Box b = crop_box(v, vs);
    if (box_out) *box_out = b;
    v->mesh_triangles = -1;
    if (!b.any) return -1;
    const MeshView m = mesh_view(v, vs, b, white);
    v->mesh_triangles = 0;
    if (m.ncells == 0) return 0;
    const long long nblocks = (m.ncells + kCellsPerBlock - 1) / kCellsPerBlock;
    v->mesh_counts.alloc((size_t)nblocks + 1);
    v->mesh_crossed.alloc((size_t)nblocks * 32);      // one bit per cell
    { Prof pf(v->stream, KID_MESH_COUNT); k_mesh_cells<false><<<(unsigned)nblocks, 256, 0, v->stream>>>(m, v->mesh_counts.p, v->mesh_crossed.p, nullptr, nullptr); }
    { Prof pf(v->stream, KID_MESH_SCAN); k_mesh_scan<<<1, 1024, 0, v->stream>>>(v->mesh_counts.p, nblocks); }
    PSG_LAUNCH_CHECK();
    unsigned total = 0;
    PSG_CUDA(cudaMemcpyAsync(&total, v->mesh_counts.p + nblocks, sizeof total, cudaMemcpyDeviceToHost, v->stream));
    PSG_CUDA(cudaStreamSynchronize(v->stream));
    if (total > 0) {
        v->mesh_vertices.alloc((size_t)total * 9);
        v->mesh_colors.alloc((size_t)total * 9);
        { Prof pf(v->stream, KID_MESH_EMIT); k_mesh_cells<true><<<(unsigned)nblocks, 256, 0, v->stream>>>(m, v->mesh_counts.p, v->mesh_crossed.p, v->mesh_vertices.p, v->mesh_colors.p); }
        PSG_LAUNCH_CHECK();
    }
    v->mesh_triangles = total;
    return total;
}

// ---- text output: n lines formatted by several host threads into per-thread strings, written in order -----------------------------
template <class Fmt>
void write_lines(FILE* f, long long n, Fmt fmt) {
    const long long batch = 1ll << 21;
    unsigned nt = std::thread::hardware_concurrency();
    nt = nt < 1 ? 1 : nt > 16 ? 16 : nt;
    for (long long b0 = 0; b0 < n; b0 += batch) {
        const long long b1 = std::min(n, b0 + batch);
        const unsigned use = (b1 - b0) < 32768 ? 1u : nt;
        std::vector<std::string> parts(use);
        auto work = [&](unsigned t) {
            const long long per = (b1 - b0 + use - 1) / use, i0 = b0 + per * t, i1 = std::min(b1, i0 + per);
            std::string& s = parts[t];
            s.reserve((size_t)(i1 > i0 ? i1 - i0 : 0) * 40);
            char buf[192];
            for (long long i = i0; i < i1; ++i) s.append(buf, (size_t)fmt(i, buf, sizeof buf));
        };
        if (use == 1) work(0);
        else {
            std::vector<std::thread> th;
            for (unsigned t = 0; t < use; ++t) th.emplace_back(work, t);
            for (auto& t : th) t.join();
        }
        for (const auto& s : parts) fwrite(s.data(), 1, s.size(), f);
    }
}

}  // namespace

namespace psg {

// MarchingCubes::savePly (MarchingCubes.cpp:653-699): ostream << float is %g.  No file when there is no triangle (the reference's
// savePly returns false on an empty vertex list).  Returns false only when the file cannot be opened.
bool write_mesh_file(psg_volume* v, float vs, const char* path, int white, long long* n_triangles) {
    const long long nt = run_marching_cubes(v, vs, white, nullptr);
    if (n_triangles) *n_triangles = nt < 0 ? 0 : nt;
    if (nt <= 0) return true;
    std::vector<float> vert((size_t)nt * 9);
    std::vector<unsigned char> col((size_t)nt * 9);
    PSG_CUDA(cudaMemcpyAsync(vert.data(), v->mesh_vertices.p, vert.size() * sizeof(float), cudaMemcpyDeviceToHost, v->stream));
    PSG_CUDA(cudaMemcpyAsync(col.data(), v->mesh_colors.p, col.size(), cudaMemcpyDeviceToHost, v->stream));
    PSG_CUDA(cudaStreamSynchronize(v->stream));
    FILE* f = fopen(path, "w");
    if (!f) return false;
    fprintf(f, "ply\nformat ascii 1.0\nelement vertex %lld\nproperty float x\nproperty float y\nproperty float z\n", 3 * nt);
    fprintf(f, "property uchar red\nproperty uchar green\nproperty uchar blue\nelement face %d\n", (int)nt);
    fprintf(f, "property list uchar int vertex_indices\nend_header\n");
    const float* vp = vert.data();
    const unsigned char* cp = col.data();
    write_lines(f, 3 * nt, [=](long long i, char* buf, size_t cap) {
        return snprintf(buf, cap, "%g %g %g %d %d %d\n", vp[3 * i], vp[3 * i + 1], vp[3 * i + 2], cp[3 * i], cp[3 * i + 1], cp[3 * i + 2]);
    });
    write_lines(f, nt, [=](long long t, char* buf, size_t cap) {
        return snprintf(buf, cap, "3 %d %d %d\n", (int)(3 * t), (int)(3 * t + 1), (int)(3 * t + 2));
    });
    fclose(f);
    return true;
}

// Optimizer::saveSDF (OptimizerAux.cpp:513-577) / VolumetricGradSdf::saveSDF: only the crop box leaves the device.
bool write_sdf_file(psg_volume* v, float vs, const char* path, bool* any) {
    const Box b = crop_box(v, vs);
    if (any) *any = b.any;
    if (!b.any) return true;
    const MeshView m = mesh_view(v, vs, b, 0);
    const long long n = (long long)m.bd[0] * m.bd[1] * m.bd[2];
    v->staging.alloc((size_t)n);
    { Prof pf(v->stream, KID_MESH_MISC); k_crop_neg_dist<<<(unsigned)((n + 255) / 256), 256, 0, v->stream>>>(m, v->staging.p, n); }
    PSG_LAUNCH_CHECK();
    std::vector<float> host((size_t)n);
    PSG_CUDA(cudaMemcpyAsync(host.data(), v->staging.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, v->stream));
    PSG_CUDA(cudaStreamSynchronize(v->stream));
    FILE* f = fopen(path, "w");
    if (!f) return false;
    fprintf(f, "%d %d %d\n", m.bd[0], m.bd[1], m.bd[2]);
    fprintf(f, "%g %g %g\n", (float)b.lo[0] * vs, (float)b.lo[1] * vs, (float)b.lo[2] * vs);
    fprintf(f, "%g\n", vs);
    const float* hp = host.data();
    write_lines(f, n, [=](long long i, char* buf, size_t cap) { return snprintf(buf, cap, "%g\n", hp[i]); });
    fclose(f);
    return true;
}

}  // namespace psg

extern "C" {

#define PSG_MESH_TRY try {
#define PSG_MESH_CATCH } catch (const psg::CudaError&) { return PSG_ERR_CUDA; } catch (const std::exception& e) { set_error("%s", e.what()); return PSG_ERR_INVALID; }

int psg_volume_mesh_box(psg_volume* v, float voxel_size, int lo[3], int hi[3], int* any) {
    if (!v || !lo || !hi || !any) { set_error("psg_volume_mesh_box: null argument"); return PSG_ERR_INVALID; }
    PSG_MESH_TRY
    const Box b = crop_box(v, voxel_size > 0.f ? voxel_size : v->vs);
    for (int a = 0; a < 3; ++a) { lo[a] = b.lo[a]; hi[a] = b.hi[a]; }
    *any = b.any ? 1 : 0;
    return PSG_OK;
    PSG_MESH_CATCH
}

int psg_volume_mesh_extract(psg_volume* v, float voxel_size, int white, long long* n_triangles) {
    if (!v || !n_triangles) { set_error("psg_volume_mesh_extract: null argument"); return PSG_ERR_INVALID; }
    PSG_MESH_TRY
    const long long nt = run_marching_cubes(v, voxel_size > 0.f ? voxel_size : v->vs, white, nullptr);
    *n_triangles = nt < 0 ? 0 : nt;
    return PSG_OK;
    PSG_MESH_CATCH
}

int psg_volume_mesh_download(psg_volume* v, float* vertices, unsigned char* colors) {
    if (!v) { set_error("psg_volume_mesh_download: null volume"); return PSG_ERR_INVALID; }
    if (v->mesh_triangles < 0) { set_error("psg_volume_mesh_download: no mesh extracted"); return PSG_ERR_STATE; }
    PSG_MESH_TRY
    PSG_CUDA(cudaSetDevice(v->device));
    const size_t n = (size_t)v->mesh_triangles * 9;
    if (n && vertices) PSG_CUDA(cudaMemcpyAsync(vertices, v->mesh_vertices.p, n * sizeof(float), cudaMemcpyDeviceToHost, v->stream));
    if (n && colors) PSG_CUDA(cudaMemcpyAsync(colors, v->mesh_colors.p, n, cudaMemcpyDeviceToHost, v->stream));
    PSG_CUDA(cudaStreamSynchronize(v->stream));
    return PSG_OK;
    PSG_MESH_CATCH
}

int psg_volume_extract_mesh(psg_volume* v, const char* filename, int white, long long* n_triangles) {
    if (!v || !filename) { set_error("psg_volume_extract_mesh: null argument"); return PSG_ERR_INVALID; }
    PSG_MESH_TRY
    if (!write_mesh_file(v, v->vs, filename, white, n_triangles)) { set_error("psg_volume_extract_mesh: cannot open %s", filename); return PSG_ERR_INVALID; }
    return PSG_OK;
    PSG_MESH_CATCH
}

int psg_volume_save_sdf(psg_volume* v, const char* filename) {
    if (!v || !filename) { set_error("psg_volume_save_sdf: null argument"); return PSG_ERR_INVALID; }
    PSG_MESH_TRY
    bool any = false;
    if (!write_sdf_file(v, v->vs, filename, &any)) { set_error("psg_volume_save_sdf: cannot open %s", filename); return PSG_ERR_INVALID; }
    return PSG_OK;
    PSG_MESH_CATCH
}

int psg_volume_extract_pc(psg_volume* v, const char* filename, long long* n_points) {
    if (!v || !filename) { set_error("psg_volume_extract_pc: null argument"); return PSG_ERR_INVALID; }
    PSG_MESH_TRY
    PSG_CUDA(cudaSetDevice(v->device));
    const double thr = std::sqrt(3.0) * (double)v->vs;
    const long long nblocks = (v->N + kCellsPerBlock - 1) / kCellsPerBlock;
    v->mesh_counts.alloc((size_t)nblocks + 1);
    { Prof pf(v->stream, KID_MESH_MISC); k_volume_points<false><<<(unsigned)nblocks, 256, 0, v->stream>>>(v->view(), thr, v->mesh_counts.p, nullptr); }
    { Prof pf(v->stream, KID_MESH_SCAN); k_mesh_scan<<<1, 1024, 0, v->stream>>>(v->mesh_counts.p, nblocks); }
    PSG_LAUNCH_CHECK();
    unsigned total = 0;
    PSG_CUDA(cudaMemcpyAsync(&total, v->mesh_counts.p + nblocks, sizeof total, cudaMemcpyDeviceToHost, v->stream));
    PSG_CUDA(cudaStreamSynchronize(v->stream));
    if (n_points) *n_points = total;
    std::vector<float> host((size_t)total * 9 + 1);
    if (total) {
        v->staging.alloc((size_t)total * 9);
        { Prof pf(v->stream, KID_MESH_MISC); k_volume_points<true><<<(unsigned)nblocks, 256, 0, v->stream>>>(v->view(), thr, v->mesh_counts.p, v->staging.p); }
        PSG_LAUNCH_CHECK();
        PSG_CUDA(cudaMemcpyAsync(host.data(), v->staging.p, (size_t)total * 9 * sizeof(float), cudaMemcpyDeviceToHost, v->stream));
        PSG_CUDA(cudaStreamSynchronize(v->stream));
    }
    FILE* f = fopen(filename, "w");
    if (!f) { set_error("psg_volume_extract_pc: cannot open %s", filename); return PSG_ERR_INVALID; }
    fprintf(f, "ply\nformat ascii 1.0\nelement vertex %u\nproperty float x\nproperty float y\nproperty float z\n", total);
    fprintf(f, "property float nx\nproperty float ny\nproperty float nz\nproperty uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n");
    const float* hp = host.data();
    write_lines(f, total, [=](long long i, char* buf, size_t cap) {
        const float* p = hp + 9 * i;
        return snprintf(buf, cap, "%g %g %g %g %g %g %d %d %d\n", p[0], p[1], p[2], p[3], p[4], p[5], int(255 * p[6]), int(255 * p[7]), int(255 * p[8]));
    });
    fclose(f);
    return PSG_OK;
    PSG_MESH_CATCH
}

int psg_opt_extract_mesh(psg_optimizer* o, const char* filename, int white, long long* n_triangles) {
    if (!o || !o->vol || !filename) { set_error("psg_opt_extract_mesh: null argument"); return PSG_ERR_INVALID; }
    PSG_MESH_TRY
    if (!write_mesh_file(o->vol, o->vs, filename, white, n_triangles)) { set_error("psg_opt_extract_mesh: cannot open %s", filename); return PSG_ERR_INVALID; }
    return PSG_OK;
    PSG_MESH_CATCH
}

int psg_opt_save_sdf(psg_optimizer* o, const char* filename) {
    if (!o || !o->vol || !filename) { set_error("psg_opt_save_sdf: null argument"); return PSG_ERR_INVALID; }
    PSG_MESH_TRY
    bool any = false;
    if (!write_sdf_file(o->vol, o->vs, filename, &any)) { set_error("psg_opt_save_sdf: cannot open %s", filename); return PSG_ERR_INVALID; }
    return PSG_OK;
    PSG_MESH_CATCH
}

}  // extern "C"
