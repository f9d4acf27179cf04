/* TEST INFRASTRUCTURE ONLY.  Never linked into, loaded by, or called from the product library.
 *
 * Plain-C restatement of the reference's surface extraction, pinned against the reference build
 * (oracle/_ref/libpsgref.so: ref_marching_cubes, ref_vol_extract_mesh, ref_opt_extract_mesh ...) by tests/test_oracle_mesh.py:
 *   or_mesh_bbox        -- the crop box of Optimizer::extract_mesh / save_white_mesh / saveSDF (OptimizerAux.cpp:278-308,
 *                          365-395, 513-541) and VolumetricGradSdf::extract_mesh (VolumetricGradSdf.cpp:234-270)
 *   or_extract_mesh     -- the crop copy (OptimizerAux.cpp:310-342) + MarchingCubes::computeIsoSurface
 *                          (third/mesh/MarchingCubes.cpp:314-541) with its helpers (:549-699)
 *   or_write_mesh_ply   -- MarchingCubes::savePly (:653-699)
 *   or_write_sdf        -- Optimizer::saveSDF (OptimizerAux.cpp:513-577)
 *   or_write_volume_pc  -- VolumetricGradSdf::extract_pc (VolumetricGradSdf.cpp:320-376)
 * Volumes are SdfVoxel arrays as psg_oracle.c keeps them: 8 floats per voxel {dist, gx, gy, gz, weight, r, g, b}.
 * Compiled with -ffp-contract=off: every float / double operation rounds where the reference's x86-64 build rounds.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../psgradientsdf_b200/csrc/mc_table.h"

static const unsigned long long kTri[256] = PSG_MC_TABLE_INIT;

/* corners of edge e as (dx, dy, dz) pairs, in the order the reference hands them to getVertex (MarchingCubes.cpp:342-506) ... */
static const int kEdgeV[12][6] = {
    {1, 1, 0, 1, 0, 0}, {1, 0, 0, 0, 0, 0}, {0, 0, 0, 0, 1, 0}, {0, 1, 0, 1, 1, 0},
    {1, 1, 1, 1, 0, 1}, {1, 0, 1, 0, 0, 1}, {0, 0, 1, 0, 1, 1}, {0, 1, 1, 1, 1, 1},
    {1, 1, 0, 1, 1, 1}, {1, 0, 0, 1, 0, 1}, {0, 0, 0, 0, 0, 1}, {0, 1, 0, 0, 1, 1}};
/* ... and to getColor: edges 2, 3, 6, 7 pass their two corners the other way round (:374,386,422,434) */
static const int kEdgeC[12][6] = {
    {1, 1, 0, 1, 0, 0}, {1, 0, 0, 0, 0, 0}, {0, 1, 0, 0, 0, 0}, {1, 1, 0, 0, 1, 0},
    {1, 1, 1, 1, 0, 1}, {1, 0, 1, 0, 0, 1}, {0, 1, 1, 0, 0, 1}, {1, 1, 1, 0, 1, 1},
    {1, 1, 0, 1, 1, 1}, {1, 0, 0, 1, 0, 1}, {0, 0, 0, 0, 0, 1}, {0, 1, 0, 0, 1, 1}};

/* The crop box: voxels with |dist| <= sqrt(3) * vs, compared in double (std::sqrt(3) is the double overload).  Returns 0 when
 * no voxel qualifies (the reference then computes garbage dimensions; callers treat it as "nothing to write"). */
int or_mesh_bbox(const float* aos, const int* dims, float vs, int* lo, int* hi) {
    const double thr = sqrt(3.0) * (double)vs;
    int any = 0;
    size_t l = 0;
    for (int a = 0; a < 3; ++a) { lo[a] = 2147483647; hi[a] = -2147483647 - 1; }
    for (int k = 0; k < dims[2]; ++k) for (int j = 0; j < dims[1]; ++j) for (int i = 0; i < dims[0]; ++i, ++l) {
        if ((double)fabsf(aos[8 * l]) > thr) continue;
        any = 1;
        if (i < lo[0]) lo[0] = i;
        if (i > hi[0]) hi[0] = i;
        if (j < lo[1]) lo[1] = j;
        if (j > hi[1]) hi[1] = j;
        if (k < lo[2]) lo[2] = k;
        if (k > hi[2]) hi[2] = k;
    }
    return any;
}

typedef struct {
    int d[3];
    float vsz[3], org[3];
    const float* tsdf; const float* w;
    const unsigned char* r; const unsigned char* g; const unsigned char* b;
} Mc;

/* MarchingCubes::interpolate (:589-609), one component at a time */
static void mc_interp(float t0, float t1, const float* v0, const float* v1, float iso, float* out) {
    if ((double)fabsf(iso - t0) < 1e-7) { memcpy(out, v0, 12); return; }
    if ((double)fabsf(iso - t1) < 1e-7) { memcpy(out, v1, 12); return; }
    if ((double)fabsf(t0 - t1) < 1e-7) { memcpy(out, v0, 12); return; }
    double mu = (double)((iso - t0) / (t1 - t0));      /* the quotient is a float */
    if (mu > 1.0) mu = 1.0; else if (mu < 0) mu = 0.0;
    for (int a = 0; a < 3; ++a) out[a] = (float)((double)v0[a] + mu * (double)(v1[a] - v0[a]));
}

static size_t mc_off(const Mc* m, int i, int j, int k) { return (size_t)k * m->d[0] * m->d[1] + (size_t)j * m->d[0] + i; }

/* MarchingCubes::getVertex (:612-619) with voxelToWorld (:646-650) */
static void mc_vertex(const Mc* m, const int* c, float iso, float* out) {
    float p1[3], p2[3];
    for (int a = 0; a < 3; ++a) { p1[a] = (float)c[a] * m->vsz[a] - m->org[a]; p2[a] = (float)c[3 + a] * m->vsz[a] - m->org[a]; }
    mc_interp(m->tsdf[mc_off(m, c[0], c[1], c[2])], m->tsdf[mc_off(m, c[3], c[4], c[5])], p1, p2, iso, out);
}

/* MarchingCubes::getColor (:622-640): green is read one voxel further along x than red, blue two (green_[idx+1], blue_[idx+2]) */
static void mc_color(const Mc* m, const int* c, float iso, unsigned char* out) {
    const size_t i1 = mc_off(m, c[0], c[1], c[2]), i2 = mc_off(m, c[3], c[4], c[5]);
    const float c1[3] = { (float)m->r[i1] / 255.0f, (float)m->g[i1 + 1] / 255.0f, (float)m->b[i1 + 2] / 255.0f };
    const float c2[3] = { (float)m->r[i2] / 255.0f, (float)m->g[i2 + 1] / 255.0f, (float)m->b[i2 + 2] / 255.0f };
    float cv[3];
    mc_interp(m->tsdf[i1], m->tsdf[i2], c1, c2, iso, cv);
    for (int a = 0; a < 3; ++a) out[a] = (unsigned char)(cv[a] * 255.0f);
}

/* MarchingCubes::computeLutIndex (:549-585) */
static int mc_case(const Mc* m, int i, int j, int k, float iso) {
    static const int cor[8][3] = {{1, 1, 0}, {1, 0, 0}, {0, 0, 0}, {0, 1, 0}, {1, 1, 1}, {1, 0, 1}, {0, 0, 1}, {0, 1, 1}};
    int idx = 0;
    for (int b = 0; b < 8; ++b) if (m->w[mc_off(m, i + cor[b][0], j + cor[b][1], k + cor[b][2])] == 0.0f) return 0;
    for (int b = 0; b < 8; ++b) if (m->tsdf[mc_off(m, i + cor[b][0], j + cor[b][1], k + cor[b][2])] > iso) idx |= 1 << b;
    return idx;
}

/* MarchingCubes::computeIsoSurface on cropped arrays.  Outputs may be null (count only).  Returns the number of triangles; vertex
 * 3t+n of triangle t goes to vertices[9t+3n..], its colour to colors[9t+3n..]; faces are (3t, 3t+1, 3t+2) (computeTriangles,
 * :613-641: three new vertices per triangle, degenerate triangles dropped). */
long long or_marching_cubes(const int* dims, const float* size3, const float* origin3, const float* tsdf, const float* weights,
                            const unsigned char* red, const unsigned char* green, const unsigned char* blue, float iso,
                            float* vertices, unsigned char* colors) {
    Mc m;
    for (int a = 0; a < 3; ++a) { m.d[a] = dims[a]; m.vsz[a] = size3[a] / (float)dims[a]; m.org[a] = origin3[a]; }
    m.tsdf = tsdf; m.w = weights; m.r = red; m.g = green; m.b = blue;
    long long nt = 0;
    for (int z = 0; z < dims[2] - 2; ++z) for (int y = 0; y < dims[1] - 2; ++y) for (int x = 0; x < dims[0] - 2; ++x) {
        const int ci = mc_case(&m, x, y, z, iso);
        if (ci == 0 || ci == 255) continue;
        float ep[12][3];
        unsigned char ec[12][3];
        unsigned have = 0;
        for (unsigned long long t = kTri[ci]; (t & 0xF) != 0xF; t >>= 4) have |= 1u << (t & 0xF);
        for (int e = 0; e < 12; ++e) if (have >> e & 1) {
            int c[6];
            for (int a = 0; a < 6; ++a) c[a] = kEdgeV[e][a] + (a % 3 == 0 ? x : a % 3 == 1 ? y : z);
            mc_vertex(&m, c, iso, ep[e]);
            for (int a = 0; a < 6; ++a) c[a] = kEdgeC[e][a] + (a % 3 == 0 ? x : a % 3 == 1 ? y : z);
            mc_color(&m, c, iso, ec[e]);
        }
        for (unsigned long long t = kTri[ci]; (t & 0xF) != 0xF; t >>= 12) {
            const int e0 = (int)(t & 0xF), e1 = (int)(t >> 4 & 0xF), e2 = (int)(t >> 8 & 0xF);
            const float *p1 = ep[e0], *p2 = ep[e1], *p3 = ep[e2];
            const int ne12 = p1[0] != p2[0] || p1[1] != p2[1] || p1[2] != p2[2];
            const int ne13 = p1[0] != p3[0] || p1[1] != p3[1] || p1[2] != p3[2];
            const int ne23 = p2[0] != p3[0] || p2[1] != p3[1] || p2[2] != p3[2];
            if (!(ne12 && ne13 && ne23)) continue;
            if (vertices) { memcpy(vertices + 9 * nt, p1, 12); memcpy(vertices + 9 * nt + 3, p2, 12); memcpy(vertices + 9 * nt + 6, p3, 12); }
            if (colors) { memcpy(colors + 9 * nt, ec[e0], 3); memcpy(colors + 9 * nt + 3, ec[e1], 3); memcpy(colors + 9 * nt + 6, ec[e2], 3); }
            ++nt;
        }
    }
    return nt;
}

/* Optimizer::extract_mesh (white = 0, OptimizerAux.cpp:278-362) / save_white_mesh (white = 1, :365-453) /
 * VolumetricGradSdf::extract_mesh (white = 0): crop, copy -dist, weight and int(255 * colour), run marching cubes with
 * size = vs * dim and origin = -vs * (xmin, ymin, zmin).  Returns the number of triangles, -1 when the crop box is empty. */
long long or_extract_mesh(const float* aos, const int* dims, float vs, int white, int* lo_out, int* hi_out, float* vertices,
                          unsigned char* colors) {
    int lo[3], hi[3];
    if (!or_mesh_bbox(aos, dims, vs, lo, hi)) return -1;
    if (lo_out) memcpy(lo_out, lo, sizeof lo);
    if (hi_out) memcpy(hi_out, hi, sizeof hi);
    const int d[3] = { hi[0] - lo[0] + 1, hi[1] - lo[1] + 1, hi[2] - lo[2] + 1 };
    const size_t n = (size_t)d[0] * d[1] * d[2];
    float* tsdf = (float*)malloc(n * 4);
    float* w = (float*)malloc(n * 4);
    unsigned char* rgb = (unsigned char*)malloc(3 * n);
    size_t pos = 0;
    for (int k = lo[2]; k <= hi[2]; ++k) for (int j = lo[1]; j <= hi[1]; ++j) for (int i = lo[0]; i <= hi[0]; ++i, ++pos) {
        const float* v = aos + 8 * ((size_t)i + (size_t)dims[0] * j + (size_t)dims[0] * dims[1] * k);
        tsdf[pos] = -v[0];
        w[pos] = v[4];
        for (int c = 0; c < 3; ++c) rgb[c * n + pos] = white ? 255 : (unsigned char)(int)(255 * v[5 + c]);
    }
    const float size3[3] = { vs * (float)d[0], vs * (float)d[1], vs * (float)d[2] };
    const float org[3] = { -vs * (float)lo[0], -vs * (float)lo[1], -vs * (float)lo[2] };
    const long long nt = or_marching_cubes(d, size3, org, tsdf, w, rgb, rgb + n, rgb + 2 * n, 0.0f, vertices, colors);
    free(tsdf); free(w); free(rgb);
    return nt;
}

/* MarchingCubes::savePly (:653-699); ostream << float prints like %g.  Returns 0 when there is nothing to write or the
 * file cannot be opened (the reference returns false in both cases). */
int or_write_mesh_ply(const char* path, long long n_tri, const float* vertices, const unsigned char* colors) {
    if (n_tri <= 0) return 0;
    FILE* f = fopen(path, "w");
    if (!f) return 0;
    fprintf(f, "ply\nformat ascii 1.0\nelement vertex %lld\nproperty float x\nproperty float y\nproperty float z\n", 3 * n_tri);
    fprintf(f, "property uchar red\nproperty uchar green\nproperty uchar blue\nelement face %d\n", (int)n_tri);
    fprintf(f, "property list uchar int vertex_indices\nend_header\n");
    for (long long i = 0; i < 3 * n_tri; ++i)
        fprintf(f, "%g %g %g %d %d %d\n", vertices[3 * i], vertices[3 * i + 1], vertices[3 * i + 2], colors[3 * i], colors[3 * i + 1],
                colors[3 * i + 2]);
    for (long long t = 0; t < n_tri; ++t) fprintf(f, "3 %d %d %d\n", (int)(3 * t), (int)(3 * t + 1), (int)(3 * t + 2));
    fclose(f);
    return 1;
}

/* Optimizer::saveSDF (OptimizerAux.cpp:513-577) / VolumetricGradSdf::saveSDF: box dimensions, its lower corner in grid
 * units * vs, the voxel size, then -dist of every voxel of the box. */
int or_write_sdf(const char* path, const float* aos, const int* dims, float vs) {
    int lo[3], hi[3];
    if (!or_mesh_bbox(aos, dims, vs, lo, hi)) return 0;
    FILE* f = fopen(path, "w");
    if (!f) return 0;
    fprintf(f, "%d %d %d\n", hi[0] - lo[0] + 1, hi[1] - lo[1] + 1, hi[2] - lo[2] + 1);
    fprintf(f, "%g %g %g\n", (float)lo[0] * vs, (float)lo[1] * vs, (float)lo[2] * vs);
    fprintf(f, "%g\n", vs);
    for (int k = lo[2]; k <= hi[2]; ++k) for (int j = lo[1]; j <= hi[1]; ++j) for (int i = lo[0]; i <= hi[0]; ++i)
        fprintf(f, "%g\n", -aos[8 * ((size_t)i + (size_t)dims[0] * j + (size_t)dims[0] * dims[1] * k)]);
    fclose(f);
    return 1;
}

/* VolumetricGradSdf::extract_pc (VolumetricGradSdf.cpp:320-376): every voxel with |dist| <= sqrt(3.0) * vs seen by at least one
 * frame (seen[l] != 0), in lin order: position vs * (i,j,k) - dist * g_hat, g_hat, int(255 * colour). */
int or_write_volume_pc(const char* path, const float* aos, const unsigned char* seen, const int* dims, float vs) {
    const double thr = sqrt(3.0) * (double)vs;
    FILE* f = fopen(path, "w");
    if (!f) return 0;
    long long count = 0;
    size_t l = 0;
    for (int k = 0; k < dims[2]; ++k) for (int j = 0; j < dims[1]; ++j) for (int i = 0; i < dims[0]; ++i, ++l)
        if (!((double)fabsf(aos[8 * l]) > thr) && seen[l]) ++count;
    fprintf(f, "ply\nformat ascii 1.0\nelement vertex %lld\nproperty float x\nproperty float y\nproperty float z\n", count);
    fprintf(f, "property float nx\nproperty float ny\nproperty float nz\nproperty uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n");
    l = 0;
    for (int k = 0; k < dims[2]; ++k) for (int j = 0; j < dims[1]; ++j) for (int i = 0; i < dims[0]; ++i, ++l) {
        const float* v = aos + 8 * l;
        if ((double)fabsf(v[0]) > thr || !seen[l]) continue;
        float g[3] = { v[1], v[2], v[3] };
        const float z = g[0] * g[0] + (g[1] * g[1] + g[2] * g[2]);     /* Eigen's 3-term reduction order */
        if (z > 0.f) { const float s = sqrtf(z); g[0] /= s; g[1] /= s; g[2] /= s; }
        const float idx[3] = { (float)i, (float)j, (float)k };
        float p[3];
        for (int a = 0; a < 3; ++a) p[a] = vs * idx[a] - v[0] * g[a];
        fprintf(f, "%g %g %g %g %g %g %d %d %d\n", p[0], p[1], p[2], g[0], g[1], g[2], (int)(255 * v[5]), (int)(255 * v[6]), (int)(255 * v[7]));
    }
    fclose(f);
    return 1;
}

/* Optimizer::save_pointcloud (OptimizerAux.cpp:456-511): the surface list's voxels with |dist| < sqrt(3) * vs (strictly, std::abs
 * in float against a double product), same record as extract_pc. */
int or_write_surface_pc(const char* path, const float* aos, const int* lin, int S, const int* dims, float vs) {
    const double thr = sqrt(3.0) * (double)vs;
    FILE* f = fopen(path, "w");
    if (!f) return 0;
    long long count = 0;
    for (int s = 0; s < S; ++s) if ((double)fabsf(aos[8 * (size_t)lin[s]]) < thr) ++count;
    fprintf(f, "ply\nformat ascii 1.0\nelement vertex %lld\nproperty float x\nproperty float y\nproperty float z\n", count);
    fprintf(f, "property float nx\nproperty float ny\nproperty float nz\nproperty uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n");
    for (int s = 0; s < S; ++s) {
        const float* v = aos + 8 * (size_t)lin[s];
        if (!((double)fabsf(v[0]) < thr)) continue;
        int rest = lin[s];
        const int k = rest / (dims[0] * dims[1]); rest -= k * dims[0] * dims[1];
        const int j = rest / dims[0], i = rest - j * dims[0];
        float g[3] = { v[1], v[2], v[3] };
        const float z = g[0] * g[0] + (g[1] * g[1] + g[2] * g[2]);
        if (z > 0.f) { const float sq = sqrtf(z); g[0] /= sq; g[1] /= sq; g[2] /= sq; }
        const float idx[3] = { (float)i, (float)j, (float)k };
        float p[3];
        for (int a = 0; a < 3; ++a) p[a] = vs * idx[a] - v[0] * g[a];
        fprintf(f, "%g %g %g %g %g %g %d %d %d\n", p[0], p[1], p[2], g[0], g[1], g[2], (int)(255 * v[5]), (int)(255 * v[6]), (int)(255 * v[7]));
    }
    fclose(f);
    return 1;
}
