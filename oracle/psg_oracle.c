/* TEST INFRASTRUCTURE ONLY -- see psg_oracle.h.
 *
 * CPU restatement of the reference's photometric-optimizer hot path.  Every function cites the
 * reference file:line it follows (paths relative to /root/reference/cpp/include).  Float evaluation
 * order is the reference's as compiled with -O2 on x86-64 (SSE2, no FMA contraction):
 *   - fixed-size Eigen 3-vector reductions (dot, squaredNorm, small matrix products) are fully
 *     unrolled as  a0 + (a1 + a2)   (third/eigen/Eigen/src/Core/Redux.h:100-116);
 *   - dynamic VectorXf inner products of length 4 / 9 use the SSE packet order
 *     (Redux.h:236-281, arch/SSE/PacketMath.h:978-988);
 *   - cv::Vec3f * double scales in double and rounds to float per channel;
 *   - the block solves follow Eigen operation by operation (eigen_dot, eigen_like_cg, rowsum2 below): against the reference
 *     build, two complete outer iterations are bit-identical for SH1, SH2 and LED (tests/test_oracle_vs_ref.py).
 * Build with -ffp-contract=off (oracle/Makefile).
 */
#include "psg_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ small helpers */
static inline float sum3(float a0, float a1, float a2) { return a0 + (a1 + a2); }
static inline float dot3(const float* a, const float* b) { return sum3(a[0] * b[0], a[1] * b[1], a[2] * b[2]); }
static inline float sqnorm3(const float* a) { return sum3(a[0] * a[0], a[1] * a[1], a[2] * a[2]); }
static inline float norm3(const float* a) { return sqrtf(sqnorm3(a)); }
/* Eigen normalized(): unchanged when squaredNorm is 0 (Core/Dot.h:121-131); per-component division */
static inline void normalized3(const float* a, float* o) {
    float z = sqnorm3(a);
    if (z > 0.f) { float s = sqrtf(z); o[0] = a[0] / s; o[1] = a[1] / s; o[2] = a[2] / s; }
    else { o[0] = a[0]; o[1] = a[1]; o[2] = a[2]; }
}
/* o = M^T v, M row-major 3x3 */
static inline void matT_vec(const float* M, const float* v, float* o) {
    for (int a = 0; a < 3; ++a) o[a] = sum3(M[0 + a] * v[0], M[3 + a] * v[1], M[6 + a] * v[2]);
}
static inline void mat_vec(const float* M, const float* v, float* o) {
    for (int a = 0; a < 3; ++a) o[a] = sum3(M[3 * a] * v[0], M[3 * a + 1] * v[1], M[3 * a + 2] * v[2]);
}
/* dynamic-size Eigen inner product, SSE packet order */
static float dotn(const float* a, const float* b, int n) {
    if (n == 4) return (a[0] * b[0] + a[2] * b[2]) + (a[1] * b[1] + a[3] * b[3]);
    if (n == 9) {
        float p0 = a[0] * b[0] + a[4] * b[4], p1 = a[1] * b[1] + a[5] * b[5];
        float p2 = a[2] * b[2] + a[6] * b[6], p3 = a[3] * b[3] + a[7] * b[7];
        return ((p0 + p2) + (p1 + p3)) + a[8] * b[8];
    }
    float r = a[0] * b[0];
    for (int i = 1; i < n; ++i) r += a[i] * b[i];
    return r;
}
/* (float) std::pow(x, 3) and 5: float argument promoted to double (C++11 <cmath>) */
static inline double powd(float x, int e) { return pow((double)x, (double)e); }

/* ------------------------------------------------------------------ volume */
struct or_volume {
    int dims[3];
    float vs, shift[3], origin[3], T, invT, zmin, zmax;
    long long N;
    int counter, nwords;
    float *dist, *gx, *gy, *gz, *w, *r, *g, *b;
    uint64_t* vis; /* frame indexed, word-major */
    int* vis_len;
};

static void vol_alloc(or_volume* v) {
    size_t N = (size_t)v->N;
    v->dist = malloc(4 * N); v->gx = malloc(4 * N); v->gy = malloc(4 * N); v->gz = malloc(4 * N);
    v->w = malloc(4 * N); v->r = malloc(4 * N); v->g = malloc(4 * N); v->b = malloc(4 * N);
    v->vis = calloc(N * (size_t)v->nwords, 8);
    v->vis_len = calloc(N, 4);
}
static void vol_free_arrays(or_volume* v) {
    free(v->dist); free(v->gx); free(v->gy); free(v->gz); free(v->w); free(v->r); free(v->g); free(v->b);
    free(v->vis); free(v->vis_len);
}

/* VoxelGrid.h:127-133 (origin), Sdf.h:83-87 (inv_T = 1./T in double), VolumetricGradSdf.cpp:14-38 (init) */
or_volume* or_volume_create(const int dims[3], float voxel_size, const float shift[3], float T, float zmin,
                            float zmax, int max_frames) {
    or_volume* v = calloc(1, sizeof(*v));
    for (int i = 0; i < 3; ++i) { v->dims[i] = dims[i]; v->shift[i] = shift[i]; }
    v->vs = voxel_size;
    const float half_vs = (float)(0.5 * (double)voxel_size);
    for (int i = 0; i < 3; ++i) v->origin[i] = shift[i] - half_vs * (float)dims[i];
    v->T = T; v->invT = (float)(1.0 / (double)T);
    v->zmin = zmin; v->zmax = zmax;
    v->N = (long long)dims[0] * dims[1] * dims[2];
    v->nwords = (max_frames + 63) / 64; if (v->nwords < 1) v->nwords = 1;
    vol_alloc(v);
    for (long long i = 0; i < v->N; ++i) {
        v->dist[i] = T; v->gx[i] = v->gy[i] = v->gz[i] = 0.f; v->w[i] = 0.f; v->r[i] = v->g[i] = v->b[i] = 0.f;
    }
    return v;
}
void or_volume_destroy(or_volume* v) { vol_free_arrays(v); free(v); }
long long or_volume_num_voxels(const or_volume* v) { return v->N; }
void or_volume_geometry(const or_volume* v, int dims[3], float* vs, float origin[3]) {
    for (int i = 0; i < 3; ++i) { dims[i] = v->dims[i]; origin[i] = v->origin[i]; }
    *vs = v->vs;
}
int or_volume_vis_words(const or_volume* v) { return v->nwords; }

/* VolumetricGradSdf::update, VolumetricGradSdf.cpp:51-138; weight/truncate Sdf.h:44-66 */
void or_volume_integrate(or_volume* v, const float* color, const float* depth, const float* nx, const float* ny,
                         const float* nz, const float* nsq, int H, int W, const float K[9], const float R[9],
                         const float t[3], int increase_counter) {
    if (increase_counter) v->counter++;
    const float fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    const int c = v->counter;
    if (c >= 64 * v->nwords) { fprintf(stderr, "or_volume_integrate: frame %d beyond max_frames\n", c); abort(); }
    const size_t N = (size_t)v->N;
    for (int k = 0; k < v->dims[2]; ++k) for (int j = 0; j < v->dims[1]; ++j) for (int i = 0; i < v->dims[0]; ++i) {
        const size_t lin = (size_t)i + (size_t)v->dims[0] * j + (size_t)v->dims[0] * v->dims[1] * k;
        float p[3] = { v->origin[0] + v->vs * (float)i, v->origin[1] + v->vs * (float)j, v->origin[2] + v->vs * (float)k };
        float q[3] = { p[0] - t[0], p[1] - t[1], p[2] - t[2] };
        float pt[3];
        matT_vec(R, q, pt);
        if (pt[2] < 0.f) continue;
        const int n = (int)((double)(cx + fx * pt[0] / pt[2]) + 0.5);
        const int m = (int)((double)(cy + fy * pt[1] / pt[2]) + 0.5);
        if (n < 0 || n >= W || m < 0 || m >= H) continue;
        const size_t pix = (size_t)m * W + n;
        const float z = depth[pix];
        if (z <= v->zmin || z >= v->zmax) continue;
        const float sdf = z - pt[2];
        float w = 0.f;
        if (sdf >= 0.f) w = 1.f; else if (sdf >= -v->T) w = 1.f + sdf * v->invT;
        if (w == 0.f) continue;
        const float nrm[3] = { nx[pix], ny[pix], nz[pix] };
        if ((double)sqnorm3(nrm) < .1) continue;
        const float zi = (float)(1.0 / (double)pt[2]);
        const float xy[3] = { zi * pt[0], zi * pt[1], zi * pt[2] };
        const float d = dot3(nrm, xy);
        if ((double)(d * d * nsq[pix]) < .25 * .25) continue;
        v->w[lin] += w;
        const float tr = fmaxf(-v->T, fminf(v->T, sdf));
        v->dist[lin] += (tr - v->dist[lin]) * w / v->w[lin];
        v->gx[lin] -= sum3((w * R[0]) * nrm[0], (w * R[1]) * nrm[1], (w * R[2]) * nrm[2]);
        v->gy[lin] -= sum3((w * R[3]) * nrm[0], (w * R[4]) * nrm[1], (w * R[5]) * nrm[2]);
        v->gz[lin] -= sum3((w * R[6]) * nrm[0], (w * R[7]) * nrm[1], (w * R[8]) * nrm[2]);
        v->r[lin] += (color[3 * pix + 2] - v->r[lin]) * w / v->w[lin];
        v->g[lin] += (color[3 * pix + 1] - v->g[lin]) * w / v->w[lin];
        v->b[lin] += (color[3 * pix + 0] - v->b[lin]) * w / v->w[lin];
        /* vis.resize(counter_); vis.push_back(true) */
        v->vis[(size_t)(c >> 6) * N + lin] |= (uint64_t)1 << (c & 63);
        v->vis_len[lin] = c + 1;
    }
}

void or_volume_download(const or_volume* v, float* a) {
    for (long long i = 0; i < v->N; ++i) {
        float* o = a + 8 * i;
        o[0] = v->dist[i]; o[1] = v->gx[i]; o[2] = v->gy[i]; o[3] = v->gz[i];
        o[4] = v->w[i]; o[5] = v->r[i]; o[6] = v->g[i]; o[7] = v->b[i];
    }
}
void or_volume_upload(or_volume* v, const float* a) {
    for (long long i = 0; i < v->N; ++i) {
        const float* o = a + 8 * i;
        v->dist[i] = o[0]; v->gx[i] = o[1]; v->gy[i] = o[2]; v->gz[i] = o[3];
        v->w[i] = o[4]; v->r[i] = o[5]; v->g[i] = o[6]; v->b[i] = o[7];
    }
}
void or_volume_vis(const or_volume* v, uint64_t* words, int* lengths) {
    memcpy(words, v->vis, (size_t)v->N * v->nwords * 8);
    if (lengths) memcpy(lengths, v->vis_len, (size_t)v->N * 4);
}
void or_volume_set_vis(or_volume* v, const uint64_t* words, const int* lengths) {
    memcpy(v->vis, words, (size_t)v->N * v->nwords * 8);
    memcpy(v->vis_len, lengths, (size_t)v->N * 4);
}

/* ------------------------------------------------------------------ optimizer state */
typedef struct { int i, j; float v; } trip;

struct or_opt {
    or_volume* vol;
    int accum64;   /* diagnostic: carry the long per-keyframe sums of the light / pose blocks in double (or_opt_set_accum64) */
    int model, order, nb, loss;
    float vs, vs_inv, K[9], lambda, lambda_sq, reg_rho, reg_n, reg_l;
    int F, H, W;
    int* frame_idx;
    float* poses;  /* F x 16 */
    float* images; /* F x H x W x 3 BGR */
    float* light;  /* SH: F x nb ; LED: 3 */
    int S;
    int* surf;          /* S lin indices ascending */
    int* rowmap;        /* N: row or -1 */
    uint64_t* visF;     /* N keyframe-indexed visibility (after select_vis) */
    /* last block */
    int n; float *b, *delta; int success, iters; float err;
    trip* Ht; long long Hn, Hcap;
};

or_opt* or_opt_create(or_volume* vol, int model, float voxel_size, const float K[9], int loss, float lambda,
                      float reg_rho, float reg_n, float reg_l) {
    or_opt* o = calloc(1, sizeof(*o));
    o->vol = vol; o->model = model; o->order = model == OR_SH2 ? 2 : 1; o->nb = model == OR_SH2 ? 9 : 4;
    o->loss = loss; o->vs = voxel_size; o->vs_inv = 1.f / voxel_size; /* Optimizer.cpp:23 */
    memcpy(o->K, K, sizeof(o->K));
    o->lambda = lambda; o->lambda_sq = lambda * lambda;
    o->reg_rho = reg_rho; o->reg_n = reg_n; o->reg_l = reg_l;
    return o;
}
void or_opt_destroy(or_opt* o) {
    free(o->frame_idx); free(o->poses); free(o->images); free(o->light); free(o->surf); free(o->rowmap);
    free(o->visF); free(o->b); free(o->delta); free(o->Ht); free(o);
}
void or_opt_set_keyframes(or_opt* o, int F, const int* frame_idx, const float* poses16, const float* images, int H,
                          int W) {
    free(o->frame_idx); free(o->poses); free(o->images);
    o->F = F; o->H = H; o->W = W;
    o->frame_idx = malloc(sizeof(int) * F); memcpy(o->frame_idx, frame_idx, sizeof(int) * F);
    o->poses = malloc(sizeof(float) * 16 * F); memcpy(o->poses, poses16, sizeof(float) * 16 * F);
    size_t n = (size_t)F * H * W * 3;
    o->images = malloc(sizeof(float) * n); memcpy(o->images, images, sizeof(float) * n);
}
static inline void pose_R(const or_opt* o, int f, float* R) {
    const float* P = o->poses + 16 * f;
    R[0] = P[0]; R[1] = P[1]; R[2] = P[2]; R[3] = P[4]; R[4] = P[5]; R[5] = P[6]; R[6] = P[8]; R[7] = P[9]; R[8] = P[10];
}
static inline void pose_t(const or_opt* o, int f, float* t) {
    const float* P = o->poses + 16 * f;
    t[0] = P[3]; t[1] = P[7]; t[2] = P[11];
}

/* PsOptimizerJa.cpp:17-28 */
static void sh_eval(const float* n, int order, float* sh) {
    sh[0] = 1.f; sh[1] = n[0]; sh[2] = n[1]; sh[3] = n[2];
    if (order == 2) {
        sh[4] = n[0] * n[1]; sh[5] = n[0] * n[2]; sh[6] = n[1] * n[2];
        sh[7] = n[0] * n[0] - n[1] * n[1]; sh[8] = n[0] * n[0] - n[2] * n[2];
    }
}

/* Optimizer::select_vis (Optimizer.cpp:30-47) + getSurfaceVoxel (OptimizerAux.cpp:237-257) */
static void select_surface(or_opt* o, int reindex_vis) {
    or_volume* v = o->vol;
    const size_t N = (size_t)v->N;
    if (reindex_vis) {
        free(o->visF);
        o->visF = calloc(N, 8);
        for (size_t lin = 0; lin < N; ++lin) {
            const int len = v->vis_len[lin];
            uint64_t out = 0; int pos = 0;
            for (int k = 0; k < o->F; ++k) {
                const int fi = o->frame_idx[k];
                if ((size_t)fi < (size_t)len) { /* unsigned compare, as frame_idx_[frame] < vis.size() */
                    if ((v->vis[(size_t)(fi >> 6) * N + lin] >> (fi & 63)) & 1) out |= (uint64_t)1 << pos;
                    ++pos;
                }
            }
            o->visF[lin] = out;
        }
    }
    free(o->surf); free(o->rowmap);
    o->rowmap = malloc(sizeof(int) * N);
    int S = 0;
    const double thr = sqrt(3.0) * (double)o->vs;
    for (size_t lin = 0; lin < N; ++lin) {
        const int on = (fabs((double)v->dist[lin]) <= thr) && (o->visF[lin] != 0);
        o->rowmap[lin] = on ? S++ : -1;
    }
    o->S = S;
    o->surf = malloc(sizeof(int) * (S > 0 ? S : 1));
    for (size_t lin = 0; lin < N; ++lin) if (o->rowmap[lin] >= 0) o->surf[o->rowmap[lin]] = (int)lin;
}

/* ---- per-voxel / per-pair arithmetic ---- */
typedef struct { float dist, g[3], rho[3]; int idx[3]; size_t lin; } voxel;

static inline voxel load_voxel(const or_opt* o, size_t lin) {
    const or_volume* v = o->vol;
    voxel x;
    x.lin = lin; x.dist = v->dist[lin]; x.g[0] = v->gx[lin]; x.g[1] = v->gy[lin]; x.g[2] = v->gz[lin];
    x.rho[0] = v->r[lin]; x.rho[1] = v->g[lin]; x.rho[2] = v->b[lin];
    const int WH = v->dims[0] * v->dims[1];
    int rest = (int)lin; /* VoxelGrid.h:88-97 */
    x.idx[2] = rest / WH; rest -= x.idx[2] * WH; x.idx[1] = rest / v->dims[0]; rest -= x.idx[1] * v->dims[0]; x.idx[0] = rest;
    return x;
}

/* Optimizer::computeDistGrad (Optimizer.cpp:287-364) with ifValidDirection (:462-474): forward neighbour if it is
 * a surface voxel, otherwise the backward neighbour unchecked.  Neighbours are addressed by linear offset exactly as
 * idx2line does (so an index one past the edge wraps like the reference).  Reads outside the array, which are
 * undefined behaviour in the reference, return the centre value here. */
static void dist_grad(const or_opt* o, const voxel* x, float* grad, float* pos, long long* nbr_lin) {
    const or_volume* v = o->vol;
    const long long stride[3] = { 1, v->dims[0], (long long)v->dims[0] * v->dims[1] };
    float n[3];
    for (int a = 0; a < 3; ++a) {
        long long fwd = (long long)x->lin + stride[a];
        int valid = (x->idx[a] + 1 <= v->dims[a]) && fwd < v->N && o->rowmap[fwd] >= 0;
        pos[a] = valid ? 1.f : -1.f;
        long long nl = valid ? fwd : (long long)x->lin - stride[a];
        float dn = (nl >= 0 && nl < v->N) ? v->dist[nl] : x->dist;
        if (nbr_lin) nbr_lin[a] = nl;
        n[a] = pos[a] * (dn - x->dist);
    }
    for (int a = 0; a < 3; ++a) grad[a] = n[a] * o->vs_inv;
}

/* Optimizer::computeDistLaplacian (Optimizer.cpp:368-393) */
static float dist_laplacian(const or_opt* o, const voxel* x) {
    const or_volume* v = o->vol;
    const long long stride[3] = { 1, v->dims[0], (long long)v->dims[0] * v->dims[1] };
    float dd[3];
    for (int a = 0; a < 3; ++a) {
        long long l1 = (long long)x->lin + stride[a], l0 = (long long)x->lin - stride[a];
        float d1 = (l1 >= 0 && l1 < v->N) ? v->dist[l1] : x->dist;
        float d0 = (l0 >= 0 && l0 < v->N) ? v->dist[l0] : x->dist;
        dd[a] = d1 + d0 - 2 * x->dist;
    }
    return (dd[0] + dd[1] + dd[2]) * o->vs_inv * o->vs_inv;
}

/* surface point in camera coordinates: R^T (voxel2world(idx) - d * g_hat - t)  (OptimizerAux.cpp:214-215) */
static inline void cam_point(const or_opt* o, const voxel* x, const float* R, const float* t, float* pt, float* ghat) {
    const or_volume* v = o->vol;
    normalized3(x->g, ghat);
    float q[3];
    for (int a = 0; a < 3; ++a) q[a] = ((v->origin[a] + v->vs * (float)x->idx[a]) - x->dist * ghat[a]) - t[a];
    matT_vec(R, q, pt);
}

static inline const float* texel(const or_opt* o, int f, int row, int col) {
    return o->images + (((size_t)f * o->H + row) * o->W + col) * 3;
}

/* interpolateImage (Auxilary.h:41-61), called as interpolateImage(n=row, m=col) -> RGB */
static void interp_image(const or_opt* o, int f, float rowf, float colf, float* rgb) {
    const int x = (int)floorf(rowf), y = (int)floorf(colf);
    float bgr[3];
    if (x + 1 < o->H && y + 1 < o->W) {
        const double wy1 = (double)y + 1.0 - (double)colf; /* (y+1.0-n) */
        const double wx1 = (double)x + 1.0 - (double)rowf; /* (x+1.0-m) */
        const float fx_ = rowf - (float)x;                 /* (m-x) */
        const float fy_ = colf - (float)y;                 /* (n-y) */
        const double a1 = wy1 * (double)fx_;
        const double a2 = wy1 * wx1;
        const float a3 = fy_ * fx_;
        const double a4 = (double)fy_ * wx1;
        const float *t10 = texel(o, f, x + 1, y), *t00 = texel(o, f, x, y), *t11 = texel(o, f, x + 1, y + 1),
                    *t01 = texel(o, f, x, y + 1);
        for (int c = 0; c < 3; ++c) {
            float s = (float)((double)t10[c] * a1) + (float)((double)t00[c] * a2);
            s = s + t11[c] * a3;
            s = s + (float)((double)t01[c] * a4);
            bgr[c] = s;
        }
    } else {
        const float* t00 = texel(o, f, x, y);
        bgr[0] = t00[0]; bgr[1] = t00[1]; bgr[2] = t00[2];
    }
    rgb[0] = bgr[2]; rgb[1] = bgr[1]; rgb[2] = bgr[0];
}

/* computeImageGradient (Auxilary.h:64-123), called with (row, col); direction 0 = along columns, 1 = along rows.
 * The two reads the reference makes one element past the image at the bottom-right corner cell are clamped. */
static void image_gradient(const or_opt* o, int f, float rowf, float colf, int direction, float* out) {
    const int x = (int)floorf(rowf), y = (int)floorf(colf);
    const int Hh = o->H, Ww = o->W;
    const float w01 = rowf - (float)x, w11 = colf - (float)y;
    const float w00 = (float)(1.0 - (double)w01), w10 = (float)(1.0 - (double)w11);
    float v0[3] = { 0, 0, 0 }, v1[3] = { 0, 0, 0 };
    float wa = 1.f, wb = 0.f; int two = 0;
    if (direction == 0) {
        if (x + 1 < Hh && y + 1 < Ww) {
            for (int c = 0; c < 3; ++c) { v0[c] = texel(o, f, x, y + 1)[c] - texel(o, f, x, y)[c];
                                          v1[c] = texel(o, f, x + 1, y + 1)[c] - texel(o, f, x + 1, y)[c]; }
            wa = w00; wb = w01; two = 1;
        } else if (x + 1 >= Hh) {
            const int yy = (y + 1 < Ww) ? y + 1 : y; /* clamp of the reference's out-of-image read */
            for (int c = 0; c < 3; ++c) v0[c] = texel(o, f, x, yy)[c] - texel(o, f, x, y)[c];
        } else {
            for (int c = 0; c < 3; ++c) { v0[c] = -texel(o, f, x, y - 1)[c] + texel(o, f, x, y)[c];
                                          v1[c] = -texel(o, f, x + 1, y - 1)[c] + texel(o, f, x + 1, y)[c]; }
            wa = w00; wb = w01; two = 1;
        }
    } else {
        if (x + 1 < Hh && y + 1 < Ww) {
            for (int c = 0; c < 3; ++c) { v0[c] = texel(o, f, x + 1, y)[c] - texel(o, f, x, y)[c];
                                          v1[c] = texel(o, f, x + 1, y + 1)[c] - texel(o, f, x, y + 1)[c]; }
            wa = w10; wb = w11; two = 1;
        } else if (x + 1 >= Hh && y + 1 < Ww) {
            for (int c = 0; c < 3; ++c) { v0[c] = -texel(o, f, x - 1, y)[c] + texel(o, f, x, y)[c];
                                          v1[c] = -texel(o, f, x - 1, y + 1)[c] + texel(o, f, x, y + 1)[c]; }
            wa = w10; wb = w11; two = 1;
        } else {
            const int xx = (x + 1 < Hh) ? x + 1 : x; /* clamp of the reference's out-of-image read */
            for (int c = 0; c < 3; ++c) v0[c] = texel(o, f, xx, y)[c] - texel(o, f, x, y)[c];
        }
    }
    /* BGR -> RGB */
    if (two) for (int c = 0; c < 3; ++c) out[c] = wa * v0[2 - c] + wb * v1[2 - c];
    else for (int c = 0; c < 3; ++c) out[c] = v0[2 - c];
}

/* test probe: interpolateImage + computeImageGradient of keyframe f at n given (row, col) positions inside the image */
void or_opt_image_probe(const or_opt* o, int f, int n, const float* rows, const float* cols, float* rgb3, float* g0, float* g1) {
    for (int i = 0; i < n; ++i) {
        if (rgb3) interp_image(o, f, rows[i], cols[i], rgb3 + 3 * i);
        image_gradient(o, f, rows[i], cols[i], 0, g0 + 3 * i);
        image_gradient(o, f, rows[i], cols[i], 1, g1 + 3 * i);
    }
}

/* Optimizer::getIntensity (OptimizerAux.cpp:207-231) */
static int get_intensity(const or_opt* o, const voxel* x, int f, const float* R, const float* t, float* rgb,
                         float* m_out, float* n_out) {
    float pt[3], gh[3];
    cam_point(o, x, R, t, pt, gh);
    const float fx = o->K[0], fy = o->K[4], cx = o->K[2], cy = o->K[5];
    const float z_inv = (float)(1.0 / (double)pt[2]);
    const float m = fx * pt[0] * z_inv + cx;
    const float n = fy * pt[1] * z_inv + cy;
    if (m_out) { *m_out = m; *n_out = n; }
    if (m < 0 || m >= (float)o->W || n < 0 || n >= (float)o->H) return 0;
    interp_image(o, f, n, m, rgb);
    return 1;
}

/* {Ps,Led}Optimizer::renderedIntensity (PsOptimizerJa.cpp:30-40, LedOptimizerJa.cpp:15-29) */
static void rendered(const or_opt* o, const voxel* x, int f, const float* R, const float* t, float* out) {
    float g[3], pos[3], n[3];
    dist_grad(o, x, g, pos, NULL);
    normalized3(g, n);
    float irr;
    if (o->model == OR_LED) {
        float pt[3], gh[3], Rp[3];
        cam_point(o, x, R, t, pt, gh);
        mat_vec(R, pt, Rp);
        /* -n^T * R * point: (-n^T R) is evaluated first as a row vector, then dotted with point */
        float nR[3];
        for (int a = 0; a < 3; ++a) nR[a] = sum3((-n[0]) * R[0 + a], (-n[1]) * R[3 + a], (-n[2]) * R[6 + a]);
        (void)Rp;
        irr = dot3(nR, pt);
        const float ld = (float)powd(norm3(pt), 3);
        irr /= ld;
        for (int c = 0; c < 3; ++c) out[c] = x->rho[c] * o->light[c] * irr;
    } else {
        float sh[9];
        sh_eval(n, o->order, sh);
        irr = dotn(o->light + (size_t)f * o->nb, sh, o->nb);
        for (int c = 0; c < 3; ++c) out[c] = x->rho[c] * irr;
    }
}

/* Optimizer::computeWeight / computeLoss (Optimizer.cpp:140-186), per 3-vector */
static void irls_weight(const or_opt* o, const float* r, float* w) {
    for (int c = 0; c < 3; ++c) {
        const float q = r[c] / o->lambda;
        switch (o->loss) {
            case OR_CAUCHY: w[c] = 1.f / (1.f + q * q); break;
            case OR_TUKEY: { float u = 1.f - q * q; w[c] = (r[c] * r[c] < o->lambda_sq) ? u * u : 0.f; } break;
            case OR_HUBER: w[c] = (r[c] * r[c] < o->lambda_sq) ? 1.f : o->lambda * fabsf(1.f / r[c]); break;
            case OR_TRUNC_L2: w[c] = (r[c] * r[c] < o->lambda_sq) ? 1.f : 0.f; break;
            default: w[c] = 1.f;
        }
    }
}
static float robust_loss(const or_opt* o, const float* r) {
    float t[3];
    for (int c = 0; c < 3; ++c) {
        const float q = r[c] / o->lambda;
        switch (o->loss) {
            case OR_CAUCHY: t[c] = logf(1.f + q * q); break;
            case OR_TUKEY: { float u = 1.f - q * q; t[c] = (r[c] * r[c] < o->lambda_sq) ? 1.f - u * u * u : 1.f; } break;
            case OR_HUBER: t[c] = (r[c] * r[c] < o->lambda_sq) ? .5f * (r[c] * r[c])
                                                                : o->lambda * (fabsf(r[c]) - 0.5f * o->lambda * 1.f); break;
            case OR_TRUNC_L2: { float cl = fminf(fmaxf(r[c], -o->lambda), o->lambda); t[c] = cl * cl; } break;
            default: t[c] = r[c] * r[c];
        }
    }
    return (t[0] + t[1]) + t[2];
}

/* ------------------------------------------------------------------ init */
/* LedOptimizer::computeLightIntensive (LedOptimizer.cpp:76-112) */
static void led_light_init(or_opt* o) {
    float I[3] = { 0, 0, 0 }, Rd[3] = { 0, 0, 0 };
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        const uint64_t vis = o->visF[x.lin];
        for (int f = 0; f < o->F; ++f) {
            if (!((vis >> f) & 1)) continue;
            float R[9], t[3], in[3], re[3];
            pose_R(o, f, R); pose_t(o, f, t);
            if (!get_intensity(o, &x, f, R, t, in, NULL, NULL)) continue;
            rendered(o, &x, f, R, t, re);
            for (int c = 0; c < 3; ++c) { I[c] += in[c]; Rd[c] += re[c]; }
        }
    }
    for (int c = 0; c < 3; ++c) o->light[c] = I[c] / Rd[c];
}

/* PsOptimizer::init (PsOptimizer.cpp:25-42), LedOptimizer::init (LedOptimizer.cpp:25-36) */
void or_opt_init(or_opt* o) {
    free(o->light);
    if (o->model == OR_LED) {
        o->light = malloc(sizeof(float) * 3);
        select_surface(o, 1);
        o->light[0] = o->light[1] = o->light[2] = 1.f;
        led_light_init(o);
    } else {
        o->light = malloc(sizeof(float) * o->nb * (o->F > 0 ? o->F : 1));
        const float s[3] = { 0.f, 0.f, -1.f };
        for (int f = 0; f < o->F; ++f) {
            float R[9], d[3];
            pose_R(o, f, R);
            mat_vec(R, s, d);
            sh_eval(d, o->order, o->light + (size_t)f * o->nb);
            o->light[(size_t)f * o->nb] = 0.02f;
        }
        select_surface(o, 1);
    }
}

int or_opt_num_surface(const or_opt* o) { return o->S; }
void or_opt_surface(const or_opt* o, int* out) { memcpy(out, o->surf, sizeof(int) * o->S); }
void or_opt_vis_keyframes(const or_opt* o, uint64_t* out) { for (int s = 0; s < o->S; ++s) out[s] = o->visF[o->surf[s]]; }
int or_opt_light_len(const or_opt* o) { return o->model == OR_LED ? 3 : o->F * o->nb; }
void or_opt_get_light(const or_opt* o, float* out) { memcpy(out, o->light, sizeof(float) * or_opt_light_len(o)); }
void or_opt_set_light(or_opt* o, const float* in) { memcpy(o->light, in, sizeof(float) * or_opt_light_len(o)); }
void or_opt_get_poses(const or_opt* o, float* out) { memcpy(out, o->poses, sizeof(float) * 16 * o->F); }
void or_opt_set_poses(or_opt* o, const float* in) { memcpy(o->poses, in, sizeof(float) * 16 * o->F); }
void or_opt_get_reg(const or_opt* o, float out[3]) { out[0] = o->reg_rho; out[1] = o->reg_n; out[2] = o->reg_l; }
void or_opt_set_reg(or_opt* o, const float in[3]) { o->reg_rho = in[0]; o->reg_n = in[1]; o->reg_l = in[2]; }
float or_opt_voxel_size(const or_opt* o) { return o->vs; }

/* Optimizer::initAlbedo (Optimizer.cpp:50-81) */
void or_opt_init_albedo(or_opt* o) {
    or_volume* v = o->vol;
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        const uint64_t vis = o->visF[x.lin];
        int count = 0; float rho[3] = { 0, 0, 0 };
        for (int f = 0; f < o->F; ++f) {
            if (!((vis >> f) & 1)) continue;
            float R[9], t[3], in[3];
            pose_R(o, f, R); pose_t(o, f, t);
            if (!get_intensity(o, &x, f, R, t, in, NULL, NULL)) continue;
            for (int c = 0; c < 3; ++c) rho[c] += in[c];
            count++;
        }
        if (count) {
            v->r[x.lin] = rho[0] / (float)count; v->g[x.lin] = rho[1] / (float)count; v->b[x.lin] = rho[2] / (float)count;
        }
    }
}

/* getPSEnergy (PsOptimizer.cpp:47-78, LedOptimizer.cpp:40-71) */
float or_opt_energy(or_opt* o, double* E64, long long* n_active) {
    float E = 0.f; double acc = 0.0; long long cnt = 0;
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        const uint64_t vis = o->visF[x.lin];
        for (int f = 0; f < o->F; ++f) {
            if (!((vis >> f) & 1)) continue;
            float R[9], t[3], in[3], re[3], r[3];
            pose_R(o, f, R); pose_t(o, f, t);
            if (!get_intensity(o, &x, f, R, t, in, NULL, NULL)) continue;
            rendered(o, &x, f, R, t, re);
            for (int c = 0; c < 3; ++c) r[c] = in[c] - re[c];
            const float l = robust_loss(o, r);
            E += l; acc += (double)l; ++cnt;
        }
    }
    if (E64) *E64 = acc / (double)o->S;
    if (n_active) *n_active = cnt;
    return E / (float)o->S;
}
/* getNormalEnergy (Optimizer.cpp:86-103) */
float or_opt_normal_energy(or_opt* o, double* E64) {
    float E = 0.f; double acc = 0.0;
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        float g[3], pos[3];
        dist_grad(o, &x, g, pos, NULL);
        const float e = norm3(g) - 1;
        E += e * e; acc += (double)(e * e);
    }
    if (E64) *E64 = acc / (double)o->S;
    return E / (float)o->S;
}
/* getLaplacianEnergy (Optimizer.cpp:106-119) */
float or_opt_laplacian_energy(or_opt* o, double* E64) {
    float E = 0.f; double acc = 0.0;
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        const float e = dist_laplacian(o, &x);
        E += e * e; acc += (double)(e * e);
    }
    if (E64) *E64 = acc / (double)o->S;
    return E / (float)o->S;
}

/* Optimizer::computeAlbedoGrad (Optimizer.cpp:396-460): 3x3 rows r,g,b; pos */
static void albedo_grad(const or_opt* o, const voxel* x, float J[9], float* pos, long long* nbr_lin) {
    const or_volume* v = o->vol;
    const long long stride[3] = { 1, v->dims[0], (long long)v->dims[0] * v->dims[1] };
    for (int a = 0; a < 3; ++a) {
        long long fwd = (long long)x->lin + stride[a];
        int valid = (x->idx[a] + 1 <= v->dims[a]) && fwd < v->N && o->rowmap[fwd] >= 0;
        pos[a] = valid ? 1.f : -1.f;
        long long nl = valid ? fwd : (long long)x->lin - stride[a];
        int ok = nl >= 0 && nl < v->N;
        if (nbr_lin) nbr_lin[a] = nl;
        float nr = ok ? v->r[nl] : x->rho[0], ng = ok ? v->g[nl] : x->rho[1], nbb = ok ? v->b[nl] : x->rho[2];
        J[0 + a] = (pos[a] * (nr - x->rho[0])) * o->vs_inv;
        J[3 + a] = (pos[a] * (ng - x->rho[1])) * o->vs_inv;
        J[6 + a] = (pos[a] * (nbb - x->rho[2])) * o->vs_inv;
    }
}
/* getAlbedoRegEnergy (Optimizer.cpp:122-136) */
float or_opt_albedo_reg_energy(or_opt* o, double* E64) {
    float E = 0.f; double acc = 0.0;
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        float J[9], pos[3];
        albedo_grad(o, &x, J, pos, NULL);
        const float e = (norm3(J) + norm3(J + 3)) + norm3(J + 6);
        E += e; acc += (double)e;
    }
    if (E64) *E64 = acc / (double)o->S;
    return E / (float)o->S;
}

void or_opt_probe_pairs(or_opt* o, float* m, float* n, unsigned char* ok, float* I3, float* rend3) {
    const int F = o->F;
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        const uint64_t vis = o->visF[x.lin];
        for (int f = 0; f < F; ++f) {
            const size_t k = (size_t)s * F + f;
            m[k] = n[k] = 0.f; ok[k] = 0;
            for (int c = 0; c < 3; ++c) I3[3 * k + c] = rend3[3 * k + c] = 0.f;
            if (!((vis >> f) & 1)) continue;
            float R[9], t[3], in[3], re[3];
            pose_R(o, f, R); pose_t(o, f, t);
            const int good = get_intensity(o, &x, f, R, t, in, &m[k], &n[k]);
            if (!good) continue;
            ok[k] = 1;
            rendered(o, &x, f, R, t, re);
            for (int c = 0; c < 3; ++c) { I3[3 * k + c] = in[c]; rend3[3 * k + c] = re[c]; }
        }
    }
}

/* computeResidual() (PsOptimizerJa.cpp:567-626, LedOptimizerJa.cpp:488-542); layout here S x F x 3 */
void or_opt_residual(or_opt* o, float* r3, float* w3) {
    const int F = o->F;
    memset(r3, 0, sizeof(float) * 3 * (size_t)o->S * F);
    memset(w3, 0, sizeof(float) * 3 * (size_t)o->S * F);
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        const uint64_t vis = o->visF[x.lin];
        for (int f = 0; f < F; ++f) {
            if (!((vis >> f) & 1)) continue;
            float R[9], t[3], in[3], re[3];
            pose_R(o, f, R); pose_t(o, f, t);
            if (!get_intensity(o, &x, f, R, t, in, NULL, NULL)) continue;
            rendered(o, &x, f, R, t, re);
            float* r = r3 + 3 * ((size_t)s * F + f);
            for (int c = 0; c < 3; ++c) r[c] = in[c] - re[c];
            irls_weight(o, r, w3 + 3 * ((size_t)s * F + f));
        }
    }
}
void or_opt_dist_grad(or_opt* o, float* out6) {
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        dist_grad(o, &x, out6 + 6 * s, out6 + 6 * s + 3, NULL);
    }
}

/* ------------------------------------------------------------------ Jacobians */
/* shared front part of poseJacobian / distJacobian (PsOptimizerJa.cpp:61-90, 160-192): division-form projection,
 * bounds test, image gradient, projection derivative, A = image_grad(3x2) * pi_grad(2x3). Returns 0 if out of image. */
static int image_pi_grad(const or_opt* o, const voxel* x, int f, const float* R, const float* t, float* pt, float* gh,
                         float A[9]) {
    cam_point(o, x, R, t, pt, gh);
    const float fx = o->K[0], fy = o->K[4], cx = o->K[2], cy = o->K[5];
    const float m = fx * pt[0] / pt[2] + cx;
    const float n = fy * pt[1] / pt[2] + cy;
    if (m < 0 || m >= (float)o->W || n < 0 || n >= (float)o->H) return 0;
    float g0[3], g1[3];
    image_gradient(o, f, n, m, 0, g0);
    image_gradient(o, f, n, m, 1, g1);
    const float z_inv = (float)(1.0 / (double)pt[2]);
    const float z_inv_sq = z_inv * z_inv;
    float pi[6] = { fx * z_inv, 0.f, -fx * pt[0] * z_inv_sq, 0.f, fy * z_inv, -fy * pt[1] * z_inv_sq };
    for (int c = 0; c < 3; ++c) for (int j = 0; j < 3; ++j) A[3 * c + j] = g0[c] * pi[j] + g1[c] * pi[3 + j];
    return 1;
}
static inline void skew3(const float* v, float* S) {
    S[0] = 0; S[1] = -v[2]; S[2] = v[1]; S[3] = v[2]; S[4] = 0; S[5] = -v[0]; S[6] = -v[1]; S[7] = v[0]; S[8] = 0;
}

/* {Ps,Led}Optimizer::poseJacobian (PsOptimizerJa.cpp:61-115, LedOptimizerJa.cpp:32-81); J row-major 3x6 */
static int pose_jacobian(const or_opt* o, const voxel* x, int f, const float* R, const float* t, float J[18]) {
    float pt[3], gh[3], A[9];
    if (!image_pi_grad(o, x, f, R, t, pt, gh, A)) return 0;
    float Sk[9];
    skew3(pt, Sk);
    for (int c = 0; c < 3; ++c) for (int j = 0; j < 3; ++j) {
        /* -A * R^T */
        J[6 * c + j] = sum3((-A[3 * c]) * R[3 * j], (-A[3 * c + 1]) * R[3 * j + 1], (-A[3 * c + 2]) * R[3 * j + 2]);
        /* A * skew(point) */
        J[6 * c + 3 + j] = sum3(A[3 * c] * Sk[j], A[3 * c + 1] * Sk[3 + j], A[3 * c + 2] * Sk[6 + j]);
    }
    if (o->model == OR_LED) {
        const float l = (float)powd(norm3(pt), 3);
        const float dl = (float)powd(norm3(pt), 5);
        float skp[3]; /* (skew(point)*point) */
        for (int a = 0; a < 3; ++a) skp[a] = sum3(Sk[3 * a] * pt[0], Sk[3 * a + 1] * pt[1], Sk[3 * a + 2] * pt[2]);
        float nR[3];
        for (int a = 0; a < 3; ++a) nR[a] = sum3(gh[0] * R[0 + a], gh[1] * R[3 + a], gh[2] * R[6 + a]);
        const float nRp = dot3(nR, pt);
        for (int c = 0; c < 3; ++c) {
            const float kt = -(x->rho[c] * o->light[c] / l);
            const float kr = -3 * x->rho[c] * o->light[c] / dl * nRp;
            for (int j = 0; j < 3; ++j) {
                J[6 * c + j] = J[6 * c + j] + kt * gh[j];
                J[6 * c + 3 + j] = J[6 * c + 3 + j] + kr * skp[j];
            }
        }
    }
    return 1;
}

/* Optimizer::normalJacobian(grad, direction, lag=false) (Optimizer.cpp:269-284) */
static void normal_jacobian(const or_opt* o, const float* grad, const float* direction, float* Jn) {
    float nd[3];
    for (int a = 0; a < 3; ++a) nd[a] = -o->vs_inv * direction[a];
    const float N_inv = (float)(1.0 / (double)fmaxf(norm3(grad), 0.001f));
    const float dN = (float)(powd(N_inv, 3) * (double)dot3(nd, grad));
    for (int a = 0; a < 3; ++a) Jn[a] = N_inv * nd[a] - dN * grad[a];
}

/* {Ps,Led}Optimizer::distJacobian (PsOptimizerJa.cpp:160-289, LedOptimizerJa.cpp:117-218).
 * Jd[k][c], k = 0 centre, 1..3 the pos-neighbours; pos returned for the column lookup. */
static int dist_jacobian(const or_opt* o, const voxel* x, int f, const float* R, const float* t, float Jd[4][3],
                         float* pos, long long* nbr_lin) {
    float pt[3], gh[3], A[9];
    if (!image_pi_grad(o, x, f, R, t, pt, gh, A)) return 0;
    float grad[3];
    dist_grad(o, x, grad, pos, nbr_lin);
    float dn[4][3], dir[3];
    normal_jacobian(o, grad, pos, dn[0]);
    const float sgn = (o->model == OR_LED) ? 1.f : -1.f; /* LedOptimizerJa.cpp:158 "+=" vs PsOptimizerJa.cpp:201 "-=" */
    for (int k = 0; k < 3; ++k) {
        dir[0] = dir[1] = dir[2] = 0.f;
        dir[k] = 0.f + sgn * pos[k];
        if (sgn < 0) dir[k] = 0.f - pos[k];
        normal_jacobian(o, grad, dir, dn[k + 1]);
    }
    float dx[4][3];
    for (int a = 0; a < 3; ++a) dx[0][a] = -gh[a] - x->dist * dn[0][a];
    for (int k = 1; k < 4; ++k) for (int a = 0; a < 3; ++a) dx[k][a] = -x->dist * dn[k][a];
    /* M = A * R^T, then dI_k = M * dx_k */
    float M[9], dI[4][3];
    for (int c = 0; c < 3; ++c) for (int j = 0; j < 3; ++j)
        M[3 * c + j] = sum3(A[3 * c] * R[3 * j], A[3 * c + 1] * R[3 * j + 1], A[3 * c + 2] * R[3 * j + 2]);
    for (int k = 0; k < 4; ++k) mat_vec(M, dx[k], dI[k]);

    if (o->model == OR_LED) {
        float gn[3];
        normalized3(grad, gn);
        float Rp[3];
        mat_vec(R, pt, Rp);
        const float radius = (float)powd(norm3(pt), 3);
        const double r5 = powd(norm3(pt), 5);
        /* grad.normalized()^T * R * point : row vector (gn^T R) first */
        float gR[3];
        for (int a = 0; a < 3; ++a) gR[a] = sum3(gn[0] * R[0 + a], gn[1] * R[3 + a], gn[2] * R[6 + a]);
        const float gRp = dot3(gR, pt);
        /* point^T * R^T */
        float pRt[3];
        for (int a = 0; a < 3; ++a) pRt[a] = sum3(pt[0] * R[3 * a], pt[1] * R[3 * a + 1], pt[2] * R[3 * a + 2]);
        for (int k = 0; k < 4; ++k) {
            float dR_[3];
            for (int a = 0; a < 3; ++a) dR_[a] = sum3(dn[k][0] * R[0 + a], dn[k][1] * R[3 + a], dn[k][2] * R[6 + a]);
            float dm = dot3(dR_, pt) + dot3(gn, dx[k]);
            const float dm2 = (float)((double)(-3 * dot3(pRt, dx[k])) / r5);
            dm = dm / radius + dm2 * gRp;
            for (int c = 0; c < 3; ++c) Jd[k][c] = dI[k][c] + (x->rho[c] * o->light[c]) * dm;
        }
    } else if (o->order == 1) {
        const float* l = o->light + (size_t)f * o->nb;
        for (int k = 0; k < 4; ++k) for (int c = 0; c < 3; ++c) {
            const float d0 = x->rho[c] * l[1], d1 = x->rho[c] * l[2], d2 = x->rho[c] * l[3];
            Jd[k][c] = dI[k][c] - sum3(d0 * dn[k][0], d1 * dn[k][1], d2 * dn[k][2]);
        }
    } else {
        const float* l = o->light + (size_t)f * o->nb;
        float n[3];
        normalized3(grad, n);
        /* dSH/dn (3 x 9), PsOptimizerJa.cpp:248-250 */
        const float D[27] = { 0, 1, 0, 0, n[1], n[2], 0, 2 * n[0], 2 * n[0],
                              0, 0, 1, 0, n[0], 0, n[2], -2 * n[1], 0,
                              0, 0, 0, 1, 0, n[0], n[1], 0, -2 * n[2] };
        for (int k = 0; k < 4; ++k) {
            float dsh[9];
            for (int i = 0; i < 9; ++i) dsh[i] = sum3(D[i] * dn[k][0], D[9 + i] * dn[k][1], D[18 + i] * dn[k][2]);
            for (int c = 0; c < 3; ++c) {
                float a[9];
                for (int i = 0; i < 9; ++i) a[i] = (x->rho[c] * l[i]) * dsh[i];
                /* fixed-size 9 reduction, fully unrolled: (a0+a1)+(a2+a3) + ((a4+a5) + (a6 + (a7+a8))) */
                const float sdot = ((a[0] + a[1]) + (a[2] + a[3])) + ((a[4] + a[5]) + (a[6] + (a[7] + a[8])));
                Jd[k][c] = dI[k][c] - sdot;
            }
        }
    }
    return 1;
}

/* b += s * (J^T * r) as Eigen evaluates it: the scalar goes into the matrix, (s * J^T) * r, and the product of a row-major
 * sparse expression with a dense vector sums every row with two accumulators, entries alternating between them in ascending
 * column order, res[i] += alpha * (tmp_a + tmp_b) with alpha = 1 (SparseCore/SparseDenseProduct.h:66-80).  Entries must
 * arrive in ascending J-row order per column. */
typedef struct { float* a; float* b; unsigned char* odd; int n; } rowsum2;
static rowsum2 rs2_new(int n) {
    rowsum2 r; r.n = n;
    r.a = calloc(n > 0 ? n : 1, sizeof(float)); r.b = calloc(n > 0 ? n : 1, sizeof(float)); r.odd = calloc(n > 0 ? n : 1, 1);
    return r;
}
static void rs2_add(rowsum2* r, int col, float scaled_j, float res) {
    const float t = scaled_j * res;
    if (r->odd[col]) r->b[col] += t; else r->a[col] += t;
    r->odd[col] ^= 1;
}
static void rs2_finish(rowsum2* r, float* b) {
    for (int i = 0; i < r->n; ++i) b[i] += r->a[i] + r->b[i];
    free(r->a); free(r->b); free(r->odd);
}

/* ------------------------------------------------------------------ sparse accumulator + CG */
typedef struct { int n; int* cnt; int* col; float* val; int cap; } rowacc;
static rowacc* ra_new(int n, int cap) {
    rowacc* a = calloc(1, sizeof(*a));
    a->n = n; a->cap = cap; a->cnt = calloc(n > 0 ? n : 1, sizeof(int));
    a->col = malloc(sizeof(int) * (size_t)(n > 0 ? n : 1) * cap); a->val = malloc(sizeof(float) * (size_t)(n > 0 ? n : 1) * cap);
    return a;
}
static void ra_free(rowacc* a) { free(a->cnt); free(a->col); free(a->val); free(a); }
static float* ra_at(rowacc* a, int i, int j) {
    int* c = a->col + (size_t)i * a->cap; float* v = a->val + (size_t)i * a->cap;
    for (int k = 0; k < a->cnt[i]; ++k) if (c[k] == j) return &v[k];
    if (a->cnt[i] >= a->cap) { fprintf(stderr, "rowacc overflow\n"); abort(); }
    c[a->cnt[i]] = j; v[a->cnt[i]] = 0.f;
    return &v[a->cnt[i]++];
}
static float ra_get(const rowacc* a, int i, int j) {
    const int* c = a->col + (size_t)i * a->cap; const float* v = a->val + (size_t)i * a->cap;
    for (int k = 0; k < a->cnt[i]; ++k) if (c[k] == j) return v[k];
    return 0.f;
}

/* sum_i a[i] * b[i] in the order Eigen's vectorised reduction adds them on the reference's x86-64 build (SSE2 packets of 4
 * floats, no FMA): two packet accumulators over strides of 8, their sum, one more packet if the length allows, the horizontal
 * sum (l0 + l2) + (l1 + l3), then the tail element by element (third/eigen/Eigen/src/Core/Redux.h:236-281, predux in
 * arch/SSE/PacketMath.h).  Vectors shorter than one packet are summed serially. */
static float eigen_dot(const float* a, const float* b, int n) {
    if (n <= 0) return 0.f;
    if (n < 4) {
        float r = a[0] * b[0];
        for (int i = 1; i < n; ++i) r += a[i] * b[i];
        return r;
    }
    const int aligned = n & ~3, aligned2 = n & ~7;
    float p0[4], p1[4];
    for (int l = 0; l < 4; ++l) p0[l] = a[l] * b[l];
    if (aligned > 4) {
        for (int l = 0; l < 4; ++l) p1[l] = a[4 + l] * b[4 + l];
        for (int i = 8; i < aligned2; i += 8) {
            for (int l = 0; l < 4; ++l) p0[l] += a[i + l] * b[i + l];
            for (int l = 0; l < 4; ++l) p1[l] += a[i + 4 + l] * b[i + 4 + l];
        }
        for (int l = 0; l < 4; ++l) p0[l] += p1[l];
        if (aligned > aligned2) for (int l = 0; l < 4; ++l) p0[l] += a[aligned2 + l] * b[aligned2 + l];
    }
    float r = (p0[0] + p0[2]) + (p0[1] + p0[3]);
    for (int i = aligned; i < n; ++i) r += a[i] * b[i];
    return r;
}

/* Eigen::ConjugateGradient<SparseMatrix<float>, Lower, DiagonalPreconditioner> statement by statement (ConjugateGradient.h:
 * 29-98, IterativeSolverBase.h, BasicPreconditioners.h:67-93): tol = epsilon, maxIters = 2n, only the lower triangle is read.
 * The product with the lower selfadjoint view walks the columns of the column-major matrix: diagonal term into res[j], then for
 * the rows i > j below it (ascending) res_j += l * p[i] and res[i] += l * p[j], then res[j] += res_j
 * (SparseCore/SparseSelfAdjointView.h:299-331); the inner products follow eigen_dot. */
static void eigen_like_cg(const rowacc* H, const float* b, float* x, int* success, int* iters_out, float* err_out) {
    const int n = H->n;
    float *res = malloc(4 * (size_t)(n + 1)), *p = malloc(4 * (size_t)(n + 1)), *z = malloc(4 * (size_t)(n + 1)),
          *tmp = malloc(4 * (size_t)(n + 1)), *invd = malloc(4 * (size_t)(n + 1)), *diag = malloc(4 * (size_t)(n + 1));
    unsigned char* has_diag = calloc((size_t)n + 1, 1);
    /* lower triangle by columns, rows ascending within a column (row i's entries (i, j < i) are appended to column j as i grows) */
    int* colptr = calloc((size_t)n + 2, sizeof(int));
    for (int i = 0; i < n; ++i) {
        const int* c = H->col + (size_t)i * H->cap;
        for (int k = 0; k < H->cnt[i]; ++k) if (c[k] < i) ++colptr[c[k] + 1];
    }
    for (int j = 0; j < n; ++j) colptr[j + 1] += colptr[j];
    const int nnz = colptr[n];
    int* rowidx = malloc(sizeof(int) * (size_t)(nnz + 1));
    float* lval = malloc(sizeof(float) * (size_t)(nnz + 1));
    int* fill = malloc(sizeof(int) * (size_t)(n + 1));
    for (int j = 0; j < n; ++j) fill[j] = colptr[j];
    for (int i = 0; i < n; ++i) {
        const int* c = H->col + (size_t)i * H->cap; const float* v = H->val + (size_t)i * H->cap;
        diag[i] = 0.f;
        for (int k = 0; k < H->cnt[i]; ++k) {
            if (c[k] == i) { diag[i] = v[k]; has_diag[i] = 1; }
            else if (c[k] < i) { rowidx[fill[c[k]]] = i; lval[fill[c[k]]++] = v[k]; }
        }
    }
    for (int i = 0; i < n; ++i) { x[i] = 0.f; res[i] = b[i]; invd[i] = (has_diag[i] && diag[i] != 0.f) ? 1.f / diag[i] : 1.f; }
    const float tol = 1.1920929e-07f;
    const int maxIters = 2 * n;
    const float rhsNorm2 = eigen_dot(b, b, n);
    int it = 0; float residualNorm2 = 0.f;
    if (rhsNorm2 == 0.f) { *success = 1; *iters_out = 0; *err_out = 0.f; goto done; }
    {
        float threshold = tol * tol * rhsNorm2;
        if (threshold < 1.17549435e-38f) threshold = 1.17549435e-38f;
        residualNorm2 = eigen_dot(res, res, n);
        if (residualNorm2 < threshold) { *iters_out = 0; *err_out = sqrtf(residualNorm2 / rhsNorm2); *success = 1; goto done; }
        for (int i = 0; i < n; ++i) p[i] = invd[i] * res[i];
        float absNew = eigen_dot(res, p, n);
        while (it < maxIters) {
            /* tmp = selfadjointView<Lower>(H) * p */
            for (int i = 0; i < n; ++i) tmp[i] = 0.f;
            for (int j = 0; j < n; ++j) {
                if (has_diag[j]) tmp[j] += diag[j] * p[j];
                const float rhs_j = p[j];
                float res_j = 0.f;
                for (int k = colptr[j]; k < colptr[j + 1]; ++k) {
                    res_j += lval[k] * p[rowidx[k]];
                    tmp[rowidx[k]] += lval[k] * rhs_j;
                }
                tmp[j] += res_j;
            }
            const float alpha = absNew / eigen_dot(p, tmp, n);
            for (int i = 0; i < n; ++i) { x[i] += alpha * p[i]; res[i] -= alpha * tmp[i]; }
            residualNorm2 = eigen_dot(res, res, n);
            if (residualNorm2 < threshold) break;
            for (int i = 0; i < n; ++i) z[i] = invd[i] * res[i];
            const float absOld = absNew;
            absNew = eigen_dot(res, z, n);
            const float beta = absNew / absOld;
            for (int i = 0; i < n; ++i) p[i] = z[i] + beta * p[i];
            ++it;
        }
        *err_out = sqrtf(residualNorm2 / rhsNorm2);
        *iters_out = it;
        *success = (*err_out <= tol) ? 1 : 0; /* IterativeSolverBase: m_info = m_error <= m_tolerance ? Success : NoConvergence */
    }
done:
    free(res); free(p); free(z); free(tmp); free(invd); free(diag); free(has_diag); free(colptr); free(rowidx); free(lval); free(fill);
}

/* ------------------------------------------------------------------ updates */
/* Optimizer::updateAlbedo(VectorXf&) (OptimizerAux.cpp:120-150) */
static void update_albedo(or_opt* o, const float* d) {
    or_volume* v = o->vol;
    for (int s = 0; s < o->S; ++s) {
        const size_t lin = (size_t)o->surf[s];
        const float r = v->r[lin] - d[3 * s], g = v->g[lin] - d[3 * s + 1], b = v->b[lin] - d[3 * s + 2];
        if ((double)r > 0.0 && (double)r < 1.0) v->r[lin] = r;
        if ((double)g > 0.0 && (double)g < 1.0) v->g[lin] = g;
        if ((double)b > 0.0 && (double)b < 1.0) v->b[lin] = b;
    }
}
/* Optimizer::updateDist + updateGrad (OptimizerAux.cpp:152-188) */
static void update_dist(or_opt* o, const float* d) {
    or_volume* v = o->vol;
    const double thr = sqrt(3.0) * (double)o->vs;
    for (int s = 0; s < o->S; ++s) {
        const size_t lin = (size_t)o->surf[s];
        if (fabs((double)d[s]) < thr) v->dist[lin] -= d[s];
    }
    for (int s = 0; s < o->S; ++s) {
        voxel x = load_voxel(o, (size_t)o->surf[s]);
        float g[3], pos[3];
        dist_grad(o, &x, g, pos, NULL);
        v->gx[x.lin] = g[0]; v->gy[x.lin] = g[1]; v->gz[x.lin] = g[2];
    }
}
/* Sophus SO3::exp (third/Sophus/sophus/so3.hpp:583-619) + Eigen Quaternion::toRotationMatrix */
static void so3_exp_matrix(const float* w, float* Rm) {
    const float theta_sq = sqnorm3(w);
    float imag, real;
    if (theta_sq < 1e-10f * 1e-10f) { /* Constants<float>::epsilon() = 1e-5 -> squared */
        const float t4 = theta_sq * theta_sq;
        imag = 0.5f - (float)(1.0 / 48.0) * theta_sq + (float)(1.0 / 3840.0) * t4;
        real = 1.f - (float)(1.0 / 8.0) * theta_sq + (float)(1.0 / 384.0) * t4;
    } else {
        const float theta = sqrtf(theta_sq);
        const float half = 0.5f * theta;
        imag = sinf(half) / theta;
        real = cosf(half);
    }
    const float qw = real, qx = imag * w[0], qy = imag * w[1], qz = imag * w[2];
    const float tx = 2.f * qx, ty = 2.f * qy, tz = 2.f * qz;
    const float twx = tx * qw, twy = ty * qw, twz = tz * qw, txx = tx * qx, txy = ty * qx, txz = tz * qx;
    const float tyy = ty * qy, tyz = tz * qy, tzz = tz * qz;
    Rm[0] = 1.f - (tyy + tzz); Rm[1] = txy - twz; Rm[2] = txz + twy;
    Rm[3] = txy + twz; Rm[4] = 1.f - (txx + tzz); Rm[5] = tyz - twx;
    Rm[6] = txz - twy; Rm[7] = tyz + twx; Rm[8] = 1.f - (txx + tyy);
}
/* Optimizer::updatePose (OptimizerAux.cpp:190-205) */
static void update_pose(or_opt* o, const float* d) {
    for (int f = 0; f < o->F; ++f) {
        float R[9], t[3], E[9], w[3] = { -d[6 * f + 3], -d[6 * f + 4], -d[6 * f + 5] };
        pose_R(o, f, R); pose_t(o, f, t);
        so3_exp_matrix(w, E);
        float* P = o->poses + 16 * f;
        P[3] = t[0] - d[6 * f]; P[7] = t[1] - d[6 * f + 1]; P[11] = t[2] - d[6 * f + 2];
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b)
            P[4 * a + b] = sum3(R[3 * a] * E[b], R[3 * a + 1] * E[3 + b], R[3 * a + 2] * E[6 + b]);
    }
}

void or_opt_set_accum64(or_opt* o, int on) { o->accum64 = on; }

/* ------------------------------------------------------------------ block step */
static void H_push(or_opt* o, int i, int j, float v) {
    if (o->Hn >= o->Hcap) { o->Hcap = o->Hcap ? 2 * o->Hcap : 1024; o->Ht = realloc(o->Ht, sizeof(trip) * (size_t)o->Hcap); }
    o->Ht[o->Hn].i = i; o->Ht[o->Hn].j = j; o->Ht[o->Hn].v = v; o->Hn++;
}

int or_opt_block(or_opt* o, int which, int apply, float damping, int albedo_reg, int normal_reg, int laplacian_reg) {
    or_volume* v = o->vol;
    const int S = o->S, F = o->F, nb = o->nb;
    int n;
    if (which == OR_ALBEDO) n = 3 * S;
    else if (which == OR_LIGHT) n = o->model == OR_LED ? 3 : nb * F;
    else if (which == OR_DIST) n = S;
    else n = 6 * F;
    rowacc* H = ra_new(n, which == OR_DIST ? 48 : (which == OR_ALBEDO ? (albedo_reg ? 48 : 2) : (which == OR_LIGHT ? (nb > 6 ? nb : 6) : 6)));
    float* b = calloc(n > 0 ? n : 1, sizeof(float));
    /* accum64: the same products, summed in double, replace the float sums of the light / pose blocks afterwards */
    const int bs64 = which == OR_POSE ? 6 : (which == OR_LIGHT ? (o->model == OR_LED ? 1 : nb) : 0);
    double* H64 = (o->accum64 && bs64) ? calloc((size_t)n * bs64, sizeof(double)) : NULL;
    double* b64 = (o->accum64 && bs64) ? calloc((size_t)n, sizeof(double)) : NULL;

    /* J^T W J and J^T W r accumulated in the reference's row order: frame-major, then surface row, then channel */
    for (int f = 0; f < F; ++f) {
        float R[9], t[3];
        pose_R(o, f, R); pose_t(o, f, t);
        for (int s = 0; s < S; ++s) {
            voxel x = load_voxel(o, (size_t)o->surf[s]);
            if (!((o->visF[x.lin] >> f) & 1)) continue;
            float in[3], re[3], r[3], w[3];
            if (!get_intensity(o, &x, f, R, t, in, NULL, NULL)) continue; /* W = 0 for this row */
            rendered(o, &x, f, R, t, re);
            for (int c = 0; c < 3; ++c) r[c] = in[c] - re[c];
            irls_weight(o, r, w);
            if (which == OR_ALBEDO) {
                float gh[3], Jr[3];
                normalized3(x.g, gh);
                if (o->model == OR_LED) { /* LedOptimizerJa.cpp:85-99 */
                    float pt[3], g2[3];
                    cam_point(o, &x, R, t, pt, g2);
                    float nR[3];
                    for (int a = 0; a < 3; ++a) nR[a] = sum3(gh[0] * R[0 + a], gh[1] * R[3 + a], gh[2] * R[6 + a]);
                    float refl = dot3(nR, pt);
                    refl /= (float)powd(norm3(pt), 3);
                    for (int c = 0; c < 3; ++c) Jr[c] = refl * o->light[c];
                } else { /* PsOptimizerJa.cpp:118-122 */
                    float sh[9];
                    sh_eval(gh, o->order, sh);
                    const float j = -dotn(o->light + (size_t)f * nb, sh, nb);
                    Jr[0] = Jr[1] = Jr[2] = j;
                }
                for (int c = 0; c < 3; ++c) {
                    *ra_at(H, 3 * s + c, 3 * s + c) += (Jr[c] * w[c]) * Jr[c];
                    b[3 * s + c] += (Jr[c] * w[c]) * r[c];
                }
            } else if (which == OR_LIGHT) {
                float gh[3];
                normalized3(x.g, gh);
                if (o->model == OR_LED) { /* LedOptimizerJa.cpp:101-115 */
                    float pt[3], g2[3];
                    cam_point(o, &x, R, t, pt, g2);
                    float nR[3];
                    for (int a = 0; a < 3; ++a) nR[a] = sum3(gh[0] * R[0 + a], gh[1] * R[3 + a], gh[2] * R[6 + a]);
                    float refl = dot3(nR, pt);
                    refl /= (float)powd(norm3(pt), 3);
                    for (int c = 0; c < 3; ++c) {
                        const float j = refl * x.rho[c];
                        *ra_at(H, c, c) += (j * w[c]) * j;
                        b[c] += (j * w[c]) * r[c];
                        if (H64) { H64[c] += (double)((j * w[c]) * j); b64[c] += (double)((j * w[c]) * r[c]); }
                    }
                } else { /* PsOptimizerJa.cpp:132-143 */
                    float sh[9];
                    sh_eval(gh, o->order, sh);
                    for (int c = 0; c < 3; ++c) {
                        float J[9];
                        for (int i = 0; i < nb; ++i) J[i] = (-x.rho[c]) * sh[i];
                        for (int i = 0; i < nb; ++i) {
                            const float jw = J[i] * w[c];
                            for (int k = 0; k < nb; ++k) *ra_at(H, nb * f + i, nb * f + k) += jw * J[k];
                            b[nb * f + i] += jw * r[c];
                            if (H64) {
                                for (int k = 0; k < nb; ++k) H64[(size_t)(nb * f + i) * nb + k] += (double)(jw * J[k]);
                                b64[nb * f + i] += (double)(jw * r[c]);
                            }
                        }
                    }
                }
            } else if (which == OR_POSE) {
                float J[18];
                if (!pose_jacobian(o, &x, f, R, t, J)) continue;
                for (int c = 0; c < 3; ++c) for (int i = 0; i < 6; ++i) {
                    const float jw = J[6 * c + i] * w[c];
                    for (int k = 0; k < 6; ++k) *ra_at(H, 6 * f + i, 6 * f + k) += jw * J[6 * c + k];
                    b[6 * f + i] += jw * r[c];
                    if (H64) {
                        for (int k = 0; k < 6; ++k) H64[(size_t)(6 * f + i) * 6 + k] += (double)(jw * J[6 * c + k]);
                        b64[6 * f + i] += (double)(jw * r[c]);
                    }
                }
            } else {
                float Jd[4][3], pos[3]; long long nl[3];
                if (!dist_jacobian(o, &x, f, R, t, Jd, pos, nl)) continue;
                int cols[4] = { s, -1, -1, -1 };
                for (int k = 0; k < 3; ++k) if (nl[k] >= 0 && nl[k] < v->N) cols[k + 1] = o->rowmap[nl[k]];
                for (int c = 0; c < 3; ++c) for (int i = 0; i < 4; ++i) {
                    if (cols[i] < 0) continue;
                    const float jw = Jd[i][c] * w[c];
                    for (int k = 0; k < 4; ++k) if (cols[k] >= 0) *ra_at(H, cols[i], cols[k]) += jw * Jd[k][c];
                    b[cols[i]] += jw * r[c];
                }
            }
        }
    }

    /* regularisers */
    if (which == OR_DIST && normal_reg) { /* Optimizer.cpp:196-218, 477-537; PsOptimizer.cpp:136-142 */
        rowacc* M = ra_new(n, 48);
        rowsum2 bb = rs2_new(n);
        for (int s = 0; s < S; ++s) {
            voxel x = load_voxel(o, (size_t)o->surf[s]);
            float g[3], pos[3]; long long nl[3];
            dist_grad(o, &x, g, pos, nl);
            float nd[3], Jr[4];
            for (int a = 0; a < 3; ++a) nd[a] = -o->vs_inv * pos[a];
            Jr[0] = dot3(g, nd);
            for (int a = 0; a < 3; ++a) Jr[a + 1] = g[a] * (0.f + o->vs_inv * pos[a]);
            const float gn = norm3(g);
            if (gn > 0.f) for (int a = 0; a < 4; ++a) Jr[a] /= gn;
            const float rr = gn - 1;
            int cols[4] = { s, -1, -1, -1 };
            for (int k = 0; k < 3; ++k) if (nl[k] >= 0 && nl[k] < v->N) cols[k + 1] = o->rowmap[nl[k]];
            for (int i = 0; i < 4; ++i) {
                if (cols[i] < 0) continue;
                for (int k = 0; k < 4; ++k) if (cols[k] >= 0) *ra_at(M, cols[i], cols[k]) += Jr[i] * Jr[k];
                rs2_add(&bb, cols[i], o->reg_n * Jr[i], rr);
            }
        }
        for (int i = 0; i < n; ++i)
            for (int k = 0; k < M->cnt[i]; ++k)
                *ra_at(H, i, M->col[(size_t)i * M->cap + k]) += o->reg_n * M->val[(size_t)i * M->cap + k];
        rs2_finish(&bb, b);
        ra_free(M);
    }
    if (which == OR_DIST && laplacian_reg) { /* Optimizer.cpp:540-590 (diagonal only); PsOptimizer.cpp:144-150 */
        const float vi2 = o->vs_inv * o->vs_inv;
        const float jd = -6 * vi2;
        for (int s = 0; s < S; ++s) {
            voxel x = load_voxel(o, (size_t)o->surf[s]);
            const float rr = dist_laplacian(o, &x);
            *ra_at(H, s, s) += o->reg_l * (jd * jd);
            b[s] += (o->reg_l * jd) * rr + 0.f;   /* (reg_l * J^T) * r, one entry per row: tmp_a + tmp_b with tmp_b = 0 */
        }
    }
    if (which == OR_ALBEDO && albedo_reg) { /* Optimizer.cpp:221-245, 593-647 (incl. the blue-channel column quirk at :618) */
        rowacc* M = ra_new(n, 48);
        rowsum2 bb = rs2_new(n);
        /* J rows: 3*s+c ; gather J first, then J^T J in row order */
        for (int s = 0; s < S; ++s) {
            voxel x = load_voxel(o, (size_t)o->surf[s]);
            float G[9], pos[3]; long long nl[3];
            albedo_grad(o, &x, G, pos, nl);
            float rd[3], gnorm[3], Jr[4][3];
            for (int a = 0; a < 3; ++a) rd[a] = -o->vs_inv * pos[a];
            for (int c = 0; c < 3; ++c) { gnorm[c] = norm3(G + 3 * c); Jr[0][c] = dot3(G + 3 * c, rd); }
            for (int a = 0; a < 3; ++a) for (int c = 0; c < 3; ++c) Jr[a + 1][c] = G[3 * c + a] * (0.f + o->vs_inv * pos[a]);
            for (int c = 0; c < 3; ++c) if (gnorm[c] != 0.f) for (int a = 0; a < 4; ++a) Jr[a][c] /= gnorm[c];
            int ncol[3] = { -1, -1, -1 };
            for (int k = 0; k < 3; ++k) if (nl[k] >= 0 && nl[k] < v->N) ncol[k] = o->rowmap[nl[k]];
            for (int c = 0; c < 3; ++c) {
                int cols[4]; float vals[4]; int m = 0;
                cols[m] = (c == 2) ? 3 * s + 1 : 3 * s + c; vals[m++] = Jr[0][c];
                for (int k = 0; k < 3; ++k) if (ncol[k] >= 0) { cols[m] = 3 * ncol[k] + c; vals[m++] = Jr[k + 1][c]; }
                /* H += reg * Jr^T * Jr parses as (reg * Jr^T) * Jr (PsOptimizer.cpp:99; the distance block writes reg * (J^T J)):
                 * the scaled transpose is evaluated first, so every term is (reg * J(k,i)) * J(k,j) */
                for (int i = 0; i < m; ++i) {
                    const float sv = o->reg_rho * vals[i];
                    for (int k = 0; k < m; ++k) *ra_at(M, cols[i], cols[k]) += sv * vals[k];
                    rs2_add(&bb, cols[i], sv, gnorm[c]);
                }
            }
        }
        for (int i = 0; i < n; ++i)
            for (int k = 0; k < M->cnt[i]; ++k)
                *ra_at(H, i, M->col[(size_t)i * M->cap + k]) += M->val[(size_t)i * M->cap + k];
        rs2_finish(&bb, b);
        ra_free(M);
    }
    if (H64) {
        for (int i = 0; i < n; ++i) {
            const int blk0 = (i / bs64) * bs64;
            for (int k = 0; k < bs64; ++k) {
                const double hv = H64[(size_t)i * bs64 + k];
                if (hv != 0.0 || ra_get(H, i, blk0 + k) != 0.f) *ra_at(H, i, blk0 + k) = (float)hv;
            }
            b[i] = (float)b64[i];
        }
        free(H64); free(b64);
    }
    /* H.diagonal() += damping * H.diagonal(); the SH light block has none (PsOptimizer.cpp:175-203) */
    const int damp = !(which == OR_LIGHT && o->model != OR_LED);
    if (damp && damping != 0.f)
        for (int i = 0; i < n; ++i) {
            int has = 0;
            for (int k = 0; k < H->cnt[i]; ++k) if (H->col[(size_t)i * H->cap + k] == i) has = 1;
            if (has) { float* d = ra_at(H, i, i); *d += damping * *d; }
        }

    free(o->b); free(o->delta);
    o->n = n; o->b = b; o->delta = calloc(n > 0 ? n : 1, sizeof(float));
    eigen_like_cg(H, b, o->delta, &o->success, &o->iters, &o->err);
    o->Hn = 0;
    for (int i = 0; i < n; ++i) for (int k = 0; k < H->cnt[i]; ++k)
        H_push(o, i, H->col[(size_t)i * H->cap + k], H->val[(size_t)i * H->cap + k]);
    ra_free(H);

    if (apply) {
        const int led = o->model == OR_LED;
        if (which == OR_ALBEDO) { if (led || o->success) update_albedo(o, o->delta); }
        else if (which == OR_LIGHT) { for (int i = 0; i < n; ++i) o->light[i] -= o->delta[i]; }
        else if (which == OR_DIST) { if (led || o->success) update_dist(o, o->delta); }
        else { if (!led || o->success) update_pose(o, o->delta); }
    }
    return n;
}
void or_opt_block_result(const or_opt* o, float* b, float* delta, int* success, int* iters, float* err) {
    if (b) memcpy(b, o->b, sizeof(float) * o->n);
    if (delta) memcpy(delta, o->delta, sizeof(float) * o->n);
    if (success) *success = o->success;
    if (iters) *iters = o->iters;
    if (err) *err = o->err;
}
long long or_opt_block_H(const or_opt* o, int* rows, int* cols, float* vals) {
    if (rows) for (long long k = 0; k < o->Hn; ++k) { rows[k] = o->Ht[k].i; cols[k] = o->Ht[k].j; vals[k] = o->Ht[k].v; }
    return o->Hn;
}

/* ------------------------------------------------------------------ upsampling */
/* Optimizer::subsampling (OptimizerAux.cpp:622-684), VolumetricGradSdf::subsample (VolumetricGradSdf.cpp:469-494),
 * VoxelGrid::grid_subsample (VoxelGrid.h:143-149) */
void or_opt_subsampling(or_opt* o) {
    or_volume* v = o->vol;
    or_volume nv = *v;
    for (int a = 0; a < 3; ++a) nv.dims[a] = 2 * v->dims[a];
    nv.N = 8 * v->N;
    vol_alloc(&nv);
    uint64_t* nvisF = calloc((size_t)nv.N, 8);
    for (long long i = 0; i < nv.N; ++i) {
        nv.dist[i] = v->T; nv.gx[i] = nv.gy[i] = nv.gz[i] = 0.f; nv.w[i] = 0.f; nv.r[i] = nv.g[i] = nv.b[i] = 0.5f;
    }
    const float vs4 = (float)(0.25 * (double)v->vs);
    const size_t N = (size_t)v->N, NN = (size_t)nv.N;
    for (int k = 0; k < v->dims[2]; ++k) for (int j = 0; j < v->dims[1]; ++j) for (int i = 0; i < v->dims[0]; ++i) {
        const size_t lin = (size_t)i + (size_t)v->dims[0] * j + (size_t)v->dims[0] * v->dims[1] * k;
        if (v->dist[lin] == v->T) continue;
        const float g0[3] = { v->gx[lin], v->gy[lin], v->gz[lin] };
        float g[3];
        normalized3(g0, g);
        for (int sub = 0; sub < 8; ++sub) {
            const int sx = sub & 1, sy = (sub >> 1) & 1, sz = (sub >> 2) & 1;
            const float a = sx ? g[0] : -g[0], bq = sy ? g[1] : -g[1], cq = sz ? g[2] : -g[2];
            const float d = v->dist[lin] + vs4 * ((a + bq) + cq); /* (+-gx +- gy) +- gz, left to right */
            const size_t nl = (size_t)(2 * i + sx) + (size_t)nv.dims[0] * (2 * j + sy) + (size_t)nv.dims[0] * nv.dims[1] * (2 * k + sz);
            nv.dist[nl] = d; nv.gx[nl] = g0[0]; nv.gy[nl] = g0[1]; nv.gz[nl] = g0[2];
            nv.w[nl] = v->w[lin]; nv.r[nl] = v->r[lin]; nv.g[nl] = v->g[lin]; nv.b[nl] = v->b[lin];
            for (int w = 0; w < v->nwords; ++w) nv.vis[(size_t)w * NN + nl] = v->vis[(size_t)w * N + lin];
            nv.vis_len[nl] = v->vis_len[lin];
            nvisF[nl] = o->visF[lin];
        }
    }
    vol_free_arrays(v);
    /* grid_subsample: voxel_size *= 0.5; origin = shift - 0.5*vs*dim - 0.5*vs */
    nv.vs = (float)((double)v->vs * 0.5);
    const float half_vs = (float)(0.5 * (double)nv.vs);
    for (int a = 0; a < 3; ++a) nv.origin[a] = (v->shift[a] - half_vs * (float)nv.dims[a]) - half_vs * 1.f;
    *v = nv;
    free(o->visF);
    o->visF = nvisF;
    o->vs = (float)((double)o->vs * 0.5);
    o->vs_inv = (float)(1.0 / (double)o->vs);
    select_surface(o, 0);
}

/* ---- rigid depth-to-SDF tracker: one pass of RigidPointOptimizer::optimize_sampled's inner loop
 * (sdf_tracker/RigidPointOptimizer.cpp:38-62) at pose (R, t), with Sdf::weights / Sdf::tsdf of the gradient-SDF volume
 * (VolumetricGradSdf.h:76-93) and VoxelGrid::nearest_index / world2voxel / voxel2world (VoxelGrid.cpp:57-72,
 * VoxelGrid.h:38-66).  Sums in one serial float each, in pixel order, like the reference. */
static int nearest_index(const or_volume* v, const float* p) {
    const float i = (p[0] - v->origin[0]) / v->vs, j = (p[1] - v->origin[1]) / v->vs, k = (p[2] - v->origin[2]) / v->vs;
    if (i <= 0 || j <= 0 || k <= 0 || i >= (v->dims[0] - 1) || j >= (v->dims[1] - 1) || k >= (v->dims[2] - 1)) return -1;
    const int im = (int)((double)i + 0.5), jm = (int)((double)j + 0.5), km = (int)((double)k + 0.5);
    return im + jm * v->dims[0] + km * v->dims[0] * v->dims[1];
}

void or_tracker_system(const or_volume* v, const float* depth, int H, int W, const float K[9], const float R[9],
                       const float t[3], int sampling, float* E_out, float g6[6], float H36[36], long long* count) {
    const float fx = K[0], fy = K[4], cx = K[2], cy = K[5];
    const float fx_inv = 1.f / fx, fy_inv = 1.f / fy;
    float E = 0.f, g[6] = {0, 0, 0, 0, 0, 0}, Hm[36];
    long long counter = 0;
    for (int i = 0; i < 36; ++i) Hm[i] = 0.f;
    for (int y = 0; y < H; y += sampling)
        for (int x = 0; x < W; x += sampling) {
            const float z = depth[y * W + x];
            if (z <= v->zmin || z >= v->zmax) continue;
            const float x0 = ((float)x - cx) * fx_inv;
            const float y0 = ((float)y - cy) * fy_inv;
            const float pc[3] = { x0 * z, y0 * z, z };
            float Rp[3], point[3];
            mat_vec(R, pc, Rp);
            for (int a = 0; a < 3; ++a) point[a] = Rp[a] + t[a];
            const int I = nearest_index(v, point);
            if (I < 0 || !(v->w[I] > 0)) continue;
            const float grad[3] = { v->gx[I], v->gy[I], v->gz[I] };
            float gn[3];
            normalized3(grad, gn);
            /* voxel2world(world2voxel(point)) - point: world2voxel adds a FLOAT 0.5 and truncates */
            float d[3];
            for (int a = 0; a < 3; ++a) {
                const float fi = (point[a] - v->origin[a]) / v->vs + 0.5f;
                const int idx = (int)fi;
                d[a] = (v->origin[a] + v->vs * (float)idx) - point[a];
            }
            const float phi0 = v->dist[I] + dot3(gn, d);
            E += phi0 * phi0;
            const float gx[6] = { gn[0], gn[1], gn[2], point[1] * gn[2] - point[2] * gn[1], point[2] * gn[0] - point[0] * gn[2],
                                  point[0] * gn[1] - point[1] * gn[0] };
            for (int a = 0; a < 6; ++a) g[a] += phi0 * gx[a];
            for (int a = 0; a < 6; ++a)
                for (int b = 0; b < 6; ++b) Hm[6 * a + b] += gx[a] * gx[b];
            ++counter;
        }
    *E_out = E;
    for (int a = 0; a < 6; ++a) g6[a] = g[a];
    for (int a = 0; a < 36; ++a) H36[a] = Hm[a];
    *count = counter;
}
