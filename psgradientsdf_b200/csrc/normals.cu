// Front end of fusion on the device: the loader's pixel conversions and the FALS normal estimator,
// so that a frame crosses PCIe as the 16-bit depth and 8-bit colour the loader read from disk instead of eight float planes.
//
// Replaces cv::NormalEstimator<float> (cpp/include/normals/NormalEstimator.h:52-176) as main_ps.cpp uses it
// (:115 construction with an 11 x 11 window, VolumetricGradSdf.cpp:63-66 compute + n_sq_inv_ptr) and
// ImageLoader::load_depth / load_color's convertTo calls (cpp/include/img_loader/ImageLoader.h:130-146, 166-188).
//
// The per-camera tables (cache(), :52-123) are computed once on the host in double, in the summation order of OpenCV's
// boxFilter (sliding row sums, then sliding column sums) because double sums of double data depend on that order.  The
// per-frame work runs in two kernels: row pass (1/z, the three products, 11 horizontal taps accumulated in double) and
// column pass (11 vertical taps, cast to float, the 3 x 3 symmetric solve with the cached Q, normalisation).  The float
// products summed here carry 24 significant bits, so their double sums are exact for any sane depth range and therefore
// independent of the order -- the result is OpenCV's bit for bit (tests/golden/normals.npz) without its serial recurrence.
#include <cmath>
#include <cstring>
#include <vector>

#include "psg_host.h"

struct psg_normals {
    int device = 0, W = 0, H = 0, window = 11;
    float K[9] = {0};
    psg::DevBuf<float> tables;            // 9 planes: x0_n, y0_n, n_sq_inv, Q11, Q12, Q13, Q22, Q23, Q33
    std::vector<float> host_tables;
    psg::DevBuf<double> rows;             // 3 planes of horizontal sums
    psg::DevBuf<float> depth;             // depth in metres
    psg::DevBuf<float> out;               // 3 planes nx ny nz (psg_normals_compute)
    psg::DevBuf<unsigned char> raw;       // H2D staging of one raw frame: depth u16 then colour u8 (or a float colour image)
    psg::DevBuf<double> partials;         // block partials of the sharpness sum
    cudaStream_t stream = nullptr;
};

namespace psg {

__host__ __device__ __forceinline__ int reflect101(int i, int n) {
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
    return i;
}

// ---- host: NormalEstimator::cache() -------------------------------------------------------------------------------
// unnormalised k x k box sums of a double image: cv::boxFilter(..., normalize = false), BORDER_DEFAULT
static void box_sum_f64(const std::vector<double>& src, int H, int W, int k, std::vector<double>& out) {
    const int r = k / 2;
    std::vector<double> hs((size_t)H * W), line((size_t)W + 2 * r), acc((size_t)W, 0.0);
    for (int y = 0; y < H; ++y) {
        for (int j = -r; j < W + r; ++j) line[(size_t)(j + r)] = src[(size_t)y * W + reflect101(j, W)];
        double s = 0.0;
        for (int i = 0; i < k; ++i) s += line[(size_t)i];
        hs[(size_t)y * W] = s;
        for (int i = 0; i + 1 < W; ++i) { s += line[(size_t)(i + k)] - line[(size_t)i]; hs[(size_t)y * W + i + 1] = s; }
    }
    for (int i = 0; i < k - 1; ++i) {
        const double* in = hs.data() + (size_t)reflect101(i - r, H) * W;
        for (int x = 0; x < W; ++x) acc[(size_t)x] += in[x];
    }
    out.resize((size_t)H * W);
    for (int y = 0; y < H; ++y) {
        const double* add = hs.data() + (size_t)reflect101(y + r, H) * W;
        const double* sub = hs.data() + (size_t)reflect101(y - r, H) * W;
        for (int x = 0; x < W; ++x) {
            const double s0 = acc[(size_t)x] + add[x];
            out[(size_t)y * W + x] = s0;
            acc[(size_t)x] = s0 - sub[x];
        }
    }
}

static void build_tables(psg_normals* ne) {
    const int W = ne->W, H = ne->H, k = ne->window;
    const size_t n = (size_t)H * W;
    const double fx_inv = 1. / (double)ne->K[0], fy_inv = 1. / (double)ne->K[4], cx = (double)ne->K[2], cy = (double)ne->K[5];
    std::vector<double> x0(n), y0(n), nsi(n), x0n(n), y0n(n), t(n), M11, M12, M13, M22, M23, M33;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const size_t i = (size_t)y * W + x;
            x0[i] = fx_inv * ((double)x - cx);
            y0[i] = fy_inv * ((double)y - cy);
            nsi[i] = 1. / (1. + x0[i] * x0[i] + y0[i] * y0[i]);
            x0n[i] = x0[i] * nsi[i];
            y0n[i] = y0[i] * nsi[i];
        }
    for (size_t i = 0; i < n; ++i) t[i] = (x0[i] * x0[i]) * nsi[i];
    box_sum_f64(t, H, W, k, M11);
    for (size_t i = 0; i < n; ++i) t[i] = (x0[i] * y0[i]) * nsi[i];
    box_sum_f64(t, H, W, k, M12);
    box_sum_f64(x0n, H, W, k, M13);
    for (size_t i = 0; i < n; ++i) t[i] = (y0[i] * y0[i]) * nsi[i];
    box_sum_f64(t, H, W, k, M22);
    box_sum_f64(y0n, H, W, k, M23);
    box_sum_f64(nsi, H, W, k, M33);
    ne->host_tables.assign(9 * n, 0.f);
    float* T = ne->host_tables.data();
    for (size_t i = 0; i < n; ++i) {
        const double det = M11[i] * (M22[i] * M33[i]) + 2 * M12[i] * (M23[i] * M13[i]) -
                           (M13[i] * (M13[i] * M22[i]) + M12[i] * (M12[i] * M33[i]) + M23[i] * (M23[i] * M11[i]));
        const double di = 1. / det;
        T[0 * n + i] = (float)x0n[i]; T[1 * n + i] = (float)y0n[i]; T[2 * n + i] = (float)nsi[i];
        T[3 * n + i] = (float)(di * (M22[i] * M33[i] - M23[i] * M23[i]));
        T[4 * n + i] = (float)(di * (M13[i] * M23[i] - M12[i] * M33[i]));
        T[5 * n + i] = (float)(di * (M12[i] * M23[i] - M13[i] * M22[i]));
        T[6 * n + i] = (float)(di * (M11[i] * M33[i] - M13[i] * M13[i]));
        T[7 * n + i] = (float)(di * (M12[i] * M13[i] - M11[i] * M23[i]));
        T[8 * n + i] = (float)(di * (M11[i] * M22[i] - M12[i] * M12[i]));
    }
}

// ---- device ---------------------------------------------------------------------------------------------------------
// ImageLoader::load_depth: convertTo(CV_32FC1, unit) = float(src) * unit
__global__ void __launch_bounds__(256) k_depth_u16(const unsigned short* __restrict__ src, float unit, float* __restrict__ dst, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = (float)src[i] * unit;
}

// row pass: z_inv = depth != 0 ? 1 / depth : 0; horizontal window sums of x0_n z_inv, y0_n z_inv, n_sq_inv z_inv
__global__ void __launch_bounds__(256) k_normals_rows(const float* __restrict__ depth, const float* __restrict__ tables, int H, int W,
                                                     int r, double* __restrict__ rows) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = H * W;
    if (i >= n) return;
    const int y = i / W, x = i - y * W;
    double s1 = 0.0, s2 = 0.0, s3 = 0.0;
    for (int d = -r; d <= r; ++d) {
        const int j = y * W + reflect101(x + d, W);
        const float z = __ldg(depth + j);
        const float zi = z != 0.f ? 1.f / z : 0.f;
        s1 += (double)(__ldg(tables + j) * zi);
        s2 += (double)(__ldg(tables + n + j) * zi);
        s3 += (double)(__ldg(tables + 2 * n + j) * zi);
    }
    rows[i] = s1; rows[n + i] = s2; rows[2 * n + i] = s3;
}

// column pass + NormalEstimator::compute's tail.  PACKED = false: planar nx, ny, nz.  PACKED = true: the 32-byte fusion
// record {depth, nx, ny, nz, n_sq_inv, b, g, r} with the colour converted from 8-bit BGR (load_color: float(src) * (1/255)).
template <bool PACKED>
__global__ void __launch_bounds__(256) k_normals_cols(const double* __restrict__ rows, const float* __restrict__ tables, int H, int W, int r,
                                                     const float* __restrict__ depth, const unsigned char* __restrict__ bgr8,
                                                     float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = H * W;
    if (i >= n) return;
    const int y = i / W, x = i - y * W;
    double s1 = 0.0, s2 = 0.0, s3 = 0.0;
    for (int d = -r; d <= r; ++d) {
        const int j = reflect101(y + d, H) * W + x;
        s1 += rows[j]; s2 += rows[n + j]; s3 += rows[2 * n + j];
    }
    const float b1 = (float)s1, b2 = (float)s2, b3 = (float)s3;
    const float Q11 = tables[3 * n + i], Q12 = tables[4 * n + i], Q13 = tables[5 * n + i];
    const float Q22 = tables[6 * n + i], Q23 = tables[7 * n + i], Q33 = tables[8 * n + i];
    float nx = (b1 * Q11 + b2 * Q12) + b3 * Q13;
    float ny = (b1 * Q12 + b2 * Q22) + b3 * Q23;
    float nz = (b1 * Q13 + b2 * Q23) + b3 * Q33;
    const float norm = sqrtf((nx * nx + ny * ny) + nz * nz);
    nx = nx / norm; ny = ny / norm; nz = nz / norm;   // 0 / 0 = NaN where the whole window is empty (OpenCV >= 4 behaviour)
    if constexpr (PACKED) {
        const float a = 1.0f / 255.0f;
        float4* o = reinterpret_cast<float4*>(out) + 2ll * i;
        o[0] = make_float4(depth[i], nx, ny, nz);
        o[1] = make_float4(tables[2 * n + i], (float)bgr8[3 * i] * a, (float)bgr8[3 * i + 1] * a, (float)bgr8[3 * i + 2] * a);
    } else {
        out[i] = nx; out[n + i] = ny; out[2 * n + i] = nz;
    }
}

// keyframe images from 8-bit BGR: texel (b, g, r, 0) = float(src) * (1/255)
__global__ void __launch_bounds__(256) k_texels_u8(const unsigned char* __restrict__ bgr8, float4* __restrict__ dst, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float a = 1.0f / 255.0f;
    dst[i] = make_float4((float)bgr8[3 * i] * a, (float)bgr8[3 * i + 1] * a, (float)bgr8[3 * i + 2] * a, 0.f);
}

// modifiedLaplacian (ps_optimizer/SharpDetector.h:22-37) of the blue channel: FM = |G_col(M_row(x))| + |M_col(G_row(x))| with
// M = (-1, 2, -1), G = (0.25, 0.5, 0.25), reflect-101 border, OpenCV's (left + right) * k1 + centre * k0 form; sum in double.
// U8 = true reads the 8-bit BGR image the loader saw and converts like load_color.
__device__ __forceinline__ float tap3(float l, float c, float r, float k1, float k0) { return (l + r) * k1 + c * k0; }
template <bool U8>
__global__ void __launch_bounds__(256) k_sharpness(const void* __restrict__ img, int H, int W, double* __restrict__ out) {
    __shared__ double s_w[8];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double fm = 0.0;
    if (i < H * W) {
        const int y = i / W, x = i - y * W;
        const int xl = reflect101(x - 1, W), xr = reflect101(x + 1, W);
        float tm[3], tg[3];
#pragma unroll
        for (int d = -1; d <= 1; ++d) {
            const long long row = (long long)reflect101(y + d, H) * W;
            float l, c, r;
            if constexpr (U8) {
                const unsigned char* p = static_cast<const unsigned char*>(img);
                const float a = 1.0f / 255.0f;
                l = (float)p[3 * (row + xl)] * a; c = (float)p[3 * (row + x)] * a; r = (float)p[3 * (row + xr)] * a;
            } else {
                const float* p = static_cast<const float*>(img);
                l = p[3 * (row + xl)]; c = p[3 * (row + x)]; r = p[3 * (row + xr)];
            }
            tm[d + 1] = tap3(l, c, r, -1.f, 2.f);
            tg[d + 1] = tap3(l, c, r, 0.25f, 0.5f);
        }
        const float Lx = tap3(tm[0], tm[1], tm[2], 0.25f, 0.5f);
        const float Ly = tap3(tg[0], tg[1], tg[2], -1.f, 2.f);
        fm = (double)(fabsf(Lx) + fabsf(Ly));
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) fm += __shfl_xor_sync(0xffffffffu, fm, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = fm;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += s_w[w];
        out[blockIdx.x] = t;          // block partials, added on the host in block order: reproducible
    }
}

static void run_normals(psg_normals* ne, cudaStream_t st, bool packed, const unsigned char* bgr8_dev, float* out) {
    const int n = ne->H * ne->W, r = ne->window / 2;
    ne->rows.alloc((size_t)3 * n);
    const unsigned g = (unsigned)((n + 255) / 256);
    { Prof pf(st, KID_NORMALS); k_normals_rows<<<g, 256, 0, st>>>(ne->depth.p, ne->tables.p, ne->H, ne->W, r, ne->rows.p); }
    PSG_LAUNCH_CHECK();
    { Prof pf(st, KID_NORMALS);
      if (packed) k_normals_cols<true><<<g, 256, 0, st>>>(ne->rows.p, ne->tables.p, ne->H, ne->W, r, ne->depth.p, bgr8_dev, out);
      else k_normals_cols<false><<<g, 256, 0, st>>>(ne->rows.p, ne->tables.p, ne->H, ne->W, r, ne->depth.p, nullptr, out); }
    PSG_LAUNCH_CHECK();
}

void launch_texels_u8(cudaStream_t st, const unsigned char* bgr8_dev, float4* dst, long long n) {
    { Prof pf(st, KID_PAD_IMAGES); k_texels_u8<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(bgr8_dev, dst, n); }
    PSG_LAUNCH_CHECK();
}

}  // namespace psg

using namespace psg;

extern "C" {

int psg_normals_create(psg_normals** out, int width, int height, const float K[9], int window, int device) {
    if (!out || !K || width <= 0 || height <= 0 || window < 1 || (window & 1) == 0 || window > 63) {
        set_error("psg_normals_create: bad argument (window must be odd, 1..63)");
        return PSG_ERR_INVALID;
    }
    if ((long long)width * height > (1ll << 28)) { set_error("psg_normals_create: image too large"); return PSG_ERR_INVALID; }
    psg_normals* ne = nullptr;
    try {
        PSG_CUDA(cudaSetDevice(device));
        ne = new psg_normals();
        ne->device = device; ne->W = width; ne->H = height; ne->window = window;
        memcpy(ne->K, K, sizeof(ne->K));
        PSG_CUDA(cudaStreamCreateWithFlags(&ne->stream, cudaStreamNonBlocking));
        build_tables(ne);
        ne->tables.alloc(ne->host_tables.size());
        PSG_CUDA(cudaMemcpyAsync(ne->tables.p, ne->host_tables.data(), ne->host_tables.size() * sizeof(float), cudaMemcpyHostToDevice, ne->stream));
        PSG_CUDA(cudaStreamSynchronize(ne->stream));
        *out = ne;
        return PSG_OK;
    } catch (const psg::CudaError&) {
        delete ne;
        return PSG_ERR_CUDA;
    }
}

int psg_normals_destroy(psg_normals* ne) {
    if (!ne) return PSG_OK;
    cudaSetDevice(ne->device);
    if (ne->stream) { cudaStreamSynchronize(ne->stream); cudaStreamDestroy(ne->stream); }
    delete ne;
    return PSG_OK;
}

int psg_normals_tables(psg_normals* ne, float* planes9) {
    if (!ne || !planes9) { set_error("psg_normals_tables: null argument"); return PSG_ERR_INVALID; }
    memcpy(planes9, ne->host_tables.data(), ne->host_tables.size() * sizeof(float));
    return PSG_OK;
}

int psg_normals_compute(psg_normals* ne, const float* depth, float* nx, float* ny, float* nz) {
    if (!ne || !depth || !nx || !ny || !nz) { set_error("psg_normals_compute: null argument"); return PSG_ERR_INVALID; }
    try {
        PSG_CUDA(cudaSetDevice(ne->device));
        const size_t n = (size_t)ne->H * ne->W;
        ne->depth.alloc(n);
        ne->out.alloc(3 * n);
        PSG_CUDA(cudaMemcpyAsync(ne->depth.p, depth, n * sizeof(float), cudaMemcpyHostToDevice, ne->stream));
        run_normals(ne, ne->stream, false, nullptr, ne->out.p);
        PSG_CUDA(cudaMemcpyAsync(nx, ne->out.p, n * sizeof(float), cudaMemcpyDeviceToHost, ne->stream));
        PSG_CUDA(cudaMemcpyAsync(ny, ne->out.p + n, n * sizeof(float), cudaMemcpyDeviceToHost, ne->stream));
        PSG_CUDA(cudaMemcpyAsync(nz, ne->out.p + 2 * n, n * sizeof(float), cudaMemcpyDeviceToHost, ne->stream));
        PSG_CUDA(cudaStreamSynchronize(ne->stream));
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

int psg_volume_integrate_raw(psg_volume* v, psg_normals* ne, const unsigned char* color_bgr_u8, const unsigned short* depth_u16,
                             float depth_unit, const float K[9], const float R[9], const float t[3], int increase_counter) {
    if (!v || !ne || !color_bgr_u8 || !depth_u16 || !K || !R || !t) { set_error("psg_volume_integrate_raw: null argument"); return PSG_ERR_INVALID; }
    if (ne->device != v->device) { set_error("psg_volume_integrate_raw: estimator and volume live on different devices"); return PSG_ERR_INVALID; }
    const int counter = v->counter + (increase_counter ? 1 : 0);
    if (counter >= 64 * v->nwords) {
        set_error("psg_volume_integrate_raw: frame %d exceeds max_frames %d given at creation", counter, v->max_frames);
        return PSG_ERR_INVALID;
    }
    try {
        PSG_CUDA(cudaSetDevice(v->device));
        const int H = ne->H, W = ne->W;
        const size_t n = (size_t)H * W;
        // raw frame: depth (2 B / pixel) then colour (3 B / pixel); everything below runs on the volume's stream
        ne->raw.alloc(n * 5);
        ne->depth.alloc(n);
        v->frame.alloc(n * 8);
        unsigned short* d16 = reinterpret_cast<unsigned short*>(ne->raw.p);
        unsigned char* c8 = ne->raw.p + n * 2;
        PSG_CUDA(cudaMemcpyAsync(d16, depth_u16, n * 2, cudaMemcpyHostToDevice, v->stream));
        PSG_CUDA(cudaMemcpyAsync(c8, color_bgr_u8, n * 3, cudaMemcpyHostToDevice, v->stream));
        { Prof pf(v->stream, KID_NORMALS); k_depth_u16<<<(unsigned)((n + 255) / 256), 256, 0, v->stream>>>(d16, depth_unit, ne->depth.p, (int)n); }
        PSG_LAUNCH_CHECK();
        run_normals(ne, v->stream, true, c8, v->frame.p);
        v->counter = counter;
        v->generation++;
        launch_integrate(v, H, W, K, R, t, counter);
        PSG_CUDA(cudaStreamSynchronize(v->stream));
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}


static int sharpness_impl(psg_normals* ne, const void* host_img, bool u8, float* measure) {
    if (!ne || !host_img || !measure) { set_error("psg_normals_sharpness: null argument"); return PSG_ERR_INVALID; }
    try {
        PSG_CUDA(cudaSetDevice(ne->device));
        const size_t n = (size_t)ne->H * ne->W;
        const size_t bytes = n * 3 * (u8 ? 1 : 4);
        ne->raw.alloc(n * 12);
        const int nblocks = (int)((n + 255) / 256);
        ne->partials.alloc((size_t)nblocks);
        PSG_CUDA(cudaMemcpyAsync(ne->raw.p, host_img, bytes, cudaMemcpyHostToDevice, ne->stream));
        { Prof pf(ne->stream, KID_NORMALS);
          if (u8) k_sharpness<true><<<nblocks, 256, 0, ne->stream>>>(ne->raw.p, ne->H, ne->W, ne->partials.p);
          else k_sharpness<false><<<nblocks, 256, 0, ne->stream>>>(ne->raw.p, ne->H, ne->W, ne->partials.p); }
        PSG_LAUNCH_CHECK();
        std::vector<double> h((size_t)nblocks);
        PSG_CUDA(cudaMemcpyAsync(h.data(), ne->partials.p, sizeof(double) * (size_t)nblocks, cudaMemcpyDeviceToHost, ne->stream));
        PSG_CUDA(cudaStreamSynchronize(ne->stream));
        double acc = 0.0;
        for (int b = 0; b < nblocks; ++b) acc += h[(size_t)b];
        *measure = (float)(acc / ((double)ne->H * (double)ne->W));
        return PSG_OK;
    } catch (const psg::CudaError&) { return PSG_ERR_CUDA; }
}

int psg_normals_sharpness(psg_normals* ne, const float* color_bgr, float* measure) { return sharpness_impl(ne, color_bgr, false, measure); }
int psg_normals_sharpness_u8(psg_normals* ne, const unsigned char* color_bgr_u8, float* measure) { return sharpness_impl(ne, color_bgr_u8, true, measure); }

// compute_centroid (voxel_ps/src/main_ps.cpp:346-375).  One serial float accumulator per component over the pixels in row
// order: the rounding of that sum fixes the volume origin for everything downstream, so it is kept serial (host, ~1 ms).
int psg_compute_centroid(const float K[9], const float* depth, int H, int W, const float pose16[16], float centroid[3]) {
    if (!K || !depth || !pose16 || !centroid || H <= 0 || W <= 0) { set_error("psg_compute_centroid: null or empty argument"); return PSG_ERR_INVALID; }
    float c[3] = { 0.f, 0.f, 0.f };
    int counter = 0;
    const float fx_inv = 1.f / K[0], fy_inv = 1.f / K[4], cx = K[2], cy = K[5];
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const float z = depth[(size_t)y * W + x];
            if (z > 0.0) {
                const float x0 = ((float)x - cx) * fx_inv, y0 = ((float)y - cy) * fy_inv;
                const float p[3] = { x0 * z, y0 * z, z };
                for (int a = 0; a < 3; ++a) {
                    const float rp = pose16[4 * a] * p[0] + (pose16[4 * a + 1] * p[1] + pose16[4 * a + 2] * p[2]);
                    c[a] += rp + pose16[4 * a + 3];
                }
                ++counter;
            }
        }
    for (int a = 0; a < 3; ++a) centroid[a] = c[a] / (float)counter;
    return PSG_OK;
}

// sampleKeyFrame (main_ps.cpp:392-421): positions (into the keyframe list) of the keyframes that are kept
int psg_sample_keyframes(int n, int max_num, int* kept, int* n_kept) {
    if (n < 0 || max_num < 2 || !kept || !n_kept) { set_error("psg_sample_keyframes: bad argument"); return PSG_ERR_INVALID; }
    int m = 0;
    if (n < max_num) {
        for (int i = 0; i < n; ++i) kept[m++] = i;
    } else {
        const int cnt = max_num - 1;
        const float step = (float)n / (float)cnt;
        float idx = 0;
        for (int k = 0; k < cnt; ++k) { kept[m++] = (int)idx; idx += step; }
        kept[m++] = n - 1;
    }
    *n_kept = m;
    return PSG_OK;
}

}  // extern "C"
