// TEST INFRASTRUCTURE ONLY -- part of the oracle build (oracle/_ref), never linked into the product.
//
// Minimal stand-in for the slice of OpenCV that the reference's hot path touches, so that the
// reference's own sources under /root/reference/cpp compile unmodified without an OpenCV install
// (OpenCV C++ is not present in this image).  Only container semantics are provided:
//   cv::Mat        row-major, reference-counted buffer with at<T>(row, col)
//   cv::Vec<T,n>   value type with OpenCV's arithmetic rules (scale by alpha computed in the
//                  promoted type of (element * alpha), then cast to the element type; add/sub in T)
// Filters (medianBlur, boxFilter ...) are NOT reimplemented: medianBlur is a no-op because the
// reference discards its result (VolumetricGradSdf.cpp:61), and normals are injected through the
// NormalEstimator stand-in (oracle/shim/normals/NormalEstimator.h).
#ifndef PSG_ORACLE_CV_SHIM_CORE_HPP
#define PSG_ORACLE_CV_SHIM_CORE_HPP

#include <cstddef>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_32F 5
#define CV_CN_SHIFT 3
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn)-1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC3 CV_MAKETYPE(CV_32F, 3)

namespace cv {

template <typename T, int N>
struct Vec {
    T val[N];
    Vec() { for (int i = 0; i < N; ++i) val[i] = T(0); }
    Vec(T a, T b, T c) { static_assert(N == 3, "3-ctor"); val[0] = a; val[1] = b; val[2] = c; }
    T& operator[](int i) { return val[i]; }
    const T& operator[](int i) const { return val[i]; }
};

template <typename T, int N>
inline Vec<T, N> operator+(const Vec<T, N>& a, const Vec<T, N>& b) {
    Vec<T, N> r; for (int i = 0; i < N; ++i) r.val[i] = static_cast<T>(a.val[i] + b.val[i]); return r;
}
template <typename T, int N>
inline Vec<T, N> operator-(const Vec<T, N>& a, const Vec<T, N>& b) {
    Vec<T, N> r; for (int i = 0; i < N; ++i) r.val[i] = static_cast<T>(a.val[i] - b.val[i]); return r;
}
template <typename T, int N>
inline Vec<T, N> operator-(const Vec<T, N>& a) {
    Vec<T, N> r; for (int i = 0; i < N; ++i) r.val[i] = static_cast<T>(-a.val[i]); return r;
}
// OpenCV: Vec * alpha -> saturate_cast<T>(val * alpha) with alpha keeping its own type.
#define PSG_CV_SCALE(ATYPE)                                                                  \
    template <typename T, int N> inline Vec<T, N> operator*(ATYPE alpha, const Vec<T, N>& a) { \
        Vec<T, N> r; for (int i = 0; i < N; ++i) r.val[i] = static_cast<T>(a.val[i] * alpha); return r; } \
    template <typename T, int N> inline Vec<T, N> operator*(const Vec<T, N>& a, ATYPE alpha) { \
        Vec<T, N> r; for (int i = 0; i < N; ++i) r.val[i] = static_cast<T>(a.val[i] * alpha); return r; }
PSG_CV_SCALE(int)
PSG_CV_SCALE(float)
PSG_CV_SCALE(double)
#undef PSG_CV_SCALE

typedef Vec<float, 3> Vec3f;
typedef Vec<unsigned char, 3> Vec3b;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Point { int x, y; Point() : x(0), y(0) {} Point(int x_, int y_) : x(x_), y(y_) {} };
struct Scalar { double val[4]; Scalar(double a = 0, double b = 0, double c = 0, double d = 0) { val[0]=a; val[1]=b; val[2]=c; val[3]=d; } };

class Mat {
public:
    int rows, cols;
    unsigned char* data;
    Mat() : rows(0), cols(0), data(nullptr), type_(0), esz_(0) {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size s, int type, const Scalar& = Scalar()) { create(s.height, s.width, type); }
    void create(int r, int c, int type) {
        rows = r; cols = c; type_ = type;
        const int depth = type & 7, cn = (type >> CV_CN_SHIFT) + 1;
        esz_ = static_cast<size_t>(cn) * (depth == CV_32F ? 4 : 1);
        buf_ = std::make_shared<std::vector<unsigned char>>(static_cast<size_t>(r) * c * esz_, 0);
        data = buf_->data();
    }
    // Wrap caller-owned memory (no copy); the caller keeps it alive.
    static Mat wrap(int r, int c, int type, void* ptr) {
        Mat m; m.rows = r; m.cols = c; m.type_ = type;
        const int depth = type & 7, cn = (type >> CV_CN_SHIFT) + 1;
        m.esz_ = static_cast<size_t>(cn) * (depth == CV_32F ? 4 : 1);
        m.data = static_cast<unsigned char*>(ptr);
        return m;
    }
    template <typename T> T& at(int r, int c) {
        return *reinterpret_cast<T*>(data + (static_cast<size_t>(r) * cols + c) * esz_);
    }
    template <typename T> const T& at(int r, int c) const {
        return *reinterpret_cast<const T*>(data + (static_cast<size_t>(r) * cols + c) * esz_);
    }
    int type() const { return type_; }
    bool empty() const { return data == nullptr; }
    void copyTo(Mat& dst) const {
        dst.create(rows, cols, type_);
        std::memcpy(dst.data, data, static_cast<size_t>(rows) * cols * esz_);
    }
    void convertTo(Mat& dst, int) const { dst = *this; }  // debug writers only; not on the hot path
private:
    int type_;
    size_t esz_;
    std::shared_ptr<std::vector<unsigned char>> buf_;
};

// The reference computes a median-filtered depth and never reads it (VolumetricGradSdf.cpp:61).
inline void medianBlur(const Mat&, Mat&, int) {}
inline bool imwrite(const std::string&, const Mat&) { return false; }

}  // namespace cv
#endif
