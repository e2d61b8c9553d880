// sr_cv_standin.hpp -- TEST INFRASTRUCTURE ONLY.  A header-only stand-in for the few OpenCV C++ facilities the
// reference's hot-path sources use (WindowSearch/BlockSearch.cpp, WindowSearch/LinearSearch.cpp,
// Reconstruction/reconstruction.cpp), so that those files compile UNCHANGED, from where they lie under
// /root/reference/src, into oracle/_ref/libsr_ref.so (recipe: oracle/Makefile, target _ref).  OpenCV's C++ headers
// and libraries do not exist in this image; Python cv2 4.13 does, and every arithmetic function below is checked
// against the cv2 function of the same name in tests/test_ref_build.py (absdiff, norm, mean, subtract, blur).
//
// Semantics restated (OpenCV 4.x, the versions the reference can be built with):
//   cv::Mat          reference-counted buffer + ROI views sharing it; operator()(Rect) throws cv::Exception when
//                    the rectangle leaves the matrix (modules/core/src/matrix.cpp, Mat::Mat(const Mat&, const Rect&))
//   cv::absdiff      per-element |a - b| on u8
//   cv::norm L2      u8: exact integer sum of squares, then sqrt in double
//   cv::mean         per-channel exact integer sum times 1/N in double
//   cv::subtract     (u8 Mat, Scalar): the scalar is converted to the source depth first (saturate_cast<uchar> of
//                    cvRound = round half to even), then subtracted with u8 saturation
//   cv::blur         CV_32F: normalised box filter, anchor k/2, BORDER_REFLECT_101, RowSum<float,double> running
//                    sums along x, ColumnSum<double,float> running sums along y, result (float)(sum * (1/(kx*ky)))
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <exception>
#include <fstream>
#include <iostream>
#include <limits>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#define CV_8U 0
#define CV_32F 5
#define CV_64F 6
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << 3))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)

#define CV_Assert(expr) do { if (!(expr)) throw cv::Exception(std::string("(-215:Assertion failed) ") + #expr); } while (0)
#define CV_Error(code, msg) throw cv::Exception(std::string(msg))

namespace cv {
typedef unsigned char uchar;
namespace Error { enum Code { StsBadSize = -201, StsBadArg = -5, StsAssert = -215 }; }

class Exception : public std::exception {
public:
    explicit Exception(const std::string& m) : msg(m) {}
    const char* what() const noexcept override { return msg.c_str(); }
    std::string msg;
};

struct Size {
    int width = 0, height = 0;
    Size() {}
    Size(int w, int h) : width(w), height(h) {}
    bool operator==(const Size& o) const { return width == o.width && height == o.height; }
};
struct Rect {
    int x = 0, y = 0, width = 0, height = 0;
    Rect() {}
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};
struct Scalar {
    double val[4] = {0, 0, 0, 0};
    Scalar() {}
    Scalar(double a, double b = 0, double c = 0, double d = 0) { val[0] = a; val[1] = b; val[2] = c; val[3] = d; }
    double& operator[](int i) { return val[i]; }
    const double& operator[](int i) const { return val[i]; }
};
template <typename T, int N>
struct Vec {
    T val[N];
    Vec() { for (int i = 0; i < N; i++) val[i] = T(); }
    Vec(T a, T b, T c) { static_assert(N == 3, "3 channels"); val[0] = a; val[1] = b; val[2] = c; }
    T& operator[](int i) { return val[i]; }
    const T& operator[](int i) const { return val[i]; }
    bool operator==(const Vec& o) const { for (int i = 0; i < N; i++) if (val[i] != o.val[i]) return false; return true; }
    bool operator!=(const Vec& o) const { return !(*this == o); }
};
typedef Vec<uchar, 3> Vec3b;

enum { NORM_L2 = 4 };

class Mat {
public:
    int rows = 0, cols = 0;
    size_t step = 0;
    uchar* data = nullptr;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    // external data, not copied, not owned (like cv::Mat(rows, cols, type, void* data, size_t step))
    Mat(int r, int c, int type, void* ext, size_t step_) : rows(r), cols(c), step(step_), data((uchar*)ext), type_(type) {}
    static Mat zeros(int r, int c, int type) {
        Mat m(r, c, type);
        if (m.data) std::memset(m.data, 0, m.step * (size_t)r);
        return m;
    }
    void create(int r, int c, int type) {
        if (r < 0 || c < 0) throw Exception("Mat::create: negative size");
        if (data && r == rows && c == cols && type == type_ && step == (size_t)c * esz(type)) return;
        rows = r; cols = c; type_ = type; step = (size_t)c * esz(type);
        buf_.reset(new uchar[step * (size_t)r + 64], std::default_delete<uchar[]>());
        data = buf_.get();
    }
    int type() const { return type_; }
    int depth() const { return type_ & 7; }
    int channels() const { return (type_ >> 3) + 1; }
    size_t elemSize() const { return esz(type_); }
    Size size() const { return Size(cols, rows); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return rows <= 1 || step == (size_t)cols * elemSize(); }
    template <typename T> T& at(int y, int x) { return *reinterpret_cast<T*>(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    template <typename T> const T& at(int y, int x) const { return *reinterpret_cast<const T*>(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    template <typename T> T* ptr(int y = 0) { return reinterpret_cast<T*>(data + (size_t)y * step); }
    template <typename T> const T* ptr(int y = 0) const { return reinterpret_cast<const T*>(data + (size_t)y * step); }
    Mat operator()(const Rect& r) const {
        if (!(0 <= r.x && 0 <= r.width && r.x + r.width <= cols && 0 <= r.y && 0 <= r.height && r.y + r.height <= rows))
            throw Exception("(-215:Assertion failed) 0 <= roi.x && 0 <= roi.width && roi.x + roi.width <= m.cols && "
                            "0 <= roi.y && 0 <= roi.height && roi.y + roi.height <= m.rows in function 'Mat'");
        Mat v(*this);
        v.data = data + (size_t)r.y * step + (size_t)r.x * elemSize();
        v.rows = r.height; v.cols = r.width;
        return v;
    }
    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int y = 0; y < rows; y++) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
        return m;
    }

private:
    static size_t esz(int type) {
        static const size_t d[8] = {1, 1, 2, 2, 4, 4, 8, 2};
        return d[type & 7] * (size_t)((type >> 3) + 1);
    }
    int type_ = 0;
    std::shared_ptr<uchar> buf_;
};

inline void absdiff(const Mat& a, const Mat& b, Mat& dst) {
    if (a.rows != b.rows || a.cols != b.cols || a.type() != b.type()) throw Exception("absdiff: size/type mismatch");
    if (a.depth() != CV_8U) throw Exception("absdiff stand-in: CV_8U only");
    dst.create(a.rows, a.cols, a.type());
    const int n = a.cols * a.channels();
    for (int y = 0; y < a.rows; y++) {
        const uchar *pa = a.ptr<uchar>(y), *pb = b.ptr<uchar>(y);
        uchar* pd = dst.ptr<uchar>(y);
        for (int i = 0; i < n; i++) pd[i] = (uchar)(pa[i] > pb[i] ? pa[i] - pb[i] : pb[i] - pa[i]);
    }
}

inline double norm(const Mat& a, int normType = NORM_L2) {
    if (normType != NORM_L2 || a.depth() != CV_8U) throw Exception("norm stand-in: NORM_L2 on CV_8U only");
    const int n = a.cols * a.channels();
    uint64_t s = 0;
    for (int y = 0; y < a.rows; y++) {
        const uchar* p = a.ptr<uchar>(y);
        uint32_t r = 0;   // a row of < 2^15 elements cannot overflow 32 bits
        for (int i = 0; i < n; i++) r += (uint32_t)p[i] * p[i];
        s += r;
    }
    return std::sqrt((double)s);
}

inline Scalar mean(const Mat& a) {
    if (a.depth() != CV_8U) throw Exception("mean stand-in: CV_8U only");
    const int cn = a.channels();
    uint64_t s[4] = {0, 0, 0, 0};
    for (int y = 0; y < a.rows; y++) {
        const uchar* p = a.ptr<uchar>(y);
        for (int x = 0; x < a.cols; x++)
            for (int c = 0; c < cn; c++) s[c] += p[x * cn + c];
    }
    const size_t total = (size_t)a.rows * a.cols;
    const double inv = total ? 1.0 / (double)total : 0.0;
    Scalar m;
    for (int c = 0; c < cn; c++) m[c] = (double)s[c] * inv;
    return m;
}

inline void subtract(const Mat& a, const Scalar& sc, Mat& dst) {
    if (a.depth() != CV_8U) throw Exception("subtract stand-in: CV_8U only");
    const int cn = a.channels();
    int q[4];
    for (int c = 0; c < 4; c++) {
        double r = std::nearbyint(sc[c]);   // cvRound: round half to even (default rounding mode)
        q[c] = r < 0 ? 0 : r > 255 ? 255 : (int)r;
    }
    Mat out(a.rows, a.cols, a.type());   // dst may alias a
    for (int y = 0; y < a.rows; y++) {
        const uchar* p = a.ptr<uchar>(y);
        uchar* d = out.ptr<uchar>(y);
        for (int x = 0; x < a.cols; x++)
            for (int c = 0; c < cn; c++) { int v = (int)p[x * cn + c] - q[c]; d[x * cn + c] = (uchar)(v < 0 ? 0 : v); }
    }
    dst = out;
}

inline int borderReflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
    return p;
}

inline void blur(const Mat& src, Mat& dst, Size ksize) {
    if (src.type() != CV_32FC1) throw Exception("blur stand-in: CV_32FC1 only");
    const int w = src.cols, h = src.rows, kx = ksize.width, ky = ksize.height, ax = kx / 2, ay = ky / 2;
    if (kx <= 0 || ky <= 0) throw Exception("blur: bad kernel size");
    Mat out(h, w, CV_32FC1);
    const double scale = 1.0 / ((double)kx * (double)ky);
    std::vector<double> rows((size_t)w * h), sum((size_t)w, 0.0);
    std::vector<float> ext((size_t)w + kx);
    for (int y = 0; y < h; y++) {   // RowSum<float, double>
        const float* s = src.ptr<float>(y);
        for (int i = 0; i < w + kx - 1; i++) ext[i] = s[borderReflect101(i - ax, w)];
        double acc = 0;
        for (int i = 0; i < kx; i++) acc += (double)ext[i];
        rows[(size_t)y * w] = acc;
        for (int x = 1; x < w; x++) { acc += (double)ext[x + kx - 1] - (double)ext[x - 1]; rows[(size_t)y * w + x] = acc; }
    }
    for (int i = 0; i < ky - 1; i++) {   // ColumnSum<double, float>
        const double* r = &rows[(size_t)borderReflect101(i - ay, h) * w];
        for (int x = 0; x < w; x++) sum[x] += r[x];
    }
    for (int y = 0; y < h; y++) {
        const double* rin = &rows[(size_t)borderReflect101(y - ay + ky - 1, h) * w];
        const double* rout = &rows[(size_t)borderReflect101(y - ay, h) * w];
        float* d = out.ptr<float>(y);
        for (int x = 0; x < w; x++) {
            double s0 = sum[x] + rin[x];
            d[x] = (float)(s0 * scale);
            sum[x] = s0 - rout[x];
        }
    }
    dst = out;
}
}  // namespace cv
