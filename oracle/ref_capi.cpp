// ref_capi.cpp -- TEST INFRASTRUCTURE ONLY.  C entry points over the reference's OWN hot-path code: the classes and
// functions below are compiled, unchanged, from /root/reference/src/{WindowSearch/BlockSearch.cpp,
// WindowSearch/LinearSearch.cpp, Reconstruction/reconstruction.cpp} against the header stand-ins in oracle/refstub
// (recipe: oracle/Makefile, target _ref -> oracle/_ref/libsr_ref.so; the .so is git-ignored and travels to the GPU
// box, the reference sources never enter this repository).  Used by tests/ to pin oracle/sr_oracle.c to the
// reference itself and by bench.py --impl reference / cpu_baseline as the timed CPU reference (kind "reference").
#include <unistd.h>

#include <climits>
#include <cstdio>
#include <cstring>
#include <sstream>
#include <string>

#include "WindowSearch/BlockSearch.h"
#include "WindowSearch/LinearSearch.h"
#include "Reconstruction/reconstruction.h"
#ifdef _OPENMP
#include <omp.h>
#endif

static_assert(sizeof(Vertex) == 32 && alignof(Vertex) == 16, "struct Vertex: Vector4f + Vector4uc = 32 bytes (reconstruction.h:13-22)");

#define REF_OK 0
#define REF_E_BADARG (-1)
#define REF_E_SHAPE (-2)   // cv::Exception (ROI outside the image)
#define REF_E_IO (-5)

namespace {
thread_local std::string g_err;
struct CoutCapture {   // the reference prints from inside its compute methods (BlockSearch.cpp:177, LinearSearch.cpp:16-17)
    std::streambuf* old;
    std::ostringstream ss;
    CoutCapture() : old(std::cout.rdbuf(ss.rdbuf())) {}
    ~CoutCapture() { std::cout.rdbuf(old); }
};
cv::Mat wrap8uc3(const uint8_t* p, int w, int h, size_t pitch) { return cv::Mat(h, w, CV_8UC3, (void*)p, pitch); }
void copy_out(const cv::Mat& m, double* out) {
    for (int y = 0; y < m.rows; y++) std::memcpy(out + (size_t)y * m.cols, m.ptr<double>(y), sizeof(double) * (size_t)m.cols);
}
}  // namespace

extern "C" {

const char* ref_last_error() { return g_err.c_str(); }
int ref_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// BlockSearch(left, right, block, minD, maxD).computeDisparityMapRight(smooth, varBlock, thres); out: h2*w2 doubles.
int ref_block_search_right(const uint8_t* left, int w1, int h1, size_t pitch1, const uint8_t* right, int w2, int h2, size_t pitch2,
                           int block, int min_disp, int max_disp, double smooth, int var_block, double thres, double* out,
                           int* max_block) {
    try {
        cv::Mat L = wrap8uc3(left, w1, h1, pitch1), R = wrap8uc3(right, w2, h2, pitch2);
        CoutCapture cap;
        BlockSearch bs(L, R, block, min_disp, max_disp);
        cv::Mat d = bs.computeDisparityMapRight(smooth, var_block != 0, thres);
        copy_out(d, out);
        if (max_block) { int mb = 0; std::sscanf(cap.ss.str().c_str(), "max block size: %d", &mb); *max_block = mb; }
        return REF_OK;
    } catch (const cv::Exception& e) { g_err = e.what(); return REF_E_SHAPE; }
}

// The same on output rows [row0, row1) only, smoothFactor == 1 (no dependency between rows then): the rows are cut
// into bands, each band is a BlockSearch of its own on a crop with the window's halo rows (the clamps up/down of
// BlockSearch.cpp:118-119 see the same borders as in the whole image), bands run on `nthreads` OpenMP threads.  The
// reference has no threads; this is the harness using "all the host threads it can" on unchanged reference code.
// Rows outside [row0,row1) of `out` are left untouched.  Returns the thread count used (> 0) or an error (< 0).
int ref_block_search_right_rows(const uint8_t* left, int w1, int h1, size_t pitch1, const uint8_t* right, int w2, int h2, size_t pitch2,
                                int block, int min_disp, int max_disp, int row0, int row1, int nthreads, double* out) {
    if (h1 != h2 || row0 < 0 || row1 > h2 || row0 >= row1) { g_err = "rows: equal heights and a non-empty row range"; return REF_E_BADARG; }
    const int half = (block - 1) / 2;
    int nt = nthreads > 0 ? nthreads : ref_max_threads();
    if (nt > row1 - row0) nt = row1 - row0;
    int rc = REF_OK, used = 1;
    CoutCapture cap;   // "max block size" once per band (BlockSearch.cpp:177)
#pragma omp parallel num_threads(nt)
    {
#ifdef _OPENMP
        const int t = omp_get_thread_num(), T = omp_get_num_threads();
#else
        const int t = 0, T = 1;
#endif
        if (t == 0) used = T;
        const int a = row0 + (int)((long)(row1 - row0) * t / T), b = row0 + (int)((long)(row1 - row0) * (t + 1) / T);
        if (a < b) {
            const int c0 = a - half < 0 ? 0 : a - half, c1 = b + half > h2 ? h2 : b + half;
            try {
                cv::Mat L = wrap8uc3(left + (size_t)c0 * pitch1, w1, c1 - c0, pitch1), R = wrap8uc3(right + (size_t)c0 * pitch2, w2, c1 - c0, pitch2);
                BlockSearch bs(L, R, block, min_disp, max_disp);
                cv::Mat d = bs.computeDisparityMapRight(1.0, false, 19.0);   // prints "max block size" once per band
                for (int y = a; y < b; y++) std::memcpy(out + (size_t)y * w2, d.ptr<double>(y - c0), sizeof(double) * (size_t)w2);
            } catch (const cv::Exception& e) {
#pragma omp critical
                { g_err = e.what(); rc = REF_E_SHAPE; }
            }
        }
    }
    return rc == REF_OK ? used : rc;
}

// BlockSearch(...).computeDisparityMapLeft(smooth); out: h1*w1 doubles.
int ref_block_search_left(const uint8_t* left, int w1, int h1, size_t pitch1, const uint8_t* right, int w2, int h2, size_t pitch2,
                          int block, int max_disp, double smooth, double* out) {
    try {
        cv::Mat L = wrap8uc3(left, w1, h1, pitch1), R = wrap8uc3(right, w2, h2, pitch2);
        BlockSearch bs(L, R, block, 0, max_disp);
        cv::Mat d = bs.computeDisparityMapLeft(smooth);
        copy_out(d, out);
        return REF_OK;
    } catch (const cv::Exception& e) { g_err = e.what(); return REF_E_SHAPE; }
}

// LinearSearch(left, right).computeDisparityMap(smooth); out: h2*w2 doubles.  The reference reads left(i, k) for k up
// to j+199 without a bound (LinearSearch.cpp:32-33): the caller must pass w1 >= w2 + 199 (and h1 >= h2).
int ref_linear_search(const uint8_t* left, int w1, int h1, size_t pitch1, const uint8_t* right, int w2, int h2, size_t pitch2,
                      double smooth, double* out, char* printed, int printed_cap) {
    if (w1 < w2 + 199 || h1 < h2) { g_err = "linear: the reference reads out of bounds for this shape"; return REF_E_SHAPE; }
    cv::Mat L = wrap8uc3(left, w1, h1, pitch1), R = wrap8uc3(right, w2, h2, pitch2);
    CoutCapture cap;
    LinearSearch ls(L, R);
    cv::Mat d = ls.computeDisparityMap(smooth);
    copy_out(d, out);
    if (printed && printed_cap > 0) { std::strncpy(printed, cap.ss.str().c_str(), (size_t)printed_cap - 1); printed[printed_cap - 1] = 0; }
    return REF_OK;
}

int ref_remove_outliers(float* disp, int w, int h, int ksize, float thr_front, float thr_back) {
    cv::Mat m(h, w, CV_32FC1, disp, sizeof(float) * (size_t)w);
    removeDisparityOutliers(m, ksize, thr_front, thr_back);
    return REF_OK;
}
int ref_scale(float* disp, int w, int h, float f) {
    cv::Mat m(h, w, CV_32FC1, disp, sizeof(float) * (size_t)w);
    scaleDisparityMap(m, f);
    return REF_OK;
}
int ref_disparity_to_depth(const float* disp, float* depth, int w, int h, float focal, float baseline) {
    cv::Mat m(h, w, CV_32FC1, (void*)disp, sizeof(float) * (size_t)w);
    cv::Mat d = convertDisparityToDepth(m, focal, baseline);
    for (int y = 0; y < h; y++) std::memcpy(depth + (size_t)y * w, d.ptr<float>(y), sizeof(float) * (size_t)w);
    return REF_OK;
}
int ref_triangle_valid(const void* vertices, unsigned a, unsigned b, unsigned c, float thr) {
    return CheckTriangularValidity((Vertex*)vertices, a, b, c, thr) ? 1 : 0;
}
int ref_write_mesh(const void* vertices, unsigned w, unsigned h, const char* path, float thr) {
    return WriteMesh((Vertex*)vertices, w, h, std::string(path), thr) ? REF_OK : REF_E_IO;
}
// reconstruction(bgr, depth, K, thr): the only output is the file "../../results/reconstruction_mesh.off" relative to
// the working directory (reconstruction.cpp:199-204), so the call runs with the working directory set to `cwd`
// (the caller provides <root>/a/b with <root>/results existing).  K: fx, fy, cx, cy into an Eigen-style Matrix3f.
int ref_reconstruction(const uint8_t* bgr, size_t pitch, const float* depth, int w, int h, float fx, float fy, float cx, float cy,
                       float thr, const char* cwd) {
    char old[PATH_MAX];
    if (!getcwd(old, sizeof old) || chdir(cwd) != 0) { g_err = "chdir failed"; return REF_E_IO; }
    cv::Mat img = wrap8uc3(bgr, w, h, pitch);
    cv::Mat dep(h, w, CV_32FC1, (void*)depth, sizeof(float) * (size_t)w);
    Matrix3f K = Matrix3f::Zero();
    K(0, 0) = fx; K(1, 1) = fy; K(0, 2) = cx; K(1, 2) = cy; K(2, 2) = 1.f;
    CoutCapture cap;
    reconstruction(img, dep, K, thr);
    int rc = cap.ss.str().find("Failed to write mesh") == std::string::npos ? REF_OK : REF_E_IO;
    if (chdir(old) != 0) rc = REF_E_IO;
    return rc;
}

// ---- the stand-in's OpenCV functions, exported so that tests can hold each against cv2 ----
int ref_cv_absdiff_norm(const uint8_t* a, const uint8_t* b, int w, int h, uint8_t* diff, double* nrm) {
    cv::Mat A = wrap8uc3(a, w, h, 3 * (size_t)w), B = wrap8uc3(b, w, h, 3 * (size_t)w), D;
    cv::absdiff(A, B, D);
    for (int y = 0; y < h; y++) std::memcpy(diff + (size_t)y * 3 * w, D.ptr<uint8_t>(y), 3 * (size_t)w);
    *nrm = cv::norm(D, cv::NORM_L2);
    return REF_OK;
}
int ref_cv_mean_subtract(const uint8_t* a, int w, int h, double mean[3], uint8_t* sub, double* nrm) {
    cv::Mat A = wrap8uc3(a, w, h, 3 * (size_t)w), S;
    cv::Scalar m = cv::mean(A);
    for (int c = 0; c < 3; c++) mean[c] = m[c];
    cv::subtract(A, m, S);
    for (int y = 0; y < h; y++) std::memcpy(sub + (size_t)y * 3 * w, S.ptr<uint8_t>(y), 3 * (size_t)w);
    *nrm = cv::norm(S, cv::NORM_L2);
    return REF_OK;
}
int ref_cv_blur(const float* src, float* dst, int w, int h, int k) {
    cv::Mat S(h, w, CV_32FC1, (void*)src, sizeof(float) * (size_t)w), D;
    cv::blur(S, D, cv::Size(k, k));
    for (int y = 0; y < h; y++) std::memcpy(dst + (size_t)y * w, D.ptr<float>(y), sizeof(float) * (size_t)w);
    return REF_OK;
}
int ref_roi_throws(int w, int h, int x, int y, int rw, int rh) {
    cv::Mat m = cv::Mat::zeros(h, w, CV_8UC3);
    try { cv::Mat v = m(cv::Rect(x, y, rw, rh)); (void)v; return 0; } catch (const cv::Exception&) { return 1; }
}
}  // extern "C"
