/*
 * sr_oracle.h -- CPU ORACLE for the stereo hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a plain-C restatement of the reference's CPU algorithm
 * (Bizilizi/stereo_reconstruction, src/WindowSearch + src/Reconstruction).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may link or call it; the product (libsr_b200.so)
 * never does.
 *
 * Pinning status: the reference's own tests hold no golden vector for this
 * path.  The restatement is pinned against a line-by-line
 * cv2 transliteration of the reference that calls the very OpenCV functions
 * the arithmetic lives in (tests/golden/make_golden.py, run in the build
 * container, vectors committed under tests/golden/).  The extension costs
 * (SAD, NCC, LR check, sub-pixel) have no reference counterpart:
 * "parity unpinned" for those.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * the reference root).
 */
#ifndef SR_ORACLE_H
#define SR_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_OK 0
#define ORC_E_BADARG (-1)
#define ORC_E_SHAPE (-2) /* the reference would throw cv::Exception / read out of bounds */
#define ORC_E_IO (-5)

enum { ORC_COST_SSD_L2 = 0, ORC_COST_SAD = 1, ORC_COST_NCC = 2 };

/* 32-byte vertex, layout of `struct Vertex` (Reconstruction/reconstruction.h:13-22):
 * Eigen Vector4f (16 B, 16-aligned) + Vector4uc (4 B) + 12 B tail padding. */
typedef struct {
    float pos[4];
    uint8_t rgba[4];
    uint8_t pad_[12];
} orc_vertex;

/* --- reference-exact functions --------------------------------------------------------- */

/* BlockSearch::computeDisparityMapRight, WindowSearch/BlockSearch.cpp:88-179.
 * out: h2*w2 doubles (CV_64F), fully written (zeros where the reference leaves zeros).
 * max_block (may be NULL) receives the "max block size" the reference prints (:177). */
int orc_block_search_right(const uint8_t* left, int w1, int h1, size_t pitch1,
                           const uint8_t* right, int w2, int h2, size_t pitch2,
                           int block, int min_disp, int max_disp, double smooth,
                           int var_block, double thres, double* out, int* max_block);

/* Same results as orc_block_search_right for var_block == 0, computed with separable
 * sliding box sums and OpenMP over row chunks ("baseline B" of BASELINE.md).  cost selects
 * ORC_COST_SSD_L2 (reference) or ORC_COST_SAD (extension; smooth is applied to
 * SAD/N the same way).  row0/row1 restrict the computed output rows (smooth pass then only
 * meaningful when row0 == 0); pass 0, h2 for the full image. */
int orc_block_search_right_fast(const uint8_t* left, int w1, int h1, size_t pitch1,
                                const uint8_t* right, int w2, int h2, size_t pitch2,
                                int block, int min_disp, int max_disp, double smooth,
                                int cost, int row0, int row1, double* out, int nthreads);

/* BlockSearch::computeDisparityMapLeft, WindowSearch/BlockSearch.cpp:24-86.  out: h1*w1. */
int orc_block_search_left(const uint8_t* left, int w1, int h1, size_t pitch1,
                          const uint8_t* right, int w2, int h2, size_t pitch2,
                          int block, int max_disp, double smooth, double* out);

/* LinearSearch::computeDisparityMap, WindowSearch/LinearSearch.cpp:10-59.  out: h2*w2.
 * Deviation: candidates are clamped to k < w1 (the reference reads out of bounds). */
int orc_linear_search(const uint8_t* left, int w1, int h1, size_t pitch1,
                      const uint8_t* right, int w2, int h2, size_t pitch2,
                      double smooth, double* out);

/* Reconstruction/reconstruction.cpp:5-18 (cv::blur semantics restated). In place. */
int orc_remove_outliers(float* disp, int w, int h, int ksize, float thr_front, float thr_back);
/* The blur alone (normalised box filter, BORDER_REFLECT_101, anchor k/2). */
int orc_box_blur(const float* src, float* dst, int w, int h, int ksize);
/* reconstruction.cpp:21-27 */
int orc_scale(float* disp, int w, int h, float f);
/* reconstruction.cpp:30-43 */
int orc_disparity_to_depth(const float* disp, float* depth, int w, int h, float focal, float baseline);
/* reconstruction.cpp:152-196 (back-projection loop only) */
int orc_backproject(const uint8_t* bgr, size_t pitch, const float* depth, int w, int h,
                    float fx, float fy, float cx, float cy, orc_vertex* out);
/* reconstruction.cpp:46-69 */
int orc_triangle_valid(const orc_vertex* v, unsigned a, unsigned b, unsigned c, float thr);
/* reconstruction.cpp:72-113: the face list in reference order.  tri: up to 6*(w-1)*(h-1). */
int orc_mesh_faces(const orc_vertex* v, int w, int h, float thr, uint32_t* tri, uint32_t* n_faces);
/* reconstruction.cpp:116-146: the COFF text file. */
int orc_write_off(const char* path, const orc_vertex* v, int w, int h, float thr);

/* utils.cpp:123-168 (evaldisp): counts are exact, avg_err is the serial float sum. */
typedef struct { int n, bad, invalid; float serr; float bad_pct, invalid_pct, total_bad_pct, avg_err; } orc_eval;
int orc_evaldisp(const float* disp, const float* gt, const uint8_t* mask, int w, int h,
                 float badthresh, float maxdisp, int rounddisp, orc_eval* out);
/* utils.cpp:8-20 */
float orc_average_disparity(const float* disp, int w, int h);

/* rectification.cpp:82-87: cv::warpPerspective(src, dst, M, size, INTER_NEAREST) on CV_64F,
 * constant-0 border.  M is the 3x3 row-major matrix PASSED to warpPerspective (H_.inv());
 * like OpenCV it is inverted internally (no WARP_INVERSE_MAP). */
int orc_warp_nearest_f64(const double* src, int sw, int sh, double* dst, int dw, int dh, const double M[9]);
int orc_warp_linear_bgr(const uint8_t* src, int sw, int sh, size_t spitch, uint8_t* dst, int dw, int dh, size_t dpitch,
                        const double M[9]);

/* main.cpp:41-50: CV_64F -> 8-bit (rint half-even, saturate) -> float32. */
int orc_quantize_u8(const double* disp, int n, float* out);

/* --- extensions (no reference counterpart; definitions of SURVEY.md 8c) ------------------ */

/* Right-referenced search with the reference's window geometry, candidate order, break rule
 * and first-minimum rule, cost in {SSD (integer compare), SAD, NCC}.
 *   disp      h2*w2 int32 winner d* (0 for black / no-candidate pixels)
 *   cost_i    (optional) winner integer cost for SSD/SAD (int64), -1 when none
 *   subpix    (optional) float d* + parabola offset (fp32), equals d* when not applicable
 *   left_ref  0: right image is the reference (d>=0 towards +x in left);
 *             1: mirrored search with the left image as reference (for the LR check) */
int orc_search_ext(const uint8_t* left, int w1, int h1, size_t pitch1,
                   const uint8_t* right, int w2, int h2, size_t pitch2,
                   int block, int min_disp, int max_disp, int cost, int left_ref,
                   int32_t* disp, int64_t* cost_i, float* subpix);

/* orc_search_ext for SSD / SAD through sliding box sums + OpenMP (whole MiddEval3 pairs in seconds); same results. */
int orc_search_ext_fast(const uint8_t* left, int w1, int h1, size_t pitch1,
                        const uint8_t* right, int w2, int h2, size_t pitch2,
                        int block, int min_disp, int max_disp, int cost, int left_ref,
                        int32_t* disp, int64_t* cost_i, float* subpix, int nthreads);

/* LR check: disp_r (h2*w2) invalidated (set to 0) where inconsistent with disp_l (h1*w1). */
int orc_lr_check(const int32_t* disp_r, int w2, int h2, const int32_t* disp_l, int w1, int h1,
                 int32_t* out);

/* Middlebury PFM (Pf, bottom-to-top rows, scale -1.0, little endian) and ASCII PLY writers. */
int orc_write_pfm(const char* path, const float* data, int w, int h);
int orc_write_ply(const char* path, const orc_vertex* v, int w, int h);

/* Deterministic synthetic pair generator of SURVEY.md 8d (splitmix64 counter hash).
 * left/right: h*w*3 BGR, tightly packed.  quant != 0 quantises R to 4 levels (tie stress). */
void orc_synth_pair(uint64_t seed, uint64_t pair, int w, int h, int min_disp, int max_disp,
                    int quant, uint8_t* left, uint8_t* right);

#ifdef __cplusplus
}
#endif
#endif
