/*
 * sr_oracle.c -- CPU ORACLE (test infrastructure, never linked into the product).
 * See sr_oracle.h for scope and pinning status.  Compile with
 *   gcc -O2 -ffp-contract=off -fopenmp -fPIC -shared
 * (-ffp-contract=off: the reference is plain C++ without FMA contraction on x86-64).
 */
#include "sr_oracle.h"
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }
static inline const uint8_t* px(const uint8_t* img, size_t pitch, int y, int x) { return img + (size_t)y * pitch + 3 * (size_t)x; }
static inline int is_black(const uint8_t* p) { return p[0] == 0 && p[1] == 0 && p[2] == 0; }

/* cv::norm(absdiff(A,B), NORM_L2)^2 over a ww x wh window of 3-channel pixels: exact integer
 * (OpenCV accumulates u8 squares in int blocks and sums them in double: exact below 2^53). */
static int64_t window_ssd(const uint8_t* a, size_t pa, int ay, int ax,
                          const uint8_t* b, size_t pb, int by, int bx, int ww, int wh) {
    int64_t s = 0;
    for (int r = 0; r < wh; r++) {
        const uint8_t* ra = px(a, pa, ay + r, ax);
        const uint8_t* rb = px(b, pb, by + r, bx);
        for (int i = 0; i < 3 * ww; i++) {
            int d = (int)ra[i] - (int)rb[i];
            s += d * d;
        }
    }
    return s;
}
static int64_t window_sad(const uint8_t* a, size_t pa, int ay, int ax,
                          const uint8_t* b, size_t pb, int by, int bx, int ww, int wh) {
    int64_t s = 0;
    for (int r = 0; r < wh; r++) {
        const uint8_t* ra = px(a, pa, ay + r, ax);
        const uint8_t* rb = px(b, pb, by + r, bx);
        for (int i = 0; i < 3 * ww; i++) s += abs((int)ra[i] - (int)rb[i]);
    }
    return s;
}

/* norm(rightWindow - mean(rightWindow)) with OpenCV's u8 semantics, BlockSearch.cpp:125-129:
 * cv::mean -> per-channel double mean; cv::subtract(u8 window, Scalar) -> as executed by
 * cv2 4.13 (the only OpenCV this image has; the reference leaves its version unpinned): the
 * scalar is converted to the source depth with saturate_cast<uchar> (cvRound, half to even)
 * and subtracted with u8 saturation, i.e. q = clamp(p - rint(mean), 0, 255) -- checked
 * value by value in tests/golden (subtract_mean.npz); it differs from a float-domain
 * rint(p - mean) only when mean is exactly k+0.5.  cv::norm L2 -> sqrt of the integer sum. */
static double window_dev_norm(const uint8_t* img, size_t pitch, int y0, int x0, int ww, int wh) {
    double mean[3] = {0, 0, 0};
    int64_t sum[3] = {0, 0, 0};
    for (int r = 0; r < wh; r++) {
        const uint8_t* p = px(img, pitch, y0 + r, x0);
        for (int i = 0; i < ww; i++) { sum[0] += p[3 * i]; sum[1] += p[3 * i + 1]; sum[2] += p[3 * i + 2]; }
    }
    int64_t n = (int64_t)ww * wh;
    for (int c = 0; c < 3; c++) mean[c] = n ? (double)sum[c] / (double)n : 0.0;
    int64_t ss = 0;
    long mi[3];
    for (int c = 0; c < 3; c++) { mi[c] = lrint(mean[c]); if (mi[c] < 0) mi[c] = 0; if (mi[c] > 255) mi[c] = 255; } /* FE_TONEAREST */
    for (int r = 0; r < wh; r++) {
        const uint8_t* p = px(img, pitch, y0 + r, x0);
        for (int i = 0; i < ww; i++)
            for (int c = 0; c < 3; c++) {
                long q = (long)p[3 * i + c] - mi[c];
                if (q < 0) q = 0;
                ss += q * q;
            }
    }
    return sqrt((double)ss);
}

/* The h1<h2 case: the left ROI may run below the left image -> cv::Exception in the reference. */
static int right_rows_fit(int h1, int h2, int half) {
    int height = imin(h1, h2);
    for (int y = imax(0, height - half - 2); y < height; y++) {
        int down = imin(h2 - y - 1, half);
        if (y + down > h1) return 0;
    }
    return 1;
}

int orc_block_search_right(const uint8_t* left, int w1, int h1, size_t pitch1,
                           const uint8_t* right, int w2, int h2, size_t pitch2,
                           int block, int min_disp, int max_disp, double smooth,
                           int var_block, double thres, double* out, int* max_block) {
    if (!left || !right || !out || w1 <= 0 || h1 <= 0 || w2 <= 0 || h2 <= 0 || block < 1) return ORC_E_BADARG;
    if (min_disp < 0) return ORC_E_SHAPE; /* left ROI starts at a negative column */
    int height = imin(h1, h2);
    if (!var_block && !right_rows_fit(h1, h2, (block - 1) / 2)) return ORC_E_SHAPE;
    memset(out, 0, sizeof(double) * (size_t)w2 * h2);             /* :97 */
    int maxBlock = block;                                          /* :98 */
    for (int y = 0; y < height; y++) {                             /* :100 */
        for (int x = 0; x < w2; x++) {                             /* :102 */
            if (is_black(px(right, pitch2, y, x))) continue;       /* :105-108 */
            int minX = 0;                                          /* :111 */
            double best = DBL_MAX;                                 /* :112 */
            int bs = block;                                        /* :114 */
            int l = imin(x, (bs - 1) / 2), r = imin(w2 - x - 1, (bs - 1) / 2);
            int u = imin(y, (bs - 1) / 2), dn = imin(h2 - y - 1, (bs - 1) / 2); /* :116-119 */
            if (var_block) {                                       /* :125-142 */
                while (window_dev_norm(right, pitch2, y - u, x - l, l + r, u + dn) < thres) {
                    int pl = l, pr = r, pu = u, pd = dn;
                    bs += 4;
                    l = imin(x, (bs - 1) / 2); r = imin(w2 - x - 1, (bs - 1) / 2);
                    u = imin(y, (bs - 1) / 2); dn = imin(h2 - y - 1, (bs - 1) / 2);
                    /* deviation: the reference never terminates once the window spans the
                     * whole image; stop at the first growth step that changes nothing. */
                    if (l == pl && r == pr && u == pu && dn == pd) break;
                }
                if (y + dn > h1) return ORC_E_SHAPE;
            }
            if (bs > maxBlock) maxBlock = bs;                      /* :144-145 */
            int ww = l + r, wh = u + dn;
            for (int cx = x + min_disp; cx < x + max_disp; cx++) { /* :147 */
                if (cx + r >= w1) break;                           /* :148-149 */
                int64_t ssd = window_ssd(left, pitch1, y - u, cx - l, right, pitch2, y - u, x - l, ww, wh);
                double dist = sqrt((double)ssd) / (double)(ww * wh); /* :156-158 (0/0 -> NaN for empty windows) */
                if (y >= 1 && out[(size_t)(y - 1) * w2 + x] == (double)(x - cx)) dist *= smooth; /* :160-162 */
                if (x >= 1 && out[(size_t)y * w2 + x - 1] == (double)(x - cx)) dist *= smooth;   /* :163-165 */
                if (dist < best) { minX = cx; best = dist; }       /* :168-171 */
            }
            out[(size_t)y * w2 + x] = (double)(minX - x);          /* :174 */
        }
    }
    if (max_block) *max_block = maxBlock;
    return ORC_OK;
}

int orc_block_search_left(const uint8_t* left, int w1, int h1, size_t pitch1,
                          const uint8_t* right, int w2, int h2, size_t pitch2,
                          int block, int max_disp, double smooth, double* out) {
    if (!left || !right || !out || w1 <= 0 || h1 <= 0 || w2 <= 0 || h2 <= 0 || block < 1) return ORC_E_BADARG;
    if ((block & 1) == 0) return ORC_E_SHAPE; /* even block: the ROI x-h+B overruns at x = w1-h-1 */
    int height = imin(h1, h2);
    int half = (block - 1) / 2;                                    /* :31 */
    memset(out, 0, sizeof(double) * (size_t)w1 * h1);              /* :33 */
    for (int y = half; y < height - half; y++) {                   /* :36 */
        for (int x = half; x < w1 - half; x++) {                   /* :38 */
            if (is_black(px(left, pitch1, y, x))) continue;        /* :41-44 */
            int minX = 0;                                          /* :50 */
            double best = DBL_MAX;
            for (int cx = x - max_disp; cx < x; cx++) {            /* :53 */
                if (cx < half || cx >= w2 - half) continue;        /* :55-57 */
                int64_t ssd = window_ssd(left, pitch1, y - half, x - half, right, pitch2, y - half, cx - half, block, block);
                double dist = sqrt((double)ssd);                   /* :64-66 */
                if (y >= 1 && out[(size_t)(y - 1) * w1 + x] == (double)(x - cx)) dist *= smooth; /* :68-70 */
                if (x >= 1 && out[(size_t)y * w1 + x - 1] == (double)(x - cx)) dist *= smooth;   /* :71-73 */
                if (dist < best) { minX = cx; best = dist; }       /* :76-79 */
            }
            out[(size_t)y * w1 + x] = (double)(x - minX);          /* :82 */
        }
    }
    return ORC_OK;
}

int orc_linear_search(const uint8_t* left, int w1, int h1, size_t pitch1,
                      const uint8_t* right, int w2, int h2, size_t pitch2,
                      double smooth, double* out) {
    if (!left || !right || !out || w1 <= 0 || h1 <= 0 || w2 <= 0 || h2 <= 0) return ORC_E_BADARG;
    if (h1 < h2 || w1 < w2) return ORC_E_SHAPE; /* leftImage.at(i,j) with i<h2, j<w2 (:24) */
    memset(out, 0, sizeof(double) * (size_t)w2 * h2);              /* :19 */
    for (int i = 0; i < h2; i++) {
        for (int j = 0; j < w2; j++) {
            if (is_black(px(left, pitch1, i, j))) continue;        /* :24-27 (LEFT pixel tested) */
            int col = 0;
            const uint8_t* pr = px(right, pitch2, i, j);
            double best = DBL_MAX;
            for (int k = j; k < j + 200 && k < w1; k++) {          /* :32, clamped to k<w1 */
                const uint8_t* pl = px(left, pitch1, i, k);
                int d0 = pr[0] - pl[0], d1 = pr[1] - pl[1], d2 = pr[2] - pl[2];
                double dist = sqrt((double)(d0 * d0) + (double)(d1 * d1) + (double)(d2 * d2)); /* :36-37 */
                if (i >= 1 && out[(size_t)(i - 1) * w2 + j] == (double)(j - k)) dist *= smooth;
                if (j >= 1 && out[(size_t)i * w2 + j - 1] == (double)(j - k)) dist *= smooth;
                if (dist < best) { col = k; best = dist; }
            }
            out[(size_t)i * w2 + j] = (double)(col - j);           /* :53 */
        }
    }
    return ORC_OK;
}

/* ---- baseline B: the same Right search with sliding box sums + OpenMP -------------------- */

/* Raster pass of the smoothFactor rule (BlockSearch.cpp:160-165) given, per pixel, the
 * unbiased cost at d=0 (c0, <0 when d=0 is not a candidate) and the first minimum over d>=1
 * (cr, dr; cr<0 when none).  The bias compares D(neighbour) with (double)(x-cx) = -d, so it
 * can only fire for d == 0 next to a stored 0 (or for the unreachable no-candidate case). */
static void smooth_pass_right(const int64_t* c0, const int64_t* cr, const int32_t* dr, const uint8_t* skip,
                              int w, int row0, int row1, const int* nwin, double smooth, double* out) {
    for (int y = row0; y < row1; y++)
        for (int x = 0; x < w; x++) {
            size_t i = (size_t)y * w + x;
            if (skip[i]) { out[i] = 0.0; continue; }
            double n = (double)nwin[i];
            if (c0[i] < 0) { out[i] = cr[i] < 0 ? (double)(0 - x) : (double)dr[i]; continue; }
            double d0 = sqrt((double)c0[i]) / n;
            if (y >= 1 && out[i - w] == 0.0) d0 *= smooth;
            if (x >= 1 && out[i - 1] == 0.0) d0 *= smooth;
            double best = DBL_MAX; int d = 0; /* minimumCorrespondX = 0 -> D = -x if nothing beats DBL_MAX */
            int any = 0;
            if (d0 < best) { best = d0; d = 0; any = 1; }
            if (cr[i] >= 0) { double dd = sqrt((double)cr[i]) / n; if (dd < best) { best = dd; d = dr[i]; any = 1; } }
            out[i] = any ? (double)d : (double)(0 - x);
        }
}

int orc_block_search_right_fast(const uint8_t* left, int w1, int h1, size_t pitch1,
                                const uint8_t* right, int w2, int h2, size_t pitch2,
                                int block, int min_disp, int max_disp, double smooth,
                                int cost, int row0, int row1, double* out, int nthreads) {
    if (!left || !right || !out || w1 <= 0 || h1 <= 0 || w2 <= 0 || h2 <= 0 || block < 1) return ORC_E_BADARG;
    if (cost != ORC_COST_SSD_L2 && cost != ORC_COST_SAD) return ORC_E_BADARG;
    if (min_disp < 0) return ORC_E_SHAPE;
    int half = (block - 1) / 2;
    if (!right_rows_fit(h1, h2, half)) return ORC_E_SHAPE;
    int height = imin(h1, h2);
    if (row0 < 0) row0 = 0;
    if (row1 > h2) row1 = h2;
    size_t npx = (size_t)w2 * h2;
    int64_t* c0 = (int64_t*)malloc(npx * 8); int64_t* cr = (int64_t*)malloc(npx * 8);
    int32_t* dr = (int32_t*)malloc(npx * 4); int* nwin = (int*)malloc(npx * 4);
    uint8_t* skip = (uint8_t*)malloc(npx);
    if (!c0 || !cr || !dr || !nwin || !skip) { free(c0); free(cr); free(dr); free(nwin); free(skip); return ORC_E_BADARG; }
    for (size_t i = 0; i < npx; i++) { c0[i] = -1; cr[i] = -1; dr[i] = 0; nwin[i] = 0; skip[i] = 1; }
    const int CH = 16;
    int r1 = imin(row1, height);
    int nchunks = (r1 - row0 + CH - 1) / CH;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
    nthreads = 1;
#endif
#pragma omp parallel num_threads(nthreads)
    {
        int64_t* col = (int64_t*)malloc(sizeof(int64_t) * (size_t)(w2 + 1));
        int64_t* pre = (int64_t*)malloc(sizeof(int64_t) * (size_t)(w2 + 1));
#pragma omp for schedule(dynamic, 1)
        for (int ch = 0; ch < nchunks; ch++) {
            int ya = row0 + ch * CH, yb = imin(ya + CH, r1);
            for (int y = ya; y < yb; y++)
                for (int x = 0; x < w2; x++) {
                    size_t i = (size_t)y * w2 + x;
                    skip[i] = (uint8_t)is_black(px(right, pitch2, y, x));
                    int l = imin(x, half), r = imin(w2 - x - 1, half), u = imin(y, half), dn = imin(h2 - y - 1, half);
                    nwin[i] = (l + r) * (u + dn);
                }
            for (int d = min_disp; d < max_disp; d++) {
                /* columns c of the right image whose left partner c+d exists */
                int cmax = imin(w2 - 1, w1 - d); /* pointwise cost defined for c in [0,cmax) */
                if (cmax <= 0) break;
                for (int y = ya; y < yb; y++) {
                    int u = imin(y, half), dn = imin(h2 - y - 1, half);
                    int ra = y - u, rb = y + dn; /* rows [ra, rb) */
                    if (y == ya) {
                        for (int c = 0; c < cmax; c++) col[c] = 0;
                        for (int r = ra; r < rb; r++) {
                            const uint8_t* pl = px(left, pitch1, r, d); const uint8_t* pr = px(right, pitch2, r, 0);
                            for (int c = 0; c < cmax; c++) {
                                int a = pl[3*c] - pr[3*c], b = pl[3*c+1] - pr[3*c+1], e = pl[3*c+2] - pr[3*c+2];
                                col[c] += cost == ORC_COST_SAD ? abs(a) + abs(b) + abs(e) : a*a + b*b + e*e;
                            }
                        }
                    } else {
                        int pu = imin(y - 1, half), pdn = imin(h2 - y, half);
                        int pra = y - 1 - pu, prb = y - 1 + pdn;
                        for (int r = pra; r < ra; r++) { /* rows leaving */
                            const uint8_t* pl = px(left, pitch1, r, d); const uint8_t* pr = px(right, pitch2, r, 0);
                            for (int c = 0; c < cmax; c++) {
                                int a = pl[3*c] - pr[3*c], b = pl[3*c+1] - pr[3*c+1], e = pl[3*c+2] - pr[3*c+2];
                                col[c] -= cost == ORC_COST_SAD ? abs(a) + abs(b) + abs(e) : a*a + b*b + e*e;
                            }
                        }
                        for (int r = prb; r < rb; r++) { /* rows entering */
                            const uint8_t* pl = px(left, pitch1, r, d); const uint8_t* pr = px(right, pitch2, r, 0);
                            for (int c = 0; c < cmax; c++) {
                                int a = pl[3*c] - pr[3*c], b = pl[3*c+1] - pr[3*c+1], e = pl[3*c+2] - pr[3*c+2];
                                col[c] += cost == ORC_COST_SAD ? abs(a) + abs(b) + abs(e) : a*a + b*b + e*e;
                            }
                        }
                    }
                    pre[0] = 0;
                    for (int c = 0; c < cmax; c++) pre[c + 1] = pre[c] + col[c];
                    for (int x = 0; x < w2; x++) {
                        int r = imin(w2 - x - 1, half);
                        if (x + d + r >= w1) break; /* :148 (monotone in x for fixed d) */
                        int l = imin(x, half);
                        size_t i = (size_t)y * w2 + x;
                        int64_t c = pre[x + r] - pre[x - l];
                        if (d == 0) c0[i] = c;
                        else if (cr[i] < 0 || c < cr[i]) { cr[i] = c; dr[i] = d; }
                    }
                }
            }
        }
        free(col); free(pre);
    }
    /* min_disp > 0: there is no d=0 candidate, cr/dr already hold the first minimum. */
    smooth_pass_right(c0, cr, dr, skip, w2, row0, r1, nwin, smooth, out);
    for (int y = imax(r1, row0); y < row1; y++) for (int x = 0; x < w2; x++) out[(size_t)y * w2 + x] = 0.0;
    free(c0); free(cr); free(dr); free(nwin); free(skip);
    return ORC_OK;
}

/* ---- Reconstruction ------------------------------------------------------------------- */

/* cv::borderInterpolate(p, len, BORDER_REFLECT_101) */
static int reflect101(int p, int len) {
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        if (p < 0) p = -p; else p = 2 * (len - 1) - p;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

/* cv::blur(src32f, dst32f, Size(k,k)): normalised box filter, anchor (k/2,k/2), REFLECT_101.
 * OpenCV (imgproc/box_filter) runs RowSum<float,double> (first window summed left to right,
 * then s += new - old) and ColumnSum<double,float> (SUM over k rows, out = (float)(SUM*scale),
 * SUM -= oldest row); double accumulators, scale = 1.0/(k*k).  Restated in that order. */
int orc_box_blur(const float* src, float* dst, int w, int h, int ksize) {
    if (!src || !dst || w <= 0 || h <= 0 || ksize <= 0) return ORC_E_BADARG;
    int a = ksize / 2;
    double scale = 1.0 / ((double)ksize * (double)ksize);
    double* rows = (double*)malloc(sizeof(double) * (size_t)w * h); /* horizontal sums per source row */
    double* sum = (double*)calloc((size_t)w, sizeof(double));
    float* ext = (float*)malloc(sizeof(float) * (size_t)(w + ksize));
    if (!rows || !sum || !ext) { free(rows); free(sum); free(ext); return ORC_E_BADARG; }
    for (int y = 0; y < h; y++) {
        for (int i = 0; i < w + ksize - 1; i++) ext[i] = src[(size_t)y * w + reflect101(i - a, w)];
        double s = 0;
        for (int i = 0; i < ksize; i++) s += (double)ext[i];
        rows[(size_t)y * w] = s;
        for (int x = 1; x < w; x++) { s += (double)ext[x + ksize - 1] - (double)ext[x - 1]; rows[(size_t)y * w + x] = s; }
    }
    /* vertical: rows y-a .. y-a+k-1 (reflected) */
    for (int i = 0; i < ksize - 1; i++) {
        const double* r = rows + (size_t)reflect101(i - a, h) * w;
        for (int x = 0; x < w; x++) sum[x] += r[x];
    }
    for (int y = 0; y < h; y++) {
        const double* rin = rows + (size_t)reflect101(y - a + ksize - 1, h) * w;
        const double* rout = rows + (size_t)reflect101(y - a, h) * w;
        for (int x = 0; x < w; x++) {
            double s0 = sum[x] + rin[x];
            dst[(size_t)y * w + x] = (float)(s0 * scale);
            sum[x] = s0 - rout[x];
        }
    }
    free(rows); free(sum); free(ext);
    return ORC_OK;
}

int orc_remove_outliers(float* disp, int w, int h, int ksize, float thr_front, float thr_back) {
    float* b = (float*)malloc(sizeof(float) * (size_t)w * h);
    if (!b) return ORC_E_BADARG;
    int rc = orc_box_blur(disp, b, w, h, ksize);                   /* reconstruction.cpp:7-8 */
    if (rc == ORC_OK)
        for (size_t i = 0; i < (size_t)w * h; i++)                 /* :11-17 */
            if (disp[i] > thr_front * b[i] || disp[i] < thr_back * b[i]) disp[i] = b[i];
    free(b);
    return rc;
}

int orc_scale(float* disp, int w, int h, float f) {
    if (!disp) return ORC_E_BADARG;
    for (size_t i = 0; i < (size_t)w * h; i++) disp[i] = f * disp[i]; /* :24 */
    return ORC_OK;
}

int orc_disparity_to_depth(const float* disp, float* depth, int w, int h, float focal, float baseline) {
    if (!disp || !depth) return ORC_E_BADARG;
    for (size_t i = 0; i < (size_t)w * h; i++)                     /* :34-39 */
        depth[i] = disp[i] == 0 ? -INFINITY : focal * baseline / disp[i];
    return ORC_OK;
}

int orc_backproject(const uint8_t* bgr, size_t pitch, const float* depth, int w, int h,
                    float fx, float fy, float cx, float cy, orc_vertex* out) {
    if (!bgr || !depth || !out) return ORC_E_BADARG;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {                              /* :169-196 */
            size_t i = (size_t)y * w + x;
            float d = depth[i];
            orc_vertex v; memset(&v, 0, sizeof v);
            if (d == -INFINITY) {
                v.pos[0] = v.pos[1] = v.pos[2] = v.pos[3] = -INFINITY; /* colour (0,0,0,0) */
            } else {
                float xc = ((float)x * d - cx * d) / fx;           /* :181 */
                float yc = ((float)y * d - cy * d) / fy;           /* :182 */
                v.pos[0] = xc; v.pos[1] = yc; v.pos[2] = d; v.pos[3] = 1.0f;
                const uint8_t* p = px(bgr, pitch, y, x);
                v.rgba[0] = p[2]; v.rgba[1] = p[1]; v.rgba[2] = p[0]; v.rgba[3] = 255; /* :188-192 */
            }
            out[i] = v;
        }
    return ORC_OK;
}

static float edge_len(const orc_vertex* a, const orc_vertex* b) {
    /* sqrtf(powf(dx,2)+powf(dy,2)+powf(dz,2)); powf(v,2) is exactly v*v (glibc), sums left to right */
    float dx = a->pos[0] - b->pos[0], dy = a->pos[1] - b->pos[1], dz = a->pos[2] - b->pos[2];
    return sqrtf(dx * dx + dy * dy + dz * dz);
}
int orc_triangle_valid(const orc_vertex* v, unsigned a, unsigned b, unsigned c, float thr) {
    if (v[a].pos[0] == -INFINITY || v[b].pos[0] == -INFINITY || v[c].pos[0] == -INFINITY) return 0; /* :48-51 */
    float l1 = edge_len(&v[a], &v[b]), l2 = edge_len(&v[a], &v[c]), l3 = edge_len(&v[b], &v[c]);
    if (l1 > thr || l2 > thr || l3 > thr) return 0;                /* :64-67 */
    return 1;
}
int orc_mesh_faces(const orc_vertex* v, int w, int h, float thr, uint32_t* tri, uint32_t* n_faces) {
    if (!v || !tri || !n_faces) return ORC_E_BADARG;
    uint32_t n = 0;
    for (unsigned y = 0; y + 1 < (unsigned)h; y++)
        for (unsigned x = 0; x + 1 < (unsigned)w; x++) {           /* :92-113 */
            unsigned a = y * w + x, b = (y + 1) * w + x, c = y * w + x + 1, d = (y + 1) * w + x + 1;
            if (orc_triangle_valid(v, a, b, c, thr)) { tri[3*n] = a; tri[3*n+1] = b; tri[3*n+2] = c; n++; }
            if (orc_triangle_valid(v, b, d, c, thr)) { tri[3*n] = b; tri[3*n+1] = d; tri[3*n+2] = c; n++; }
        }
    *n_faces = n;
    return ORC_OK;
}
int orc_write_off(const char* path, const orc_vertex* v, int w, int h, float thr) {
    size_t cells = (size_t)imax(w - 1, 0) * imax(h - 1, 0);
    uint32_t* tri = (uint32_t*)malloc(sizeof(uint32_t) * 6 * (cells ? cells : 1));
    uint32_t nf = 0;
    if (!tri) return ORC_E_BADARG;
    orc_mesh_faces(v, w, h, thr, tri, &nf);
    FILE* f = fopen(path, "w");
    if (!f) { free(tri); return ORC_E_IO; }                        /* :117 */
    fprintf(f, "COFF\n%u %u 0\n", (unsigned)(w * h), nf);          /* :120-121 */
    for (size_t n = 0; n < (size_t)w * h; n++) {                   /* :124-137; ostream<<float == %g */
        if (v[n].pos[0] == -INFINITY)
            fprintf(f, "0 0 0 %u %u %u %u\n", v[n].rgba[0], v[n].rgba[1], v[n].rgba[2], v[n].rgba[3]);
        else
            fprintf(f, "%g %g %g %u %u %u %u\n", (double)v[n].pos[0], (double)v[n].pos[1], (double)v[n].pos[2],
                    v[n].rgba[0], v[n].rgba[1], v[n].rgba[2], v[n].rgba[3]);
    }
    for (uint32_t n = 0; n < nf; n++) fprintf(f, "3 %u %u %u\n", tri[3*n], tri[3*n+1], tri[3*n+2]); /* :140-143 */
    fclose(f); free(tri);
    return ORC_OK;
}

int orc_evaldisp(const float* disp, const float* gt, const uint8_t* mask, int w, int h,
                 float badthresh, float maxdisp, int rounddisp, orc_eval* o) {
    if (!disp || !gt || !mask || !o) return ORC_E_BADARG;
    int n = 0, bad = 0, invalid = 0; float serr = 0;
    for (size_t i = 0; i < (size_t)w * h; i++) {                   /* utils.cpp:135-159 */
        float g = gt[i];
        if (g == INFINITY) continue;
        float d = disp[i];
        int valid = d != 0;
        if (valid) d = fmaxf(0.0f, fminf(maxdisp, d));
        if (valid && rounddisp) d = roundf(d);
        float err = fabsf(d - g);
        if (mask[i] != 255) continue;
        n++;
        if (valid) { serr += err; if (err > badthresh) bad++; } else invalid++;
    }
    o->n = n; o->bad = bad; o->invalid = invalid; o->serr = serr;
    o->bad_pct = (float)(100.0 * bad / n); o->invalid_pct = (float)(100.0 * invalid / n);
    o->total_bad_pct = (float)(100.0 * (bad + invalid) / n); o->avg_err = serr / (float)(n - invalid);
    return ORC_OK;
}
float orc_average_disparity(const float* disp, int w, int h) {
    float sum = 0; int cnt = 0;
    for (size_t i = 0; i < (size_t)w * h; i++) if (!isinf(disp[i])) { sum += disp[i]; cnt++; }
    return sum / cnt;
}

static int invert3(const double m[9], double o[9]) {
    double c0 = m[4]*m[8] - m[5]*m[7], c1 = m[5]*m[6] - m[3]*m[8], c2 = m[3]*m[7] - m[4]*m[6];
    double det = m[0]*c0 + m[1]*c1 + m[2]*c2;
    if (det == 0) return 0;
    double id = 1.0 / det;
    o[0] = c0*id; o[1] = (m[2]*m[7] - m[1]*m[8])*id; o[2] = (m[1]*m[5] - m[2]*m[4])*id;
    o[3] = c1*id; o[4] = (m[0]*m[8] - m[2]*m[6])*id; o[5] = (m[2]*m[3] - m[0]*m[5])*id;
    o[6] = c2*id; o[7] = (m[1]*m[6] - m[0]*m[7])*id; o[8] = (m[0]*m[4] - m[1]*m[3])*id;
    return 1;
}
/* cv::warpPerspective nearest (modules/imgproc/src/imgwarp.cpp, WarpPerspectiveInvoker): dst(x,y) =
 * src(round(X/W), round(Y/W)) with [X Y W] = M^-1 [x y 1] in double, W==0 -> 0, saturate_cast<int> (rint, half to
 * even), constant border 0.  OpenCV walks the destination in blocks of bw0 x bh0 pixels (bh0 = min(16, dh),
 * bw0 = min(1024 / bh0, dw), i.e. 64 columns for any map of 16 rows or more) and evaluates a pixel as
 * ((M0*bx + M1*y + M2) + M0*x1) * (1 / ((M6*bx + M7*y + M8) + M6*x1)) with bx the block's first column and x1 the
 * offset inside it -- not as M0*x + M1*y + M2.  The two differ in the last bit, which decides the source pixel when
 * the landing point is k + 0.5 (golden case warp_strong_perspective: one such pixel).  cv::invert of a 3x3 uses the
 * cofactor formula below (modules/core/src/lapack.cpp), bit for bit.  Checked against cv2.warpPerspective on the
 * tests/golden warp_* cases. */
int orc_warp_nearest_f64(const double* src, int sw, int sh, double* dst, int dw, int dh, const double M[9]) {
    double m[9];
    if (!invert3(M, m)) memset(m, 0, sizeof m);
    int bh0 = dh < 16 ? dh : 16;
    int bw0 = 1024 / bh0 < dw ? 1024 / bh0 : dw;
    for (int y = 0; y < dh; y++)
        for (int x = 0; x < dw; x++) {
            int bx = x - x % bw0, x1 = x - bx;
            double X0 = m[0]*bx + m[1]*y + m[2], Y0 = m[3]*bx + m[4]*y + m[5], W0 = m[6]*bx + m[7]*y + m[8];
            double W = W0 + m[6]*x1;
            W = W != 0 ? 1.0 / W : 0;
            double fx = fmax(-2147483648.0, fmin(2147483647.0, (X0 + m[0]*x1) * W));
            double fy = fmax(-2147483648.0, fmin(2147483647.0, (Y0 + m[3]*x1) * W));
            long ix = lrint(fx), iy = lrint(fy);
            dst[(size_t)y * dw + x] = (ix >= 0 && ix < sw && iy >= 0 && iy < sh) ? src[(size_t)iy * sw + ix] : 0.0;
        }
    return ORC_OK;
}

/* cv::warpPerspective(image, dst, M, size) with the default flags (INTER_LINEAR, BORDER_CONSTANT 0) on CV_8UC3 --
 * rectification.cpp:486-493, the rectifying warp of both input images.  OpenCV's fixed-point bilinear path
 * (imgwarp.cpp: WarpPerspectiveInvoker + remapBilinear<FixedPtCast<int, uchar, 15>>): the landing point is computed
 * block-wise like the nearest warp above but scaled by INTER_TAB_SIZE = 32, rounded half-to-even to integers X, Y;
 * source pixel (X >> 5, Y >> 5) saturated to short, fractions X & 31, Y & 31; the weights come from
 * BilinearTab_i[fy * 32 + fx] = saturate_cast<short>((1-fy/32 | fy/32) * (1-fx/32 | fx/32) * 32768) -- exact integers
 * (32-fy)(32-fx)*32 ...; the one entry that saturates (fx = fy = 0: 32767, OpenCV then adds the missing 1 to another
 * entry) gives the same pixel value as 32768 for every input, so the closed form below is used; value =
 * (sum of weight * neighbour + 2^14) >> 15, a neighbour outside the image counts as the border value 0.
 * Pinned to cv2.warpPerspective by the tests/golden warplin_* cases. */
int orc_warp_linear_bgr(const uint8_t* src, int sw, int sh, size_t spitch, uint8_t* dst, int dw, int dh, size_t dpitch,
                        const double M[9]) {
    double m[9];
    if (!invert3(M, m)) memset(m, 0, sizeof m);
    int bh0 = dh < 16 ? dh : 16;
    int bw0 = 1024 / bh0 < dw ? 1024 / bh0 : dw;
    for (int y = 0; y < dh; y++)
        for (int x = 0; x < dw; x++) {
            int bx = x - x % bw0, x1 = x - bx;
            double X0 = m[0]*bx + m[1]*y + m[2], Y0 = m[3]*bx + m[4]*y + m[5], W0 = m[6]*bx + m[7]*y + m[8];
            double W = W0 + m[6]*x1;
            W = W != 0 ? 32.0 / W : 0;
            double fX = fmax(-2147483648.0, fmin(2147483647.0, (X0 + m[0]*x1) * W));
            double fY = fmax(-2147483648.0, fmin(2147483647.0, (Y0 + m[3]*x1) * W));
            long X = lrint(fX), Y = lrint(fY);
            long sx = X >> 5, sy = Y >> 5;
            if (sx < -32768) sx = -32768; if (sx > 32767) sx = 32767;
            if (sy < -32768) sy = -32768; if (sy > 32767) sy = 32767;
            int fx = (int)(X & 31), fy = (int)(Y & 31);
            int w[4] = {(32 - fy) * (32 - fx) * 32, (32 - fy) * fx * 32, fy * (32 - fx) * 32, fy * fx * 32};
            for (int c = 0; c < 3; c++) {
                int acc = 1 << 14;
                for (int k = 0; k < 4; k++) {
                    long px = sx + (k & 1), py = sy + (k >> 1);
                    int v = (px >= 0 && px < sw && py >= 0 && py < sh) ? src[(size_t)py * spitch + 3 * px + c] : 0;
                    acc += v * w[k];
                }
                int o = acc >> 15;
                dst[(size_t)y * dpitch + 3 * x + c] = (uint8_t)(o < 0 ? 0 : o > 255 ? 255 : o);
            }
        }
    return ORC_OK;
}

int orc_quantize_u8(const double* disp, int n, float* out) {
    for (int i = 0; i < n; i++) {
        double v = disp[i];
        long q = isnan(v) ? 0 : lrint(fmax(-1e9, fmin(1e9, v)));
        if (q < 0) q = 0;
        if (q > 255) q = 255;
        out[i] = (float)(uint8_t)q;
    }
    return ORC_OK;
}

/* ---- extensions ------------------------------------------------------------------------- */

typedef struct { int64_t sl, sr, sll, srr, slr, ssd, sad; } wstats;
static void window_stats(const uint8_t* a, size_t pa, int ay, int ax, const uint8_t* b, size_t pb, int by, int bx,
                         int ww, int wh, wstats* s) {
    memset(s, 0, sizeof *s);
    for (int r = 0; r < wh; r++) {
        const uint8_t* ra = px(a, pa, ay + r, ax); const uint8_t* rb = px(b, pb, by + r, bx);
        for (int i = 0; i < 3 * ww; i++) {
            int l = ra[i], q = rb[i];
            s->sl += l; s->sr += q; s->sll += l * l; s->srr += q * q; s->slr += l * q;
            s->ssd += (l - q) * (l - q); s->sad += abs(l - q);
        }
    }
}
/* NCC cost (extension; fp32 with a fixed operation order):
 *   num = n*slr - sl*sr,  vl = n*sll - sl^2,  vr = n*srr - sr^2      (exact integers)
 *   rl = 1/sqrt((float)vl), rr = 1/sqrt((float)vr)                   (per-window factors, d-independent per image)
 *   cost = -(((float)num * rl) * rr)                                 (+inf when a variance is not positive)
 * i.e. -ZNCC of the two windows.  The factored form costs two multiplications per candidate once the
 * per-window factors are tabulated, which is what the sliding-sum CUDA kernel does; every operation is a
 * correctly rounded fp32 operation, so the GPU reproduces it bit for bit. */
static float ncc_cost(const wstats* s, int64_t n) {
    int64_t num = n * s->slr - s->sl * s->sr, vl = n * s->sll - s->sl * s->sl, vr = n * s->srr - s->sr * s->sr;
    if (vl <= 0 || vr <= 0) return INFINITY;
    volatile float sl_ = sqrtf((float)vl), sr_ = sqrtf((float)vr);
    volatile float rl = 1.0f / sl_, rr = 1.0f / sr_;
    volatile float a = (float)num * rl;
    volatile float c = a * rr;
    return -c;
}

int orc_search_ext(const uint8_t* left, int w1, int h1, size_t pitch1,
                   const uint8_t* right, int w2, int h2, size_t pitch2,
                   int block, int min_disp, int max_disp, int cost, int left_ref,
                   int32_t* disp, int64_t* cost_i, float* subpix) {
    if (!left || !right || !disp || block < 1 || min_disp < 0) return ORC_E_BADARG;
    int half = (block - 1) / 2;
    /* reference image = the one whose pixels we assign a disparity to */
    const uint8_t* ref = left_ref ? left : right; const uint8_t* oth = left_ref ? right : left;
    int wr = left_ref ? w1 : w2, hr = left_ref ? h1 : h2, wo = left_ref ? w2 : w1, ho = left_ref ? h2 : h1;
    size_t pr_ = left_ref ? pitch1 : pitch2, po = left_ref ? pitch2 : pitch1;
    int height = imin(hr, ho);
    if (!right_rows_fit(ho, hr, half)) return ORC_E_SHAPE;
    int sgn = left_ref ? -1 : 1; /* partner column = x + sgn*d */
    int nd = imax(max_disp - min_disp, 1);
#pragma omp parallel
    {
        double* cst = (double*)malloc(sizeof(double) * (size_t)nd);
        int64_t* csti = (int64_t*)malloc(sizeof(int64_t) * (size_t)nd);
#pragma omp for schedule(dynamic, 4)
        for (int y = 0; y < hr; y++)
            for (int x = 0; x < wr; x++) {
                size_t i = (size_t)y * wr + x;
                disp[i] = 0; if (cost_i) cost_i[i] = -1; if (subpix) subpix[i] = 0.0f;
                if (y >= height || is_black(px(ref, pr_, y, x))) continue;
                int l = imin(x, half), r = imin(wr - x - 1, half), u = imin(y, half), dn = imin(hr - y - 1, half);
                int ww = l + r, wh = u + dn;
                int64_t n = (int64_t)ww * wh * 3;
                int nv = 0;
                for (int d = min_disp; d < max_disp; d++) {
                    int cx = x + sgn * d;
                    if (left_ref ? (cx - l < 0) : (cx + r >= wo)) break;
                    if (left_ref && cx + r > wo) { /* partner window beyond the right image's width (w1>w2) */
                        cst[nv] = INFINITY; csti[nv] = INT64_MAX; nv++; continue;
                    }
                    wstats s;
                    /* stats with "l" = left image sample, "q" = right image sample */
                    if (left_ref) window_stats(ref, pr_, y - u, x - l, oth, po, y - u, cx - l, ww, wh, &s);
                    else window_stats(oth, po, y - u, cx - l, ref, pr_, y - u, x - l, ww, wh, &s);
                    if (cost == ORC_COST_NCC) { cst[nv] = n ? ncc_cost(&s, n) : INFINITY; csti[nv] = 0; }
                    else { csti[nv] = cost == ORC_COST_SAD ? s.sad : s.ssd; cst[nv] = (double)csti[nv]; }
                    nv++;
                }
                int bi = -1; double best = INFINITY;
                for (int k = 0; k < nv; k++) if (cst[k] < best) { best = cst[k]; bi = k; }
                if (bi < 0) { /* no candidate, or all costs +inf: d* = first candidate if any (mirrors "min stays"), else 0 */
                    disp[i] = 0; continue;
                }
                disp[i] = min_disp + bi;
                if (cost_i && cost != ORC_COST_NCC) cost_i[i] = csti[bi];
                if (subpix) {
                    float dsub = (float)(min_disp + bi);
                    if (bi >= 1 && bi + 1 < nv && isfinite(cst[bi - 1]) && isfinite(cst[bi + 1])) {
                        float cm = (float)cst[bi - 1], c0 = (float)cst[bi], cp = (float)cst[bi + 1];
                        float den = (cm - 2.0f * c0) + cp;
                        if (den > 0.0f) dsub = dsub + (cm - cp) / (2.0f * den);
                    }
                    subpix[i] = dsub;
                }
            }
        free(cst); free(csti);
    }
    return ORC_OK;
}

/* The same search as orc_search_ext for the integer costs (SSD, SAD), with vertical sliding column sums + prefix sums
 * per row and OpenMP over row chunks, so that whole MiddEval3 pairs finish in seconds (the literal version above is
 * O(window) per candidate).  Same definition, candidate order, break rule, first-minimum rule and parabola; held equal
 * to orc_search_ext on random pairs and real crops by tests/test_oracle_golden.py. */
int orc_search_ext_fast(const uint8_t* left, int w1, int h1, size_t pitch1,
                        const uint8_t* right, int w2, int h2, size_t pitch2,
                        int block, int min_disp, int max_disp, int cost, int left_ref,
                        int32_t* disp, int64_t* cost_i, float* subpix, int nthreads) {
    if (!left || !right || !disp || block < 1 || min_disp < 0) return ORC_E_BADARG;
    if (cost != ORC_COST_SSD_L2 && cost != ORC_COST_SAD) return ORC_E_BADARG;
    int half = (block - 1) / 2;
    const uint8_t* ref = left_ref ? left : right; const uint8_t* oth = left_ref ? right : left;
    int wr = left_ref ? w1 : w2, hr = left_ref ? h1 : h2, wo = left_ref ? w2 : w1, ho = left_ref ? h2 : h1;
    size_t pr_ = left_ref ? pitch1 : pitch2, po = left_ref ? pitch2 : pitch1;
    int height = imin(hr, ho);
    if (!right_rows_fit(ho, hr, half)) return ORC_E_SHAPE;
    int sgn = left_ref ? -1 : 1;
    size_t npx = (size_t)wr * hr;
    const int64_t NONE = -1, INF = INT64_MAX;
    int64_t* best = (int64_t*)malloc(npx * 8); int64_t* prev = (int64_t*)malloc(npx * 8);
    int64_t* cm = (int64_t*)malloc(npx * 8); int64_t* cp = (int64_t*)malloc(npx * 8);
    if (!best || !prev || !cm || !cp) { free(best); free(prev); free(cm); free(cp); return ORC_E_BADARG; }
    for (size_t i = 0; i < npx; i++) { best[i] = NONE; prev[i] = NONE; cm[i] = NONE; cp[i] = NONE; disp[i] = 0; }
    const int CH = 16;
    int nchunks = (height + CH - 1) / CH;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
    nthreads = 1;
#endif
#pragma omp parallel num_threads(nthreads)
    {
        int64_t* col = (int64_t*)malloc(sizeof(int64_t) * (size_t)(wr + 1));
        int64_t* pre = (int64_t*)malloc(sizeof(int64_t) * (size_t)(wr + 1));
#pragma omp for schedule(dynamic, 1)
        for (int ch = 0; ch < nchunks; ch++) {
            int ya = ch * CH, yb = imin(ya + CH, height);
            for (int d = min_disp; d < max_disp; d++) {
                /* reference columns c whose partner c + sgn*d exists; the window never reaches column wr-1 */
                int clo = left_ref ? d : 0, chi = imin(wr - 1, left_ref ? wo + d : wo - d);
                if (chi <= clo) { if (left_ref) continue; else break; }
                for (int y = ya; y < yb; y++) {
                    int u = imin(y, half), dn = imin(hr - y - 1, half);
                    int ra = y - u, rb = y + dn;
                    int rfirst, rlast, sub_a = 0, sub_b = 0;
                    if (y == ya) { for (int c = clo; c < chi; c++) col[c] = 0; rfirst = ra; rlast = rb; }
                    else {
                        int pu = imin(y - 1, half), pdn = imin(hr - y, half);
                        sub_a = y - 1 - pu; sub_b = ra; rfirst = y - 1 + pdn; rlast = rb;
                    }
                    for (int pass = 0; pass < 2; pass++) {
                        int r0 = pass ? rfirst : sub_a, r1 = pass ? rlast : sub_b;
                        for (int r = r0; r < r1; r++) {
                            const uint8_t* pa = px(ref, pr_, r, 0); const uint8_t* pb = px(oth, po, r, sgn * d);
                            for (int c = clo; c < chi; c++) {
                                int a = pa[3*c] - pb[3*c], b = pa[3*c+1] - pb[3*c+1], e = pa[3*c+2] - pb[3*c+2];
                                int64_t v = cost == ORC_COST_SAD ? abs(a) + abs(b) + abs(e) : a*a + b*b + e*e;
                                col[c] += pass ? v : -v;
                            }
                        }
                    }
                    pre[clo] = 0;
                    for (int c = clo; c < chi; c++) pre[c + 1] = pre[c] + col[c];
                    for (int x = 0; x < wr; x++) {
                        size_t i = (size_t)y * wr + x;
                        int l = imin(x, half), r = imin(wr - x - 1, half);
                        int cx = x + sgn * d;
                        int64_t c;
                        if (left_ref) {
                            if (cx - l < 0) continue;                 /* break: no candidate from here on for this pixel */
                            c = cx + r > wo ? INF : pre[x + r] - pre[x - l];
                        } else {
                            if (cx + r >= wo) break;                  /* monotone in x for a fixed d */
                            c = pre[x + r] - pre[x - l];
                        }
                        if (is_black(px(ref, pr_, y, x))) continue;
                        /* parabola neighbours: cost right after the current winner */
                        if (best[i] != NONE && disp[i] == d - 1 && cp[i] == NONE) cp[i] = c;
                        if (c != INF && (best[i] == NONE || c < best[i])) { best[i] = c; disp[i] = d; cm[i] = prev[i]; cp[i] = NONE; }
                        prev[i] = c;
                    }
                }
            }
        }
        free(col); free(pre);
    }
    for (size_t i = 0; i < npx; i++) {
        int has = best[i] != NONE;
        if (!has) disp[i] = 0;
        if (cost_i) cost_i[i] = has ? best[i] : -1;
        if (subpix) {
            float dsub = has ? (float)disp[i] : 0.0f;
            if (has && cm[i] != NONE && cp[i] != NONE && cm[i] != INF && cp[i] != INF) {
                float a = (float)(double)cm[i], c0 = (float)(double)best[i], b = (float)(double)cp[i];
                float den = (a - 2.0f * c0) + b;
                if (den > 0.0f) dsub = dsub + (a - b) / (2.0f * den);
            }
            subpix[i] = dsub;
        }
    }
    free(best); free(prev); free(cm); free(cp);
    return ORC_OK;
}

int orc_lr_check(const int32_t* disp_r, int w2, int h2, const int32_t* disp_l, int w1, int h1, int32_t* out) {
    for (int y = 0; y < h2; y++)
        for (int x = 0; x < w2; x++) {
            int d = disp_r[(size_t)y * w2 + x];
            int u = x + d;
            int ok = y < h1 && u >= 0 && u < w1 && abs(disp_l[(size_t)y * w1 + u] - d) <= 1;
            out[(size_t)y * w2 + x] = ok ? d : 0;
        }
    return ORC_OK;
}

int orc_write_pfm(const char* path, const float* data, int w, int h) {
    FILE* f = fopen(path, "wb");
    if (!f) return ORC_E_IO;
    fprintf(f, "Pf\n%d %d\n-1.0\n", w, h);
    for (int y = h - 1; y >= 0; y--) fwrite(data + (size_t)y * w, sizeof(float), (size_t)w, f);
    fclose(f);
    return ORC_OK;
}
int orc_write_ply(const char* path, const orc_vertex* v, int w, int h) {
    size_t n = 0;
    for (size_t i = 0; i < (size_t)w * h; i++) if (v[i].pos[0] != -INFINITY) n++;
    FILE* f = fopen(path, "w");
    if (!f) return ORC_E_IO;
    fprintf(f, "ply\nformat ascii 1.0\nelement vertex %zu\nproperty float x\nproperty float y\nproperty float z\n"
               "property uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n", n);
    for (size_t i = 0; i < (size_t)w * h; i++)
        if (v[i].pos[0] != -INFINITY)
            fprintf(f, "%.9g %.9g %.9g %u %u %u\n", (double)v[i].pos[0], (double)v[i].pos[1], (double)v[i].pos[2],
                    v[i].rgba[0], v[i].rgba[1], v[i].rgba[2]);
    fclose(f);
    return ORC_OK;
}

static inline uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline unsigned hN(uint64_t seed, uint64_t p, uint64_t y, uint64_t x, uint64_t c) {
    return (unsigned)(splitmix64(seed ^ (p * 0x9E3779B97F4A7C15ull) ^ (y << 40) ^ (x << 8) ^ c) & 0xFF);
}
void orc_synth_pair(uint64_t seed, uint64_t pair, int w, int h, int min_disp, int max_disp,
                    int quant, uint8_t* left, uint8_t* right) {
    int span = imax(max_disp - min_disp - 1, 1);
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; y++) {
        uint8_t* R = right + (size_t)y * w * 3; uint8_t* L = left + (size_t)y * w * 3;
        for (int x = 0; x < w; x++)
            for (int c = 0; c < 3; c++) {
                unsigned v = 16 + ((((3 * hN(seed, pair, (uint64_t)(y >> 2), (uint64_t)(x >> 2), (uint64_t)c) + hN(seed, pair, (uint64_t)y, (uint64_t)x, (uint64_t)c)) >> 2) * 7) >> 3);
                if (quant) v = 16 + (v >> 6) * 64;
                R[3 * x + c] = (uint8_t)v;
                L[3 * x + c] = (uint8_t)(1 | hN(seed + 1, pair, (uint64_t)y, (uint64_t)x, (uint64_t)c));
            }
        for (int x = 0; x < w; x++) {
            int g = min_disp + (((x >> 6) * 5 + (y >> 5) * 3) % span);
            if (x + g < w) for (int c = 0; c < 3; c++) L[3 * (x + g) + c] = R[3 * x + c];
        }
        if (!quant)
            for (int x = 0; x < w; x++)
                for (int c = 0; c < 3; c++) {
                    int v = (int)L[3 * x + c] + (int)(hN(seed + 2, pair, (uint64_t)y, (uint64_t)x, (uint64_t)c) % 5) - 2;
                    L[3 * x + c] = (uint8_t)(v < 1 ? 1 : v > 255 ? 255 : v);
                }
    }
}
