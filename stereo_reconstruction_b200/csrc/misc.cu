// misc.cu -- neighbours of the hot path: nearest-neighbour homography warp of the
// CV_64F disparity map (rectification.cpp:82-87) and evaldisp (utils.cpp:123-168).
#include <cmath>
#include "sr_internal.cuh"

// cv::warpPerspective(src, dst, M, size, INTER_NEAREST), constant-0 border: dst(x,y) = src(rint(X/W), rint(Y/W)) with
// [X Y W] = M^-1 [x y 1], evaluated in double in OpenCV's order: the destination is walked in blocks of bw0 columns
// (64 for maps of 16 rows or more) and a pixel is ((M0*bx + M1*y + M2) + M0*x1) * (1 / ((M6*bx + M7*y + M8) + M6*x1)),
// bx = first column of the block, x1 = offset inside it (WarpPerspectiveInvoker) -- the last bit matters where the
// landing point is k + 0.5.
__global__ void warp_nearest_kernel(const double* __restrict__ src, int sw, int sh, double* __restrict__ dst, int dw, int dh,
                                    int bw0, double m0, double m1, double m2, double m3, double m4, double m5, double m6,
                                    double m7, double m8) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= dw) return;
    const int bx = x - x % bw0, x1 = x - bx;
    double X0 = __dadd_rn(__dadd_rn(__dmul_rn(m0, bx), __dmul_rn(m1, y)), m2);
    double Y0 = __dadd_rn(__dadd_rn(__dmul_rn(m3, bx), __dmul_rn(m4, y)), m5);
    double W0 = __dadd_rn(__dadd_rn(__dmul_rn(m6, bx), __dmul_rn(m7, y)), m8);
    double W = __dadd_rn(W0, __dmul_rn(m6, x1));
    W = W != 0 ? __ddiv_rn(1.0, W) : 0;
    double fx = fmax(-2147483648.0, fmin(2147483647.0, __dmul_rn(__dadd_rn(X0, __dmul_rn(m0, x1)), W)));
    double fy = fmax(-2147483648.0, fmin(2147483647.0, __dmul_rn(__dadd_rn(Y0, __dmul_rn(m3, x1)), W)));
    long long ix = __double2ll_rn(fx), iy = __double2ll_rn(fy);
    dst[(size_t)y * dw + x] = (ix >= 0 && ix < sw && iy >= 0 && iy < sh) ? src[(size_t)iy * sw + ix] : 0.0;
}

static bool invert3(const double m[9], double o[9]) {
    double c0 = m[4] * m[8] - m[5] * m[7], c1 = m[5] * m[6] - m[3] * m[8], c2 = m[3] * m[7] - m[4] * m[6];
    double det = m[0] * c0 + m[1] * c1 + m[2] * c2;
    if (det == 0) return false;
    double id = 1.0 / det;
    o[0] = c0 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    o[3] = c1 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    o[6] = c2 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
    return true;
}

// cv::warpPerspective(image, dst, M, size) with the default INTER_LINEAR / BORDER_CONSTANT(0) on CV_8UC3
// (rectification.cpp:486-493: the rectifying warp of both input images).  OpenCV's fixed-point bilinear path:
// the landing point is computed like the nearest warp's but scaled by INTER_TAB_SIZE = 32 (W = 32 / W), rounded to
// an integer X: source pixel sx = X >> 5 (saturated to short), fraction fx = X & 31; the four neighbours are blended
// with the 15-bit weights (32-fy)(32-fx)*32, (32-fy)fx*32, fy(32-fx)*32, fy*fx*32 (the table BilinearTab_i: products
// of exact float fractions, so its entries are these integers; its one saturated entry, fx = fy = 0 -> 32767 with the
// missing 1 added elsewhere, yields the same pixel value as 32768), result = (sum + 2^14) >> 15; a neighbour outside
// the image is the border value 0.  Four pixels per thread: 12 output bytes as three words.
__device__ __forceinline__ uint32_t warp_linear_px(const uint8_t* __restrict__ src, int sw, int sh, size_t spitch, int bw0, int x, int y,
                                                   double m0, double m1, double m2, double m3, double m4, double m5, double m6,
                                                   double m7, double m8) {
    const int bx = x - x % bw0, x1 = x - bx;
    double X0 = __dadd_rn(__dadd_rn(__dmul_rn(m0, bx), __dmul_rn(m1, y)), m2);
    double Y0 = __dadd_rn(__dadd_rn(__dmul_rn(m3, bx), __dmul_rn(m4, y)), m5);
    double W0 = __dadd_rn(__dadd_rn(__dmul_rn(m6, bx), __dmul_rn(m7, y)), m8);
    double W = __dadd_rn(W0, __dmul_rn(m6, x1));
    W = W != 0 ? __ddiv_rn(32.0, W) : 0;
    double fX = fmax(-2147483648.0, fmin(2147483647.0, __dmul_rn(__dadd_rn(X0, __dmul_rn(m0, x1)), W)));
    double fY = fmax(-2147483648.0, fmin(2147483647.0, __dmul_rn(__dadd_rn(Y0, __dmul_rn(m3, x1)), W)));
    const int X = __double2int_rn(fX), Y = __double2int_rn(fY);
    const int sx = max(-32768, min(32767, X >> 5)), sy = max(-32768, min(32767, Y >> 5));
    const int fx = X & 31, fy = Y & 31;
    const int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32, w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
    const bool x0 = (unsigned)sx < (unsigned)sw, x1ok = (unsigned)(sx + 1) < (unsigned)sw;
    const bool y0 = (unsigned)sy < (unsigned)sh, y1ok = (unsigned)(sy + 1) < (unsigned)sh;
    const uint8_t* r0 = src + (size_t)(y0 ? sy : 0) * spitch;
    const uint8_t* r1 = src + (size_t)(y1ok ? sy + 1 : 0) * spitch;
    const int c0 = 3 * (x0 ? sx : 0), c1 = 3 * (x1ok ? sx + 1 : 0);
    uint32_t out = 0;
#pragma unroll
    for (int c = 0; c < 3; c++) {
        int p00 = (y0 && x0) ? r0[c0 + c] : 0, p01 = (y0 && x1ok) ? r0[c1 + c] : 0;
        int p10 = (y1ok && x0) ? r1[c0 + c] : 0, p11 = (y1ok && x1ok) ? r1[c1 + c] : 0;
        int v = (p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11 + (1 << 14)) >> 15;
        out |= (uint32_t)min(255, max(0, v)) << (8 * c);
    }
    return out;
}

__global__ void __launch_bounds__(128) warp_linear_bgr_kernel(const uint8_t* __restrict__ src, int sw, int sh, size_t spitch,
                                                              uint8_t* __restrict__ dst, int dw, int dh, size_t dpitch, int bw0,
                                                              double m0, double m1, double m2, double m3, double m4, double m5,
                                                              double m6, double m7, double m8, int words) {
    const int x = 4 * (blockIdx.x * blockDim.x + threadIdx.x), y = blockIdx.y;
    if (x >= dw) return;
    uint32_t p[4];
#pragma unroll
    for (int i = 0; i < 4; i++)
        p[i] = x + i < dw ? warp_linear_px(src, sw, sh, spitch, bw0, x + i, y, m0, m1, m2, m3, m4, m5, m6, m7, m8) : 0u;
    uint8_t* o = dst + (size_t)y * dpitch + 3 * (size_t)x;
    if (words && x + 3 < dw) {                                   // 4-byte aligned rows: three words
        uint32_t* ow = reinterpret_cast<uint32_t*>(o);
        ow[0] = p[0] | (p[1] << 24);
        ow[1] = (p[1] >> 8) | (p[2] << 16);
        ow[2] = (p[2] >> 16) | (p[3] << 8);
    } else {
        for (int i = 0; i < 4 && x + i < dw; i++) { o[3 * i] = p[i] & 255; o[3 * i + 1] = (p[i] >> 8) & 255; o[3 * i + 2] = (p[i] >> 16) & 255; }
    }
}

// evaldisp: one thread per pixel, integer counts by atomics, serr as a double sum
__global__ void evaldisp_kernel(const float* __restrict__ disp, const float* __restrict__ gt, const uint8_t* __restrict__ mask,
                                size_t n, float badthresh, float maxdisp, int rounddisp, int* __restrict__ counts,
                                double* __restrict__ serr) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    int cn = 0, cbad = 0, cinv = 0;
    double e = 0;
    if (i < n) {
        float g = gt[i];
        if (g != INFINITY && mask[i] == 255) {   // utils.cpp:139, :148
            float d = disp[i];
            bool valid = d != 0;
            if (valid) d = fmaxf(0.0f, fminf(maxdisp, d));
            if (valid && rounddisp) d = roundf(d);
            float err = fabsf(d - g);
            cn = 1;
            if (valid) { e = (double)err; cbad = err > badthresh; } else cinv = 1;
        }
    }
    for (int o = 16; o; o >>= 1) {
        cn += __shfl_down_sync(0xffffffffu, cn, o);
        cbad += __shfl_down_sync(0xffffffffu, cbad, o);
        cinv += __shfl_down_sync(0xffffffffu, cinv, o);
        e += __shfl_down_sync(0xffffffffu, e, o);
    }
    if ((threadIdx.x & 31) == 0 && cn) {
        atomicAdd(counts, cn); atomicAdd(counts + 1, cbad); atomicAdd(counts + 2, cinv); atomicAdd(serr, e);
    }
}

// device buffers; M is the matrix PASSED to warpPerspective (inverted here like OpenCV does, cofactor formula of
// cv::invert for 3x3); asynchronous on the context stream
int sr_launch_warp_nearest(sr_ctx* ctx, const double* d_src, int sw, int sh, double* d_dst, int dw, int dh, const double M[9]) {
    double m[9];
    if (!invert3(M, m)) for (double& v : m) v = 0;
    const int bh0 = dh < 16 ? dh : 16;
    const int bw0 = 1024 / bh0 < dw ? 1024 / bh0 : dw;
    warp_nearest_kernel<<<dim3(cdiv(dw, 128), dh), 128, 0, ctx->stream>>>(d_src, sw, sh, d_dst, dw, dh, bw0, m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8]);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// the same for the bilinear warp of a CV_8UC3 image (device buffers, row pitches in bytes)
int sr_launch_warp_linear_bgr(sr_ctx* ctx, const uint8_t* d_src, int sw, int sh, size_t spitch, uint8_t* d_dst, int dw, int dh,
                              size_t dpitch, const double M[9]) {
    double m[9];
    if (!invert3(M, m)) for (double& v : m) v = 0;
    const int bh0 = dh < 16 ? dh : 16;
    const int bw0 = 1024 / bh0 < dw ? 1024 / bh0 : dw;
    const int words = (((uintptr_t)d_dst | dpitch) & 3) == 0;
    warp_linear_bgr_kernel<<<dim3(cdiv(cdiv(dw, 4), 128), dh), 128, 0, ctx->stream>>>(d_src, sw, sh, spitch, d_dst, dw, dh, dpitch, bw0, m[0], m[1],
                                                                                       m[2], m[3], m[4], m[5], m[6], m[7], m[8], words);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

extern "C" {

int sr_warp_linear_bgr(sr_ctx* ctx, const uint8_t* src, int sw, int sh, size_t src_pitch, uint8_t* dst, int dw, int dh,
                       size_t dst_pitch, const double M[9]) {
    if (!ctx || !src || !dst || !M || sw <= 0 || sh <= 0 || dw <= 0 || dh <= 0 || src_pitch < (size_t)sw * 3 || dst_pitch < (size_t)dw * 3)
        return sr_fail(ctx, SR_E_BADARG, "sr_warp_linear_bgr: bad argument");
    if (sw >= 32768 || sh >= 32768) return sr_fail(ctx, SR_E_SHAPE, "sr_warp_linear_bgr: source beyond 32767 pixels (cv::remap's short coordinates)");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t srow = (size_t)sw * 3, drow = ((size_t)dw * 3 + 3) & ~(size_t)3;
    uint8_t* ds = (uint8_t*)sr_scratch(ctx, 11, srow * sh);
    uint8_t* dd = (uint8_t*)sr_scratch(ctx, 12, drow * dh);
    if (!ds || !dd) return SR_E_CUDA;
    SR_CUDA(ctx, cudaMemcpy2DAsync(ds, srow, src, src_pitch, srow, sh, cudaMemcpyHostToDevice, ctx->stream));
    int rc = sr_launch_warp_linear_bgr(ctx, ds, sw, sh, srow, dd, dw, dh, drow, M);
    if (rc != SR_OK) return rc;
    SR_CUDA(ctx, cudaMemcpy2DAsync(dst, dst_pitch, dd, drow, (size_t)dw * 3, dh, cudaMemcpyDeviceToHost, ctx->stream));
    return sr_ctx_sync(ctx);
}

int sr_warp_nearest_f64(sr_ctx* ctx, const double* src, int sw, int sh, double* dst, int dw, int dh, const double M[9]) {
    if (!ctx || !src || !dst || !M || sw <= 0 || sh <= 0 || dw <= 0 || dh <= 0) return sr_fail(ctx, SR_E_BADARG, "sr_warp_nearest_f64: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    double* ds = (double*)sr_scratch(ctx, 11, (size_t)sw * sh * 8);
    double* dd = (double*)sr_scratch(ctx, 12, (size_t)dw * dh * 8);
    if (!ds || !dd) return SR_E_CUDA;
    SR_CUDA(ctx, cudaMemcpyAsync(ds, src, (size_t)sw * sh * 8, cudaMemcpyHostToDevice, ctx->stream));
    int rc = sr_launch_warp_nearest(ctx, ds, sw, sh, dd, dw, dh, M);
    if (rc != SR_OK) return rc;
    SR_CUDA(ctx, cudaMemcpyAsync(dst, dd, (size_t)dw * dh * 8, cudaMemcpyDeviceToHost, ctx->stream));
    return sr_ctx_sync(ctx);
}

int sr_evaldisp(sr_ctx* ctx, const float* disp, const float* gt, const uint8_t* mask, int w, int h, float badthresh,
                float maxdisp, int rounddisp, sr_eval* out) {
    if (!ctx || !disp || !gt || !mask || !out || w <= 0 || h <= 0) return sr_fail(ctx, SR_E_BADARG, "sr_evaldisp: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)w * h;
    float* dd = (float*)sr_scratch(ctx, 4, n * 4);
    float* dg = (float*)sr_scratch(ctx, 5, n * 4);
    uint8_t* dm = (uint8_t*)sr_scratch(ctx, 13, n + 64);
    if (!dd || !dg || !dm) return SR_E_CUDA;
    int* counts = (int*)(dm + ((n + 15) & ~(size_t)15));
    double* serr = (double*)(counts + 4);
    SR_CUDA(ctx, cudaMemcpyAsync(dd, disp, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    SR_CUDA(ctx, cudaMemcpyAsync(dg, gt, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    SR_CUDA(ctx, cudaMemcpyAsync(dm, mask, n, cudaMemcpyHostToDevice, ctx->stream));
    SR_CUDA(ctx, cudaMemsetAsync(counts, 0, 32, ctx->stream));
    evaldisp_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(dd, dg, dm, n, badthresh, maxdisp, rounddisp, counts, serr);
    SR_LAUNCHED(ctx);
    int c[4];
    double s;
    SR_CUDA(ctx, cudaMemcpyAsync(c, counts, 16, cudaMemcpyDeviceToHost, ctx->stream));
    SR_CUDA(ctx, cudaMemcpyAsync(&s, serr, 8, cudaMemcpyDeviceToHost, ctx->stream));
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    out->n = c[0]; out->bad = c[1]; out->invalid = c[2]; out->serr = s;
    out->bad_pct = (float)(100.0 * c[1] / c[0]); out->invalid_pct = (float)(100.0 * c[2] / c[0]);
    out->total_bad_pct = (float)(100.0 * (c[1] + c[2]) / c[0]);
    out->avg_err = (float)(s / (double)(c[0] - c[2]));
    return SR_OK;
}

}  // extern "C"
