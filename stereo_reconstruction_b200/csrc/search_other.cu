// search_other.cu -- the two secondary reference searches:
//   LinearSearch::computeDisparityMap        (LinearSearch.cpp:10-59)   per-pixel colour distance
//   BlockSearch::computeDisparityMapLeft     (BlockSearch.cpp:24-86)    left-referenced, true raster
//                                                                       dependency on the smoothing rule
#include "sr_internal.cuh"

// ---- LinearSearch ------------------------------------------------------------------------------------
// Window = one pixel, 200 hard-coded candidates k = j .. j+199 (clamped to k < w1: the reference
// reads out of bounds there), skip test on the LEFT pixel, bias vs (double)(j-k) = -d (fires for
// d == 0 only), strict < => smallest d.  Produces the c0 / (cr,dr) records of resolve.cu with N = 1
// (sqrt(ssd)/1 == sqrt(ssd) exactly).
__global__ void __launch_bounds__(128) linear_search_kernel(const uint32_t* __restrict__ Lp, int pitchL,
                                                            const uint32_t* __restrict__ Rp, int pitchR, int w1, int w2, int h2,
                                                            unsigned long long* __restrict__ c0,
                                                            unsigned long long* __restrict__ cr, int* __restrict__ dr,
                                                            int* __restrict__ nwin) {
    int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= w2) return;
    size_t o = (size_t)i * w2 + j;
    c0[o] = SR_NONE64; cr[o] = SR_NONE64; dr[o] = 0; nwin[o] = -1;
    const uint32_t* lrow = Lp + (size_t)i * pitchL;
    if ((__ldg(lrow + j) & 0xFFFFFFu) == 0) return;            // :24-27
    uint32_t r = __ldg(Rp + (size_t)i * pitchR + j);
    unsigned long long zc = SR_NONE64, bc = SR_NONE64;
    int bd = 0;
    int kend = min(j + 200, w1);                                // :32
    for (int k = j; k < kend; k++) {
        uint32_t ad = __vabsdiffu4(r, __ldg(lrow + k));
        unsigned long long c = __dp4a(ad, ad, 0u);
        if (k == j) zc = c;
        else if (c < bc) { bc = c; bd = k - j; }
    }
    c0[o] = zc; cr[o] = bc; dr[o] = bd; nwin[o] = 1;
}

static int ensure_direct_buffers(sr_ctx* ctx, sr_pair* pair) {
    if (!pair->cr) {
        SR_CUDA(ctx, cudaMalloc((void**)&pair->cr, pair->npx_cap * 8));
        SR_CUDA(ctx, cudaMalloc((void**)&pair->dr, pair->npx_cap * 4));
        SR_CUDA(ctx, cudaMalloc((void**)&pair->nwin, pair->npx_cap * 4));
    }
    return SR_OK;
}

int sr_launch_search_linear(sr_ctx* ctx, sr_pair* pair, double smooth) {
    int rc = ensure_direct_buffers(ctx, pair);
    if (rc != SR_OK) return rc;
    linear_search_kernel<<<dim3(cdiv(pair->w2, 128), pair->h2), 128, 0, ctx->stream>>>(
        pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, pair->w1, pair->w2, pair->h2, pair->c0, pair->cr, pair->dr, pair->nwin);
    SR_LAUNCHED(ctx);
    SearchGeom g;
    g.w1 = pair->w1; g.h1 = pair->h1; g.w2 = pair->w2; g.h2 = pair->h2; g.height = pair->h2;
    g.block = 1; g.half = 0; g.dlo = 0; g.dhi = 200; g.row0 = 0; g.row1 = pair->h2;
    return sr_launch_resolve_right(ctx, pair, g, smooth, 1, 1);
}

// ---- Left variant ---------------------------------------------------------------------------------------
// Pixels y in [h, min(h1,h2)-h), x in [h, w1-h); full symmetric BxB window; candidates
// cx = x-maxD .. x-1 ascending (d = maxD .. 1 descending; `continue` unless h <= cx < w2-h);
// dist = sqrt(SSD); the bias compares the stored disparity of the upper / left neighbour with +d, a
// genuine raster dependency on arbitrary d; strict < => the LARGEST d wins ties; D = x - cx*, and
// D = x when no candidate is valid (minimumCorrespondX stays 0, :50,:82).
//
// Phase 1 (parallel): per pixel the three best unbiased candidates in (cost, iteration order).
// At most two candidates (d_up, d_left) are biased, so the winner is always among
// top-3 U {d_up, d_left}.  Phase 2 (wavefront): one warp per row, row y lags row y-1 by one pixel;
// the two extra window costs are evaluated cooperatively by the warp.
struct Top3 { unsigned long long c[3]; int d[3]; };

__device__ __forceinline__ unsigned long long left_window_ssd_thread(const uint32_t* __restrict__ Lp, int pitchL,
                                                                     const uint32_t* __restrict__ Rp, int pitchR, int y0, int lx0,
                                                                     int rx0, int B) {
    unsigned long long total = 0;
    for (int r = 0; r < B; r++) {
        const uint32_t* lrow = Lp + (size_t)(y0 + r) * pitchL + lx0;
        const uint32_t* rrow = Rp + (size_t)(y0 + r) * pitchR + rx0;
        uint32_t acc = 0;
        for (int i = 0; i < B; i++) {
            uint32_t ad = __vabsdiffu4(__ldg(lrow + i), __ldg(rrow + i));
            acc = __dp4a(ad, ad, acc);
        }
        total += acc;
    }
    return total;
}

__global__ void __launch_bounds__(128) left_top3_kernel(const uint32_t* __restrict__ Lp, int pitchL,
                                                        const uint32_t* __restrict__ Rp, int pitchR, int w1, int w2, int height,
                                                        int B, int max_disp, unsigned long long* __restrict__ tc,
                                                        int* __restrict__ td, size_t npx) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    int h = (B - 1) / 2;
    if (x >= w1) return;
    size_t o = (size_t)y * w1 + x;
    Top3 t;
    for (int i = 0; i < 3; i++) { t.c[i] = SR_NONE64; t.d[i] = -1; }
    bool inside = y >= h && y < height - h && x >= h && x < w1 - h;
    bool skip = !inside || (__ldg(Lp + (size_t)y * pitchL + x) & 0xFFFFFFu) == 0;   // :41-44
    if (!skip) {
        for (int cx = x - max_disp; cx < x; cx++) {              // :53
            if (cx < h || cx >= w2 - h) continue;               // :55-57
            unsigned long long c = left_window_ssd_thread(Lp, pitchL, Rp, pitchR, y - h, x - h, cx - h, B);
            int d = x - cx;
            // sorted insertion, strict <: on equal cost the earlier candidate stays ahead
            if (c < t.c[2]) {
                if (c < t.c[1]) {
                    t.c[2] = t.c[1]; t.d[2] = t.d[1];
                    if (c < t.c[0]) { t.c[1] = t.c[0]; t.d[1] = t.d[0]; t.c[0] = c; t.d[0] = d; }
                    else { t.c[1] = c; t.d[1] = d; }
                } else { t.c[2] = c; t.d[2] = d; }
            }
        }
    }
    for (int i = 0; i < 3; i++) { tc[i * npx + o] = t.c[i]; td[i * npx + o] = skip ? -2 : t.d[i]; }
}

#define LEFT_WARPS 32
__global__ void __launch_bounds__(LEFT_WARPS * 32) left_raster_kernel(const uint32_t* __restrict__ Lp, int pitchL,
                                                                      const uint32_t* __restrict__ Rp, int pitchR, int w1, int w2,
                                                                      int height, int B, int max_disp, double s,
                                                                      const unsigned long long* __restrict__ tc,
                                                                      const int* __restrict__ td, size_t npx,
                                                                      double* out, volatile int* done) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int h = (B - 1) / 2;
    const int nwin = B * B;
    for (int y = h + warp; y < height - h; y += LEFT_WARPS) {
        double dleft = 0.0;                                       // D(y, h-1) is border: 0
        for (int x = h; x < w1 - h; x++) {
            size_t o = (size_t)y * w1 + x;
            // wait until D(y-1, x) is final (rows above h are border rows: always 0)
            double dup = 0.0;
            if (y - 1 >= h) {
                if (lane == 0) while (done[y - 1] <= x) { }
                __threadfence();                                  // acquire side of the writer's store -> fence -> flag
                __syncwarp();
                dup = __ldcg(out + o - w1);
            }
            int d3[3]; unsigned long long c3[3];
            for (int i = 0; i < 3; i++) { d3[i] = td[i * npx + o]; c3[i] = tc[i * npx + o]; }
            double result = 0.0;
            if (d3[0] != -2) {                                    // not skipped
                // candidate list in iteration order (d descending): top-3 plus the two biased ones
                int cd[5]; unsigned long long cc[5]; int n = 0;
                for (int i = 0; i < 3; i++) if (d3[i] > 0) { cd[n] = d3[i]; cc[n] = c3[i]; n++; }
                for (int e = 0; e < 2; e++) {
                    double dv = e == 0 ? dup : dleft;
                    int d = (int)dv;
                    if ((double)d != dv || d < 1 || d > max_disp) continue;
                    int cx = x - d;
                    if (cx < h || cx >= w2 - h) continue;
                    bool have = false;
                    for (int i = 0; i < n; i++) have |= cd[i] == d;
                    if (have) continue;
                    // warp-cooperative window SSD
                    unsigned long long part = 0;
                    for (int idx = lane; idx < nwin; idx += 32) {
                        int r = idx / B, c = idx - r * B;
                        uint32_t ad = __vabsdiffu4(__ldg(Lp + (size_t)(y - h + r) * pitchL + x - h + c),
                                                   __ldg(Rp + (size_t)(y - h + r) * pitchR + cx - h + c));
                        part += __dp4a(ad, ad, 0u);
                    }
                    for (int of = 16; of; of >>= 1) part += __shfl_xor_sync(0xffffffffu, part, of);
                    cd[n] = d; cc[n] = part; n++;
                }
                // strict-< scan in iteration order (largest d first)
                double best = 1.7976931348623157e308;
                int bestd = -1;
                for (int it = 0; it < n; it++) {
                    int pick = -1;
                    for (int i = 0; i < n; i++) if (cd[i] >= 0 && (pick < 0 || cd[i] > cd[pick])) pick = i;
                    double dist = sqrt((double)cc[pick]);
                    if (dup == (double)cd[pick]) dist = __dmul_rn(dist, s);     // :68-70 (needs y >= 1: always here)
                    if (dleft == (double)cd[pick]) dist = __dmul_rn(dist, s);   // :71-73
                    if (dist < best) { best = dist; bestd = cd[pick]; }
                    cd[pick] = -1;
                }
                result = bestd >= 0 ? (double)bestd : (double)x;  // no valid candidate: x - 0
            }
            if (lane == 0) {
                out[o] = result;
                __threadfence();
                done[y] = x + 1;
            }
            dleft = result;
        }
    }
}

int sr_launch_search_left(sr_ctx* ctx, sr_pair* pair, int block, int max_disp, double smooth) {
    int w1 = pair->w1, h1 = pair->h1, w2 = pair->w2, h2 = pair->h2;
    int height = h1 < h2 ? h1 : h2;
    size_t npx = (size_t)w1 * h1;
    unsigned long long* tc = (unsigned long long*)sr_scratch(ctx, 14, npx * 3 * 8);
    int* td = (int*)sr_scratch(ctx, 15, npx * 3 * 4 + (size_t)h1 * 4);
    if (!tc || !td) return SR_E_CUDA;
    int* done = td + npx * 3;
    if (max_disp < 0) max_disp = 0;
    SR_CUDA(ctx, cudaMemsetAsync(pair->out, 0, npx * sizeof(double), ctx->stream));   // :33
    SR_CUDA(ctx, cudaMemsetAsync(done, 0, (size_t)h1 * 4, ctx->stream));
    left_top3_kernel<<<dim3(cdiv(w1, 128), h1), 128, 0, ctx->stream>>>(pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, w1, w2,
                                                                        height, block, max_disp, tc, td, npx);
    SR_LAUNCHED(ctx);
    left_raster_kernel<<<1, LEFT_WARPS * 32, 0, ctx->stream>>>(pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, w1, w2, height,
                                                               block, max_disp, smooth, tc, td, npx, pair->out, done);
    SR_LAUNCHED(ctx);
    return SR_OK;
}
