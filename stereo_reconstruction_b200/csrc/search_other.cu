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
                                                        int B, int max_disp, double* __restrict__ tq,
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
    // the raster pass compares distances: sqrt once, here, off its dependency chain
    for (int i = 0; i < 3; i++) { tq[i * npx + o] = sqrt((double)t.c[i]); td[i * npx + o] = skip ? -2 : t.d[i]; }
}

// Phase 1 through the tile kernel (search_tile.cu, sr_launch_search_tile_leftvariant): for smoothFactor <= 1 a biased
// distance never exceeds the unbiased one, so the winner is among {unbiased winner, d_up, d_left} and ONE entry per pixel
// is enough.  Key planes over the index i = off - d (ascending i = the reference's iteration order) -> (cost, d).
__global__ void __launch_bounds__(128) left_top1_from_planes_kernel(const uint32_t* __restrict__ Lp, int pitchL, int w1, int height, int h,
                                                                    const uint32_t* __restrict__ planes, size_t stride, int npass,
                                                                    int dbase, int shift, int off,
                                                                    double* __restrict__ tq, int* __restrict__ td,
                                                                    double* __restrict__ out_direct) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w1) return;
    size_t o = (size_t)y * w1 + x;
    bool inside = y >= h && y < height - h && x >= h && x < w1 - h;
    bool skip = !inside || (__ldg(Lp + (size_t)y * pitchL + x) & 0xFFFFFFu) == 0;   // :41-44
    unsigned long long bc = SR_NONE64;
    int bd = -1;
    if (!skip) {
        for (int p = 0; p < npass; p++) {
            uint32_t k = __ldg(planes + (size_t)p * stride + o);
            if (k != 0xFFFFFFFFu && (unsigned long long)(k >> shift) < bc) {
                bc = k >> shift;
                bd = off - (dbase + (p << shift) + (int)(k & ((1u << shift) - 1u)));
            }
        }
    }
    if (out_direct) {
        // smoothFactor == 1: the two multiplications (:68-73) change nothing, the unbiased winner IS the result
        // (x - 0 where no candidate was valid, :50,:82); no raster pass
        out_direct[o] = skip ? 0.0 : (bd > 0 ? (double)bd : (double)x);
        return;
    }
    tq[o] = sqrt((double)bc);                                     // the raster pass compares distances (:65)
    td[o] = skip ? -2 : bd;
}

// Phase 2: the raster rule as a wavefront.  One warp per row; the warps of ALL CTAs (grid <= SM count, one CTA per SM,
// so every CTA is resident) take the rows round-robin, row y runs K pixels behind row y-1.  The run time is the length
// of the dependency chain, (w / K + rows) chunk steps, so everything inside a step is arranged for latency:
//  * per chunk, in parallel (G = 32 / K lanes per pixel): the chunk's entries and upper results, sqrt of the entries'
//    costs, and the window cost of d_up where the entries do not hold it -- each lane takes whole window rows with all
//    loads of a row in flight at once;
//  * then the K pixels one after the other: the left neighbour's result is a register; its window cost, where needed,
//    is one round of <= 10 independent loads per lane (window offsets precomputed per lane) and a u32 butterfly;
//    the winner is the minimum of (distance, then larger d) over at most five register slots, which is what the
//    reference's strict-< scan in descending d computes;
//  * one store of the chunk, fence, progress flag.
#define LEFT_WARPS 16
#define LEFT_RING 64           // results a leader keeps in shared memory for its follower (power of two, > chunk)
#define LEFT_SLOTS 10          // window pixels per lane of the warp-cooperative cost: 32 * 10 >= 17 * 17

// one window row (B pixels, B <= 17 per round) of the SSD, all loads issued before the first use
__device__ __forceinline__ uint32_t left_row_ssd(const uint32_t* __restrict__ lrow, const uint32_t* __restrict__ rrow, int B) {
    uint32_t acc = 0;
    for (int c0 = 0; c0 < B; c0 += 17) {
        uint32_t l[17], r[17];
#pragma unroll
        for (int c = 0; c < 17; c++) {
            const bool ok = c0 + c < B;
            l[c] = ok ? __ldg(lrow + c0 + c) : 0u;
            r[c] = ok ? __ldg(rrow + c0 + c) : 0u;
        }
#pragma unroll
        for (int c = 0; c < 17; c++) { uint32_t ad = __vabsdiffu4(l[c], r[c]); acc = __dp4a(ad, ad, acc); }
    }
    return acc;                                                   // <= 17 * 195075 per round; rows are summed in 64 bits
}

template <int K>
__global__ void __launch_bounds__(LEFT_WARPS * 32, 1) left_raster_kernel(const uint32_t* __restrict__ Lp, int pitchL,
                                                                         const uint32_t* __restrict__ Rp, int pitchR, int w1, int w2,
                                                                         int height, int B, int max_disp, double s,
                                                                         const double* __restrict__ tq,
                                                                         const int* __restrict__ td, size_t npx, int ntop,
                                                                         double* out, volatile int* done) {
    constexpr int G = 32 / K;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gw = blockIdx.x * LEFT_WARPS + warp, nw = gridDim.x * LEFT_WARPS;
    const int h = (B - 1) / 2;
    const int nwin = B * B;
    const int pi = lane / G, q = lane % G;
    // this lane's window pixels of the warp-cooperative cost
    const bool small_win = nwin <= 32 * LEFT_SLOTS;
    int offL[LEFT_SLOTS], offR[LEFT_SLOTS];
#pragma unroll
    for (int i = 0; i < LEFT_SLOTS; i++) {
        int idx = lane + 32 * i;
        int r = idx / B, c = idx - r * B;
        offL[i] = idx < nwin ? r * pitchL + c : -1;
        offR[i] = r * pitchR + c;
    }
    auto warp_ssd = [&](int y, int xi, int d) -> unsigned long long {
        const uint32_t* lb = Lp + (size_t)(y - h) * pitchL + xi - h;
        const uint32_t* rb = Rp + (size_t)(y - h) * pitchR + xi - d - h;
        if (small_win) {
            uint32_t l[LEFT_SLOTS], r[LEFT_SLOTS];
#pragma unroll
            for (int i = 0; i < LEFT_SLOTS; i++) {
                const bool ok = offL[i] >= 0;
                l[i] = ok ? __ldg(lb + offL[i]) : 0u;
                r[i] = ok ? __ldg(rb + offR[i]) : 0u;
            }
            uint32_t acc = 0;                                     // <= 289 * 195075 < 2^32 over the whole warp
#pragma unroll
            for (int i = 0; i < LEFT_SLOTS; i++) { uint32_t ad = __vabsdiffu4(l[i], r[i]); acc = __dp4a(ad, ad, acc); }
#pragma unroll
            for (int of = 16; of; of >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, of);
            return acc;
        }
        unsigned long long part = 0;
        for (int idx = lane; idx < nwin; idx += 32) {
            int r = idx / B, c = idx - r * B;
            uint32_t ad = __vabsdiffu4(__ldg(lb + (size_t)r * pitchL + c), __ldg(rb + (size_t)r * pitchR + c));
            part += __dp4a(ad, ad, 0u);
        }
        for (int of = 16; of; of >>= 1) part += __shfl_xor_sync(0xffffffffu, part, of);
        return part;
    };
    // Rows y-1 and y belong to neighbouring warps of one CTA (except for warp 0): their hand-over goes through shared
    // memory -- a ring of the leader's last LEFT_RING results and a progress word (round * 2^20 + pixels done) per warp --
    // instead of store / fence / flag / poll through L2.  The leader stays within the ring of its follower (it waits
    // before it overwrites a slot the follower has not consumed).  Warp 0 follows the last warp of the previous CTA
    // (or of the previous round) through global memory.  Every result is an integer (a disparity, or x): the state of
    // the chain is kept in ints, doubles appear only where two different candidates compete.
    __shared__ int s_ring[LEFT_WARPS][LEFT_RING];
    __shared__ int s_prog[LEFT_WARPS];
    volatile int* vprog = s_prog;
    if (lane == 0) s_prog[warp] = 0;
    __syncthreads();
    const bool smem_leader = warp > 0;                            // the row above comes from warp - 1 of this CTA
    const bool smem_follower = warp < LEFT_WARPS - 1;             // warp + 1 of this CTA takes the row below
    const int xend = w1 - h;
    int round = 0;
    for (int y = h + gw; y < height - h; y += nw, round++) {
        int dleft = 0;                                            // D(y, h-1) is border: 0
        const bool has_follower = smem_follower && y + 1 < height - h;
        // entries of a chunk are fetched one chunk ahead (they do not depend on the raster)
        int nd3[3] = {-2, -1, -1};
        double nq3[3] = {0.0, 0.0, 0.0};
        const int* td_row = td + (size_t)y * w1;                  // row bases: no 64-bit index arithmetic per chunk
        const double* tq_row = tq + (size_t)y * w1;
        double* out_row = out + (size_t)y * w1;
        auto fetch = [&](int xf) {
            nd3[0] = -2; nd3[1] = -1; nd3[2] = -1;
            if (xf < xend) {
#pragma unroll
                for (int i = 0; i < 3; i++)
                    if (i < ntop) { nd3[i] = __ldg(td_row + i * npx + xf); nq3[i] = __ldg(tq_row + i * npx + xf); }
            }
        };
        fetch(h + pi);
        for (int x0 = h; x0 < xend; x0 += K) {
            const int n = min(K, xend - x0);
            const int x = x0 + pi;
            const bool on = pi < n;
            int d3[3]; double q3[3];
#pragma unroll
            for (int i = 0; i < 3; i++) { d3[i] = nd3[i]; q3[i] = nq3[i]; }
            fetch(x + K);
            int dup = 0;
            if (y - 1 >= h) {                                     // rows above h are border rows: always 0
                // every lane polls the same word (one broadcast access, no divergent branch to reconverge)
                if (smem_leader) {
                    while (vprog[warp - 1] < (round << 20) + x0 + n) { }
                    __threadfence_block();
                    if (on) dup = s_ring[warp - 1][x & (LEFT_RING - 1)];
                } else {
                    while (done[y - 1] < x0 + n) { }
                    __threadfence();                              // acquire side of the writer's store -> fence -> flag
                    if (on) dup = (int)__ldcg(out_row - w1 + x);
                }
            }
            // window cost of the upper neighbour's disparity, where it is a valid candidate the entries do not hold
            int dupc = -1;
            double qup = 0.0;
            {
                const int d = dup;
                const bool need = on && d3[0] != -2 && d >= 1 && d <= max_disp && x - d >= h && x - d < w2 - h &&
                                  d != d3[0] && d != d3[1] && d != d3[2];
                if (__any_sync(0xffffffffu, need)) {
                    unsigned long long part = 0;
                    if (need)
                        for (int r = q; r < B; r += G)
                            part += left_row_ssd(Lp + (size_t)(y - h + r) * pitchL + x - h, Rp + (size_t)(y - h + r) * pitchR + x - d - h, B);
#pragma unroll
                    for (int of = G / 2; of; of >>= 1) part += __shfl_xor_sync(0xffffffffu, part, of);
                    if (need) { dupc = d; qup = sqrt((double)part); }
                }
            }
            int mine = 0;
            for (int i = 0; i < n; i++) {
                const int src = i * G;
                const int xi = x0 + i;
                const int e0 = __shfl_sync(0xffffffffu, d3[0], src);
                const int du = __shfl_sync(0xffffffffu, dupc, src);
                int result = 0;
                if (e0 != -2) {                                   // not skipped (warp-uniform)
                    int seen1 = -1, seen2 = -1;
                    if (ntop > 1) { seen1 = __shfl_sync(0xffffffffu, d3[1], src); seen2 = __shfl_sync(0xffffffffu, d3[2], src); }
                    const int dl = dleft;
                    const bool need_left = dl >= 1 && dl <= max_disp && xi - dl >= h && xi - dl < w2 - h &&
                                           dl != e0 && dl != seen1 && dl != seen2 && dl != du;
                    if (ntop == 1 && du < 0 && !need_left) {
                        result = e0 > 0 ? e0 : xi;                // one candidate (or none: x - 0): nothing to compare
                    } else {
                        const int dupi = __shfl_sync(0xffffffffu, dup, src);
                        double best = 1.7976931348623157e308;
                        int bestd = -1;
                        // a candidate enters when it is strictly better, or equal with a larger d: the strict-< scan in
                        // descending d order (:53-79)
                        auto offer = [&](int d, double dist) {
                            if (dupi == d) dist = __dmul_rn(dist, s);    // :68-70 (needs y >= 1: always here)
                            if (dl == d) dist = __dmul_rn(dist, s);      // :71-73
                            if (bestd < 0 ? dist < best : (dist < best || (dist == best && d > bestd))) { best = dist; bestd = d; }
                        };
                        if (e0 > 0) offer(e0, __shfl_sync(0xffffffffu, q3[0], src));
                        if (ntop > 1) {
                            if (seen1 > 0) offer(seen1, __shfl_sync(0xffffffffu, q3[1], src));
                            if (seen2 > 0) offer(seen2, __shfl_sync(0xffffffffu, q3[2], src));
                        }
                        if (du >= 0) offer(du, __shfl_sync(0xffffffffu, qup, src));
                        if (need_left) offer(dl, sqrt((double)warp_ssd(y, xi, dl)));
                        result = bestd >= 0 ? bestd : xi;         // no valid candidate: x - 0
                    }
                }
                if (pi == i) mine = result;
                dleft = result;
            }
            if (on && q == 0) out_row[x] = (double)mine;
            if (smem_follower) {
                // the follower must have consumed what this chunk overwrites: the same row RING pixels back, or (start
                // of a row) its whole previous row
                const int back = x0 + n - LEFT_RING;
                const int need = has_follower && back > h ? (round << 20) + back : (round - 1) * (1 << 20) + xend;
                while (vprog[warp + 1] < need) { }
                if (on && q == 0) s_ring[warp][x & (LEFT_RING - 1)] = mine;
            } else {
                // the follower is in another CTA: store -> fence -> flag through L2.  (Publishing once per 32 pixels
                // instead was slower, 4.0 -> 6.4 ms on 900x750: the extra lag of those rows costs more than the fences.)
                __threadfence();
                __syncwarp();
                if (lane == 0) { __threadfence(); done[y] = x0 + n; }
            }
            __syncwarp();
            if (lane == 0) { __threadfence_block(); vprog[warp] = (round << 20) + x0 + n; }
        }
    }
}

int sr_launch_search_left(sr_ctx* ctx, sr_pair* pair, int block, int max_disp, double smooth) {
    int w1 = pair->w1, h1 = pair->h1, w2 = pair->w2, h2 = pair->h2;
    int height = h1 < h2 ? h1 : h2;
    size_t npx = (size_t)w1 * h1;
    // smoothFactor <= 1 (the way the reference is called): unbiased winners from the tile kernel, one entry per pixel;
    // else (a penalty: the two biased candidates may both lose) the three best by the direct kernel
    static const char* direct_env = getenv("SR_LEFT_DIRECT");     // test hook: 1 = always the direct top-3 kernel
    int off = 0, handled = 0;
    if (smooth <= 1.0 && !(direct_env && atoi(direct_env) != 0)) {
        handled = sr_launch_search_tile_leftvariant(ctx, pair, block, max_disp, &off);
        if (handled < 0) return handled;
    }
    const int ntop = handled ? 1 : 3;
    // (slots 14 / 13: the tile kernel's key planes live in slot 0, its work lists are not scratch)
    double* tc = (double*)sr_scratch(ctx, 14, npx * 3 * 8);    // sqrt of the entries' costs
    int* td = (int*)sr_scratch(ctx, 15, npx * 3 * 4 + (size_t)h1 * 4);
    if (!tc || !td) return SR_E_CUDA;
    int* done = td + npx * 3;
    SR_CUDA(ctx, cudaMemsetAsync(pair->out, 0, npx * sizeof(double), ctx->stream));   // :33
    SR_CUDA(ctx, cudaMemsetAsync(done, 0, (size_t)h1 * 4, ctx->stream));
    const int h = (block - 1) / 2;
    if (handled) {
        left_top1_from_planes_kernel<<<dim3(cdiv(w1, 128), h1), 128, 0, ctx->stream>>>(
            pair->L.px, pair->L.pitch, w1, height, h, pair->planes, pair->npx_cap, pair->npass, pair->plane_dbase, pair->key_shift, off, tc, td,
            smooth == 1.0 ? pair->out : nullptr);
        SR_LAUNCHED(ctx);
        if (smooth == 1.0) return SR_OK;
    } else {
        left_top3_kernel<<<dim3(cdiv(w1, 128), h1), 128, 0, ctx->stream>>>(pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, w1, w2,
                                                                            height, block, max_disp, tc, td, npx);
        SR_LAUNCHED(ctx);
    }
    const int rows = height - 2 * h;
    if (rows <= 0 || w1 - 2 * h <= 0) return SR_OK;
    int grid = cdiv(rows, LEFT_WARPS);
    if (grid > ctx->sms) grid = ctx->sms;
    if (const char* g_env = getenv("SR_LEFT_GRID")) {             // test hook (read per call): fewer CTAs = several rounds of rows per warp
        int g = atoi(g_env);
        if (g >= 1 && g < grid) grid = g;
    }
    static const char* chunk_env = getenv("SR_LEFT_CHUNK");       // tuning: 2 / 4 / 8 / 16 pixels per progress step
    // the chain is (w / K + rows) chunk steps of (fixed + K x per-pixel) time: measured 4.05 / 4.48 / 5.64 / 7.79 ms for
    // K = 2 / 4 / 8 / 16 on 900x750 (block 17, 128 candidates)
    const int chunk = chunk_env ? atoi(chunk_env) : (w1 - 2 * h < 2 * rows ? 2 : 4);
#define SR_LEFT_RASTER(K)                                                                                                    \
    left_raster_kernel<K><<<grid, LEFT_WARPS * 32, 0, ctx->stream>>>(pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, w1, w2, \
                                                                      height, block, max_disp, smooth, tc, td, npx, ntop, pair->out, done)
    if (chunk == 2) SR_LEFT_RASTER(2);
    else if (chunk == 8) SR_LEFT_RASTER(8);
    else if (chunk == 16) SR_LEFT_RASTER(16);
    else SR_LEFT_RASTER(4);
#undef SR_LEFT_RASTER
    SR_LAUNCHED(ctx);
    return SR_OK;
}
