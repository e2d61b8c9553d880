// search_direct.cu -- universal (slow) right-referenced search: one thread per output pixel,
// O(window) work per candidate exactly like the reference loop (BlockSearch.cpp:100-175).
// Used for: varBlock (per-pixel windows, BlockSearch.cpp:125-145), window sizes the tile
// kernel is not instantiated for, and the d == 0 cost of the smoothing rule.
#include "sr_internal.cuh"

template <int COST>
__device__ __forceinline__ unsigned long long window_cost(const uint32_t* __restrict__ Lp, int pitchL,
                                                          const uint32_t* __restrict__ Rp, int pitchR,
                                                          int y0, int lx0, int rx0, int ww, int wh) {
    unsigned long long total = 0;
    for (int r = 0; r < wh; r++) {
        const uint32_t* lrow = Lp + (size_t)(y0 + r) * pitchL + lx0;
        const uint32_t* rrow = Rp + (size_t)(y0 + r) * pitchR + rx0;
        uint32_t acc = 0;  // a row of <= 2^14 pixels cannot overflow 32 bits (3*65025 per pixel)
        for (int i = 0; i < ww; i++) {
            uint32_t ad = __vabsdiffu4(__ldg(lrow + i), __ldg(rrow + i));
            acc = (COST == SR_COST_SAD) ? __dp4a(ad, 0x01010101u, acc) : __dp4a(ad, ad, acc);
        }
        total += acc;
    }
    return total;
}

// norm(rightWindow - mean(rightWindow)) with OpenCV's u8 semantics (BlockSearch.cpp:125-129):
// per-channel double mean, scalar saturate-cast to u8 (round half to even), saturating u8
// subtract, L2 norm = sqrt of the exact integer sum.
__device__ double window_dev_norm(const uint32_t* __restrict__ Rp, int pitchR, int y0, int x0, int ww, int wh) {
    unsigned long long s0 = 0, s1 = 0, s2 = 0;
    for (int r = 0; r < wh; r++) {
        const uint32_t* row = Rp + (size_t)(y0 + r) * pitchR + x0;
        for (int i = 0; i < ww; i++) {
            uint32_t p = __ldg(row + i);
            s0 += p & 255u; s1 += (p >> 8) & 255u; s2 += (p >> 16) & 255u;
        }
    }
    long long n = (long long)ww * wh;
    double m0 = n ? __ddiv_rn((double)s0, (double)n) : 0.0, m1 = n ? __ddiv_rn((double)s1, (double)n) : 0.0,
           m2 = n ? __ddiv_rn((double)s2, (double)n) : 0.0;
    int q0 = min(255, max(0, (int)__double2ll_rn(m0))), q1 = min(255, max(0, (int)__double2ll_rn(m1))),
        q2 = min(255, max(0, (int)__double2ll_rn(m2)));
    uint32_t mpk = (uint32_t)q0 | ((uint32_t)q1 << 8) | ((uint32_t)q2 << 16);
    unsigned long long ss = 0;
    for (int r = 0; r < wh; r++) {
        const uint32_t* row = Rp + (size_t)(y0 + r) * pitchR + x0;
        uint32_t acc = 0;
        for (int i = 0; i < ww; i++) {
            uint32_t p = __ldg(row + i);
            // saturating per-byte p - m : max(p, m) - m
            uint32_t q = __vmaxu4(p, mpk) - mpk;  // no borrow between bytes since max >= m
            acc = __dp4a(q, q, acc);
        }
        ss += acc;
    }
    return sqrt((double)ss);
}

template <int COST>
__global__ void __launch_bounds__(128) search_direct_right_kernel(
    const uint32_t* __restrict__ Lp, int pitchL, const uint32_t* __restrict__ Rp, int pitchR, SearchGeom g,
    int block, int var_block, double thres, int want_c0, unsigned long long* __restrict__ c0,
    unsigned long long* __restrict__ cr, int* __restrict__ dr, int* __restrict__ nwin, int* __restrict__ maxblock,
    int* __restrict__ status) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = g.row0 + blockIdx.y;
    if (x >= g.w2 || y >= g.row1) return;
    size_t i = (size_t)y * g.w2 + x;
    c0[i] = SR_NONE64; cr[i] = SR_NONE64; dr[i] = 0; nwin[i] = -1;
    if (y >= g.height) return;  // rows the reference never visits stay 0
    if ((__ldg(Rp + (size_t)y * pitchR + x) & 0xFFFFFFu) == 0) return;  // black pixel (:105-108)
    int bs = block;
    int l = min(x, (bs - 1) / 2), r = min(g.w2 - x - 1, (bs - 1) / 2);
    int u = min(y, (bs - 1) / 2), dn = min(g.h2 - y - 1, (bs - 1) / 2);
    if (var_block) {
        while (window_dev_norm(Rp, pitchR, y - u, x - l, l + r, u + dn) < thres) {
            int pl = l, pr = r, pu = u, pd = dn;
            bs += 4;
            l = min(x, (bs - 1) / 2); r = min(g.w2 - x - 1, (bs - 1) / 2);
            u = min(y, (bs - 1) / 2); dn = min(g.h2 - y - 1, (bs - 1) / 2);
            // the reference never terminates once the window spans the image; stop at the
            // first growth step that changes nothing (documented deviation)
            if (l == pl && r == pr && u == pu && dn == pd) break;
        }
        if (y + dn > g.h1) { atomicExch(status, SR_E_SHAPE); return; }
        if (bs > block) atomicMax(maxblock, bs);
    }
    int ww = l + r, wh = u + dn;
    nwin[i] = ww * wh;
    unsigned long long bc = SR_NONE64, zc = SR_NONE64;
    int bd = 0;
    for (int d = g.dlo; d < g.dhi; d++) {
        if (x + d + r >= g.w1) break;  // :148-149
        unsigned long long c = window_cost<COST>(Lp, pitchL, Rp, pitchR, y - u, x + d - l, x - l, ww, wh);
        if (d == 0 && want_c0) zc = c;
        else if (c < bc) { bc = c; bd = d; }
    }
    c0[i] = zc; cr[i] = bc; dr[i] = bd;
}

int sr_launch_search_direct(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, int cost, int var_block,
                            double thres, int want_c0) {
    size_t npx = pair->npx_cap;
    if (!pair->cr) {
        SR_CUDA(ctx, cudaMalloc((void**)&pair->cr, npx * 8));
        SR_CUDA(ctx, cudaMalloc((void**)&pair->dr, npx * 4));
        SR_CUDA(ctx, cudaMalloc((void**)&pair->nwin, npx * 4));
    }
    dim3 grid(cdiv(g.w2, 128), g.row1 - g.row0);
    if (grid.y == 0) return SR_OK;
    int* status = pair->maxblock + 1;
    if (cost == SR_COST_SAD)
        search_direct_right_kernel<SR_COST_SAD><<<grid, 128, 0, ctx->stream>>>(
            pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, g, g.block, var_block, thres, want_c0, pair->c0,
            pair->cr, pair->dr, pair->nwin, pair->maxblock, status);
    else
        search_direct_right_kernel<SR_COST_SSD_L2><<<grid, 128, 0, ctx->stream>>>(
            pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, g, g.block, var_block, thres, want_c0, pair->c0,
            pair->cr, pair->dr, pair->nwin, pair->maxblock, status);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// Cost at d == 0 for the fixed (border-clamped) window -> u32 plane in pair->c0 (the smoothing rule's
// "stay at 0" candidate; whether d == 0 is a candidate at all -- black pixel, break rule -- is decided
// by the resolve kernels).  Separable box sum, O(1) per pixel: the window [x-l, x+r) x [y-u, y+dn) is the
// fixed window [x-h, x+h) x [y-h, y+h) with the pointwise cost forced to 0 outside columns [0, w2-1)
// and rows [0, h2-1) (same identity as the tile kernel).
template <int COST>
__device__ __forceinline__ uint32_t pcost0(const uint32_t* __restrict__ Lp, int pitchL, const uint32_t* __restrict__ Rp,
                                           int pitchR, int y, int x) {
    uint32_t l = __ldg(Lp + (size_t)y * pitchL + x), r = __ldg(Rp + (size_t)y * pitchR + x);
    uint32_t ad = __vabsdiffu4(l, r);
    return COST == SR_COST_SAD ? __dp4a(ad, 0x01010101u, 0u) : __dp4a(ad, ad, 0u);
}

#define C0_BAND 32
template <int COST>
__global__ void __launch_bounds__(128) cost0_vertical_kernel(const uint32_t* __restrict__ Lp, int pitchL,
                                                             const uint32_t* __restrict__ Rp, int pitchR, SearchGeom g,
                                                             uint32_t* __restrict__ vsum) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y0 = g.row0 + C0_BAND * blockIdx.y, y1 = min(y0 + C0_BAND, g.row1);
    if (x >= g.w2) return;
    const int h = g.half;
    const bool col_ok = x < g.w2 - 1;
    auto row_ok = [&](int r) { return r >= 0 && r < g.h2 - 1; };
    uint32_t s = 0;
    if (col_ok)
        for (int r = y0 - h; r < y0 + h - 1; r++)
            if (row_ok(r)) s += pcost0<COST>(Lp, pitchL, Rp, pitchR, r, x);
    for (int y = y0; y < y1; y++) {
        if (col_ok && row_ok(y + h - 1)) s += pcost0<COST>(Lp, pitchL, Rp, pitchR, y + h - 1, x);
        vsum[(size_t)y * g.w2 + x] = s;
        if (col_ok && row_ok(y - h)) s -= pcost0<COST>(Lp, pitchL, Rp, pitchR, y - h, x);
    }
}

__global__ void __launch_bounds__(128) cost0_horizontal_kernel(const uint32_t* __restrict__ vsum, SearchGeom g,
                                                               uint32_t* __restrict__ c0) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = g.row0 + blockIdx.y;
    if (x >= g.w2 || y >= g.row1) return;
    const uint32_t* row = vsum + (size_t)y * g.w2;
    uint32_t s = 0;
    int xa = max(x - g.half, 0), xb = min(x + g.half, g.w2);
    for (int i = xa; i < xb; i++) s += __ldg(row + i);
    c0[(size_t)y * g.w2 + x] = s;
}

int sr_launch_cost0(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, int cost) {
    int rows = g.row1 - g.row0;
    if (rows <= 0) return SR_OK;
    uint32_t* vsum = (uint32_t*)sr_scratch(ctx, 1, pair->npx_cap * sizeof(uint32_t));
    if (!vsum) return SR_E_CUDA;
    dim3 gridv(cdiv(g.w2, 128), cdiv(rows, C0_BAND));
    if (cost == SR_COST_SAD)
        cost0_vertical_kernel<SR_COST_SAD><<<gridv, 128, 0, ctx->stream>>>(pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, g, vsum);
    else
        cost0_vertical_kernel<SR_COST_SSD_L2><<<gridv, 128, 0, ctx->stream>>>(pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, g, vsum);
    SR_LAUNCHED(ctx);
    cost0_horizontal_kernel<<<dim3(cdiv(g.w2, 128), rows), 128, 0, ctx->stream>>>(vsum, g, (uint32_t*)pair->c0);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// Effective evaluation count of the reference semantics: for every non-black
// pixel of the computed rows, the candidates before the break.
__global__ void count_evals_kernel(const uint32_t* __restrict__ Rp, int pitchR, SearchGeom g,
                                   unsigned long long* __restrict__ total) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = g.row0 + blockIdx.y;
    unsigned long long n = 0;
    if (x < g.w2 && y < g.row1 && y < g.height && (__ldg(Rp + (size_t)y * pitchR + x) & 0xFFFFFFu) != 0) {
        int r = min(g.w2 - x - 1, g.half);
        int lim = min(g.dhi, g.w1 - x - r);
        if (lim > g.dlo) n = (unsigned long long)(lim - g.dlo);
    }
    for (int o = 16; o; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(total, n);
}

int sr_launch_count_evals(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g) {
    SR_CUDA(ctx, cudaMemsetAsync(pair->evals, 0, 8, ctx->stream));
    dim3 grid(cdiv(g.w2, 128), g.row1 - g.row0);
    if (grid.y == 0) return SR_OK;
    count_evals_kernel<<<grid, 128, 0, ctx->stream>>>(pair->R.px, pair->R.pitch, g, pair->evals);
    SR_LAUNCHED(ctx);
    return SR_OK;
}
