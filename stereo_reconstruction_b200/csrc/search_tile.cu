// search_tile.cu -- the fused cost + winner-take-all kernel (BlockSearch.cpp:100-175, fixed window).
//
// Work decomposition (B200: 148 SMs, 64K registers and up to 227 KB shared memory per SM):
//   CTA    = one column tile (496 output columns) x one band of rows; NW warps, loops over passes
//            of 4*NW candidate disparities.
//   warp   = 4 consecutive disparities (one aligned group) of the pass, all 512 tile columns.
//   thread = XT=16 consecutive image columns x DT=4 disparities: 64 vertical column sums in
//            registers (the cost volume never leaves the register file).
// Per output row:  (1) vertical sliding update of the 64 column sums: add the pointwise cost of the
//   row entering the window, subtract the row leaving it (packed BGRx pixels: VABSDIFF4 + IDP.4A);
//   (2) horizontal window sums by an in-register sliding sum over the thread's 16 columns plus
//   2h-1 halo columns fetched from the next lane with SHFL; (3) first-minimum WTA on the key
//   cost*64 + d_local (smallest d wins ties, BlockSearch.cpp:168), min over the thread's 4
//   disparities in registers, over the CTA's warps through shared memory, over passes in a
//   CTA-private read-modify-write of the (cost<<32 | d) record in HBM/L2.
// Image rows are staged once per CTA and pass into a shared-memory ring (cp.async, 16 B chunks).
// Shared-memory rows are padded by one 16 B chunk after every 4 chunks (one thread's 16 pixels):
// the 64 B lane stride of the LDS.128 reads becomes 80 B, which is bank-conflict free, and every
// read address is "lane base + warp-uniform offset".
//
// Window geometry of the reference: rows [y-u, y+dn), cols [x-l, x+r) with l=min(x,h),
// r=min(w2-x-1,h), u=min(y,h), dn=min(h2-y-1,h)  ==  the fixed window [x-h,x+h) x [y-h,y+h) with
// the pointwise cost forced to 0 outside columns [0,w2-1) and rows [0,h2-1).
#include "sr_internal.cuh"

namespace {

constexpr int XT = 16;             // columns per thread
constexpr int DT = 4;              // disparities per thread
constexpr int TILE_OUT = 31 * XT;  // 496 output columns per CTA (lane 31 only feeds lane 30's halo)
constexpr int RCH = 128;           // right-image 16 B chunks per ring row (512 pixels)
constexpr int LCH = 144;           // left-image chunks per ring row (512 + 4*16 pixels)
constexpr int RPOS = RCH + RCH / 4;  // padded chunk positions
constexpr int LPOS = LCH + LCH / 4;
constexpr int SLOT_U4 = RPOS + LPOS;  // uint4 per ring slot
constexpr int KEY_U4 = 160;           // uint4 per warp key row (128 chunks padded)
constexpr int PF = 2;                 // rows prefetched ahead
#ifndef SR_TILE_FMA_FROM
#define SR_TILE_FMA_FROM 2             // disparity pairs >= this index slide on the FMA pipe (4 = none)
#endif
constexpr int FMA_FROM = SR_TILE_FMA_FROM;

__device__ __forceinline__ int padpos(int c) { return c + (c >> 2); }

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

template <int COST>
__device__ __forceinline__ uint32_t pcost_add(uint32_t l, uint32_t r, uint32_t acc) {
    if (COST == SR_COST_SAD) {
        uint32_t o;
        asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(o) : "r"(l), "r"(r), "r"(acc));
        return o;
    }
    uint32_t ad = __vabsdiffu4(l, r);
    return __dp4a(ad, ad, acc);
}

// 16-byte shared-memory load at a 32-bit byte offset from the dynamic shared base
__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
    return v;
}

// vertical update of the 64 column sums with one staged image row (ADD: row enters the window).
// slot: shared address of the ring slot (warp-uniform); rb: this lane's byte offset inside the right
// segment; lb[i]: byte offsets (lane base + warp-uniform chunk) of the 5 left chunks.
template <int COST, bool ADD>
__device__ __forceinline__ void vertical_update(uint32_t (&col)[XT][DT], uint32_t slot, uint32_t rb,
                                                const uint32_t (&lb)[5], uint32_t colmask) {
    uint4 la = lds128(slot + lb[0]);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint4 lbv = lds128(slot + lb[k + 1]);
        uint4 rr = lds128(slot + rb + 16 * k);
        uint32_t l7[7] = {la.x, la.y, la.z, la.w, lbv.x, lbv.y, lbv.z};
        uint32_t r4[4] = {rr.x, rr.y, rr.z, rr.w};
#pragma unroll
        for (int jj = 0; jj < 4; jj++)
#pragma unroll
            for (int t = 0; t < DT; t++) {
                if (ADD) col[4 * k + jj][t] = pcost_add<COST>(l7[jj + t], r4[jj], col[4 * k + jj][t]);
                else col[4 * k + jj][t] -= pcost_add<COST>(l7[jj + t], r4[jj], 0u);
            }
        la = lbv;
    }
    if (colmask != 0xFFFFu) {
#pragma unroll
        for (int j = 0; j < XT; j++)
            if (!((colmask >> j) & 1u)) {
#pragma unroll
                for (int t = 0; t < DT; t++) col[j][t] = 0;
            }
    }
}

struct TileArgs {
    const uint32_t* L; int pitchL;
    const uint32_t* R; int pitchR;
    uint32_t* planes;    // one u32 key plane (cost*64 + d_local) per pass of 64 disparities
    size_t plane_stride; // elements per plane
    int w1, h1, w2, h2;
    int half;
    int dlo, dhi;        // candidate range
    int dbase;           // dlo rounded down to a multiple of 4
    int npass;
    int row0, row1;      // output rows (already clipped to min(h1,h2))
    int band;            // rows per CTA band
    uint32_t k64;        // == 64, passed as data so that key = S*64+d stays one IMAD (FMA pipe)
    uint32_t kp1, km1;   // == 1, == 0xFFFFFFFF (-1): IMAD-based sliding sums
};

// horizontal window sums + in-thread WTA for one output row.  Disparities are processed in pairs so
// that the two keys fold into one 3-input min (VIMNMX3).  The sliding update S += in - out of the
// second pair runs as two IMADs (FMA pipe: multipliers kp1 = 1, km1 = -1 are kernel arguments so that
// they are not strength-reduced) to balance the two integer pipes.  MASKED: candidates outside
// [dlo, dlim(x)) are discarded (tiles that touch the break rule or the range ends).
template <int H2, bool MASKED>
__device__ __forceinline__ void window_wta(const uint32_t (&col)[XT][DT], uint32_t (&keys)[XT], uint32_t k64,
                                           uint32_t kp1, uint32_t km1, uint32_t dl0, int d0, int x0, const TileArgs& a) {
    constexpr int HALO = H2 - 1;
#pragma unroll
    for (int tp = 0; tp < DT; tp += 2) {
        uint32_t w0[XT + HALO], w1[XT + HALO];
#pragma unroll
        for (int i = 0; i < XT; i++) { w0[i] = col[i][tp]; w1[i] = col[i][tp + 1]; }
#pragma unroll
        for (int i = 0; i < HALO; i++) {
            w0[XT + i] = __shfl_down_sync(0xffffffffu, col[i][tp], 1);
            w1[XT + i] = __shfl_down_sync(0xffffffffu, col[i][tp + 1], 1);
        }
        uint32_t S0 = 0, S1 = 0;
#pragma unroll
        for (int i = 0; i < H2; i++) { S0 += w0[i]; S1 += w1[i]; }
        const uint32_t dla = dl0 + tp, dlb = dl0 + tp + 1;
#pragma unroll
        for (int j = 0; j < XT; j++) {
            uint32_t ka = S0 * k64 + dla, kb = S1 * k64 + dlb;
            if (MASKED) {
                int x = x0 + j;
                int lim = min(a.dhi, a.w1 - x - min(a.w2 - x - 1, a.half));
                if (d0 + tp < a.dlo || d0 + tp >= lim) ka = 0xFFFFFFFFu;
                if (d0 + tp + 1 < a.dlo || d0 + tp + 1 >= lim) kb = 0xFFFFFFFFu;
            }
            keys[j] = min(min(keys[j], ka), kb);
            if (j + 1 < XT) {
                if (tp >= FMA_FROM) {
                    S0 = w0[j + H2] * kp1 + S0; S0 = w0[j] * km1 + S0;
                    S1 = w1[j + H2] * kp1 + S1; S1 = w1[j] * km1 + S1;
                } else {
                    S0 += w0[j + H2] - w0[j];
                    S1 += w1[j + H2] - w1[j];
                }
            }
        }
    }
}

template <int H2, int COST, int NW>
__global__ void __launch_bounds__(NW * 32, 1) search_tile_kernel(TileArgs a) {
    constexpr int NSLOT = H2 + PF + 1;   // +1: half of the warps subtract the leaving row one iteration late
    static_assert(RCH + LCH <= NW * 32, "one staged chunk per thread");
    extern __shared__ __align__(128) uint4 smem4[];
    uint4* ring = smem4;                          // NSLOT x SLOT_U4
    uint4* keysm = smem4 + NSLOT * SLOT_U4;       // 2 x NW x KEY_U4

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // warp-uniform for the compiler
    const int h = a.half;
    const int CB = -XT + TILE_OUT * (int)blockIdx.x;          // first own column of lane 0
    const int yb0 = a.row0 + a.band * (int)blockIdx.y;
    const int yb1 = min(yb0 + a.band, a.row1);
    if (yb0 >= yb1) return;
    const int cb = CB + XT * lane;                             // this thread's own columns [cb, cb+16)
    const uint32_t k64 = a.k64;

    // own columns that contribute (inside [0, w2-1)); others are forced to 0
    uint32_t colmask = 0;
#pragma unroll
    for (int j = 0; j < XT; j++) colmask |= (uint32_t)(cb + j >= 0 && cb + j < a.w2 - 1) << j;
    // per-output candidate limit: d < min(dhi, w1 - x - r(x))  (break rule, BlockSearch.cpp:148)
    int dlim_min = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < XT; j++) {
        int x = cb + h + j;
        if (x >= 0 && x < a.w2) dlim_min = min(dlim_min, min(a.dhi, a.w1 - x - min(a.w2 - x - 1, h)));
    }
    // lane bases inside a slot / key row (5 chunk positions per lane: 4 data + 1 pad), as 32-bit
    // shared-memory byte offsets: every hot-loop load is "uniform slot base + per-thread constant"
    const int lane_u4 = 5 * lane;
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem4);
    const uint32_t rb = 16u * lane_u4;
    uint32_t lb[5];
#pragma unroll
    for (int i = 0; i < 5; i++) lb[i] = 16u * (RPOS + lane_u4 + padpos(warp + i));
    const int y_start = yb0 - (H2 - 1);
    // this thread's chunk of every staged row (RCH + LCH <= NW*32 chunks per row)
    const int ld_dst = tid < RCH ? padpos(tid) : RPOS + padpos(tid - RCH);
    const int ld_src = CB + 4 * (tid < RCH ? tid : tid - RCH);
    auto row_ok = [&](int r) { return r >= 0 && r < a.h2 - 1; };
    auto slot_of = [&](int r) { return ((r % NSLOT) + NSLOT) % NSLOT; };

    {
        const int pass = (int)blockIdx.z;
        const int dcta = a.dbase + pass * (DT * NW);           // first disparity of the pass (multiple of 4)
        const int d0 = dcta + DT * warp;                        // this warp's 4 disparities
        const int nactive = min(NW, (a.dhi - dcta + DT - 1) / DT);  // warps with at least one d < dhi
        const bool active = warp < nactive;
        const bool late_sub = (warp >> 2) & 1;
        const bool needmask = __any_sync(0xffffffffu, d0 < a.dlo || d0 + DT - 1 >= dlim_min);

        uint32_t col[XT][DT];
#pragma unroll
        for (int j = 0; j < XT; j++)
#pragma unroll
            for (int t = 0; t < DT; t++) col[j][t] = 0;

        auto issue_row = [&](int r) {
            if (row_ok(r) && tid < RCH + LCH) {
                uint4* dst = ring + slot_of(r) * SLOT_U4 + ld_dst;
                const uint32_t* src = (tid < RCH ? a.R + (size_t)r * a.pitchR : a.L + (size_t)r * a.pitchL + dcta) + ld_src;
                cp_async16(dst, src);
            }
            cp_async_commit();
        };
        // CTA-wide reduction of the keys of output row y -> this pass's key plane
        uint32_t* plane = a.planes + (size_t)pass * a.plane_stride;
        auto reduce_row = [&](int y) {
            const uint32_t* kb = reinterpret_cast<const uint32_t*>(keysm + (y & 1) * NW * KEY_U4);
            for (int q = tid; q < TILE_OUT; q += NW * 32) {
                int x = CB + h + q;
                if (x < 0 || x >= a.w2) continue;
                int word = q + 4 * (q >> 4);
                uint32_t kk[NW];
#pragma unroll
                for (int w = 0; w < NW; w++) kk[w] = kb[w * (KEY_U4 * 4) + word];
                uint32_t k = 0xFFFFFFFFu;
#pragma unroll
                for (int w = 0; w < NW; w++) k = min(k, kk[w]);
                plane[(size_t)y * a.w2 + x] = k;
            }
        };

        // prologue: rows of the first PF iterations
#pragma unroll
        for (int i = 0; i < PF; i++) issue_row(y_start + i + h - 1);

        int s_in = slot_of(y_start + h - 1), s_out = slot_of(y_start - h);
        for (int y = y_start; y < yb1; y++) {
            cp_async_wait<PF - 1>();
            __syncthreads();
            issue_row(y + PF + h - 1);
            if (y - 1 >= yb0) reduce_row(y - 1);
            if (active) {
                // Phase staggering: the warps of one scheduler alternate between "subtract now" and
                // "subtract at the start of the next iteration", so that ALU-only phases (pointwise
                // costs) of some warps overlap the FMA-heavy window/WTA phase of the others.
                if (late_sub && y - 1 >= yb0 && row_ok(y - 1 - h) && colmask != 0) {
                    const int sp = s_out == 0 ? NSLOT - 1 : s_out - 1;
                    vertical_update<COST, false>(col, smem_base + (uint32_t)sp * (SLOT_U4 * 16), rb, lb, colmask);
                }
                const int rin = y + h - 1;
                if (row_ok(rin) && colmask != 0) {
                    vertical_update<COST, true>(col, smem_base + (uint32_t)s_in * (SLOT_U4 * 16), rb, lb, colmask);
                }
                if (y >= yb0) {
                    uint32_t keys[XT];
#pragma unroll
                    for (int j = 0; j < XT; j++) keys[j] = 0xFFFFFFFFu;
                    if (needmask) window_wta<H2, true>(col, keys, k64, a.kp1, a.km1, (uint32_t)(DT * warp), d0, cb + h, a);
                    else window_wta<H2, false>(col, keys, k64, a.kp1, a.km1, (uint32_t)(DT * warp), d0, cb + h, a);
                    uint4* kw = keysm + ((y & 1) * NW + warp) * KEY_U4 + lane_u4;
#pragma unroll
                    for (int k = 0; k < 4; k++)
                        kw[k] = make_uint4(keys[4 * k], keys[4 * k + 1], keys[4 * k + 2], keys[4 * k + 3]);
                    const int rout = y - h;
                    if (!late_sub && y + 1 < yb1 && row_ok(rout) && colmask != 0) {
                        vertical_update<COST, false>(col, smem_base + (uint32_t)s_out * (SLOT_U4 * 16), rb, lb, colmask);
                    }
                }
            }
            else if (y >= yb0) {
                uint4* kw = keysm + ((y & 1) * NW + warp) * KEY_U4 + lane_u4;
#pragma unroll
                for (int k = 0; k < 4; k++) kw[k] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
            }
            s_in = s_in + 1 == NSLOT ? 0 : s_in + 1;
            s_out = s_out + 1 == NSLOT ? 0 : s_out + 1;
        }
        cp_async_wait<0>();
        __syncthreads();
        reduce_row(yb1 - 1);
    }
}

template <int H2, int COST>
int launch(sr_ctx* ctx, const TileArgs& a, int ntx, int nbands) {
    constexpr int NW = 16;
    size_t smem = ((size_t)(H2 + PF + 1) * SLOT_U4 + 2 * NW * KEY_U4) * sizeof(uint4);
    static bool configured = false;
    if (!configured) {
        SR_CUDA(ctx, cudaFuncSetAttribute(search_tile_kernel<H2, COST, NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = true;
    }
    SR_CUDA(ctx, cudaEventRecord(ctx->kev0, ctx->stream));
    search_tile_kernel<H2, COST, NW><<<dim3(ntx, nbands, a.npass), NW * 32, smem, ctx->stream>>>(a);
    SR_LAUNCHED(ctx);
    SR_CUDA(ctx, cudaEventRecord(ctx->kev1, ctx->stream));
    ctx->kev_valid = true;
    return 1;
}

template <int COST>
int dispatch_h2(sr_ctx* ctx, const TileArgs& a, int ntx, int nbands) {
    switch (2 * a.half) {
        case 2: return launch<2, COST>(ctx, a, ntx, nbands);
        case 4: return launch<4, COST>(ctx, a, ntx, nbands);
        case 6: return launch<6, COST>(ctx, a, ntx, nbands);
        case 8: return launch<8, COST>(ctx, a, ntx, nbands);
        case 10: return launch<10, COST>(ctx, a, ntx, nbands);
        case 12: return launch<12, COST>(ctx, a, ntx, nbands);
        case 14: return launch<14, COST>(ctx, a, ntx, nbands);
        case 16: return launch<16, COST>(ctx, a, ntx, nbands);
        default: return 0;
    }
}

}  // namespace

int sr_launch_search_tile(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, int cost, int left_ref) {
    if (left_ref) return 0;
    if (g.half < 1 || g.half > 8) return 0;
    if (cost != SR_COST_SSD_L2 && cost != SR_COST_SAD) return 0;
    int r1 = g.row1 < g.height ? g.row1 : g.height;
    if (r1 <= g.row0) { pair->npass = 0; return 1; }  // nothing to compute: the resolve pass fills the rows
    TileArgs a;
    a.L = pair->L.px; a.pitchL = pair->L.pitch; a.R = pair->R.px; a.pitchR = pair->R.pitch;
    a.w1 = g.w1; a.h1 = g.h1; a.w2 = g.w2; a.h2 = g.h2; a.half = g.half;
    a.dlo = g.dlo; a.dhi = g.dhi; a.dbase = g.dlo & ~3;
    a.npass = (g.dhi - a.dbase + 63) / 64;
    a.row0 = g.row0; a.row1 = r1; a.k64 = 64; a.kp1 = 1u; a.km1 = 0xFFFFFFFFu;
    int ntx = (g.w2 + XT - g.half + TILE_OUT - 1) / TILE_OUT;
    int rows = r1 - g.row0;
    if (a.npass <= 0) {
        // empty candidate range: every record is "none"
        pair->npass = 0;
        return 1;
    }
    // one u32 key plane per pass (merged by the resolve kernels): no read-modify-write in the hot loop
    pair->npass = a.npass; pair->plane_dbase = a.dbase;
    pair->planes = (uint32_t*)sr_scratch(ctx, 0, (size_t)a.npass * pair->npx_cap * sizeof(uint32_t));
    if (!pair->planes) return SR_E_CUDA;
    a.planes = pair->planes; a.plane_stride = pair->npx_cap;
    // bands: CTAs = ntx * npass * nbands, one CTA per SM at a time; minimise waves * (band + warm-up rows)
    const int sms = 148;
    int maxnb = rows / (2 * g.half + 2);
    if (maxnb < 1) maxnb = 1;
    if (maxnb > 256) maxnb = 256;
    long bestcost = -1;
    int nbands = 1;
    for (int nb = 1; nb <= maxnb; nb++) {
        int band = (rows + nb - 1) / nb;
        int nbe = (rows + band - 1) / band;
        long ctas = (long)ntx * a.npass * nbe;
        long waves = (ctas + sms - 1) / sms;
        long c = waves * (band + 2 * g.half - 1 + 2);
        if (bestcost < 0 || c < bestcost) { bestcost = c; nbands = nbe; a.band = band; }
    }
    if (cost == SR_COST_SAD) return dispatch_h2<SR_COST_SAD>(ctx, a, ntx, nbands);
    return dispatch_h2<SR_COST_SSD_L2>(ctx, a, ntx, nbands);
}
