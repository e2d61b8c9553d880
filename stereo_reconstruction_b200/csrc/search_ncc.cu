// search_ncc.cu -- NCC (zero-mean normalised cross-correlation, fp32) through sliding sums.
// Extension cost of north_star / SURVEY.md 8c (no reference counterpart); the definition is
//   num = n*S_lr - S_l*S_r,  cost = -(((float)num * rl) * rr),  rl = 1/sqrt((float)(n*S_ll - S_l^2)), rr likewise,
// with exact integer window sums over the reference's window [x-l, x+r) x [y-u, y+dn).
//
// Interior pixels (full 2h x 2h window) are searched here; the frame of border pixels with clamped windows
// goes through the per-pixel kernel of search_ext.cu.  Structure = the fused tile kernel of search_tile.cu
// with the dot product as pointwise "cost":
//   * thread = 8 image columns x 4 disparities: 32 column sums of L(x+d).R(x) (IDP.4A) in registers,
//     updated per output row by adding the row that enters the window and subtracting the one leaving;
//   * horizontal window sums S_lr by an in-register sliding sum, halo columns from the next lane(s) by SHFL;
//   * the d-independent window moments come from tables (ncc_tables: -S_l and rl per left window,
//     S_r and rr per right window, indexed by the window's first column), staged per row with cp.async;
//   * per candidate: num as one 64-bit IMAD.WIDE chain, (float)num by the hardware I2F.S64 (round to nearest, the
//     conversion the CPU performs; measured faster than an exact hi/lo magic-number split + FFMA, which costs six
//     issue slots), two FMULs, compare/select against the running best (strict >: the first disparity wins ties);
//   * (score, d) pairs of the 16 warps are merged through shared memory, one pair per pass of 64
//     disparities goes to HBM, the passes are merged by ncc_merge_kernel (search_ext.cu).
// Shared-memory rows use "2 data chunks + 1 pad chunk" per lane (48 B lane stride): LDS.128 conflict-free.
#include <cstdlib>
#include "sr_internal.cuh"

namespace {

constexpr int XT = 8, DT = 4, NW = 16;
constexpr int RCH = 64;                 // right-image 16 B chunks per staged row (256 pixels)
constexpr int LCH = 82;                 // left-image chunks (256 + 64 disparities + 8 pixels)
constexpr int RPOS = RCH + RCH / 2;     // padded chunk positions
constexpr int LPOS = LCH + LCH / 2;
constexpr int SLOT_U4 = RPOS + LPOS;    // image ring slot
constexpr int TSLOT_U4 = 2 * RPOS + 2 * LPOS;   // table slot: S_r, rr, -S_l, rl
constexpr int NT = 3;                   // table slots (PF + 1)
constexpr int KEY_U4 = 160;             // uint4 per warp key row: 32 lanes x (4 data + 1 pad)
constexpr int PF = 2;

__device__ __forceinline__ int padpos(int c) { return c + (c >> 1); }
__device__ __forceinline__ void cp_async16_s(uint32_t saddr, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(saddr), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
    return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t saddr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t saddr, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(saddr), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}

struct NccArgs {
    const uint32_t* L; int pitchL;
    const uint32_t* R; int pitchR;
    const int32_t* nsl; const float* rl;    // left tables (pitchL), indexed by window start column
    const int32_t* sr; const float* rr;     // right tables (pitchR)
    float* pscore; int* pdisp;              // per-pass planes
    size_t plane_stride;
    int w1, h1, w2, h2, half;
    int dlo, dhi, dbase, npass;
    int row0, row1, band;
};

// col[j][t] +/-= sum over the 3 channels of L(x_j + d_t) * R(x_j) for one staged image row
template <bool ADD>
__device__ __forceinline__ void vertical_update(uint32_t (&col)[XT][DT], uint32_t slot, uint32_t lane_b, const uint32_t (&uo)[3]) {
    uint4 r0 = lds128(slot + lane_b), r1 = lds128(slot + lane_b + 16);
    uint4 l0 = lds128(slot + uo[0] + lane_b), l1 = lds128(slot + uo[1] + lane_b), l2 = lds128(slot + uo[2] + lane_b);
    uint32_t r[XT] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    uint32_t l[XT + DT] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w, l2.x, l2.y, l2.z, l2.w};
#pragma unroll
    for (int j = 0; j < XT; j++)
#pragma unroll
        for (int t = 0; t < DT; t++) {
            if (ADD) col[j][t] = __dp4a(l[j + t], r[j], col[j][t]);
            else col[j][t] -= __dp4a(l[j + t], r[j], 0u);
        }
}

// Window sums + scores + running best for one output row.  tb: shared address of this row's table slot.
template <int H2, bool MASKED>
__device__ __forceinline__ void window_best(const uint32_t (&col)[XT][DT], float (&best)[XT], int (&bd)[XT], uint32_t tb,
                                            uint32_t lane_b, const uint32_t (&uo)[3], int d0, int x0, const NccArgs& a) {
    constexpr int HALO = H2 - 1;
    constexpr uint32_t N = (uint32_t)(H2 * H2 * 3);
    // tables: S_r / rr of the thread's 8 right windows, -S_l / rl of its 11 left windows
    uint4 q0 = lds128(tb + lane_b), q1 = lds128(tb + lane_b + 16);
    uint4 f0 = lds128(tb + 16u * RPOS + lane_b), f1 = lds128(tb + 16u * RPOS + lane_b + 16);
    const uint32_t tl = tb + 32u * RPOS, tr = tl + 16u * LPOS;
    uint4 a0 = lds128(tl + uo[0] + lane_b), a1 = lds128(tl + uo[1] + lane_b), a2 = lds128(tl + uo[2] + lane_b);
    uint4 b0 = lds128(tr + uo[0] + lane_b), b1 = lds128(tr + uo[1] + lane_b), b2 = lds128(tr + uo[2] + lane_b);
    const int sr[XT] = {(int)q0.x, (int)q0.y, (int)q0.z, (int)q0.w, (int)q1.x, (int)q1.y, (int)q1.z, (int)q1.w};
    const float rr[XT] = {__uint_as_float(f0.x), __uint_as_float(f0.y), __uint_as_float(f0.z), __uint_as_float(f0.w),
                          __uint_as_float(f1.x), __uint_as_float(f1.y), __uint_as_float(f1.z), __uint_as_float(f1.w)};
    const int nsl[XT + DT] = {(int)a0.x, (int)a0.y, (int)a0.z, (int)a0.w, (int)a1.x, (int)a1.y, (int)a1.z, (int)a1.w,
                              (int)a2.x, (int)a2.y, (int)a2.z, (int)a2.w};
    const float rl[XT + DT] = {__uint_as_float(b0.x), __uint_as_float(b0.y), __uint_as_float(b0.z), __uint_as_float(b0.w),
                               __uint_as_float(b1.x), __uint_as_float(b1.y), __uint_as_float(b1.z), __uint_as_float(b1.w),
                               __uint_as_float(b2.x), __uint_as_float(b2.y), __uint_as_float(b2.z), __uint_as_float(b2.w)};
#pragma unroll
    for (int t = 0; t < DT; t++) {
        uint32_t w[XT + HALO];
#pragma unroll
        for (int i = 0; i < XT; i++) w[i] = col[i][t];
#pragma unroll
        for (int i = 0; i < HALO; i++)
            w[XT + i] = i < XT ? __shfl_down_sync(0xffffffffu, col[i][t], 1) : __shfl_down_sync(0xffffffffu, col[i - XT][t], 2);
        uint32_t S = 0;
#pragma unroll
        for (int i = 0; i < H2; i++) S += w[i];
        const int d = d0 + t;
#pragma unroll
        for (int j = 0; j < XT; j++) {
            // num = n * S_lr - S_l * S_r  (|num| < 2^36)
            long long num = (long long)((unsigned long long)S * N) + (long long)nsl[j + t] * (long long)sr[j];
            float score = __fmul_rn(__fmul_rn(__ll2float_rn(num), rl[j + t]), rr[j]);
            bool better = score > best[j];
            if (MASKED) {
                int x = x0 + j;
                better = better && d >= a.dlo && d < min(a.dhi, a.w1 - x - a.half);
            }
            if (better) { best[j] = score; bd[j] = d; }
            if (j + 1 < XT) S += w[j + H2] - w[j];
        }
    }
}

template <int H2>
__global__ void __launch_bounds__(NW * 32, 1) search_ncc_kernel(NccArgs a) {
    constexpr int HALO = H2 - 1;
    constexpr int NLH = HALO > XT ? 2 : 1;              // lanes that only feed their neighbours' halo
    constexpr int TILE_OUT = (32 - NLH) * XT;
    constexpr int NSLOT = H2 + PF + 1;
    constexpr uint32_t SLOT_B = SLOT_U4 * 16, RING_B = NSLOT * SLOT_B, TSLOT_B = TSLOT_U4 * 16, TAB_B = NT * TSLOT_B;
    constexpr uint32_t KEYROW_B = KEY_U4 * 16, KEYBUF_B = NW * KEYROW_B;
    extern __shared__ __align__(128) uint4 smem4[];     // image ring | table slots | 2 key buffers

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int h = a.half;
    const int CB = -XT + TILE_OUT * (int)blockIdx.x;     // first own column of lane 0
    const int yb0 = a.row0 + a.band * (int)blockIdx.y;
    const int yb1 = min(yb0 + a.band, a.row1);
    if (yb0 >= yb1) return;
    const int cb = CB + XT * lane;
    const int pass = (int)(gridDim.z - 1 - blockIdx.z);
    const int dcta = a.dbase + pass * (DT * NW);
    const int d0 = dcta + DT * warp;
    const int x0 = cb + h;                                // first output column of this thread

    // candidate limits (interior columns only: h <= x <= w2-1-h): d < min(dhi, w1 - x - h), non-increasing in x
    const int x_first = max(CB + h, h);
    const int dlim_tile = min(a.dhi, a.w1 - x_first - h);
    if (dcta >= dlim_tile) return;                        // planes are pre-filled with "no candidate"
    const bool active = d0 < dlim_tile && d0 + DT > a.dlo;
    int dlim_min = 0x7fffffff;
#pragma unroll
    for (int j = 0; j < XT; j++) {
        int x = x0 + j;
        if (x >= h && x <= a.w2 - 1 - h) dlim_min = min(dlim_min, min(a.dhi, a.w1 - x - h));
    }
    const bool needmask = __any_sync(0xffffffffu, d0 < a.dlo || d0 + DT - 1 >= dlim_min);

    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem4);
    const uint32_t tab_base = smem_base + RING_B, keys_base = tab_base + TAB_B;
    const uint32_t lane_b = 48u * lane;
    uint32_t uo[3], uoL[3];
#pragma unroll
    for (int i = 0; i < 3; i++) { uo[i] = 16u * padpos(warp + i); uoL[i] = 16u * (RPOS + padpos(warp + i)); }
    auto slot_of = [&](int r) { return (uint32_t)(((r % NSLOT) + NSLOT) % NSLOT) * SLOT_B; };
    auto next_slot = [&](uint32_t sb) { return sb + SLOT_B == RING_B ? 0u : sb + SLOT_B; };
    auto row_ok = [&](int r) { return r >= 0 && r < a.h2 - 1; };

    // staging: image chunk `tid` of the row entering PF rows ahead, table chunk `tid` of the output row PF ahead
    const int y_start = yb0 - (H2 - 1);
    int r_pf = y_start + h - 1;                           // image row staged next
    int t_pf = y_start;                                   // table row staged next
    const bool im_on = tid < RCH + LCH, im_right = tid < RCH;
    const uint32_t im_dst = smem_base + 16u * (im_right ? padpos(tid) : RPOS + padpos(tid - RCH));
    const int im_pitch = im_right ? a.pitchR : a.pitchL;
    const uint32_t* im_ptr = (im_right ? a.R : a.L + dcta) + (ptrdiff_t)r_pf * im_pitch + CB + 4 * (im_right ? tid : tid - RCH);
    // table segments: [0,RCH) S_r, [RCH,2RCH) rr, [2RCH, 2RCH+LCH) -S_l, [2RCH+LCH, 2RCH+2LCH) rl
    const bool tb_on = tid < 2 * RCH + 2 * LCH;
    const int seg = tid < RCH ? 0 : tid < 2 * RCH ? 1 : tid < 2 * RCH + LCH ? 2 : 3;
    const int tch = tid - (seg == 0 ? 0 : seg == 1 ? RCH : seg == 2 ? 2 * RCH : 2 * RCH + LCH);
    const uint32_t tb_dst = tab_base + 16u * ((seg == 0 ? 0 : seg == 1 ? RPOS : seg == 2 ? 2 * RPOS : 2 * RPOS + LPOS) + padpos(tch));
    const int tb_pitch = seg < 2 ? a.pitchR : a.pitchL;
    const uint32_t* tb_src0 = seg == 0 ? (const uint32_t*)a.sr : seg == 1 ? (const uint32_t*)a.rr
                            : seg == 2 ? (const uint32_t*)a.nsl : (const uint32_t*)a.rl;
    const uint32_t* tb_ptr = tb_src0 + (seg < 2 ? 0 : dcta) + (ptrdiff_t)t_pf * tb_pitch + CB + 4 * tch;
    uint32_t slot_pf = slot_of(r_pf);
    auto issue_row = [&]() {
        if (row_ok(r_pf) && im_on) cp_async16_s(slot_pf + im_dst, im_ptr);
        if (tb_on && t_pf >= yb0 && t_pf < yb1) cp_async16_s(tb_dst + (uint32_t)(t_pf % NT) * TSLOT_B, tb_ptr);
        cp_async_commit();
        r_pf++; t_pf++; im_ptr += im_pitch; tb_ptr += tb_pitch; slot_pf = next_slot(slot_pf);
    };

    // merge of the 16 warps' (score, d) pairs: warps 0..7 serve even rows, warps 8..15 odd rows; each
    // quarter-warp owns one producer lane's 8 columns, the two quarter-warps of a half-warp are 4
    // producer lanes apart (disjoint bank ranges -> conflict-free LDS.64)
    const int wr = warp & 7;
    const int rgrp = (wr >> 2) * 16 + (wr & 3) + 4 * (lane >> 3);
    const int xq = CB + XT * rgrp + h + (lane & 7);
    const bool red_col = rgrp < 32 - NLH && xq >= h && xq <= a.w2 - 1 - h;
    const uint32_t red_b = keys_base + 80u * rgrp + 8u * (lane & 7);
    auto reduce_row = [&](int y, uint32_t buf) {
        if ((warp >> 3) != (y & 1) || !red_col) return;
        float best = -INFINITY;
        int bd = -1;
#pragma unroll
        for (int w = 0; w < NW; w++) {
            uint2 k = lds64(red_b + buf + w * KEYROW_B);
            float s = __uint_as_float(k.x);
            if (s > best) { best = s; bd = (int)k.y; }
        }
        size_t o = (size_t)pass * a.plane_stride + (size_t)y * a.w2 + xq;
        a.pscore[o] = best;
        a.pdisp[o] = bd;
    };

    uint32_t col[XT][DT];
#pragma unroll
    for (int j = 0; j < XT; j++)
#pragma unroll
        for (int t = 0; t < DT; t++) col[j][t] = 0;
#pragma unroll
    for (int i = 0; i < PF; i++) issue_row();

    uint32_t s_in = slot_of(y_start + h - 1), s_out = slot_of(y_start - h);
    const uint32_t key_st = keys_base + warp * KEYROW_B + 80u * lane;
    for (int y = y_start; y < yb1; y++) {
        const uint32_t buf = (y & 1) ? KEYBUF_B : 0u;
        cp_async_wait<PF - 1>();
        __syncthreads();
        issue_row();
        if (y - 1 >= yb0) reduce_row(y - 1, buf ^ KEYBUF_B);
        if (active) {
            if (row_ok(y + h - 1)) vertical_update<true>(col, smem_base + s_in, lane_b, uoL);
            if (y >= yb0) {
                float best[XT];
                int bd[XT];
#pragma unroll
                for (int j = 0; j < XT; j++) { best[j] = -INFINITY; bd[j] = -1; }
                const uint32_t tb = tab_base + (uint32_t)(y % NT) * TSLOT_B;
                if (needmask) window_best<H2, true>(col, best, bd, tb, lane_b, uo, d0, x0, a);
                else window_best<H2, false>(col, best, bd, tb, lane_b, uo, d0, x0, a);
#pragma unroll
                for (int k = 0; k < 4; k++)
                    sts128(key_st + buf + 16 * k, __float_as_uint(best[2 * k]), (uint32_t)bd[2 * k],
                           __float_as_uint(best[2 * k + 1]), (uint32_t)bd[2 * k + 1]);
            }
            // the row leaving the window; during the warm-up rows (y < yb0) nothing has to leave yet
            if (y >= yb0 && y + 1 < yb1 && row_ok(y - h)) vertical_update<false>(col, smem_base + s_out, lane_b, uoL);
        } else if (y >= yb0) {
#pragma unroll
            for (int k = 0; k < 4; k++) sts128(key_st + buf + 16 * k, 0xFF800000u, 0xFFFFFFFFu, 0xFF800000u, 0xFFFFFFFFu);
        }
        s_out = next_slot(s_out); s_in = next_slot(s_in);
    }
    cp_async_wait<0>();
    __syncthreads();
    reduce_row(yb1 - 1, ((yb1 - 1) & 1) ? KEYBUF_B : 0u);
}

template <int H2>
int launch_ncc(sr_ctx* ctx, const NccArgs& a, int ntx, int nbands) {
    size_t smem = ((size_t)(H2 + PF + 1) * SLOT_U4 + (size_t)NT * TSLOT_U4 + 2 * NW * KEY_U4) * sizeof(uint4);
    static DeviceOnce configured;
    if (configured.need(ctx->device)) {
        SR_CUDA(ctx, cudaFuncSetAttribute(search_ncc_kernel<H2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured.done(ctx->device);
    }
    SR_CUDA(ctx, cudaEventRecord(ctx->kev0, ctx->stream));
    search_ncc_kernel<H2><<<dim3(ntx, nbands, a.npass), NW * 32, smem, ctx->stream>>>(a);
    SR_LAUNCHED(ctx);
    SR_CUDA(ctx, cudaEventRecord(ctx->kev1, ctx->stream));
    ctx->kev_valid = true;
    return 1;
}

}  // namespace

// ---- window-moment tables ---------------------------------------------------------------------------
// V1/V2[y][x] = sum over rows [y-h, y+h) of the 3-channel sum / sum of squares of pixel (., x);
// tables[y][xs] over columns [xs, xs+2h): S (negated for the left image) and 1/sqrt((float)(n*S2 - S^2)).
#define NCC_BAND 32
__global__ void __launch_bounds__(128) ncc_vertical_kernel(const uint32_t* __restrict__ img, int pitch, int w, int himg, int half,
                                                           int ya, int yb, uint32_t* __restrict__ v1, uint32_t* __restrict__ v2) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y0 = ya + NCC_BAND * blockIdx.y, y1 = min(y0 + NCC_BAND, yb);
    if (x >= w || y0 >= y1) return;
    uint32_t s1 = 0, s2 = 0;
    for (int r = y0 - half; r < y0 + half - 1; r++) {
        uint32_t p = __ldg(img + (size_t)r * pitch + x);
        s1 = __dp4a(p, 0x01010101u, s1); s2 = __dp4a(p, p, s2);
    }
    for (int y = y0; y < y1; y++) {
        uint32_t p = __ldg(img + (size_t)(y + half - 1) * pitch + x);
        s1 = __dp4a(p, 0x01010101u, s1); s2 = __dp4a(p, p, s2);
        v1[(size_t)y * w + x] = s1; v2[(size_t)y * w + x] = s2;
        uint32_t q = __ldg(img + (size_t)(y - half) * pitch + x);
        s1 -= __dp4a(q, 0x01010101u, 0u); s2 -= __dp4a(q, q, 0u);
    }
}

__global__ void __launch_bounds__(128) ncc_horizontal_kernel(const uint32_t* __restrict__ v1, const uint32_t* __restrict__ v2, int w,
                                                             int half, int ya, int negate, int pitch, int32_t* __restrict__ ts,
                                                             float* __restrict__ tr) {
    int xs = blockIdx.x * blockDim.x + threadIdx.x;
    int y = ya + blockIdx.y;
    if (xs > w - 2 * half) return;
    const uint32_t* r1 = v1 + (size_t)y * w + xs;
    const uint32_t* r2 = v2 + (size_t)y * w + xs;
    uint32_t s1 = 0, s2 = 0;
    for (int i = 0; i < 2 * half; i++) { s1 += __ldg(r1 + i); s2 += __ldg(r2 + i); }
    long long n = (long long)(2 * half) * (2 * half) * 3;
    long long var = n * (long long)s2 - (long long)s1 * (long long)s1;
    float inv = var > 0 ? __fdiv_rn(1.0f, __fsqrt_rn(__ll2float_rn(var))) : __int_as_float(0x7FC00000);   // NaN: never the best
    ts[(size_t)y * pitch + xs] = negate ? -(int32_t)s1 : (int32_t)s1;
    tr[(size_t)y * pitch + xs] = inv;
}

static int build_tables(sr_ctx* ctx, const PackedImage& img, int half, int ya, int yb, int negate, int32_t* ts, float* tr) {
    if (yb <= ya) return SR_OK;
    size_t n = (size_t)img.w * img.h;
    uint32_t* v1 = (uint32_t*)sr_scratch(ctx, 2, n * sizeof(uint32_t));
    uint32_t* v2 = (uint32_t*)sr_scratch(ctx, 3, n * sizeof(uint32_t));
    if (!v1 || !v2) return SR_E_CUDA;
    ncc_vertical_kernel<<<dim3(cdiv(img.w, 128), cdiv(yb - ya, NCC_BAND)), 128, 0, ctx->stream>>>(img.px, img.pitch, img.w, img.h, half,
                                                                                                 ya, yb, v1, v2);
    SR_LAUNCHED(ctx);
    ncc_horizontal_kernel<<<dim3(cdiv(img.w, 128), yb - ya), 128, 0, ctx->stream>>>(v1, v2, img.w, half, ya, negate, img.pitch,
                                                                                    ts + SR_PADL, tr + SR_PADL);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// Interior pixels of the right-referenced NCC search -> per-pass (score, d) planes in ctx scratch 0.
// Returns 1 when it ran, 0 when the shape is not supported (caller falls back), < 0 on error.
int sr_launch_search_ncc(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, float** pscore, int** pdisp, void** tab) {
    const int h = g.half;
    if (h < 1 || h > 8) return 0;
    if (g.w2 < 2 * h + 2 || g.h2 < 2 * h + 2 || g.w1 < 2 * h + 2 || g.h1 < 2 * h + 2) return 0;
    NccArgs a;
    a.dlo = g.dlo; a.dhi = g.dhi; a.dbase = g.dlo & ~3;
    a.npass = (g.dhi - a.dbase + 63) / 64;
    if (a.npass <= 0) return 0;
    // interior output rows: full window inside both images
    int r0 = g.row0 > h ? g.row0 : h;
    int r1 = g.row1 < g.height ? g.row1 : g.height;
    if (r1 > g.h2 - h) r1 = g.h2 - h;
    if (r1 > g.h1 - h) r1 = g.h1 - h;
    size_t npx = pair->npx_cap;
    uint32_t* planes = (uint32_t*)sr_scratch(ctx, 0, (size_t)a.npass * npx * 2 * sizeof(uint32_t));
    if (!planes) return SR_E_CUDA;
    *pscore = (float*)planes; *pdisp = (int*)(planes + (size_t)a.npass * npx);
    pair->npass = a.npass; pair->plane_dbase = a.dbase;
    // "no candidate" everywhere: score NaN (never > best), d = -1
    SR_CUDA(ctx, cudaMemsetAsync(planes, 0xFF, (size_t)a.npass * npx * 2 * sizeof(uint32_t), ctx->stream));
    if (r1 <= r0) return 1;
    // tables of the images currently bound to pair->L / pair->R (the mirrored search binds its own pair and table)
    if (!*tab) {
        size_t nl = (size_t)pair->L.pitch * pair->L.h, nr = (size_t)pair->R.pitch * pair->R.h;
        SR_CUDA(ctx, cudaMalloc(tab, (2 * nl + 2 * nr) * 4));
        SR_CUDA(ctx, cudaMemsetAsync(*tab, 0, (2 * nl + 2 * nr) * 4, ctx->stream));
    }
    size_t nl = (size_t)pair->L.pitch * pair->L.h, nr = (size_t)pair->R.pitch * pair->R.h;
    int32_t* nsl = (int32_t*)*tab; float* rl = (float*)(nsl + nl);
    int32_t* srt = (int32_t*)(rl + nl); float* rrt = (float*)(srt + nr);
    int rc;
    if ((rc = build_tables(ctx, pair->L, h, r0, r1, 1, nsl, rl)) != SR_OK) return rc;
    if ((rc = build_tables(ctx, pair->R, h, r0, r1, 0, srt, rrt)) != SR_OK) return rc;
    a.L = pair->L.px; a.pitchL = pair->L.pitch; a.R = pair->R.px; a.pitchR = pair->R.pitch;
    a.nsl = nsl + SR_PADL; a.rl = rl + SR_PADL; a.sr = srt + SR_PADL; a.rr = rrt + SR_PADL;
    a.pscore = *pscore; a.pdisp = *pdisp; a.plane_stride = npx;
    a.w1 = g.w1; a.h1 = g.h1; a.w2 = g.w2; a.h2 = g.h2; a.half = h;
    a.row0 = r0; a.row1 = r1;
    const int tile_out = (2 * h - 1 > XT ? 30 : 31) * XT;
    int ntx = cdiv(g.w2 - 2 * h + XT, tile_out);
    int rows = r1 - r0;
    const int sms = ctx->sms;
    int maxnb = rows / (2 * h + 2); if (maxnb < 1) maxnb = 1; if (maxnb > 256) maxnb = 256;
    long bestcost = -1; int nbands = 1; a.band = rows;
    for (int nb = 1; nb <= maxnb; nb++) {
        int band = cdiv(rows, nb), nbe = cdiv(rows, band);
        long ctas = (long)ntx * a.npass * nbe;
        long c = (long)((double)(ctas + sms) / sms * (band + 2 * h + 1) * 16.0);
        if (bestcost < 0 || c < bestcost) { bestcost = c; nbands = nbe; a.band = band; }
    }
    switch (2 * h) {
        case 2: return launch_ncc<2>(ctx, a, ntx, nbands);
        case 4: return launch_ncc<4>(ctx, a, ntx, nbands);
        case 6: return launch_ncc<6>(ctx, a, ntx, nbands);
        case 8: return launch_ncc<8>(ctx, a, ntx, nbands);
        case 10: return launch_ncc<10>(ctx, a, ntx, nbands);
        case 12: return launch_ncc<12>(ctx, a, ntx, nbands);
        case 14: return launch_ncc<14>(ctx, a, ntx, nbands);
        case 16: return launch_ncc<16>(ctx, a, ntx, nbands);
        default: return 0;
    }
}
