// search_ext.cu -- extension searches (no reference counterpart; definitions of SURVEY.md 8c):
// SAD / SSD (integer compare) / NCC (fp32) costs with the reference's window geometry, candidate
// order, break rule and first-minimum rule; optional sub-pixel parabola fit; optional left-right
// consistency check.  Output convention of the extensions: black / no-candidate pixels -> 0.
//
// SSD / SAD right-referenced searches reuse the fused tile kernel for the integer winner d*; the
// two neighbouring costs of the parabola are then recomputed directly (O(window) per pixel).
// NCC and the left-referenced (mirrored) search of the LR check run in the per-pixel direct kernel.
#include <cmath>
#include "sr_internal.cuh"

struct WStats { long long sl, sr, sll, srr, slr, ssd, sad; };

// stats over a ww x wh window; "l" = left-image sample, "q" = right-image sample
__device__ __forceinline__ WStats window_stats(const uint32_t* __restrict__ Lp, int pitchL, int lx0,
                                               const uint32_t* __restrict__ Rp, int pitchR, int rx0, int y0, int ww, int wh,
                                               bool want_moments) {
    WStats s = {0, 0, 0, 0, 0, 0, 0};
    for (int r = 0; r < wh; r++) {
        const uint32_t* lrow = Lp + (size_t)(y0 + r) * pitchL + lx0;
        const uint32_t* rrow = Rp + (size_t)(y0 + r) * pitchR + rx0;
        uint32_t sl = 0, sr = 0, sll = 0, srr = 0, slr = 0, ssd = 0, sad = 0;  // a row of < 2^14 pixels fits 32 bits
        for (int i = 0; i < ww; i++) {
            uint32_t l = __ldg(lrow + i), q = __ldg(rrow + i);
            uint32_t ad = __vabsdiffu4(l, q);
            ssd = __dp4a(ad, ad, ssd);
            sad = __dp4a(ad, 0x01010101u, sad);
            if (want_moments) {
                sl = __dp4a(l, 0x01010101u, sl); sr = __dp4a(q, 0x01010101u, sr);
                sll = __dp4a(l, l, sll); srr = __dp4a(q, q, srr); slr = __dp4a(l, q, slr);
            }
        }
        s.sl += sl; s.sr += sr; s.sll += sll; s.srr += srr; s.slr += slr; s.ssd += ssd; s.sad += sad;
    }
    return s;
}

// NCC cost (extension definition of SURVEY.md 8c; fp32, fixed operation order): exact int64 moments,
// per-window factors rl = 1/sqrt((float)vl), rr = 1/sqrt((float)vr), cost = -(((float)num * rl) * rr).
__device__ __forceinline__ float ncc_inv_std(long long v) {
    return __fdiv_rn(1.0f, __fsqrt_rn(__ll2float_rn(v)));
}
__device__ __forceinline__ float ncc_cost(const WStats& s, long long n) {
    long long num = n * s.slr - s.sl * s.sr, vl = n * s.sll - s.sl * s.sl, vr = n * s.srr - s.sr * s.sr;
    if (vl <= 0 || vr <= 0) return INFINITY;
    return -__fmul_rn(__fmul_rn(__ll2float_rn(num), ncc_inv_std(vl)), ncc_inv_std(vr));
}

__device__ __forceinline__ float parabola(float dsub, float cm, float c0, float cp) {
    float den = __fadd_rn(__fsub_rn(cm, __fmul_rn(2.0f, c0)), cp);
    if (den > 0.0f) dsub = __fadd_rn(dsub, __fdiv_rn(__fsub_rn(cm, cp), __fmul_rn(2.0f, den)));
    return dsub;
}

struct ExtGeom {
    int wr, hr, wo, ho;   // reference image (gets the disparities) and the other image
    int height, half, dlo, dhi, left_ref, cost;
    int row0;             // first output row of the launch (grid.y rows from here)
    int ix0, ix1, iy0, iy1;   // pixels inside [ix0,ix1) x [iy0,iy1) are someone else's (NCC interior): skipped
    int row1;                 // end row of the launch when the border frame is enumerated compactly
    int no_black;             // 1: the black-pixel rule is applied by a later kernel (mirror-packed searches)
};

// Universal extension kernel: one thread per reference pixel.
__global__ void __launch_bounds__(128) ext_direct_kernel(const uint32_t* __restrict__ Lp, int pitchL,
                                                         const uint32_t* __restrict__ Rp, int pitchR, ExtGeom g,
                                                         int* __restrict__ disp, float* __restrict__ subpix) {
    int x, y;
    if (g.ix1 > g.ix0) {
        // border frame around the interior rectangle, enumerated compactly (1-D grid) so that every lane works:
        // full rows [row0, iy0), the two side strips of rows [iy0, iy1), full rows [iy1, row1)
        long idx = (long)blockIdx.x * blockDim.x + threadIdx.x;
        const int W = g.wr, side = g.ix0 + (W - g.ix1);
        const long top = (long)(g.iy0 - g.row0) * W, mid = (long)(g.iy1 - g.iy0) * side;
        if (idx < top) { y = g.row0 + (int)(idx / W); x = (int)(idx % W); }
        else if (idx < top + mid) {
            long k = idx - top;
            int c = (int)(k % side);
            y = g.iy0 + (int)(k / side); x = c < g.ix0 ? c : g.ix1 + (c - g.ix0);
        } else {
            long k = idx - top - mid;
            y = g.iy1 + (int)(k / W); x = (int)(k % W);
            if (y >= g.row1) return;
        }
    } else {
        x = blockIdx.x * blockDim.x + threadIdx.x; y = g.row0 + blockIdx.y;
        if (x >= g.wr) return;
    }
    size_t o = (size_t)y * g.wr + x;
    int dbest = 0;
    float dsub = 0.0f;
    const uint32_t* refp = g.left_ref ? Lp : Rp;
    int refpitch = g.left_ref ? pitchL : pitchR;
    if (y < g.height && (__ldg(refp + (size_t)y * refpitch + x) & 0xFFFFFFu) != 0) {
        int l = min(x, g.half), r = min(g.wr - x - 1, g.half), u = min(y, g.half), dn = min(g.hr - y - 1, g.half);
        int ww = l + r, wh = u + dn;
        long long n = (long long)ww * wh * 3;
        int sgn = g.left_ref ? -1 : 1;
        bool ncc = g.cost == SR_COST_NCC;
        // running first minimum with its two neighbours (costs compared as double like the oracle)
        double best = INFINITY, prev = INFINITY, cm = INFINITY, cp = INFINITY, c0 = INFINITY;
        int bi = -1, k = 0;
        bool want_next = false;
        for (int d = g.dlo; d < g.dhi; d++, k++) {
            int cx = x + sgn * d;
            if (g.left_ref ? (cx - l < 0) : (cx + r >= g.wo)) break;
            double c;
            if (g.left_ref && cx + r > g.wo) c = INFINITY;       // partner window beyond the right image
            else {
                WStats s = g.left_ref ? window_stats(Lp, pitchL, x - l, Rp, pitchR, cx - l, y - u, ww, wh, ncc)
                                      : window_stats(Lp, pitchL, cx - l, Rp, pitchR, x - l, y - u, ww, wh, ncc);
                if (ncc) c = n ? (double)ncc_cost(s, n) : (double)INFINITY;
                else c = (double)(g.cost == SR_COST_SAD ? s.sad : s.ssd);
            }
            if (want_next) { cp = c; want_next = false; }
            if (c < best) { best = c; bi = k; cm = prev; c0 = c; cp = INFINITY; want_next = true; }
            prev = c;
        }
        if (bi >= 0) {
            dbest = g.dlo + bi;
            dsub = (float)dbest;
            // both neighbours evaluated and finite (bi >= 1 and bi+1 < nv in the oracle)
            if (bi >= 1 && !want_next && isfinite(cm) && isfinite(cp)) dsub = parabola(dsub, (float)cm, (float)c0, (float)cp);
        }
    }
    disp[o] = dbest;
    if (subpix) subpix[o] = dsub;
}

// NCC on the border frame around the interior rectangle (clamped windows): one WARP per pixel, lanes
// stride over the candidates, (cost, d) first-minimum reduction by shuffles.  Same arithmetic and
// candidate rules as ext_direct_kernel (right-referenced), which spends one thread per pixel and is
// latency-bound on a thin frame.
__global__ void __launch_bounds__(128) ncc_border_warp_kernel(const uint32_t* __restrict__ Lp, int pitchL,
                                                              const uint32_t* __restrict__ Rp, int pitchR, ExtGeom g,
                                                              long npix, int* __restrict__ disp, float* __restrict__ subpix) {
    __shared__ uint32_t rwin[4][256];                        // one right window per warp
    const int lane = threadIdx.x & 31;
    long idx = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (idx >= npix) return;
    int x, y;
    {
        const int W = g.wr, side = g.ix0 + (W - g.ix1);
        const long top = (long)(g.iy0 - g.row0) * W, mid = (long)(g.iy1 - g.iy0) * side;
        if (idx < top) { y = g.row0 + (int)(idx / W); x = (int)(idx % W); }
        else if (idx < top + mid) {
            long k = idx - top;
            int c = (int)(k % side);
            y = g.iy0 + (int)(k / side); x = c < g.ix0 ? c : g.ix1 + (c - g.ix0);
        } else {
            long k = idx - top - mid;
            y = g.iy1 + (int)(k / W); x = (int)(k % W);
        }
    }
    size_t o = (size_t)y * g.wr + x;
    int dbest = 0;
    float dsub = 0.0f;
    if (y < g.height && (g.no_black || (__ldg(Rp + (size_t)y * pitchR + x) & 0xFFFFFFu) != 0)) {
        int l = min(x, g.half), r = min(g.wr - x - 1, g.half), u = min(y, g.half), dn = min(g.hr - y - 1, g.half);
        int ww = l + r, wh = u + dn;
        long long n = (long long)ww * wh * 3;
        int dmax = min(g.dhi, g.wo - x - r);                 // candidates d in [dlo, dmax): before the break
        // the right window is the same for every candidate: staged once per warp, its moments summed once
        uint32_t* rw = rwin[threadIdx.x >> 5];
        const int npw = ww * wh;                              // <= 16 x 16 (half <= 8 on this path)
        uint32_t s_r = 0, s_rr = 0;
        for (int i = lane; i < npw; i += 32) {
            uint32_t q = __ldg(Rp + (size_t)(y - u + i / ww) * pitchR + (x - l + i % ww));
            rw[i] = q;
            s_r = __dp4a(q, 0x01010101u, s_r); s_rr = __dp4a(q, q, s_rr);
        }
        s_r = __reduce_add_sync(0xffffffffu, s_r); s_rr = __reduce_add_sync(0xffffffffu, s_rr);
        __syncwarp();
        auto cost_of = [&](int d) -> float {
            if (!n) return INFINITY;
            uint32_t sl = 0, sll = 0, slr = 0;                // window totals fit 32 bits (<= 256 pixels)
            for (int rr_ = 0; rr_ < wh; rr_++) {
                const uint32_t* lrow = Lp + (size_t)(y - u + rr_) * pitchL + (x + d - l);
                for (int i = 0; i < ww; i++) {
                    uint32_t lp = __ldg(lrow + i), q = rw[rr_ * ww + i];
                    sl = __dp4a(lp, 0x01010101u, sl); sll = __dp4a(lp, lp, sll); slr = __dp4a(lp, q, slr);
                }
            }
            WStats st = {(long long)sl, (long long)s_r, (long long)sll, (long long)s_rr, (long long)slr, 0, 0};
            return ncc_cost(st, n);
        };
        float best = INFINITY; int bd = -1;
        for (int d = g.dlo + lane; d < dmax; d += 32) {
            float c = cost_of(d);
            if (c < best) { best = c; bd = d; }
        }
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            float oc = __shfl_down_sync(0xffffffffu, best, off);
            int od = __shfl_down_sync(0xffffffffu, bd, off);
            if (oc < best || (oc == best && od >= 0 && od < bd)) { best = oc; bd = od; }
        }
        bd = __shfl_sync(0xffffffffu, bd, 0);
        best = __shfl_sync(0xffffffffu, best, 0);
        if (bd >= 0) {
            dbest = bd;
            dsub = (float)bd;
            if (subpix && bd - 1 >= g.dlo && bd + 1 < dmax) {
                // lanes 0 / 1 recompute the two neighbours of the winner
                float cn = INFINITY;
                if (lane < 2) cn = cost_of(bd - 1 + 2 * lane);
                float cm = __shfl_sync(0xffffffffu, cn, 0), cp = __shfl_sync(0xffffffffu, cn, 1);
                if (isfinite(cm) && isfinite(cp)) dsub = parabola(dsub, cm, best, cp);
            }
        }
    }
    if (lane == 0) {
        disp[o] = dbest;
        if (subpix) subpix[o] = dsub;
    }
}

// Winner from the tile kernel's key planes + direct recomputation of the parabola neighbours.
__global__ void __launch_bounds__(128) ext_from_planes_kernel(const uint32_t* __restrict__ Lp, int pitchL,
                                                              const uint32_t* __restrict__ Rp, int pitchR, ExtGeom g,
                                                              const uint32_t* __restrict__ planes, size_t stride, int npass,
                                                              int dbase, int shift, int want_subpix, int* __restrict__ disp,
                                                              float* __restrict__ subpix) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = g.row0 + blockIdx.y;
    if (x >= g.wr) return;
    size_t o = (size_t)y * g.wr + x;
    int dbest = 0;
    float dsub = 0.0f;
    if (y < g.height && (__ldg(Rp + (size_t)y * pitchR + x) & 0xFFFFFFu) != 0) {
        uint32_t bc = 0xFFFFFFFFu; int bd = -1;
        for (int p = 0; p < npass; p++) {
            uint32_t k = __ldg(planes + (size_t)p * stride + o);
            if (k != 0xFFFFFFFFu && (k >> shift) < bc) { bc = k >> shift; bd = dbase + (p << shift) + (int)(k & ((1u << shift) - 1u)); }
        }
        if (bd >= 0) {
            dbest = bd;
            dsub = (float)bd;
            if (want_subpix) {
                int l = min(x, g.half), r = min(g.wr - x - 1, g.half), u = min(y, g.half), dn = min(g.hr - y - 1, g.half);
                // candidates d*-1 and d*+1 must both exist: d*-1 >= dlo, d*+1 < dhi and before the break
                if (bd - 1 >= g.dlo && bd + 1 < g.dhi && x + bd + 1 + r < g.wo) {
                    WStats sm = window_stats(Lp, pitchL, x + bd - 1 - l, Rp, pitchR, x - l, y - u, l + r, u + dn, false);
                    WStats sp = window_stats(Lp, pitchL, x + bd + 1 - l, Rp, pitchR, x - l, y - u, l + r, u + dn, false);
                    long long cm = g.cost == SR_COST_SAD ? sm.sad : sm.ssd, cp = g.cost == SR_COST_SAD ? sp.sad : sp.ssd;
                    dsub = parabola(dsub, __ll2float_rn(cm), __uint2float_rn(bc), __ll2float_rn(cp));
                }
            }
        }
    }
    disp[o] = dbest;
    if (subpix) subpix[o] = dsub;
}

// The same for the integer costs when nothing else follows (no LR check): winner from the key plane(s), the two
// parabola neighbours recomputed in ONE walk over the window (the right-image pixel is loaded once for both, packed
// BGRx words, VABSDIFF4.ACC / IDP.4A), sub-pixel value written straight into the CV_64F map -- one kernel instead of
// ext_from_planes_kernel (generic seven-moment window statistics, twice) + ext_finish_kernel.
// Several pairs of one shape in one launch (blockIdx.z = pair): batches of small pairs.
struct SubpixPairs { const uint32_t* L[16]; const uint32_t* R[16]; const uint32_t* planes[16]; double* out[16]; };

// HALF > 0: instantiated for that window; interior pixels (full 2*HALF x 2*HALF window) take a fully unrolled walk --
// all loads of a window row are independent and in flight together (the generic loops below are latency-bound:
// 27 -> 15 us on a 1280x720 pair with a 7x7 block).
template <int COST, int HALF>
__global__ void __launch_bounds__(128) ext_subpix_out_kernel(SubpixPairs pp, int pitchL, int pitchR, ExtGeom g,
                                                             size_t stride, int npass, int dbase, int shift) {
    const uint32_t* __restrict__ Lp = pp.L[blockIdx.z];
    const uint32_t* __restrict__ Rp = pp.R[blockIdx.z];
    const uint32_t* __restrict__ planes = pp.planes[blockIdx.z];
    double* __restrict__ out = pp.out[blockIdx.z];
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = g.row0 + blockIdx.y;
    if (x >= g.wr) return;
    size_t o = (size_t)y * g.wr + x;
    float dsub = 0.0f;
    if (y < g.height && (__ldg(Rp + (size_t)y * pitchR + x) & 0xFFFFFFu) != 0) {
        uint32_t bc = 0xFFFFFFFFu; int bd = -1;
        for (int p = 0; p < npass; p++) {
            uint32_t k = __ldg(planes + (size_t)p * stride + o);
            if (k != 0xFFFFFFFFu && (k >> shift) < bc) { bc = k >> shift; bd = dbase + (p << shift) + (int)(k & ((1u << shift) - 1u)); }
        }
        if (bd >= 0) {
            dsub = (float)bd;
            const int l = min(x, g.half), r = min(g.wr - x - 1, g.half), u = min(y, g.half), dn = min(g.hr - y - 1, g.half);
            if (bd - 1 >= g.dlo && bd + 1 < g.dhi && x + bd + 1 + r < g.wo) {
                uint32_t cm = 0, cp = 0;   // a window of < 2^14 pixels fits 32 bits for both costs
                if (HALF > 0 && l == HALF && r == HALF && u == HALF && dn == HALF) {
#pragma unroll
                    for (int rr = 0; rr < 2 * HALF; rr++) {
                        const uint32_t* rrow = Rp + (size_t)(y - HALF + rr) * pitchR + (x - HALF);
                        const uint32_t* lrow = Lp + (size_t)(y - HALF + rr) * pitchL + (x - HALF + bd);
                        uint32_t lv[2 * HALF + 2], q[HALF > 0 ? 2 * HALF : 1];
#pragma unroll
                        for (int i = 0; i < 2 * HALF + 2; i++) lv[i] = __ldg(lrow + i - 1);
#pragma unroll
                        for (int i = 0; i < 2 * HALF; i++) q[i] = __ldg(rrow + i);
#pragma unroll
                        for (int i = 0; i < 2 * HALF; i++) {
                            if (COST == SR_COST_SAD) {
                                asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(cm) : "r"(lv[i]), "r"(q[i]));
                                asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(cp) : "r"(lv[i + 2]), "r"(q[i]));
                            } else {
                                const uint32_t da = __vabsdiffu4(lv[i], q[i]), db = __vabsdiffu4(lv[i + 2], q[i]);
                                cm = __dp4a(da, da, cm); cp = __dp4a(db, db, cp);
                            }
                        }
                    }
                } else {
                    for (int rr = y - u; rr < y + dn; rr++) {
                        const uint32_t* rrow = Rp + (size_t)rr * pitchR + (x - l);
                        const uint32_t* lrow = Lp + (size_t)rr * pitchL + (x - l + bd);
                        // the two windows are two columns apart: one left-image load per column feeds both (a three-word shift)
                        uint32_t a = __ldg(lrow - 1), mid = __ldg(lrow), b;
                        for (int i = 0; i < l + r; i++, a = mid, mid = b) {
                            const uint32_t q = __ldg(rrow + i);
                            b = __ldg(lrow + i + 1);
                            if (COST == SR_COST_SAD) {
                                asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(cm) : "r"(a), "r"(q));
                                asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(cp) : "r"(b), "r"(q));
                            } else {
                                const uint32_t da = __vabsdiffu4(a, q), db = __vabsdiffu4(b, q);
                                cm = __dp4a(da, da, cm); cp = __dp4a(db, db, cp);
                            }
                        }
                    }
                }
                dsub = parabola(dsub, __uint2float_rn(cm), __uint2float_rn(bc), __uint2float_rn(cp));
            }
        }
    }
    out[o] = (double)dsub;
}

// NCC interior pixels: first maximum of the per-pass (score, d) planes of search_ncc.cu (passes ascend in d,
// strict > keeps the smaller d), plus the parabola through the directly recomputed neighbour costs.
__global__ void __launch_bounds__(128) ncc_merge_kernel(const uint32_t* __restrict__ Lp, int pitchL,
                                                        const uint32_t* __restrict__ Rp, int pitchR, ExtGeom g,
                                                        const float* __restrict__ pscore, const int* __restrict__ pdisp,
                                                        size_t stride, int npass, int want_subpix, int* __restrict__ disp,
                                                        float* __restrict__ subpix) {
    int x = g.ix0 + blockIdx.x * blockDim.x + threadIdx.x, y = g.iy0 + blockIdx.y;
    if (x >= g.ix1) return;
    size_t o = (size_t)y * g.wr + x;
    int dbest = 0;
    float dsub = 0.0f;
    if (g.no_black || (__ldg(Rp + (size_t)y * pitchR + x) & 0xFFFFFFu) != 0) {
        float best = -INFINITY; int bd = -1;
        for (int p = 0; p < npass; p++) {
            float s = __ldg(pscore + (size_t)p * stride + o);
            if (s > best) { best = s; bd = __ldg(pdisp + (size_t)p * stride + o); }
        }
        if (bd >= 0) {
            dbest = bd;
            dsub = (float)bd;
            const int h = g.half;
            if (want_subpix && bd - 1 >= g.dlo && bd + 1 < g.dhi && x + bd + 1 + h < g.wo) {
                long long n = (long long)(2 * h) * (2 * h) * 3;
                float cm = ncc_cost(window_stats(Lp, pitchL, x + bd - 1 - h, Rp, pitchR, x - h, y - h, 2 * h, 2 * h, true), n);
                float cp = ncc_cost(window_stats(Lp, pitchL, x + bd + 1 - h, Rp, pitchR, x - h, y - h, 2 * h, 2 * h, true), n);
                if (isfinite(cm) && isfinite(cp)) dsub = parabola(dsub, cm, -best, cp);
            }
        }
    }
    disp[o] = dbest;
    if (subpix) subpix[o] = dsub;
}

// Left-referenced winners from the key planes of the tile search on the mirror-packed pair:
// output column xc of that search is left-image column w1-1-xc (see sr_search_run_ext).
__global__ void __launch_bounds__(128) left_from_mirror_planes_kernel(const uint32_t* __restrict__ Lp, int pitchL, int w1, int h1,
                                                                      int height, const uint32_t* __restrict__ planes, size_t stride,
                                                                      int npass, int dbase, int shift, int* __restrict__ disp) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w1) return;
    int dbest = 0;
    if (y < height && (__ldg(Lp + (size_t)y * pitchL + x) & 0xFFFFFFu) != 0) {
        size_t o = (size_t)y * w1 + (w1 - 1 - x);
        uint32_t bc = 0xFFFFFFFFu;
        for (int p = 0; p < npass; p++) {
            uint32_t k = __ldg(planes + (size_t)p * stride + o);
            if (k != 0xFFFFFFFFu && (k >> shift) < bc) { bc = k >> shift; dbest = dbase + (p << shift) + (int)(k & ((1u << shift) - 1u)); }
        }
    }
    disp[(size_t)y * w1 + x] = dbest;
}

// mirrored-coordinate map (column w1-1-x) of a search on the mirror-packed pair -> left map, black rule on L
__global__ void __launch_bounds__(128) unflip_left_kernel(const uint32_t* __restrict__ Lp, int pitchL, int w1, int h1, int height,
                                                          const int* __restrict__ disp_m, int* __restrict__ disp) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= w1) return;
    bool on = y < height && (__ldg(Lp + (size_t)y * pitchL + x) & 0xFFFFFFu) != 0;
    disp[(size_t)y * w1 + x] = on ? disp_m[(size_t)y * w1 + (w1 - 1 - x)] : 0;
}

// LR check + conversion to the CV_64F output map.
__global__ void ext_finish_kernel(const int* __restrict__ disp_r, const float* __restrict__ subpix, int w2, int h2,
                                  const int* __restrict__ disp_l, int w1, int h1, int row0, double* __restrict__ out) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, y = row0 + blockIdx.y;
    if (x >= w2) return;
    size_t o = (size_t)y * w2 + x;
    int d = disp_r[o];
    bool ok = true;
    if (disp_l) {
        int u = x + d;
        int dl = (y < h1 && u >= 0 && u < w1) ? disp_l[(size_t)y * w1 + u] : 0x40000000;
        ok = y < h1 && u >= 0 && u < w1 && abs(dl - d) <= 1;
    }
    out[o] = ok ? (subpix ? (double)subpix[o] : (double)d) : 0.0;
}

// winner + parabola straight into pair->out from the key planes the tile kernel left in pair->planes (SSD / SAD, no LR check)
bool sr_ext_subpix_out_applies(const sr_search_params* p) {
    return p->subpixel && !p->lr_check && (p->cost == SR_COST_SAD || p->cost == SR_COST_SSD_L2) &&
           3ull * p->block * p->block * 65025ull < 0xFFFFFFFFull;
}
template <int COST>
static void launch_subpix(sr_ctx* ctx, dim3 grid, const SubpixPairs& pp, sr_pair* pair, const ExtGeom& g) {
#define SR_SUBPIX(H) ext_subpix_out_kernel<COST, H><<<grid, 128, 0, ctx->stream>>>(pp, pair->L.pitch, pair->R.pitch, g, pair->npx_cap, \
                                                                                    pair->npass, pair->plane_dbase, pair->key_shift)
    switch (g.half) {                                    // the common small windows get the unrolled interior walk
        case 1: SR_SUBPIX(1); break;
        case 2: SR_SUBPIX(2); break;
        case 3: SR_SUBPIX(3); break;
        case 4: SR_SUBPIX(4); break;
        default: SR_SUBPIX(0); break;
    }
#undef SR_SUBPIX
}

// n pairs of one shape and one search (same planes layout): one launch
int sr_launch_ext_subpix_out_multi(sr_ctx* ctx, sr_pair* const* pairs, int n, const sr_search_params* p, int row0, int row1) {
    if (n < 1 || n > 16) return sr_fail(ctx, SR_E_BADARG, "sr_launch_ext_subpix_out_multi: 1..16 pairs");
    sr_pair* pair = pairs[0];
    const int w1 = pair->w1, h1 = pair->h1, w2 = pair->w2, h2 = pair->h2, half = (p->block - 1) / 2;
    int dhi = p->max_disp, dlo = p->min_disp;
    if (dhi < dlo) dhi = dlo;
    ExtGeom g{w2, h2, w1, h1, h1 < h2 ? h1 : h2, half, dlo, dhi, 0, p->cost, row0};
    SubpixPairs pp;
    for (int q = 0; q < n; q++) {
        sr_pair* pq = pairs[q];
        if (pq->L.pitch != pair->L.pitch || pq->R.pitch != pair->R.pitch || pq->npass != pair->npass || pq->key_shift != pair->key_shift ||
            pq->plane_dbase != pair->plane_dbase || pq->npx_cap != pair->npx_cap)
            return sr_fail(ctx, SR_E_BADARG, "sr_launch_ext_subpix_out_multi: pairs of different layout");
        pp.L[q] = pq->L.px; pp.R[q] = pq->R.px; pp.planes[q] = pq->planes; pp.out[q] = pq->out;
    }
    dim3 grid(cdiv(w2, 128), row1 - row0, n);
    if (p->cost == SR_COST_SAD) launch_subpix<SR_COST_SAD>(ctx, grid, pp, pair, g);
    else launch_subpix<SR_COST_SSD_L2>(ctx, grid, pp, pair, g);
    SR_LAUNCHED(ctx);
    return SR_OK;
}
int sr_launch_ext_subpix_out(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p, int row0, int row1) {
    return sr_launch_ext_subpix_out_multi(ctx, &pair, 1, p, row0, row1);
}

static int rows_fit_ext(int ho, int hr, int half) {
    int height = ho < hr ? ho : hr;
    for (int y = (height - half - 2 > 0 ? height - half - 2 : 0); y < height; y++) {
        int down = hr - y - 1 < half ? hr - y - 1 : half;
        if (y + down > ho) return 0;
    }
    return 1;
}

int sr_search_run_ext(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p, int row0, int row1) {
    int w1 = pair->w1, h1 = pair->h1, w2 = pair->w2, h2 = pair->h2;
    int half = (p->block - 1) / 2;
    if (p->var_block) return sr_fail(ctx, SR_E_BADARG, "var_block is a reference-only option (SSD_L2 cost, no extensions)");
    if (p->smooth != 1.0) return sr_fail(ctx, SR_E_BADARG, "the smoothing rule is defined for the reference cost only; extensions need smooth == 1");
    if (!rows_fit_ext(h1, h2, half) || (p->lr_check && !rows_fit_ext(h2, h1, half)))
        return sr_fail(ctx, SR_E_SHAPE, "image heights do not fit the window rows (h1=%d, h2=%d)", h1, h2);
    row0 = row0 < 0 ? 0 : row0;
    row1 = row1 > h2 ? h2 : row1;
    if (p->lr_check && (row0 != 0 || row1 != h2)) return sr_fail(ctx, SR_E_BADARG, "lr_check needs the full image (no row bands)");
    pair->ow = w2; pair->oh = h2; pair->last_row0 = row0; pair->last_row1 = row1;
    if (row1 <= row0) return SR_OK;
    size_t npx = pair->npx_cap;
    int* disp_r = (int*)sr_scratch(ctx, 14, npx * 4 * 2);
    float* subpix = p->subpixel ? (float*)sr_scratch(ctx, 15, npx * 4) : nullptr;
    if (!disp_r || (p->subpixel && !subpix)) return SR_E_CUDA;
    // the left-referenced map outlives this call (sr_search_left_disparity): it lives in a buffer of the pair, not in
    // the context's shared scratch, which the next search on ANY pair of the context may overwrite or regrow
    if (p->lr_check && !pair->disp_left_own) SR_CUDA(ctx, cudaMalloc((void**)&pair->disp_left_own, npx * sizeof(int)));
    int* disp_l = p->lr_check ? pair->disp_left_own : nullptr;
    int dhi = p->max_disp, dlo = p->min_disp;
    if (dhi < dlo) dhi = dlo;
    ExtGeom g{w2, h2, w1, h1, h1 < h2 ? h1 : h2, half, dlo, dhi, 0, p->cost, row0};
    bool tiled = false;
    if (p->cost != SR_COST_NCC && half >= 1 && half <= 8 && w2 > 1 && h2 > 1) {
        SearchGeom sg;
        sg.w1 = w1; sg.h1 = h1; sg.w2 = w2; sg.h2 = h2; sg.height = g.height; sg.block = p->block; sg.half = half;
        sg.dlo = dlo; sg.dhi = dhi > w1 ? w1 : dhi; if (sg.dhi < sg.dlo) sg.dhi = sg.dlo;
        sg.row0 = row0; sg.row1 = row1;
        int handled = sr_launch_search_tile(ctx, pair, sg, p->cost, 0);
        if (handled < 0) return handled;
        if (handled && sr_ext_subpix_out_applies(p)) {
            ctx->last_kernel = "tile+ext";
            return sr_launch_ext_subpix_out(ctx, pair, p, row0, row1);
        }
        if (handled) {
            tiled = true;
            ext_from_planes_kernel<<<dim3(cdiv(w2, 128), row1 - row0), 128, 0, ctx->stream>>>(
                pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, g, pair->planes, pair->npx_cap,
                pair->npass, pair->plane_dbase, pair->key_shift, p->subpixel, disp_r, subpix);
            SR_LAUNCHED(ctx);
        }
    }
    bool ncc_tiled = false;
    if (p->cost == SR_COST_NCC && half >= 1 && half <= 8) {
        // interior pixels (full window) through the sliding-sum NCC kernel, the frame of clamped windows directly
        SearchGeom sg;
        sg.w1 = w1; sg.h1 = h1; sg.w2 = w2; sg.h2 = h2; sg.height = g.height; sg.block = p->block; sg.half = half;
        sg.dlo = dlo; sg.dhi = dhi > w1 ? w1 : dhi; if (sg.dhi < sg.dlo) sg.dhi = sg.dlo;
        sg.row0 = row0; sg.row1 = row1;
        float* ps = nullptr; int* pd = nullptr;
        int handled = sg.dhi > sg.dlo ? sr_launch_search_ncc(ctx, pair, sg, &ps, &pd, &pair->ncc_tab) : 0;
        if (handled < 0) return handled;
        if (handled) {
            ncc_tiled = true;
            g.ix0 = half; g.ix1 = w2 - half;
            g.iy0 = row0 > half ? row0 : half;
            g.iy1 = row1 < g.height ? row1 : g.height;
            if (g.iy1 > h2 - half) g.iy1 = h2 - half;
            if (g.iy1 > h1 - half) g.iy1 = h1 - half;
            if (g.iy1 < g.iy0) g.iy1 = g.iy0;
            if (g.iy1 > g.iy0 && g.ix1 > g.ix0) {
                ncc_merge_kernel<<<dim3(cdiv(g.ix1 - g.ix0, 128), g.iy1 - g.iy0), 128, 0, ctx->stream>>>(
                    pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, g, ps, pd, pair->npx_cap, pair->npass, p->subpixel,
                    disp_r, subpix);
                SR_LAUNCHED(ctx);
            } else { g.ix1 = g.ix0; }
        }
    }
    if (!tiled) {   // all pixels, or -- after the NCC kernel -- the border frame outside [ix0,ix1) x [iy0,iy1)
        dim3 grid(cdiv(w2, 128), row1 - row0);
        if (g.ix1 > g.ix0) {
            g.row1 = row1;
            long nborder = (long)(row1 - row0) * w2 - (long)(g.iy1 - g.iy0) * (g.ix1 - g.ix0);
            grid = dim3((unsigned)((nborder + 127) / 128), 1);
        }
        if (g.ix1 > g.ix0 && p->cost == SR_COST_NCC) {
            long nborder = (long)(row1 - row0) * w2 - (long)(g.iy1 - g.iy0) * (g.ix1 - g.ix0);
            if (nborder > 0) {
                ncc_border_warp_kernel<<<(unsigned)((nborder * 32 + 127) / 128), 128, 0, ctx->stream>>>(
                    pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, g, nborder, disp_r, subpix);
                SR_LAUNCHED(ctx);
            }
        } else if (grid.x > 0) {
            ext_direct_kernel<<<grid, 128, 0, ctx->stream>>>(pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, g, disp_r, subpix);
            SR_LAUNCHED(ctx);
        }
    }
    ctx->last_kernel = tiled ? "tile+ext" : ncc_tiled ? "ncc-tile+ext" : "ext-direct";
    if (p->lr_check) {
        // Left-referenced search (window [x-l, x+r) on the left image, partner at x - d, break when the
        // partner window leaves the right image on the left).  With both images mirror-packed about the
        // left width (stored[s] = image[w1-2-s]) and the roles swapped, the right-referenced tile kernel
        // computes exactly this search: its window [xc-h, xc+h) covers left columns (x-h-1, x+h-1] ->
        // [x-h, x+h) for x = w1-1-xc, its column range [0, w1-1) is the left image's, and its break rule
        // d < w1 - xc - min(w1-1-xc, h) is d <= x - min(x, h).  Needs w1 <= w2 (no partner beyond R).
        bool mirrored = false;
        if (half >= 1 && half <= 8 && w1 > 1 && h1 > 1 && w1 <= w2) {
            int rc;
            if (!pair->Lm.base) {
                if ((rc = sr_alloc_packed(ctx, &pair->Lm, w1, h1, w1 + 704)) != SR_OK) return rc;
                if ((rc = sr_alloc_packed(ctx, &pair->Rm, w1, h2, w1 + 704)) != SR_OK) return rc;
            }
            if ((rc = sr_launch_pack_mirror(ctx, pair->rawL, (size_t)w1 * 3, w1, h1, w1, &pair->Lm)) != SR_OK) return rc;
            if ((rc = sr_launch_pack_mirror(ctx, pair->rawR, (size_t)w2 * 3, w2, h2, w1, &pair->Rm)) != SR_OK) return rc;
            SearchGeom sg;
            sg.w1 = w1; sg.h1 = h2; sg.w2 = w1; sg.h2 = h1; sg.height = g.height; sg.block = p->block; sg.half = half;
            sg.dlo = dlo; sg.dhi = dhi > w1 ? w1 : dhi; if (sg.dhi < sg.dlo) sg.dhi = sg.dlo;
            sg.row0 = 0; sg.row1 = h1;
            PackedImage keepL = pair->L, keepR = pair->R;
            pair->L = pair->Rm; pair->R = pair->Lm;   // tile roles: "left" = searched image, "right" = reference image
            int handled = 0;
            float* ps = nullptr; int* pd = nullptr;
            if (p->cost != SR_COST_NCC) handled = sr_launch_search_tile(ctx, pair, sg, p->cost, 0);
            else if (sg.dhi > sg.dlo) handled = sr_launch_search_ncc(ctx, pair, sg, &ps, &pd, &pair->ncc_tab_m);
            PackedImage mL = pair->L, mR = pair->R;
            pair->L = keepL; pair->R = keepR;
            if (handled < 0) return handled;
            if (handled && p->cost != SR_COST_NCC) {
                mirrored = true;
                left_from_mirror_planes_kernel<<<dim3(cdiv(w1, 128), h1), 128, 0, ctx->stream>>>(
                    pair->L.px, pair->L.pitch, w1, h1, g.height, pair->planes, pair->npx_cap, pair->npass, pair->plane_dbase, pair->key_shift, disp_l);
                SR_LAUNCHED(ctx);
                ctx->last_kernel = tiled ? "tile+ext+mirror" : "ext-direct+mirror";
            } else if (handled) {
                // NCC: interior from the (score, d) planes, border frame by the warp-per-pixel kernel, both in mirrored
                // coordinates without the black rule; the un-flip applies it on the left image
                mirrored = true;
                int* disp_m = (int*)sr_scratch(ctx, 13, pair->npx_cap * sizeof(int));
                if (!disp_m) return SR_E_CUDA;
                ExtGeom gm{w1, h1, w1, h2, g.height, half, sg.dlo, sg.dhi, 0, p->cost, 0};
                gm.no_black = 1; gm.row1 = h1;
                gm.ix0 = half; gm.ix1 = w1 - half; gm.iy0 = half;
                gm.iy1 = h1 < g.height ? h1 : g.height;
                if (gm.iy1 > h1 - half) gm.iy1 = h1 - half;
                if (gm.iy1 > h2 - half) gm.iy1 = h2 - half;
                if (gm.iy1 < gm.iy0) gm.iy1 = gm.iy0;
                if (gm.iy1 > gm.iy0 && gm.ix1 > gm.ix0) {
                    ncc_merge_kernel<<<dim3(cdiv(gm.ix1 - gm.ix0, 128), gm.iy1 - gm.iy0), 128, 0, ctx->stream>>>(
                        mL.px, mL.pitch, mR.px, mR.pitch, gm, ps, pd, pair->npx_cap, pair->npass, 0, disp_m, nullptr);
                    SR_LAUNCHED(ctx);
                } else { gm.ix1 = gm.ix0 + 1; gm.iy1 = gm.iy0; }   // no interior: every pixel is "border" (ix1 > ix0 keeps the compact form)
                long nborder = (long)h1 * w1 - (long)(gm.iy1 - gm.iy0) * (gm.ix1 - gm.ix0);
                if (nborder > 0) {
                    ncc_border_warp_kernel<<<(unsigned)((nborder * 32 + 127) / 128), 128, 0, ctx->stream>>>(
                        mL.px, mL.pitch, mR.px, mR.pitch, gm, nborder, disp_m, nullptr);
                    SR_LAUNCHED(ctx);
                }
                unflip_left_kernel<<<dim3(cdiv(w1, 128), h1), 128, 0, ctx->stream>>>(pair->L.px, pair->L.pitch, w1, h1, g.height, disp_m, disp_l);
                SR_LAUNCHED(ctx);
                ctx->last_kernel = ncc_tiled ? "ncc-tile+ext+mirror" : "ext-direct+mirror";
            }
        }
        if (!mirrored) {
            ExtGeom gl{w1, h1, w2, h2, h1 < h2 ? h1 : h2, half, dlo, dhi, 1, p->cost, 0};
            ext_direct_kernel<<<dim3(cdiv(w1, 128), h1), 128, 0, ctx->stream>>>(pair->L.px, pair->L.pitch, pair->R.px, pair->R.pitch, gl,
                                                                                 disp_l, nullptr);
            SR_LAUNCHED(ctx);
        }
    }
    pair->disp_left = disp_l;
    ext_finish_kernel<<<dim3(cdiv(w2, 128), row1 - row0), 128, 0, ctx->stream>>>(disp_r, subpix, w2, h2, disp_l, w1, h1, row0, pair->out);
    SR_LAUNCHED(ctx);
    return SR_OK;
}
