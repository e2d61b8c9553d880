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
#include <cstdlib>
#include <functional>
#include "sr_internal.cuh"
#include <stdio.h>
#include <string.h>

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
#define SR_TILE_OVERSHOOT_PAD 768     // right padding of the shifted right image of the Left variant: one tile of overshoot + halo

__device__ __forceinline__ int padpos(int c) { return c + (c >> 2); }

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async16_s(uint32_t saddr, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(saddr), "l"(gmem));
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

// col - (pointwise cost): the row leaving the window.  Both integer pipes issue 64 lanes/clk/SM (ALU: VABSDIFF4,
// IADD3, VIMNMX3; FMA: IDP.4A, IMAD), so the subtraction goes to the FMA pipe, which the rest of the loop leaves
// idle: SAD = VABSDIFF4 (ALU) + IDP.4A.U8.S8 with the byte multipliers -1 (FMA) instead of VABSDIFF4 + IADD
// (ALU + ALU); SSD = VABSDIFF4 + IDP.4A + IADD (an IMAD by -1 in its place -- FSUB, km1 = 0xFFFFFFFF as a kernel argument
// so that it is not strength-reduced -- measured 1 % faster on 4K / 9x9 and slower on small shapes, where the second
// copy of the subtract phase in the instruction stream costs more than it wins: not instantiated).
template <int COST, bool FSUB>
__device__ __forceinline__ uint32_t pcost_sub(uint32_t l, uint32_t r, uint32_t acc, uint32_t km1) {
    uint32_t ad = __vabsdiffu4(l, r);
    if (COST == SR_COST_SAD) {
        int o;
        asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(o) : "r"(ad), "r"(0xFFFFFFFFu), "r"((int)acc));
        return (uint32_t)o;
    }
    uint32_t sq = __dp4a(ad, ad, 0u);
    return FSUB ? sq * km1 + acc : acc - sq;
}

// vertical update of the 64 column sums with one staged image row (ADD: row enters the window).
// slot: shared address of the ring slot (warp-uniform); rb: this lane's byte offset inside the right
// segment; lb[i]: byte offsets (lane base + warp-uniform chunk) of the 5 left chunks.
template <int COST, bool ADD, bool FSUB>
__device__ __forceinline__ void vertical_update(uint32_t (&col)[XT][DT], uint32_t slot, uint32_t lane_b,
                                                const uint32_t (&uo)[5], uint32_t colmask, uint32_t km1) {
    // all nine LDS.128 of the row first (36 registers: the WTA's key / halo registers are dead here), then the
    // 64 pointwise updates: the load latency is paid once per phase instead of once per 16-column group
    uint4 lv[5], rv[4];
#pragma unroll
    for (int k = 0; k < 5; k++) lv[k] = lds128((slot + uo[k]) + lane_b);
#pragma unroll
    for (int k = 0; k < 4; k++) rv[k] = lds128(slot + lane_b + 16 * k);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        uint32_t l7[7] = {lv[k].x, lv[k].y, lv[k].z, lv[k].w, lv[k + 1].x, lv[k + 1].y, lv[k + 1].z};
        uint32_t r4[4] = {rv[k].x, rv[k].y, rv[k].z, rv[k].w};
#pragma unroll
        for (int jj = 0; jj < 4; jj++)
#pragma unroll
            for (int t = 0; t < DT; t++) {
                if (ADD) col[4 * k + jj][t] = pcost_add<COST>(l7[jj + t], r4[jj], col[4 * k + jj][t]);
                else col[4 * k + jj][t] = pcost_sub<COST, FSUB>(l7[jj + t], r4[jj], col[4 * k + jj][t], km1);
            }
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

constexpr int MAXP = 16;  // pairs of one launch (sr_launch_search_tile_multi)
struct PairPtrs { const uint32_t* L; const uint32_t* R; uint32_t* planes; };
struct TileArgs {
    int npairs;          // > 0: the work items span several pairs of one shape (item.x = tile | pair << 16), images and
    PairPtrs ppv[MAXP];  //      key planes from ppv[pair]; 0: the single pair below.  (Kept inside the parameter block: the same
                         //      table behind a device pointer compiled to a 3 % slower SAD kernel -- 1.779 vs 1.725 ms per 4K step.)
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
    const int4* work;    // work items {tile, first row, end row, pass}, one per CTA, in schedule order
    uint32_t k64, nk64;  // == 64, == -64: in-kernel key = S*64 + d_local (d_local = candidate index within the pass);
                         // kernel arguments so that the multiplications stay IMADs on the FMA pipe
    int merged;          // 1: one plane for all passes: the key reduction rewrites the winner as
    int shift;           //    cost << shift | (pass*64 + d_local) and folds it in with RED.MIN
    uint32_t km1;        // == 0xFFFFFFFF (-1): IMAD-based subtract
    int red_mode;        // 1: half of the warps merge the previous row's keys at the end of the iteration
    uint32_t zmask;      // bit w: warp w slides its window sums on the FMA pipe (window_wta ZMODE)
    // Left-variant geometry (kernel template LM, BlockSearch.cpp:24-86): full window [x-h, x-h+H2) x [y-h, y-h+H2) with
    // H2 = blockSize odd, every image column < cmax / row < rmax contributes (no clamped windows), candidate index
    // i = off - d ascending (= d descending: the smallest i of equal costs is the LARGEST d, :76) with per-column
    // bounds max(dlo, lo_a - x) <= i < min(dhi, lim_a - x).  w2 = width of the reference (output) image.
    int lmode, hwin;     // hwin: window size H2 (2*half, or 2*half + 1 in lmode)
    int pw;              // candidates per pass: 64 (16 warps, one CTA per SM) or 32 (8 warps, two CTAs per SM)
    uint32_t* c0;        // != NULL (needs dbase == 0): the raw window cost at d == 0 of every pixel goes to this u32 plane -- the
                         // smoothing rule treats d == 0 apart (BlockSearch.cpp:160-165) and the sums are warp 0's anyway
    int cmax, rmax, lim_a, lo_a;
};

// horizontal window sums + in-thread WTA for one output row.  Disparities are processed in pairs so
// that the two keys fold into one 3-input min (VIMNMX3).  Two instruction mixes, chosen per warp so that the two
// integer pipes (64 lanes/clk/SM each) fill evenly:
//   ZMODE = false: S += in - out as one IADD3 (ALU), key = S*64 + d_local as one IMAD (FMA);
//   ZMODE = true:  the window sum is kept as the key itself, K = 64*S + d_local, and slides with two IMADs
//                  (K += in*64; K += out*(-64); FMA pipe) -- no ALU instruction but the min.
// With the leaving-row subtract on the FMA pipe (pcost_sub) a SAD evaluation is 3.5 ALU + 2 FMA instructions
// without ZMODE and 2.5 + 3 with it; one warp in four without ZMODE balances the pipes at 2.75 each.
// MASKED: candidates outside [dlo, dlim(x)) are discarded (tiles that touch the break rule or the range ends); which
// ones does not depend on the row, so the kernel forms the 64 validity bits of a thread once per CTA (vm).
template <int H2, bool MASKED, bool ZMODE, bool LM = false>
__device__ __forceinline__ void window_wta(const uint32_t (&col)[XT][DT], uint32_t (&keys)[XT], uint32_t k64,
                                           uint32_t nk64, uint32_t dl0, const uint32_t (&vm)[2]) {
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
        if (ZMODE) { S0 = S0 * k64 + dla; S1 = S1 * k64 + dlb; }
#pragma unroll
        for (int j = 0; j < XT; j++) {
            uint32_t ka, kb;
            if (ZMODE) { ka = S0; kb = S1; }
            else { ka = S0 * k64 + dla; kb = S1 * k64 + dlb; }
            if (MASKED) {                                   // validity bits of (column j, disparity tp / tp + 1): see the kernel
                if (!((vm[tp >> 1] >> j) & 1u)) ka = 0xFFFFFFFFu;
                if (!((vm[tp >> 1] >> (16 + j)) & 1u)) kb = 0xFFFFFFFFu;
            }
            keys[j] = min(min(keys[j], ka), kb);
            if (j + 1 < XT) {
                if (ZMODE) {
                    S0 = w0[j + H2] * k64 + S0; S0 = w0[j] * nk64 + S0;
                    S1 = w1[j + H2] * k64 + S1; S1 = w1[j] * nk64 + S1;
                } else {
                    S0 += w0[j + H2] - w0[j];
                    S1 += w1[j + H2] - w1[j];
                }
            }
        }
    }
}

// Window sums of this warp's first disparity for the thread's 16 output columns -> one row of the d == 0 cost plane
// (warp 0 of pass 0 only: ~50 instructions per row for one warp in sixteen of one pass; replaces the separable box-sum
// kernels cost0_vertical / cost0_horizontal, 40 us on a 1436x992 pair).
template <int H2>
__device__ __forceinline__ void emit_c0(const uint32_t (&col)[XT][DT], uint32_t* __restrict__ c0row, int x0, int w2, int lane) {
    // no halo array: the neighbour lane's columns are fetched where the window reaches them (the kernel sits at its
    // register limit)
    auto colv = [&](int i) { return i < XT ? col[i < XT ? i : 0][0] : __shfl_down_sync(0xffffffffu, col[i >= XT ? i - XT : 0][0], 1); };
    uint32_t S = 0;
#pragma unroll
    for (int i = 0; i < H2; i++) S += colv(i);
#pragma unroll
    for (int j = 0; j < XT; j++) {
        const int x = x0 + j;
        if (lane < 31 && x >= 0 && x < w2) c0row[x] = S;
        if (j + 1 < XT) S += colv(j + H2) - col[j][0];
    }
}

// The SSD form of the unmasked window phase: all four disparities of the thread at once -- four independent sliding
// chains instead of two -- and no halo arrays: the neighbour lane's column is fetched where the window reaches it.
// Measured against window_wta (tools/ab_tile.py): SSD 1.5 - 2.5 % faster (4K / 512: 1.889 -> 1.854 ms, 1436x992 block 11:
// 0.169 -> 0.165), SAD unchanged within +-1 % -- so SAD keeps window_wta and its per-warp pipe mixes.
template <int H2, bool ZMODE>
__device__ __forceinline__ void window_wta4(const uint32_t (&col)[XT][DT], uint32_t (&keys)[XT], uint32_t k64, uint32_t nk64, uint32_t dl0) {
    uint32_t S[DT];
#pragma unroll
    for (int t = 0; t < DT; t++) {
        S[t] = 0;
#pragma unroll
        for (int i = 0; i < H2; i++) S[t] += i < XT ? col[i < XT ? i : 0][t] : __shfl_down_sync(0xffffffffu, col[i >= XT ? i - XT : 0][t], 1);
        if (ZMODE) S[t] = S[t] * k64 + (dl0 + t);
    }
#pragma unroll
    for (int j = 0; j < XT; j++) {
        uint32_t k[DT];
#pragma unroll
        for (int t = 0; t < DT; t++) k[t] = ZMODE ? S[t] : S[t] * k64 + (dl0 + t);
        keys[j] = min(min(keys[j], k[0]), k[1]);
        keys[j] = min(min(keys[j], k[2]), k[3]);
        if (j + 1 < XT) {
#pragma unroll
            for (int t = 0; t < DT; t++) {
                const int i = j + H2;
                const uint32_t in = i < XT ? col[i < XT ? i : 0][t] : __shfl_down_sync(0xffffffffu, col[i >= XT ? i - XT : 0][t], 1);
                if (ZMODE) { S[t] = in * k64 + S[t]; S[t] = col[j][t] * nk64 + S[t]; }
                else S[t] += in - col[j][t];
            }
        }
    }
}

__device__ __forceinline__ void sts128(uint32_t saddr, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(saddr), "r"(x), "r"(y), "r"(z), "r"(w));
}
__device__ __forceinline__ uint32_t lds32(uint32_t saddr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}

template <int H2, int COST, int NW, bool LM = false, bool C0 = false>
__global__ void __launch_bounds__(NW * 32, NW <= 8 ? 2 : 1) search_tile_kernel(TileArgs a) {
    constexpr int NSLOT = H2 + PF + 1;   // +1: half of the warps subtract the leaving row one iteration late
    constexpr uint32_t SLOT_B = SLOT_U4 * 16, RING_B = NSLOT * SLOT_B, KEYROW_B = KEY_U4 * 16, KEYBUF_B = NW * KEYROW_B;
    // NW = 16: one CTA per SM, 64 candidates per pass.  NW = 8: two CTAs per SM that are not synchronised with each other
    // (their phases drift apart and fill each other's barrier / single-pipe phases), 32 candidates per pass; every thread
    // then stages up to two chunks per row and owns two columns of the key reduction.
    constexpr int NST = (RCH + LCH + NW * 32 - 1) / (NW * 32);   // staged chunks per thread and row
    constexpr int NRG = 32 / (2 * NW) > 0 ? 32 / (2 * NW) : 1;   // 16-column groups per half-warp in the key reduction
    static_assert(NST * NW * 32 >= RCH + LCH && NRG * 2 * NW * XT >= TILE_OUT, "staging / key reduction cover the tile");
    extern __shared__ __align__(128) uint4 smem4[];   // ring: NSLOT x SLOT_U4, then keys: 2 x NW x KEY_U4

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // warp-uniform for the compiler
    const int h = a.half;
    const int e_in = LM ? H2 - 1 - h : h - 1;                  // row y + e_in enters the window of output row y (H2 == 2h unless LM)
    // One CTA per work item {tile, first row, end row, pass}.  The list is built on the host (tile_work_list):
    // the hardware hands out CTAs in list order, so the order of the list is the schedule.
    const int4 wk = __ldg(a.work + blockIdx.x);
    const int CB = -XT + TILE_OUT * (wk.x & 0xFFFF);           // first own column of lane 0
    const uint32_t* imgL = a.L; const uint32_t* imgR = a.R; uint32_t* planes = a.planes;
    if (a.npairs > 0) { const PairPtrs pp = a.ppv[wk.x >> 16]; imgL = pp.L; imgR = pp.R; planes = pp.planes; }
    const int yb0 = wk.y, yb1 = wk.z;
    if (yb0 >= yb1) return;
    const int cb = CB + XT * lane;                             // this thread's own columns [cb, cb+16)
    const uint32_t k64 = a.k64, nk64 = a.nk64, km1 = a.km1;
    const int pass = wk.w;
    const int dcta = a.dbase + pass * (DT * NW);               // first disparity of the pass (multiple of 4)
    const int d0 = dcta + DT * warp;                            // this warp's 4 disparities
    const uint32_t dkey0 = (uint32_t)(DT * warp);               // candidate index within the pass: the key's low 6 bits
    const bool zmode = (a.zmask >> warp) & 1u;
    const bool late_sub = (warp >> 2) & 1;
    const bool red_late = a.red_mode && ((warp >> 3) & 1);   // spreads the LDS burst of the key merge over the iteration

    // own columns that contribute (inside [0, w2-1)); others are forced to 0
    uint32_t colmask = 0;
#pragma unroll
    for (int j = 0; j < XT; j++) colmask |= (uint32_t)(cb + j >= 0 && cb + j < (LM ? a.cmax : a.w2 - 1)) << j;
    // per-output candidate limit: d < min(dhi, w1 - x - r(x))  (break rule, BlockSearch.cpp:148)
    auto lim_of = [&](int x) { return LM ? min(a.dhi, a.lim_a - x) : min(a.dhi, a.w1 - x - min(a.w2 - x - 1, h)); };
    int dlim_min = 0x7fffffff, dlo_max = a.dlo;
#pragma unroll
    for (int j = 0; j < XT; j++) {
        int x = cb + h + j;
        if (x >= 0 && x < a.w2) {
            dlim_min = min(dlim_min, lim_of(x));
            if (LM) dlo_max = max(dlo_max, a.lo_a - x);
        }
    }
    const bool needmask = __any_sync(0xffffffffu, d0 < dlo_max || d0 + DT - 1 >= dlim_min);
    // bit (t & 1) * 16 + j of vm[t >> 1]: candidate d0 + t is valid for output column cb + h + j (break rule :148, range
    // ends; the Left variant's per-column lower bound) -- row-independent, formed once
    uint32_t vm[2] = {0u, 0u};
    if (needmask) {
#pragma unroll
        for (int j = 0; j < XT; j++) {
            const int x = cb + h + j;
            const int lim = lim_of(x), lo = LM ? max(a.dlo, a.lo_a - x) : a.dlo;
#pragma unroll
            for (int t = 0; t < DT; t++)
                if (d0 + t >= lo && d0 + t < lim) vm[t >> 1] |= 1u << ((t & 1) * 16 + j);
        }
    }
    // The break rule cuts a triangle off the (x, d) plane at the right image edge: the largest limit of
    // the tile (its first output column) decides which warps / whole CTAs have no candidate at all.
    const int x_first = max(CB + h, 0);                         // first output column of the tile inside the image (the limit never grows with x)
    const int dlim_tile = lim_of(x_first);
    const bool active = d0 < dlim_tile;                         // warps with at least one candidate
    if (dcta >= dlim_tile) {
        // nothing to search in this (tile, pass): the planes still get their "no candidate" keys
        // (the merged plane is preset to them)
        if (a.merged) return;
        for (int y = yb0 + warp; y < yb1; y += NW)
            for (int q = lane; q < TILE_OUT; q += 32) {
                int x = CB + h + q;
                if (x >= 0 && x < a.w2) planes[(size_t)pass * a.plane_stride + (size_t)y * a.w2 + x] = 0xFFFFFFFFu;
            }
        return;
    }

    // Shared-memory addressing, all 32-bit byte addresses: a hot-loop load is
    // "uniform (slot base + chunk offset of this warp) + per-thread lane offset".
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(smem4);
    const uint32_t keys_base = smem_base + RING_B;
    const uint32_t lane_b = 80u * lane;                        // 5 chunk positions per lane: 4 data + 1 pad
    uint32_t uo[5];                                            // warp-uniform offsets of the 5 left chunks
#pragma unroll
    for (int i = 0; i < 5; i++) uo[i] = 16u * (RPOS + padpos(warp + i));
    auto slot_of = [&](int r) { return (uint32_t)(((r % NSLOT) + NSLOT) % NSLOT) * SLOT_B; };
    auto next_slot = [&](uint32_t sb) { return sb + SLOT_B == RING_B ? 0u : sb + SLOT_B; };
    auto row_ok = [&](int r) { return r >= 0 && r < (LM ? a.rmax : a.h2 - 1); };

    // staging: thread t copies chunk t of every row (right chunks first, then left chunks)
    const int y_start = yb0 - (H2 - 1);
    int r_pf = y_start + e_in;                                 // next row to stage
    bool ld_on[NST]; uint32_t ld_dst[NST]; int ld_pitch[NST]; const uint32_t* ld_ptr[NST];
#pragma unroll
    for (int c = 0; c < NST; c++) {
        const int ch = tid + c * NW * 32;
        const bool ld_right = ch < RCH;
        ld_on[c] = ch < RCH + LCH;
        ld_dst[c] = smem_base + 16u * (ld_right ? padpos(ch) : RPOS + padpos(ch - RCH));
        ld_pitch[c] = ld_right ? a.pitchR : a.pitchL;
        ld_ptr[c] = (ld_right ? imgR : imgL + dcta) + (ptrdiff_t)r_pf * ld_pitch[c] + CB + 4 * (ld_right ? ch : ch - RCH);
    }
    uint32_t slot_pf = slot_of(r_pf);
    auto issue_row = [&]() {
        const bool ok = row_ok(r_pf);
#pragma unroll
        for (int c = 0; c < NST; c++) {
            if (ok && ld_on[c]) cp_async16_s(slot_pf + ld_dst[c], ld_ptr[c]);
            ld_ptr[c] += ld_pitch[c];
        }
        cp_async_commit();
        r_pf++; slot_pf = next_slot(slot_pf);
    };

    // key reduction: each half-warp owns one 16-column group (= one thread's columns in the key row,
    // 20 words apart because of the padding).  The two groups of a warp are 4 apart: their bank
    // ranges [20g, 20g+16) and [20g+80, 20g+96) mod 32 are disjoint, so the LDS is conflict-free.
    const int rgrp = (warp >> 2) * 8 + (warp & 3) + ((lane >> 4) << 2);   // + 16 for the second group (NW = 8)
    const int rq = XT * rgrp + (lane & 15);
    const int xq = CB + h + rq;
    bool red_on[NRG];
#pragma unroll
    for (int r = 0; r < NRG; r++) red_on[r] = rgrp + 16 * r < 31 && xq + 16 * XT * r >= 0 && xq + 16 * XT * r < a.w2;
    const uint32_t red_b = keys_base + 4u * (20 * rgrp + (lane & 15));
    // merged: 2^(shift-6) consecutive passes share a plane (key = cost << shift | index within the plane's range)
    uint32_t* pl = planes + (size_t)(a.merged ? (pass * (DT * NW)) >> a.shift : pass) * a.plane_stride + (size_t)yb0 * a.w2 + xq;
    const bool merged = a.merged != 0;
    const int mshift = a.shift - 6;
    const uint32_t mpass = (uint32_t)((pass * (DT * NW)) & ((64 << mshift) - 1));   // first candidate of the pass within its plane
    auto reduce_row = [&](uint32_t buf) {
#pragma unroll
        for (int r = 0; r < NRG; r++)
            if (red_on[r]) {
                uint32_t kk[NW];
#pragma unroll
                for (int w = 0; w < NW; w++) kk[w] = lds32(red_b + 4u * (20 * 16 * r) + buf + w * KEYROW_B);
                uint32_t k = 0xFFFFFFFFu;
#pragma unroll
                for (int w = 0; w < NW; w++) k = min(k, kk[w]);
                // merged: the key carries the global candidate index, so the first minimum over all passes is one
                // fire-and-forget RED.MIN in L2 (the plane is preset to "none")
                if (merged) { if (k != 0xFFFFFFFFu) atomicMin(pl + 16 * XT * r, ((k & ~63u) << mshift) + (mpass + (k & 63u))); }
                else pl[16 * XT * r] = k;
            }
        pl += a.w2;
    };

    uint32_t col[XT][DT];
#pragma unroll
    for (int j = 0; j < XT; j++)
#pragma unroll
        for (int t = 0; t < DT; t++) col[j][t] = 0;

#pragma unroll
    for (int i = 0; i < PF; i++) issue_row();                  // rows of the first PF iterations

    uint32_t s_in = slot_of(y_start + e_in), s_out = slot_of(y_start - h), s_prev = slot_of(y_start - h - 1);
    const uint32_t key_st = keys_base + warp * KEYROW_B + lane_b;
    for (int y = y_start; y < yb1; y++) {
        const uint32_t buf = (y & 1) ? KEYBUF_B : 0u;
        cp_async_wait<PF - 1>();
        __syncthreads();
        issue_row();
        if (!red_late && y - 1 >= yb0) reduce_row(buf ^ KEYBUF_B);
        if (active) {
            // Phase staggering: the warps of one scheduler alternate between "subtract now" and
            // "subtract at the start of the next iteration", so that ALU-only phases (pointwise
            // costs) of some warps overlap the FMA-heavy window/WTA phase of the others.
            if (late_sub && y - 1 >= yb0 && row_ok(y - 1 - h) && colmask != 0)
                vertical_update<COST, false, false>(col, smem_base + s_prev, lane_b, uo, colmask, km1);
            if (row_ok(y + e_in) && colmask != 0) vertical_update<COST, true, false>(col, smem_base + s_in, lane_b, uo, colmask, km1);
            if (y >= yb0) {
                uint32_t keys[XT];
#pragma unroll
                for (int j = 0; j < XT; j++) keys[j] = 0xFFFFFFFFu;
                if (needmask) window_wta<H2, true, false, LM>(col, keys, k64, nk64, dkey0, vm);
                else if (COST == SR_COST_SAD && zmode) window_wta<H2, false, true>(col, keys, k64, nk64, dkey0, vm);
                else if (COST == SR_COST_SAD) window_wta<H2, false, false>(col, keys, k64, nk64, dkey0, vm);
                else window_wta4<H2, false>(col, keys, k64, nk64, dkey0);
#pragma unroll
                for (int k = 0; k < 4; k++)
                    sts128(key_st + buf + 16 * k, keys[4 * k], keys[4 * k + 1], keys[4 * k + 2], keys[4 * k + 3]);
                // (a template flag and kernels of its own: with this line in the common kernel every shape ran 0.5 - 3 % slower)
                if (C0 && d0 == 0) emit_c0<H2>(col, a.c0 + (size_t)y * a.w2, cb + h, a.w2, lane);
                if (!late_sub && y + 1 < yb1 && row_ok(y - h) && colmask != 0)
                    vertical_update<COST, false, false>(col, smem_base + s_out, lane_b, uo, colmask, km1);
            }
        } else if (y >= yb0) {
#pragma unroll
            for (int k = 0; k < 4; k++) sts128(key_st + buf + 16 * k, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
        }
        if (red_late && y - 1 >= yb0) reduce_row(buf ^ KEYBUF_B);
        s_prev = s_out; s_out = next_slot(s_out); s_in = next_slot(s_in);
    }
    cp_async_wait<0>();
    __syncthreads();
    reduce_row(((yb1 - 1) & 1) ? KEYBUF_B : 0u);
}

template <int H2, int COST, bool LM = false, int NW = 16, bool C0 = false>
int launch(sr_ctx* ctx, const TileArgs& a, int nwork) {
    size_t smem = ((size_t)(H2 + PF + 1) * SLOT_U4 + 2 * NW * KEY_U4) * sizeof(uint4);
    static DeviceOnce configured;
    if (configured.need(ctx->device)) {
        SR_CUDA(ctx, cudaFuncSetAttribute(search_tile_kernel<H2, COST, NW, LM, C0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured.done(ctx->device);
    }
    SR_CUDA(ctx, cudaEventRecord(ctx->kev0, ctx->stream));
    search_tile_kernel<H2, COST, NW, LM, C0><<<nwork, NW * 32, smem, ctx->stream>>>(a);
    SR_LAUNCHED(ctx);
    SR_CUDA(ctx, cudaEventRecord(ctx->kev1, ctx->stream));
    ctx->kev_valid = true;
    return 1;
}

template <int COST>
int dispatch_h2(sr_ctx* ctx, const TileArgs& a, int nwork) {
    if (a.c0 != nullptr) {                                // d == 0 cost plane wanted (smoothing rule): kernels of their own
        switch (2 * a.half) {
            case 2: return launch<2, COST, false, 16, true>(ctx, a, nwork);
            case 4: return launch<4, COST, false, 16, true>(ctx, a, nwork);
            case 6: return launch<6, COST, false, 16, true>(ctx, a, nwork);
            case 8: return launch<8, COST, false, 16, true>(ctx, a, nwork);
            case 10: return launch<10, COST, false, 16, true>(ctx, a, nwork);
            case 12: return launch<12, COST, false, 16, true>(ctx, a, nwork);
            case 14: return launch<14, COST, false, 16, true>(ctx, a, nwork);
            case 16: return launch<16, COST, false, 16, true>(ctx, a, nwork);
            default: return 0;
        }
    }
    switch (2 * a.half) {
        case 2: return a.pw == 32 ? launch<2, COST, false, 8>(ctx, a, nwork) : launch<2, COST>(ctx, a, nwork);
        case 4: return a.pw == 32 ? launch<4, COST, false, 8>(ctx, a, nwork) : launch<4, COST>(ctx, a, nwork);
        case 6: return a.pw == 32 ? launch<6, COST, false, 8>(ctx, a, nwork) : launch<6, COST>(ctx, a, nwork);
        case 8: return a.pw == 32 ? launch<8, COST, false, 8>(ctx, a, nwork) : launch<8, COST>(ctx, a, nwork);
        case 10: return a.pw == 32 ? launch<10, COST, false, 8>(ctx, a, nwork) : launch<10, COST>(ctx, a, nwork);
        case 12: return launch<12, COST>(ctx, a, nwork);
        case 14: return launch<14, COST>(ctx, a, nwork);
        case 16: return launch<16, COST>(ctx, a, nwork);
        default: return 0;
    }
}

// Left variant: odd windows 3 ... 17, SSD only (the reference's cost)
int dispatch_left(sr_ctx* ctx, const TileArgs& a, int nwork) {
    switch (a.hwin) {
        case 3: return launch<3, SR_COST_SSD_L2, true>(ctx, a, nwork);
        case 5: return launch<5, SR_COST_SSD_L2, true>(ctx, a, nwork);
        case 7: return launch<7, SR_COST_SSD_L2, true>(ctx, a, nwork);
        case 9: return launch<9, SR_COST_SSD_L2, true>(ctx, a, nwork);
        case 11: return launch<11, SR_COST_SSD_L2, true>(ctx, a, nwork);
        case 13: return launch<13, SR_COST_SSD_L2, true>(ctx, a, nwork);
        case 15: return launch<15, SR_COST_SSD_L2, true>(ctx, a, nwork);
        case 17: return launch<17, SR_COST_SSD_L2, true>(ctx, a, nwork);
        default: return 0;
    }
}

// The schedule of one launch: work items {tile, first row, end row, pass} in the order the CTAs should start.
//  * Bands, then tiles, then passes (heaviest -- masked / ragged -- pass first): the passes of one (tile, band)
//    run at the same time and share its image rows in L2, so DRAM reads stay near one sweep of the two images
//    (pass-major order re-reads them once per wave: 295 MB instead of 106 MB on the 4K pair).
//  * The rows are cut into `nb` equal bands followed by a tail of bands 1/div as tall over the last `pct` percent of
//    the rows: the last CTAs to start are short, so the SMs run dry together at the end of the kernel.
//  * With the merged key plane, (tile, pass) combinations without a single candidate (break rule at the right
//    image edge) are left out altogether.
// Lists are cached per geometry on the context (the row-chunk pipeline repeats the same few shapes).
// `trial(list, n)` (may be empty) runs the kernel on a candidate list and returns its time in ms: see "autotuned" below.
int tile_work_list(sr_ctx* ctx, const SearchGeom& g, const TileArgs& a, int ntx, int rows, const int4** work, int* nwork,
                   const std::function<float(const int4*, int)>& trial) {
    const int npairs = a.npairs > 0 ? a.npairs : 1;
    static const char* ord_env = getenv("SR_TILE_ORDER");   // 0: pass-major (development comparison)
    static const char* tl_env = getenv("SR_TILE_TAIL");     // "percent,divisor" tuning override
    static const char* nb_env = getenv("SR_TILE_NBANDS");   // tuning override
    const int order = ord_env ? atoi(ord_env) : 1;
    int f_pct = -1, f_div = -1, f_nb = nb_env ? atoi(nb_env) : 0;
    if (tl_env) sscanf(tl_env, "%d,%d", &f_pct, &f_div);
    const int no_tail = ctx->more_chunks_follow ? 1 : 0;   // row-chunk pipeline: the next chunk's CTAs fill the end of this launch
    TileWork key;
    const int kv[TileWork::NKEY] = {a.lmode ? a.lim_a : g.w1, g.w2, g.half, a.dlo, a.dhi, a.row0, a.row1, a.merged, order, f_pct * 64 + f_div, f_nb,
                                    no_tail + 2 * npairs + 64 * a.lmode + 128 * (a.pw == 32)};
    for (int i = 0; i < TileWork::NKEY; i++) key.key[i] = kv[i];
    for (TileWork& t : ctx->tile_work)
        if (!memcmp(t.key, key.key, sizeof key.key)) { t.stamp = ++ctx->tile_work_clock; *work = t.dev; *nwork = t.n; return SR_OK; }

    const int warm = a.hwin - 1, minband = a.hwin + 2;
    auto dlim_at = [&](int x) {
        if (a.lmode) return a.lim_a - x < a.dhi ? a.lim_a - x : a.dhi;
        int r = g.w2 - x - 1 < g.half ? g.w2 - x - 1 : g.half;
        return g.w1 - x - r < a.dhi ? g.w1 - x - r : a.dhi;
    };
    auto x_first_of = [&](int tile) { int x = -XT + TILE_OUT * tile + g.half; return x < 0 ? 0 : x; };
    auto has_candidates = [&](int tile, int pass) { return a.dbase + a.pw * pass < dlim_at(x_first_of(tile & 0xFFFF)); };
    // (tile, pass) columns in launch order within a band, with their relative cost per row
    struct Col { int tile, pass; };
    std::vector<Col> cols;
    for (int q = 0; q < npairs; q++)                           // several pairs: more columns per band, taller bands
        for (int t = 0; t < ntx; t++)
            for (int p = a.npass - 1; p >= 0; p--) {
                if (a.merged && !has_candidates(t, p)) continue;   // planes mode: the CTA still writes its "none" keys
                cols.push_back({t | (q << 16), p});
            }
    auto make_bands = [&](int nb, int pct, int div) {
        std::vector<int2> bands;
        if (nb < 1) nb = 1;
        int band = (rows + nb - 1) / nb;
        if (band < minband) band = minband;
        int band2 = div > 1 ? band / div : band;
        if (band2 < minband) band2 = minband;
        int rows_tail = (div > 1 && pct > 0) ? rows * pct / 100 : 0;
        int nb1 = (rows - rows_tail + band - 1) / band;
        int split = a.row0 + (nb1 * band < rows ? nb1 * band : rows);
        for (int y = a.row0; y < split; y += band) bands.push_back(make_int2(y, y + band < split ? y + band : split));
        for (int y = split; y < a.row1; y += band2) bands.push_back(make_int2(y, y + band2 < a.row1 ? y + band2 : a.row1));
        return bands;
    };
    auto make_items = [&](const std::vector<int2>& bands) {
        std::vector<int4> items;
        if (order) {
            for (int2 b : bands) for (const Col& c : cols) items.push_back(make_int4(c.tile, b.x, b.y, c.pass));
        } else {
            for (int p = a.npass - 1; p >= 0; p--) for (int2 b : bands) for (const Col& c : cols) if (c.pass == p) items.push_back(make_int4(c.tile, b.x, b.y, c.pass));
        }
        return items;
    };
    // Heuristic: CTAs are list-scheduled one per SM and their durations vary (masked passes on the break-rule diagonal,
    // image edges), so the end of the launch costs about one CTA duration on top of the average load: nb minimises
    // (CTAs + SMs) / SMs x (band + warm-up rows); the last 10 % of the rows go in bands a quarter as tall (20 % in thirds
    // for launches of 600 rows or fewer).
    // AUTOTUNED on first use of a geometry: the kernel's time is very sensitive to how the CTA count falls on the 148 SMs
    // (1436x992 / 145 disparities / block 17: 0.198 ms with 34 bands, 0.244 ms with 37; 4K: 1.73 ms with 20-24 bands + tail,
    // 1.86-2.15 ms without a tail, profiles/r02_tile_schedule_sweep.log), and no formula I tried tracks it -- a
    // list-scheduling simulation with per-CTA cost estimates chose 12 tall bands for the 4K pair (1.85 ms).  So the first
    // launch of a geometry on a context runs up to 12 candidate schedules (0.5 ... 1.5 x the heuristic's band count, two
    // tail shapes) twice each on the real data, keeps the fastest and caches it (every candidate computes the same, correct
    // result; a few ms to a few tens of ms once per shape; SR_TILE_AUTOTUNE=0 or the row-chunk pipeline: heuristic only).
    std::vector<int4> items;
    std::vector<int2> best_bands;
    if (!cols.empty()) {
        const int maxnb = rows / minband < 1 ? 1 : (rows / minband > 256 ? 256 : rows / minband);
        const int slots = ctx->sms * (a.pw == 32 ? 2 : 1);     // CTAs resident at a time
        int nb_best = 1;
        double best = -1;
        for (int nb = 1; nb <= maxnb; nb++) {
            int bnd = (rows + nb - 1) / nb, nbe = (rows + bnd - 1) / bnd;
            double c = ((double)cols.size() * nbe + slots) / slots * (bnd + warm + 2);
            if (best < 0 || c < best) { best = c; nb_best = nbe; }
        }
        if (f_nb > 0) nb_best = f_nb < maxnb ? f_nb : maxnb;
        int pct = rows <= 600 ? 20 : 10, div = rows <= 600 ? 3 : 4;
        if (f_pct >= 0) { pct = f_pct; div = f_div > 0 ? f_div : 2; }
        if (no_tail || !order || nb_best < 4) pct = f_pct >= 0 && !no_tail ? pct : 0;
        best_bands = make_bands(nb_best, pct, div);
        items = make_items(best_bands);
        static const char* at_env = getenv("SR_TILE_AUTOTUNE");
        if (trial && !(at_env && atoi(at_env) == 0) && f_nb <= 0 && f_pct < 0 && !no_tail && order) {
            struct Cand { int nb, pct, div; };
            std::vector<Cand> cands;
            static const double scale[] = {0.5, 0.7, 0.85, 1.0, 1.2, 1.5};
            for (double sc : scale) {
                int nb = (int)(nb_best * sc + 0.5);
                nb = nb < 1 ? 1 : (nb > maxnb ? maxnb : nb);
                bool dup = false;
                for (const Cand& c : cands) dup = dup || c.nb == nb;
                if (dup) continue;
                if (nb >= 4) { cands.push_back({nb, 10, 4}); cands.push_back({nb, 20, 3}); }
                else cands.push_back({nb, 0, 1});
            }
            int4* tmp = nullptr;
            size_t tmp_cap = 0;
            float best_ms = -1;
            // the heuristic's own plan first: a candidate replaces it only when it is at least 2 % faster (timing noise
            // must not trade a good plan for an equal one)
            cands.insert(cands.begin(), Cand{nb_best, pct, div});
            bool first = true;
            for (const Cand& c : cands) {
                std::vector<int2> bands = make_bands(c.nb, c.pct, c.div);
                std::vector<int4> it = make_items(bands);
                if (it.empty() || it.size() > 60000) continue;
                if (it.size() > tmp_cap) {
                    cudaFree(tmp);
                    tmp_cap = it.size() + 1024;
                    SR_CUDA(ctx, cudaMalloc((void**)&tmp, tmp_cap * sizeof(int4)));
                }
                SR_CUDA(ctx, cudaMemcpy(tmp, it.data(), it.size() * sizeof(int4), cudaMemcpyHostToDevice));
                float ms = -1;
                for (int rep = 0; rep < 3; rep++) {
                    float m = trial(tmp, (int)it.size());
                    if (m < 0) { ms = -1; break; }
                    if (ms < 0 || m < ms) ms = m;
                }
                if (ms < 0) break;
                if (getenv("SR_TILE_PLAN_DEBUG")) fprintf(stderr, "[tile tune] nb %d tail %d/%d: %zu bands, %zu items, %.4f ms\n", c.nb, c.pct, c.div, bands.size(), it.size(), ms);
                if (first || ms < 0.98f * best_ms) { best_ms = ms; best_bands.swap(bands); items.swap(it); }
                first = false;
            }
            cudaFree(tmp);
        }
    }
    static const char* dbg_env = getenv("SR_TILE_PLAN_DEBUG");
    if (dbg_env) {
        fprintf(stderr, "[tile plan] rows %d..%d cols %zu bands %zu (first %d, last %d rows) items %zu\n", a.row0, a.row1,
                cols.size(), best_bands.size(), best_bands.empty() ? 0 : best_bands.front().y - best_bands.front().x,
                best_bands.empty() ? 0 : best_bands.back().y - best_bands.back().x, items.size());
    }
    key.n = (int)items.size();
    key.dev = nullptr;
    if (key.n) {
        if (ctx->tile_work.size() >= 64) {   // drop the least recently used list (no launch may still read it)
            size_t lru = 0;
            for (size_t i = 1; i < ctx->tile_work.size(); i++) if (ctx->tile_work[i].stamp < ctx->tile_work[lru].stamp) lru = i;
            SR_CUDA(ctx, cudaDeviceSynchronize());
            cudaFree(ctx->tile_work[lru].dev);
            ctx->tile_work.erase(ctx->tile_work.begin() + lru);
        }
        SR_CUDA(ctx, cudaMalloc((void**)&key.dev, items.size() * sizeof(int4)));
        SR_CUDA(ctx, cudaMemcpy(key.dev, items.data(), items.size() * sizeof(int4), cudaMemcpyHostToDevice));
    }
    key.stamp = ++ctx->tile_work_clock;
    if (key.n) ctx->tile_work.push_back(key);
    *work = key.dev; *nwork = key.n;
    return SR_OK;
}

}  // namespace

// One launch over `n` pairs of one shape (n == 1: the plain call).  With several pairs the work list carries the pair
// index: a 1280x720 / 128-disparity pair alone is 6 (tile, pass) columns for 148 SMs -- 50 bands of 15 rows, a third of
// every CTA's rows warm-up; eight pairs make 48 columns and bands of ~100 rows.  Key planes of the pairs lie one after
// the other in the context's scratch slot 0 (pair q: planes + q * npass * npx_cap).
namespace {
// Left variant (lmode): images and bounds of the launch (see TileArgs)
struct LeftMode { const uint32_t* searched; int pitch; int cmax, rmax, lim_a, lo_a; };
int tile_launch_impl(sr_ctx* ctx, sr_pair* const* pairs, int n, const SearchGeom& g, int cost, const LeftMode* lm, bool want_c0 = false) {
    if (n < 1 || n > MAXP) return 0;
    sr_pair* pair = pairs[0];
    if (g.half < 1 || g.half > 8) return 0;
    if (cost != SR_COST_SSD_L2 && cost != SR_COST_SAD) return 0;
    int r1 = g.row1 < g.height ? g.row1 : g.height;
    if (r1 <= g.row0) { for (int q = 0; q < n; q++) pairs[q]->npass = 0; return 1; }  // nothing to compute: the resolve pass fills the rows
    TileArgs a;
    a.npairs = n > 1 ? n : 0;
    a.L = pair->L.px; a.pitchL = pair->L.pitch; a.R = pair->R.px; a.pitchR = pair->R.pitch;
    for (int q = 1; q < n; q++)
        if (pairs[q]->L.pitch != a.pitchL || pairs[q]->R.pitch != a.pitchR || pairs[q]->npx_cap != pair->npx_cap) return 0;
    a.w1 = g.w1; a.h1 = g.h1; a.w2 = g.w2; a.h2 = g.h2; a.half = g.half;
    a.dlo = g.dlo; a.dhi = g.dhi; a.dbase = g.dlo & ~3;
    // 8 warps per CTA, 32 candidates per pass and two CTAs per SM (possible where two rings + key buffers fit the SM's shared
    // memory: windows up to 10 rows).  Measured 4-5 % SLOWER than 16 warps / one CTA per SM at equal work (4K / 512 SAD:
    // kernel 1.776 vs 1.693 ms; the two unsynchronised CTAs do not make up for twice the passes), so it is taken only where
    // the narrower pass saves a tenth of the candidates or more (145 candidates: 160 instead of 192 -- 0.182 -> 0.173 ms
    // on 1436x992, block 11).  SR_TILE_NW=16 / 8 overrides (A/B measurements, tools/ab_tile.py).
    static const char* nw_env = getenv("SR_TILE_NW");
    const int span = g.dhi - a.dbase;
    int nw = 16;
    if (!lm && !want_c0 && 2 * g.half <= 10 && span > 0) {
        if (10 * ((span + 31) / 32 * 32) <= 9 * ((span + 63) / 64 * 64)) nw = 8;
        if (nw_env) nw = atoi(nw_env) == 8 ? 8 : 16;
    }
    a.pw = 4 * nw;
    a.npass = (g.dhi - a.dbase + a.pw - 1) / a.pw;
    a.row0 = g.row0; a.row1 = r1; a.k64 = 64u; a.nk64 = (uint32_t)-64; a.merged = 0; a.shift = 6; a.km1 = 0xFFFFFFFFu;
    a.lmode = 0; a.hwin = 2 * g.half; a.cmax = a.rmax = a.lim_a = a.lo_a = 0;
    a.c0 = (want_c0 && !lm && n == 1 && a.dbase == 0) ? (uint32_t*)pair->c0 : nullptr;
    pair->c0_from_tile = a.c0 != nullptr;
    if (lm) {
        // reference image (window centre, output) = the LEFT image, searched image = the shifted copy of the right one
        a.lmode = 1; a.hwin = 2 * g.half + 1;
        a.L = lm->searched; a.pitchL = lm->pitch; a.R = pair->L.px; a.pitchR = pair->L.pitch;
        a.cmax = lm->cmax; a.rmax = lm->rmax; a.lim_a = lm->lim_a; a.lo_a = lm->lo_a;
    }
    // per-warp instruction mix of the SAD window phase (window_wta ZMODE); SR_TILE_ZMASK overrides for tuning
    static const char* zm_env = getenv("SR_TILE_ZMASK");
    a.zmask = zm_env ? (uint32_t)strtoul(zm_env, nullptr, 0) : 0xFFF0u;   // measured: tools/tune_tile.sh
    static const char* rm_env = getenv("SR_TILE_REDMODE");
    a.red_mode = rm_env ? atoi(rm_env) : 1;
    int ntx = (g.w2 + XT - g.half + TILE_OUT - 1) / TILE_OUT;
    int rows = r1 - g.row0;
    if (a.npass <= 0) {
        // empty candidate range: every record is "none"
        for (int q = 0; q < n; q++) pairs[q]->npass = 0;
        return 1;
    }
    // Key planes.  Where cost and candidate index fit one u32 together (cost << shift | index), 2^(shift-6) consecutive
    // passes fold into ONE plane with RED.MIN -- all of them when the cost is small enough (SAD always; SSD on small
    // windows / ranges), else as many as fit (SSD 9x9 at 512 disparities: 2 planes of 4 passes instead of 8 planes -- the
    // kernel's DRAM writes drop from 209 to ~55 MB and the resolve pass reads 2 planes instead of 8).  The planes stay in
    // L2; a plane's candidate range is [dbase + (p << shift), +2^shift).  shift == 6: one plane per pass, plain stores.
    int full = 6;
    while ((64 << (full - 6)) < a.npass * a.pw) full++;
    unsigned long long cmax = (unsigned long long)g.block * g.block * 3ull * (cost == SR_COST_SAD ? 255ull : 65025ull);
    static const char* mg_env = getenv("SR_TILE_MERGE");    // tuning / test override: 0 = never merge
    int shift = 6;
    if (!(mg_env && atoi(mg_env) == 0))
        while (shift < full && ((cmax + 1) << (shift + 1)) < 0xFFFFFFFFull) shift++;
    a.merged = shift > 6 || a.pw < 64;                  // several passes per plane: RED.MIN into a preset plane
    a.shift = shift;
    const int planes_per_pair = (a.npass * a.pw + (1 << shift) - 1) >> shift;
    uint32_t* base = (uint32_t*)sr_scratch(ctx, 0, (size_t)n * planes_per_pair * pair->npx_cap * sizeof(uint32_t));
    if (!base) return SR_E_CUDA;
    PairPtrs hp[MAXP];
    for (int q = 0; q < n; q++) {
        sr_pair* pq = pairs[q];
        pq->npass = planes_per_pair; pq->plane_dbase = a.dbase; pq->key_shift = shift;
        pq->planes = base + (size_t)q * planes_per_pair * pair->npx_cap;
        hp[q].L = pq->L.px; hp[q].R = pq->R.px; hp[q].planes = pq->planes;
        if (a.merged)
            for (int pp = 0; pp < planes_per_pair; pp++)
                SR_CUDA(ctx, cudaMemsetAsync(pq->planes + (size_t)pp * pair->npx_cap + (size_t)g.row0 * g.w2, 0xFF,
                                             (size_t)(r1 - g.row0) * g.w2 * sizeof(uint32_t), ctx->stream));
    }
    a.planes = pair->planes; a.plane_stride = pair->npx_cap;
    for (int q = 0; q < n; q++) a.ppv[q] = hp[q];
    int nwork = 0;
    // a trial run for the autotuner: the same launch on a candidate list, timed with the launch's own events
    std::function<float(const int4*, int)> trial = [&](const int4* list, int cnt) -> float {
        TileArgs t = a;
        t.work = list;
        int ok = t.lmode ? dispatch_left(ctx, t, cnt) : cost == SR_COST_SAD ? dispatch_h2<SR_COST_SAD>(ctx, t, cnt) : dispatch_h2<SR_COST_SSD_L2>(ctx, t, cnt);
        float ms = -1.f;
        if (ok != 1 || cudaEventSynchronize(ctx->kev1) != cudaSuccess || cudaEventElapsedTime(&ms, ctx->kev0, ctx->kev1) != cudaSuccess) return -1.f;
        return ms;
    };
    int rc = tile_work_list(ctx, g, a, ntx, rows, &a.work, &nwork, trial);
    if (rc != SR_OK) return rc;
    if (nwork == 0) return 1;                                // merged plane, no (tile, pass) has a candidate: all "none"
    if (a.lmode) return dispatch_left(ctx, a, nwork);
    if (cost == SR_COST_SAD) return dispatch_h2<SR_COST_SAD>(ctx, a, nwork);
    return dispatch_h2<SR_COST_SSD_L2>(ctx, a, nwork);
}
}  // namespace

int sr_launch_search_tile_multi(sr_ctx* ctx, sr_pair* const* pairs, int n, const SearchGeom& g, int cost) {
    return tile_launch_impl(ctx, pairs, n, g, cost, nullptr);
}

// Cost + first-minimum search of the Left variant (BlockSearch.cpp:24-86) through the tile kernel: per left pixel the
// unbiased winner over cx = x - maxD ... x - 1.  The right image is packed a second time, shifted right by
// off = maxD rounded up to 4 (pair->Rs), so that candidate index i = off - d addresses it at x + i like the Right
// search addresses the left image: ascending i is the reference's iteration order (d descending) and the first minimum
// is the reference's winner.  Key planes as usual (pair->planes, index i); *off_out = off.  Returns 1 when handled.
int sr_launch_search_tile_leftvariant(sr_ctx* ctx, sr_pair* pair, int block, int max_disp, int* off_out) {
    const int w1 = pair->w1, h1 = pair->h1, w2 = pair->w2, h2 = pair->h2;
    const int height = h1 < h2 ? h1 : h2, h = (block - 1) / 2;
    if ((block & 1) == 0 || h < 1 || h > 8 || max_disp < 1 || max_disp > 4096) return 0;
    if (height - 2 * h <= 0 || w1 - 2 * h <= 0) return 0;
    const int off = (max_disp + 3) & ~3;
    const int need_w = (w1 > w2 ? w1 : w2) + off;
    if (!pair->Rs.base || pair->Rs.w < need_w || pair->Rs.h != h2) {
        if (pair->Rs.base) { SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); cudaFree(pair->Rs.base); pair->Rs = PackedImage(); }
        int rc = sr_alloc_packed(ctx, &pair->Rs, need_w, h2, SR_TILE_OVERSHOOT_PAD);
        if (rc != SR_OK) return rc;
    }
    // the shift differs from call to call: clear, then pack the image columns at px + off
    SR_CUDA(ctx, cudaMemsetAsync(pair->Rs.base, 0, (size_t)pair->Rs.pitch * h2 * sizeof(uint32_t), ctx->stream));
    PackedImage view = pair->Rs;
    view.px = pair->Rs.px + off; view.w = w2;
    int rc = sr_launch_pack(ctx, pair->rawR, (size_t)w2 * 3, w2, h2, &view);
    if (rc != SR_OK) return rc;
    SearchGeom g;
    g.w1 = w2; g.h1 = h2; g.w2 = w1; g.h2 = h1; g.height = height; g.block = block; g.half = h;
    g.dlo = off - max_disp; g.dhi = off; g.row0 = h; g.row1 = height - h;
    LeftMode lm{pair->Rs.px, pair->Rs.pitch, w1, height, w2 - h + off, h + off};
    *off_out = off;
    return tile_launch_impl(ctx, &pair, 1, g, SR_COST_SSD_L2, &lm);
}

// want_c0: also leave the raw d == 0 window costs in pair->c0 as a u32 plane (pair->c0_from_tile says whether it did)
int sr_launch_search_tile(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, int cost, int left_ref, int want_c0) {
    if (left_ref) return 0;
    return tile_launch_impl(ctx, &pair, 1, g, cost, nullptr, want_c0 != 0);
}
