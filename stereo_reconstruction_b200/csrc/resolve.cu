// resolve.cu -- turns the per-pixel search records into the reference's CV_64F disparity map,
// including the smoothFactor rule of BlockSearch.cpp:160-165.
//
// In the Right variant the rule compares the neighbour's stored disparity with
// (double)(pointX - correspondX) = -d, so it only ever fires for d == 0 next to a stored 0
// ("hard parts").  Per pixel that leaves three booleans a_k = "d = 0 wins when its
// cost was multiplied k times by smoothFactor" (k = 0,1,2), evaluated in fp64 exactly as the
// reference does (sqrt(ssd)/N, then *s, *s), and one alternative value (the first minimum over
// d >= 1, or -x when nothing is a candidate).  The raster dependency is then a carry chain
//     Z(y,x) = Z(y,x-1) ? F1(y,x) : F0(y,x),   F0 = Zup ? a1 : a0,  F1 = Zup ? a2 : a1
// resolved row by row with bit-parallel carry propagation (one 32-pixel word per thread).
#include "sr_internal.cuh"

struct Planes { const uint32_t* planes; size_t stride; int npass, dbase, shift; };

struct PixRec {
    bool skip;                  // black pixel / unvisited row -> 0.0
    unsigned long long c0, cr;  // SR_NONE64 when absent
    int dr;
    int n;                      // window pixel count N
};

__device__ __forceinline__ PixRec load_rec(int from_direct, int have_c0, size_t i, int x, int y,
                                           Planes pl, const unsigned long long* c0,
                                           const unsigned long long* cr, const int* dr, const int* nwin,
                                           const uint32_t* Rp, int pitchR, const SearchGeom& g) {
    PixRec r;
    if (from_direct) {
        r.n = nwin[i];
        r.skip = r.n < 0;
        r.c0 = c0[i]; r.cr = cr[i]; r.dr = dr[i];
    } else {
        r.skip = y >= g.height || (__ldg(Rp + (size_t)y * pitchR + x) & 0xFFFFFFu) == 0;
        int l = min(x, g.half), rr = min(g.w2 - x - 1, g.half), u = min(y, g.half), dn = min(g.h2 - y - 1, g.half);
        r.n = (l + rr) * (u + dn);
        // first minimum over the key planes (cost << shift | candidate index; one plane per pass of 64
        // candidates, or one plane for all when the tile kernel folded them; passes ascend in d, strict <
        // keeps the smaller d on ties)
        uint32_t bc = 0xFFFFFFFFu; int bd = 0;
        for (int p = 0; p < pl.npass; p++) {
            uint32_t k = __ldg(pl.planes + (size_t)p * pl.stride + i);
            if (k != 0xFFFFFFFFu && (k >> pl.shift) < bc) { bc = k >> pl.shift; bd = pl.dbase + (p << pl.shift) + (int)(k & ((1u << pl.shift) - 1u)); }
        }
        r.cr = bc == 0xFFFFFFFFu ? SR_NONE64 : (unsigned long long)bc;
        r.dr = bd;
        if (have_c0 == 2) {
            // u32 plane written by the tile kernel: raw window cost at d == 0; it is a candidate when the
            // break rule lets d == 0 through (x + r < w1)
            r.c0 = (!r.skip && x + rr < g.w1) ? (unsigned long long)((const uint32_t*)c0)[i] : SR_NONE64;
        } else {
            r.c0 = have_c0 ? c0[i] : SR_NONE64;
        }
    }
    return r;
}

// smooth == 1 (or no d == 0 candidate in range): pointwise.
__global__ void resolve_plain_kernel(SearchGeom g, int from_direct, int have_c0, Planes best,
                                     const unsigned long long* __restrict__ c0, const unsigned long long* __restrict__ cr,
                                     const int* __restrict__ dr, const int* __restrict__ nwin,
                                     const uint32_t* __restrict__ Rp, int pitchR, double* __restrict__ out) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = g.row0 + blockIdx.y;
    if (x >= g.w2 || y >= g.row1) return;
    size_t i = (size_t)y * g.w2 + x;
    PixRec r = load_rec(from_direct, have_c0, i, x, y, best, c0, cr, dr, nwin, Rp, pitchR, g);
    double v;
    if (r.skip) v = 0.0;
    else if (r.n == 0) v = (double)(0 - x);  // 0/0 = NaN never beats DBL_MAX (:168)
    else if (r.c0 != SR_NONE64 && (r.cr == SR_NONE64 || r.c0 <= r.cr)) v = 0.0;  // d=0 first, strict <
    else if (r.cr != SR_NONE64) v = (double)r.dr;
    else v = (double)(0 - x);  // no candidate survived the break (:111,:174)
    out[i] = v;
}

// The same for the tile kernel's key planes, two pixels per thread (8-byte loads of the packed pixels, the key
// plane(s) and the u32 d = 0 plane, one 16-byte store): twice the bytes in flight per thread.  Needs an even
// width and even pitches (checked by the launcher); per pixel the logic is load_rec(from_direct = 0) +
// resolve_plain_kernel, line for line.
__global__ void resolve_plain2_kernel(SearchGeom g, int have_c0 /* 0 or 2 */, Planes pl, const uint32_t* __restrict__ c0u,
                                      const uint32_t* __restrict__ Rp, int pitchR, double* __restrict__ out) {
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int y = g.row0 + blockIdx.y;
    if (x >= g.w2 || y >= g.row1) return;
    const size_t i = (size_t)y * g.w2 + x;
    const bool yin = y < g.height;
    const uint2 px2 = yin ? __ldg(reinterpret_cast<const uint2*>(Rp + (size_t)y * pitchR + x)) : make_uint2(0u, 0u);
    uint32_t bc[2] = {0xFFFFFFFFu, 0xFFFFFFFFu};
    int bd[2] = {0, 0};
    for (int p = 0; p < pl.npass; p++) {
        const uint2 k2 = __ldg(reinterpret_cast<const uint2*>(pl.planes + (size_t)p * pl.stride + i));
        const uint32_t k[2] = {k2.x, k2.y};
#pragma unroll
        for (int e = 0; e < 2; e++)
            if (k[e] != 0xFFFFFFFFu && (k[e] >> pl.shift) < bc[e]) {
                bc[e] = k[e] >> pl.shift;
                bd[e] = pl.dbase + (p << pl.shift) + (int)(k[e] & ((1u << pl.shift) - 1u));
            }
    }
    const uint2 c02 = have_c0 == 2 ? __ldg(reinterpret_cast<const uint2*>(c0u + i)) : make_uint2(0u, 0u);
    const uint32_t px[2] = {px2.x, px2.y}, c0raw[2] = {c02.x, c02.y};
    double v[2];
#pragma unroll
    for (int e = 0; e < 2; e++) {
        const int xe = x + e;
        const bool skip = !yin || (px[e] & 0xFFFFFFu) == 0;
        const int l = min(xe, g.half), rr = min(g.w2 - xe - 1, g.half), u = min(y, g.half), dn = min(g.h2 - y - 1, g.half);
        const int n = (l + rr) * (u + dn);
        const unsigned long long cr = bc[e] == 0xFFFFFFFFu ? SR_NONE64 : (unsigned long long)bc[e];
        const unsigned long long c0 = (have_c0 == 2 && !skip && xe + rr < g.w1) ? (unsigned long long)c0raw[e] : SR_NONE64;
        if (skip) v[e] = 0.0;
        else if (n == 0) v[e] = (double)(0 - xe);
        else if (c0 != SR_NONE64 && (cr == SR_NONE64 || c0 <= cr)) v[e] = 0.0;
        else if (cr != SR_NONE64) v[e] = (double)bd[e];
        else v[e] = (double)(0 - xe);
    }
    *reinterpret_cast<double2*>(out + i) = make_double2(v[0], v[1]);
}

// The same two-pixel resolve with the int16 wire format of the band gather as its output ("fuse by letting
// the WTA epilogue write straight into the registered send buffer"): ranks > 0 of a row-band sharded pair never form
// the CV_64F band -- the disparities go as short2 stores into the root's gather window over NVLink, and the last block
// publishes the step in the root's flag word of this rank (the acknowledgement of the previous step was awaited by a
// one-thread kernel earlier on the stream, sr_peer_wait_ack).  Values outside int16 (an image wider than 32767 pixels
// whose "no candidate" value -x does not fit) set the pair's sticky status instead of the flag.
__global__ void __launch_bounds__(128) resolve_wire16_kernel(SearchGeom g, Planes pl, const uint32_t* __restrict__ Rp, int pitchR,
                                                             short* win16, volatile unsigned* flags, int rank, unsigned seq,
                                                             unsigned* counter, int* status) {
    // a few resident CTAs per SM walk the (row, 256-pixel segment) tiles: ONE system-scope fence per CTA at the end (a
    // fence waits for the NVLink write acknowledgements; one per 128-thread tile made the kernel latency-bound)
    const int segs = (g.w2 + 255) / 256, rows = g.row1 - g.row0;
    int bad = 0;
    for (int tile = blockIdx.x; tile < segs * rows; tile += gridDim.x) {
        const int x = 2 * ((tile % segs) * 128 + threadIdx.x);
        const int y = g.row0 + tile / segs;
        if (x >= g.w2) continue;
        const size_t i = (size_t)y * g.w2 + x;
        const bool yin = y < g.height;
        const uint2 px2 = yin ? __ldg(reinterpret_cast<const uint2*>(Rp + (size_t)y * pitchR + x)) : make_uint2(0u, 0u);
        uint32_t bc[2] = {0xFFFFFFFFu, 0xFFFFFFFFu};
        int bd[2] = {0, 0};
        for (int p = 0; p < pl.npass; p++) {
            const uint2 k2 = __ldg(reinterpret_cast<const uint2*>(pl.planes + (size_t)p * pl.stride + i));
            const uint32_t k[2] = {k2.x, k2.y};
#pragma unroll
            for (int e = 0; e < 2; e++)
                if (k[e] != 0xFFFFFFFFu && (k[e] >> pl.shift) < bc[e]) {
                    bc[e] = k[e] >> pl.shift;
                    bd[e] = pl.dbase + (p << pl.shift) + (int)(k[e] & ((1u << pl.shift) - 1u));
                }
        }
        const uint32_t px[2] = {px2.x, px2.y};
        int v[2];
#pragma unroll
        for (int e = 0; e < 2; e++) {
            const int xe = x + e;
            const bool skip = !yin || (px[e] & 0xFFFFFFu) == 0;
            const int l = min(xe, g.half), rr = min(g.w2 - xe - 1, g.half), u = min(y, g.half), dn = min(g.h2 - y - 1, g.half);
            const int n = (l + rr) * (u + dn);
            if (skip) v[e] = 0;
            else if (n == 0) v[e] = 0 - xe;
            else if (bc[e] != 0xFFFFFFFFu) v[e] = bd[e];
            else v[e] = 0 - xe;
            if (v[e] < -32768 || v[e] > 32767) { bad = 1; v[e] = 0; }
        }
        *reinterpret_cast<short2*>(win16 + i) = make_short2((short)v[0], (short)v[1]);
    }
    if (bad) *status = SR_E_BADARG;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(counter, 1u) == gridDim.x - 1) {   // last block: every block's stores are fenced
            *counter = 0;
            __threadfence_system();
            if (*(volatile int*)status == SR_OK) flags[rank] = seq;
        }
    }
}

// Phase A of the smoothing pass: bit planes a0,a1,a2 (ballot words) and the alternative value.
__global__ void smooth_prepare_kernel(SearchGeom g, int from_direct, int have_c0, double s, Planes best,
                                      const unsigned long long* __restrict__ c0, const unsigned long long* __restrict__ cr,
                                      const int* __restrict__ dr, const int* __restrict__ nwin,
                                      const uint32_t* __restrict__ Rp, int pitchR, uint32_t* __restrict__ bits,
                                      int words, double* __restrict__ alt) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;  // blockDim.x multiple of 32, grid covers words*32
    int y = g.row0 + blockIdx.y;
    bool a0 = false, a1 = false, a2 = false;
    if (x < g.w2) {
        size_t i = (size_t)y * g.w2 + x;
        PixRec r = load_rec(from_direct, have_c0, i, x, y, best, c0, cr, dr, nwin, Rp, pitchR, g);
        double v;  // value when d = 0 does not win
        if (r.skip) { v = 0.0; }
        else if (r.n == 0 || r.cr == SR_NONE64) v = (r.n != 0 && r.c0 != SR_NONE64) ? 0.0 : (double)(0 - x);
        else v = (double)r.dr;
        if (!r.skip && r.n != 0 && r.c0 != SR_NONE64 && r.cr != SR_NONE64) {
            double n = (double)r.n;
            double d0 = __ddiv_rn(sqrt((double)r.c0), n);
            double dd = __ddiv_rn(sqrt((double)r.cr), n);
            double d1 = __dmul_rn(d0, s), d2 = __dmul_rn(d1, s);
            // reference order: d=0 sets best unless NaN; the rest replaces it only if dd < best
            a0 = !(dd < d0); a1 = !(dd < d1); a2 = !(dd < d2);
            // a NaN d0*s^k (s = NaN/inf corner) never becomes best: then the rest wins
            if (d0 != d0) a0 = false;
            if (d1 != d1) a1 = false;
            if (d2 != d2) a2 = false;
        }
        if (v == 0.0) a0 = a1 = a2 = true;  // the stored value is 0 either way
        alt[i] = v;
    }
    uint32_t w0 = __ballot_sync(0xffffffffu, a0), w1 = __ballot_sync(0xffffffffu, a1), w2 = __ballot_sync(0xffffffffu, a2);
    if ((threadIdx.x & 31) == 0) {
        int wi = x >> 5;
        if (wi < words) {
            size_t o = ((size_t)y * words + wi) * 3;     // planes are laid out by absolute output row (row chunks share them)
            bits[o] = w0; bits[o + 1] = w1; bits[o + 2] = w2;
        }
    }
}

// carry chain inside one 32-bit word: z_x = cin_x ? F1_x : F0_x with cin_x = z_{x-1}.
__device__ __forceinline__ uint32_t chain_word(uint32_t F0, uint32_t F1, uint32_t cin, bool monotone) {
    if (monotone) {
        // F0 <= F1 bitwise: generate = F0, propagate = F1 & ~F0 -> binary adder carries
        unsigned long long sum = (unsigned long long)F0 + (unsigned long long)F1 + cin;
        uint32_t carry_in = (uint32_t)sum ^ F0 ^ F1;  // carry into each bit
        return (F0 & F1) | ((F0 ^ F1) & carry_in);    // carry out of each bit
    }
    uint32_t z = 0, c = cin;
    for (int b = 0; b < 32; b++) {
        c = c ? (F1 >> b) & 1u : (F0 >> b) & 1u;
        z |= c << b;
    }
    return z;
}

// Phase B: one CTA walks the rows; thread t owns word t (+ k*blockDim) of every row.
// Requires row0 == 0 semantics: the row above the first computed row is taken from zprev
// (all-ones = "stored 0" never applies at y == 0 since the rule needs y >= 1).
#define SMOOTH_THREADS 1024
__global__ void __launch_bounds__(SMOOTH_THREADS) smooth_raster_kernel(int rows, int words, int first_row_has_up,
                                                                       const uint32_t* __restrict__ bits,
                                                                       uint32_t* __restrict__ zbits, int monotone) {
    __shared__ uint32_t sg[SMOOTH_THREADS / 32], sp[SMOOTH_THREADS / 32], scin[SMOOTH_THREADS / 32];
    __shared__ uint32_t chunk_carry;
    int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    int nchunks = (words + SMOOTH_THREADS - 1) / SMOOTH_THREADS;
    for (int y = 0; y < rows; y++) {
        if (t == 0) chunk_carry = 0;  // x == 0 has no left neighbour: treated as "not zero"
        __syncthreads();
        for (int ch = 0; ch < nchunks; ch++) {
            int wi = ch * SMOOTH_THREADS + t;
            uint32_t F0 = 0, F1 = 0;
            if (wi < words) {
                const uint32_t* b = bits + ((size_t)y * words + wi) * 3;
                uint32_t a0 = b[0], a1 = b[1], a2 = b[2];
                // first_row_has_up: zbits points at a later row chunk, the row above its first row is the previous chunk's last
                uint32_t up = (y > 0 || first_row_has_up) ? zbits[((ptrdiff_t)y - 1) * words + wi] : 0u;
                F0 = (up & a1) | (~up & a0);
                F1 = (up & a2) | (~up & a1);
            }
            // word summary: carry out (bit 31 of z) for cin = 0 / 1
            uint32_t g0 = chain_word(F0, F1, 0, monotone) >> 31, g1 = chain_word(F0, F1, 1, monotone) >> 31;
            uint32_t G = __ballot_sync(0xffffffffu, g0), P = __ballot_sync(0xffffffffu, g1);  // per-warp 32-word functions
            // warp summary over its 32 words (functions compose like words of F0/F1 bits)
            if (lane == 0) {
                sg[warp] = chain_word(G, P, 0, monotone) >> 31;
                sp[warp] = chain_word(G, P, 1, monotone) >> 31;
            }
            __syncthreads();
            if (warp == 0) {
                uint32_t G2 = __ballot_sync(0xffffffffu, sg[lane]), P2 = __ballot_sync(0xffffffffu, sp[lane]);
                uint32_t zz = chain_word(G2, P2, chunk_carry, monotone);  // carry out of each warp
                uint32_t cin_w = (zz << 1) | chunk_carry;                  // carry into each warp
                scin[lane] = (cin_w >> lane) & 1u;
                __syncwarp();
                if (lane == 0) chunk_carry = zz >> 31;
            }
            __syncthreads();
            uint32_t zwarp = chain_word(G, P, scin[warp], monotone);  // carry out of each word of this warp
            uint32_t cin_words = (zwarp << 1) | scin[warp];
            uint32_t cin = (cin_words >> lane) & 1u;
            if (wi < words) zbits[(size_t)y * words + wi] = chain_word(F0, F1, cin, monotone);
            __syncthreads();
        }
        __threadfence_block();
    }
}

// Phase B, fast form: ONE warp walks the rows (rows are strictly sequential, and within a row the only
// cross-word dependency is a one-bit carry).  Lane l owns the KW consecutive words l*KW .. l*KW+KW-1 of
// every row, so the "row above" bits are its own registers; the carry crosses lanes through two ballots
// (per-lane transfer function: carry-out for carry-in 0 / 1) and one adder-style chain on the ballot
// words.  The a0/a1/a2 bit planes of the coming rows are prefetched into a shared-memory ring with
// 4-byte cp.async (they do not depend on earlier rows), which hides the L2/HBM latency that dominated
// the one-CTA version (~1.1 us/row -> ~0.1 us/row).
#define RASTER_PF 8
template <int KW>
__global__ void __launch_bounds__(32) smooth_raster_warp_kernel(int rows, int words, const uint32_t* __restrict__ bits,
                                                                uint32_t* __restrict__ zbits, int monotone, int first_row_has_up) {
    extern __shared__ uint32_t ring[];   // RASTER_PF x (32*KW*3) words
    const int lane = threadIdx.x;
    const int w0 = lane * KW;
    const int row_words = 32 * KW * 3;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(ring);
    auto prefetch = [&](int y) {
        if (y < rows) {
            uint32_t dst = sbase + 4u * ((y % RASTER_PF) * row_words + w0 * 3);
            const uint32_t* src = bits + ((size_t)y * words + w0) * 3;
#pragma unroll
            for (int i = 0; i < 3 * KW; i++)
                if (w0 + i / 3 < words) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(dst + 4u * i), "l"(src + i) : "memory");
        }
        asm volatile("cp.async.commit_group;\n" ::: "memory");
    };
#pragma unroll 1
    for (int y = 0; y < RASTER_PF - 1; y++) prefetch(y);
    uint32_t up[KW];
#pragma unroll
    for (int k = 0; k < KW; k++)   // y == 0 has no row above (the rule needs y >= 1); a later row chunk starts from the row before it
        up[k] = (first_row_has_up && w0 + k < words) ? zbits[(ptrdiff_t)(w0 + k) - words] : 0u;
#pragma unroll 1
    for (int y = 0; y < rows; y++) {
        prefetch(y + RASTER_PF - 1);
        asm volatile("cp.async.wait_group %0;\n" ::"n"(RASTER_PF - 1) : "memory");
        const uint32_t* r = ring + (y % RASTER_PF) * row_words + w0 * 3;
        uint32_t F0[KW], F1[KW];
#pragma unroll
        for (int k = 0; k < KW; k++) {
            uint32_t a0 = 0, a1 = 0, a2 = 0;
            if (w0 + k < words) { a0 = r[3 * k]; a1 = r[3 * k + 1]; a2 = r[3 * k + 2]; }
            F0[k] = (up[k] & a1) | (~up[k] & a0);
            F1[k] = (up[k] & a2) | (~up[k] & a1);
        }
        // this lane's KW words resolved for both possible carries into the lane; their last carries are the
        // lane's transfer function
        uint32_t z0[KW], z1[KW];
        uint32_t c0 = 0, c1 = 1;
#pragma unroll
        for (int k = 0; k < KW; k++) {
            z0[k] = chain_word(F0[k], F1[k], c0, monotone); c0 = z0[k] >> 31;
            z1[k] = chain_word(F0[k], F1[k], c1, monotone); c1 = z1[k] >> 31;
        }
        uint32_t G = __ballot_sync(0xffffffffu, c0), P = __ballot_sync(0xffffffffu, c1);
        uint32_t zz = chain_word(G, P, 0u, monotone);         // carry out of every lane (x == 0 has carry-in 0)
        const bool cin = ((zz << 1) >> lane) & 1u;            // carry into this lane: picks one of the two solutions
#pragma unroll
        for (int k = 0; k < KW; k++) {
            uint32_t z = cin ? z1[k] : z0[k];
            up[k] = z;
            if (w0 + k < words) zbits[(size_t)y * words + w0 + k] = z;
        }
    }
}

// The same walk for the monotone case (0 <= s <= 1, the way the reference is called: a0 <= a1 <= a2 bitwise), trimmed for
// the one thing that bounds it -- the dependent chain from one row's result to the next (300 -> ~150 cycles per row):
//  * the lane's KW words are ONE multi-word addition per possible carry-in (add-with-carry from word to word instead of
//    extracting bit 31 and starting a new 64-bit addition per word): z = carry-out of every bit of F0 + F1 + cin;
//  * F0 = a0 ^ (up & (a0 ^ a1)), F1 = a1 ^ (up & (a1 ^ a2)): one LOP3 per word on the chain;
//  * no shared-memory ring: the planes of the next 16 / KW rows are ordinary loads into registers, issued while the
//    current rows are resolved; results through a running pointer; words beyond the row are zero (all-kill: they
//    neither generate nor propagate).
template <int KW>
__global__ void __launch_bounds__(32) smooth_raster_mono_kernel(int rows, int words, const uint32_t* __restrict__ bits,
                                                                uint32_t* __restrict__ zbits, int first_row_has_up) {
    // The planes of the next U rows are loaded into registers (plain loads, nothing volatile: the compiler is free to
    // spread them over the stalls of the chain) while the current U rows are resolved.
    constexpr int U = 16 / KW;
    const int lane = threadIdx.x;
    const int w0 = lane * KW;
    int nvalid = words - w0;                          // words of this lane inside the row; the others are all-kill (zero)
    nvalid = nvalid < 0 ? 0 : (nvalid > KW ? KW : nvalid);
    const size_t rowstride = (size_t)words * 3;
    const uint32_t* src = bits + (size_t)w0 * 3;
    uint32_t cur[U][KW * 3], nxt[U][KW * 3];
    auto load_block = [&](uint32_t (&dst)[U][KW * 3], int y0) {
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int i = 0; i < KW * 3; i++) dst[u][i] = (y0 + u < rows && i < 3 * nvalid) ? __ldg(src + (size_t)u * rowstride + i) : 0u;
        src += (size_t)U * rowstride;
    };
    load_block(cur, 0);
    uint32_t up[KW];
#pragma unroll
    for (int k = 0; k < KW; k++)   // y == 0 has no row above (the rule needs y >= 1); a later row chunk starts from the row before it
        up[k] = (first_row_has_up && k < nvalid) ? zbits[(ptrdiff_t)(w0 + k) - words] : 0u;
    uint32_t* zo = zbits + w0;
#pragma unroll 1
    for (int yb = 0; yb < rows; yb += U) {
        load_block(nxt, yb + U);
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (yb + u < rows) {                      // (uniform)
                uint32_t F0[KW], F1[KW], s0[KW], s1[KW];
                uint32_t c0 = 0, c1 = 1;
#pragma unroll
                for (int k = 0; k < KW; k++) {
                    const uint32_t a0 = cur[u][3 * k], a1 = cur[u][3 * k + 1], a2 = cur[u][3 * k + 2];
                    F0[k] = a0 ^ (up[k] & (a0 ^ a1));
                    F1[k] = a1 ^ (up[k] & (a1 ^ a2));
                    const unsigned long long t0 = (unsigned long long)F0[k] + F1[k] + c0, t1 = (unsigned long long)F0[k] + F1[k] + c1;
                    s0[k] = (uint32_t)t0; c0 = (uint32_t)(t0 >> 32);
                    s1[k] = (uint32_t)t1; c1 = (uint32_t)(t1 >> 32);
                }
                const uint32_t G = __ballot_sync(0xffffffffu, c0), P = __ballot_sync(0xffffffffu, c1);
                const unsigned long long sum = (unsigned long long)G + (unsigned long long)P;      // lane x == 0 has carry-in 0
                const uint32_t zz = (G & P) | ((G ^ P) & ((uint32_t)sum ^ G ^ P));                // carry out of every lane
                const bool cin = ((zz << 1) >> lane) & 1u;                                         // carry into this lane
#pragma unroll
                for (int k = 0; k < KW; k++) {
                    const uint32_t sm = cin ? s1[k] : s0[k];
                    const uint32_t z = (F0[k] & F1[k]) | ((F0[k] ^ F1[k]) & (sm ^ F0[k] ^ F1[k]));   // carry out of every bit
                    up[k] = z;
                    if (k < nvalid) zo[k] = z;
                }
                zo += words;
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int i = 0; i < KW * 3; i++) cur[u][i] = nxt[u][i];
    }
}

template <int KW>
static void launch_raster_warp(cudaStream_t st, int rows, int words, const uint32_t* bits, uint32_t* zb, int monotone, int has_up) {
    const size_t smem = RASTER_PF * 32 * KW * 3 * sizeof(uint32_t);
    static const char* old_env = getenv("SR_RASTER_GENERIC");   // test hook: 1 = the generic kernel for monotone factors too
    if (monotone && !(old_env && atoi(old_env) != 0)) smooth_raster_mono_kernel<KW><<<1, 32, 0, st>>>(rows, words, bits, zb, has_up);
    else smooth_raster_warp_kernel<KW><<<1, 32, smem, st>>>(rows, words, bits, zb, monotone, has_up);
}

__global__ void smooth_expand_kernel(SearchGeom g, int words, const uint32_t* __restrict__ zbits,
                                     const double* __restrict__ alt, double* __restrict__ out) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = g.row0 + blockIdx.y;
    if (x >= g.w2 || y >= g.row1) return;
    size_t i = (size_t)y * g.w2 + x;
    uint32_t z = (zbits[(size_t)y * words + (x >> 5)] >> (x & 31)) & 1u;
    out[i] = z ? 0.0 : alt[i];
}

int sr_launch_resolve_right(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, double smooth, int from_direct, int have_c0) {
    int rows = g.row1 - g.row0;
    if (rows <= 0) return SR_OK;
    Planes pl{pair->planes, pair->npx_cap, pair->npass, pair->plane_dbase, pair->key_shift};
    dim3 grid(cdiv(g.w2, 128), rows);
    bool plain = smooth == 1.0 || !have_c0;
    pair->wire_done = false;
    if (pair->wire_seq && plain && !from_direct && have_c0 == 0 && (g.w2 & 1) == 0 && (pl.stride & 1) == 0 && (pair->R.pitch & 1) == 0 &&
        (((uintptr_t)pl.planes | (uintptr_t)pair->R.px) & 7) == 0 && pair->peer_win && (pair->peer_mode == 2 || pair->peer_mode == 3)) {
        int rc = sr_peer_wait_ack(ctx, pair, pair->wire_seq);
        if (rc != SR_OK) return rc;
        int wgrid = cdiv(g.w2, 256) * rows;
        if (wgrid > ctx->sms * 16) wgrid = ctx->sms * 16;
        resolve_wire16_kernel<<<wgrid, 128, 0, ctx->stream>>>(g, pl, pair->R.px, pair->R.pitch, sr_peer_window16(pair),
                                                                                   sr_peer_flags(pair), pair->wire_rank, pair->wire_seq,
                                                                                   pair->peer_counter, pair->peer_status_d);
        SR_LAUNCHED(ctx);
        pair->wire_done = true;
        return SR_OK;
    }
    if (plain && !from_direct && have_c0 != 1 && (g.w2 & 1) == 0 && (pl.stride & 1) == 0 && (pair->R.pitch & 1) == 0 &&
        (((uintptr_t)pl.planes | (uintptr_t)pair->R.px | (uintptr_t)pair->c0) & 7) == 0 && ((uintptr_t)pair->out & 15) == 0) {
        resolve_plain2_kernel<<<dim3(cdiv(g.w2, 256), rows), 128, 0, ctx->stream>>>(g, have_c0, pl, (const uint32_t*)pair->c0,
                                                                                   pair->R.px, pair->R.pitch, pair->out);
        SR_LAUNCHED(ctx);
        return SR_OK;
    }
    if (plain) {
        resolve_plain_kernel<<<grid, 128, 0, ctx->stream>>>(g, from_direct, have_c0, pl, pair->c0, pair->cr, pair->dr,
                                                            pair->nwin, pair->R.px, pair->R.pitch, pair->out);
        SR_LAUNCHED(ctx);
        return SR_OK;
    }
    int words = cdiv(g.w2, 32);
    if (!pair->bits) {
        size_t maxwords = (size_t)cdiv(pair->w1 > pair->w2 ? pair->w1 : pair->w2, 32) * (pair->h1 > pair->h2 ? pair->h1 : pair->h2);
        SR_CUDA(ctx, cudaMalloc((void**)&pair->bits, maxwords * 4 * sizeof(uint32_t)));
        SR_CUDA(ctx, cudaMalloc((void**)&pair->alt, pair->npx_cap * sizeof(double)));
    }
    // planes and z bits by absolute output row: the rows of one call (a row chunk of the one-shot pipeline) are a slice
    const int oh_total = pair->oh > 0 ? pair->oh : g.h2;
    uint32_t* zb_all = pair->bits + (size_t)oh_total * words * 3;
    const uint32_t* bits_c = pair->bits + (size_t)g.row0 * words * 3;
    uint32_t* zb = zb_all + (size_t)g.row0 * words;
    const int has_up = g.row0 > 0 ? 1 : 0;             // the chunk above has been resolved (stream order / ctx->raster_wait)
    dim3 gridw(words, rows);
    smooth_prepare_kernel<<<dim3(cdiv(words * 32, 128), rows), 128, 0, ctx->stream>>>(
        g, from_direct, have_c0, smooth, pl, pair->c0, pair->cr, pair->dr, pair->nwin, pair->R.px, pair->R.pitch, pair->bits,
        words, pair->alt);
    SR_LAUNCHED(ctx);
    // s <= 1 makes a0 <= a1 <= a2 (more multiplications -> smaller cost): adder-style carries
    int monotone = (smooth >= 0.0 && smooth <= 1.0) ? 1 : 0;
    if (ctx->raster_wait) SR_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->raster_wait, 0));   // previous chunk's raster (other stream)
    if (words <= 32) launch_raster_warp<1>(ctx->stream, rows, words, bits_c, zb, monotone, has_up);
    else if (words <= 64) launch_raster_warp<2>(ctx->stream, rows, words, bits_c, zb, monotone, has_up);
    else if (words <= 128) launch_raster_warp<4>(ctx->stream, rows, words, bits_c, zb, monotone, has_up);
    else if (words <= 256) launch_raster_warp<8>(ctx->stream, rows, words, bits_c, zb, monotone, has_up);
    else smooth_raster_kernel<<<1, SMOOTH_THREADS, 0, ctx->stream>>>(rows, words, has_up, bits_c, zb, monotone);
    SR_LAUNCHED(ctx);
    smooth_expand_kernel<<<grid, 128, 0, ctx->stream>>>(g, words, zb_all, pair->alt, pair->out);
    SR_LAUNCHED(ctx);
    return SR_OK;
}
