// recon.cu -- Reconstruction kernels (src/Reconstruction/reconstruction.cpp) on device buffers.
// Every kernel is O(1) work per pixel and HBM-bound; arithmetic follows the reference's operation
// order exactly (no FMA contraction: the library is compiled with --fmad=false; IEEE division
// and square root are nvcc defaults), so results are bit-identical to the CPU code.
#include <cstdlib>
#include "sr_internal.cuh"
#include "recon.cuh"

// cv::borderInterpolate(p, len, BORDER_REFLECT_101)
__host__ __device__ static inline int reflect101(int p, int len) {
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        if (p < 0) p = -p; else p = 2 * (len - 1) - p;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

// ---- cv::blur(32F, Size(k,k)) (reconstruction.cpp:8) ----------------------------------------------
// OpenCV's box filter = RowSum<float,double> (first window summed left to right, then
// s += new - old) followed by ColumnSum<double,float> (running column sums over k rows,
// out = (float)(sum * 1/(k*k))).  Both are sequential double-precision recurrences; to stay
// bit-exact for arbitrary float input the recurrences are kept and parallelised across the
// other dimension: one thread per row for the horizontal pass, one per column for the vertical.
__global__ void blur_rows_kernel(const float* __restrict__ src, int w, int h, int k, double* __restrict__ rows) {
    int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= h) return;
    const float* s = src + (size_t)y * w;
    double* o = rows + (size_t)y * w;
    int a = k / 2;
    double acc = 0;
    for (int i = 0; i < k; i++) acc += (double)s[reflect101(i - a, w)];
    o[0] = acc;
    for (int x = 1; x < w; x++) {
        acc += (double)s[reflect101(x + k - 1 - a, w)] - (double)s[reflect101(x - 1 - a, w)];
        o[x] = acc;
    }
}

__global__ void blur_cols_kernel(const double* __restrict__ rows, int w, int h, int k, float* __restrict__ dst) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= w) return;
    int a = k / 2;
    double scale = 1.0 / ((double)k * (double)k);
    double sum = 0;
    for (int i = 0; i < k - 1; i++) sum += rows[(size_t)reflect101(i - a, h) * w + x];
    for (int y = 0; y < h; y++) {
        double s0 = sum + rows[(size_t)reflect101(y - a + k - 1, h) * w + x];
        dst[(size_t)y * w + x] = (float)(s0 * scale);
        sum = s0 - rows[(size_t)reflect101(y - a, h) * w + x];
    }
}

// ---- the same blur, order-free --------------------------------------------------------------------
// The recurrences above are sequential (one thread per row / column: latency-bound, 1.1 ms for a
// 1436x992 map with k = 500).  When every partial sum is EXACTLY representable in fp64 the order of
// the additions does not matter and the window sums can come from parallel prefix sums with the
// identical result.  blur_exact_check_kernel proves that on the device: all nonzero inputs are multiples
// of 2^q with |v| < 2^E, any signed sum of N of them is a multiple of 2^q below N*2^E, exact iff
// E - q + ceil(log2 N) <= 53, with N = the longest prefix any kernel below forms: (h + k - 1) * k inputs in a column
// prefix (>= the w + k - 1 of a row prefix for every map at least as tall as it is wide / k, the host takes the
// maximum).  (Always true for the reference's 8-bit hand-off, main.cpp:41-50, and for sub-pixel maps within a 2^10
// dynamic range.)  Both kernel families are launched; each returns at once when the flag says the other one
// applies -- no host round trip.
__device__ __forceinline__ void exact_acc(uint32_t b, int& E, int& q, int& bad) {
    uint32_t ex = (b >> 23) & 0xFFu, man = b & 0x7FFFFFu;
    if (ex == 0xFFu || (ex == 0 && man != 0)) { bad = 1; return; }   // inf / NaN / denormal: keep the recurrence
    if (ex == 0) return;                                              // zero
    uint32_t m = man | 0x800000u;
    int e = (int)ex - 127;
    E = max(E, e + 1);
    q = min(q, e - 23 + (__ffs(m) - 1));
}
// nvec = number of 16-byte words read with vector loads (two per thread in flight); the tail is read by element
__global__ void blur_exact_check_kernel(const float* __restrict__ src, size_t n, size_t nvec, int* __restrict__ flag /* [emax+1, qmin, bad] */) {
    const size_t stride = (size_t)gridDim.x * blockDim.x, tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    int E = -1000, q = 1000, bad = 0;
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    for (size_t i = tid; i < nvec; i += 2 * stride) {
        uint4 a = s4[i];
        uint4 b = i + stride < nvec ? s4[i + stride] : make_uint4(0u, 0u, 0u, 0u);   // zeros are neutral
        exact_acc(a.x, E, q, bad); exact_acc(a.y, E, q, bad); exact_acc(a.z, E, q, bad); exact_acc(a.w, E, q, bad);
        exact_acc(b.x, E, q, bad); exact_acc(b.y, E, q, bad); exact_acc(b.z, E, q, bad); exact_acc(b.w, E, q, bad);
    }
    for (size_t i = 4 * nvec + tid; i < n; i += stride) exact_acc(__float_as_uint(src[i]), E, q, bad);
    for (int off = 16; off > 0; off >>= 1) {
        E = max(E, __shfl_xor_sync(0xffffffffu, E, off));
        q = min(q, __shfl_xor_sync(0xffffffffu, q, off));
        bad |= __shfl_xor_sync(0xffffffffu, bad, off);
    }
    // one set of atomics per CTA (blockDim.x <= 1024)
    __shared__ int sE[32], sq[32], sbad[32];
    const int warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
    if ((threadIdx.x & 31) == 0) { sE[warp] = E; sq[warp] = q; sbad[warp] = bad; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < nwarp; i++) { E = max(E, sE[i]); q = min(q, sq[i]); bad |= sbad[i]; }
        if (E > -1000) atomicMax(flag, E);
        if (q < 1000) atomicMin(flag + 1, q);
        if (bad) atomicOr(flag + 2, 1);
    }
}
// logn = ceil(log2 N) of the comment above (host: blur_logn)
__device__ __forceinline__ bool blur_is_exact(const int* flag, int logn) {
    if (flag[2]) return false;
    if (flag[0] <= -1000) return true;                                      // all zeros
    return flag[0] - flag[1] + logn <= 53;
}
static int blur_logn(int w, int h, int k) {
    long long nn = (long long)(h + k - 1) * k, rw = (long long)w + k - 1;
    if (rw > nn) nn = rw;
    if (2LL * k * k > nn) nn = 2LL * k * k;
    int logn = 1;
    while ((1LL << logn) < nn) logn++;
    return logn;
}

// One CTA per row: the reflect-101 extended row is staged in shared memory with coalesced loads, turned into
// inclusive prefix sums there (per-thread chunks of odd length: conflict-free 64-bit accesses, then a scan of
// the chunk totals), and rowsum[x] = PE[x + k - 1] - PE[x - 1].
#define BLUR_T 512
__global__ void __launch_bounds__(BLUR_T) blur_rows_fast_kernel(const float* __restrict__ src, int w, int h, int k, int logn,
                                                                double* __restrict__ rows, const int* __restrict__ flag) {
    extern __shared__ double pe[];               // w + k - 1 prefix values
    __shared__ double wsum[BLUR_T / 32];
    if (!blur_is_exact(flag, logn)) return;
    const int y = blockIdx.x, t = threadIdx.x, a = k / 2, we = w + k - 1;
    const float* s = src + (size_t)y * w;
    for (int j = t; j < we; j += BLUR_T) pe[j] = (double)s[reflect101(j - a, w)];
    __syncthreads();
    const int C = ((we + BLUR_T - 1) / BLUR_T) | 1, j0 = min(t * C, we), j1 = min(j0 + C, we);
    double acc = 0;
    for (int j = j0; j < j1; j++) { acc += pe[j]; pe[j] = acc; }
    // exclusive scan of the per-thread totals
    double v = acc;
    for (int off = 1; off < 32; off <<= 1) { double u = __shfl_up_sync(0xffffffffu, v, off); if ((t & 31) >= off) v += u; }
    if ((t & 31) == 31) wsum[t >> 5] = v;
    __syncthreads();
    double base = v - acc;
    for (int i = 0; i < (t >> 5); i++) base += wsum[i];
    for (int j = j0; j < j1; j++) pe[j] += base;
    __syncthreads();
    double* o = rows + (size_t)y * w;
    for (int x = t; x < w; x += BLUR_T) o[x] = pe[x + k - 1] - (x > 0 ? pe[x - 1] : 0.0);
}

// Columns: segment sums -> per-column exclusive scan of the segments -> prefix array PE (h + k - 1 rows)
// -> dst = (float)((PE[y + k - 1] - PE[y - 1]) * scale).  Threads run along x (coalesced).
#define BLUR_NSEG 64
__global__ void blur_cols_seg_kernel(const double* __restrict__ rows, int w, int h, int k, int logn, double* __restrict__ seg,
                                     const int* __restrict__ flag) {
    if (!blur_is_exact(flag, logn)) return;
    int x = blockIdx.x * blockDim.x + threadIdx.x, sgi = blockIdx.y;
    if (x >= w) return;
    const int a = k / 2, he = h + k - 1, SL = (he + BLUR_NSEG - 1) / BLUR_NSEG;
    const int j0 = sgi * SL, j1 = min(j0 + SL, he);
    // four loads in flight; any order of the additions gives the same (exact) sum
    double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    int j = j0;
    for (; j + 3 < j1; j += 4) {
        double r0 = rows[(size_t)reflect101(j - a, h) * w + x], r1 = rows[(size_t)reflect101(j + 1 - a, h) * w + x];
        double r2 = rows[(size_t)reflect101(j + 2 - a, h) * w + x], r3 = rows[(size_t)reflect101(j + 3 - a, h) * w + x];
        a0 += r0; a1 += r1; a2 += r2; a3 += r3;
    }
    for (; j < j1; j++) a0 += rows[(size_t)reflect101(j - a, h) * w + x];
    seg[(size_t)sgi * w + x] = (a0 + a1) + (a2 + a3);
}
__global__ void blur_cols_prefix_kernel(const double* __restrict__ rows, int w, int h, int k, int logn,
                                        const double* __restrict__ seg, double* __restrict__ pe, const int* __restrict__ flag) {
    if (!blur_is_exact(flag, logn)) return;
    int x = blockIdx.x * blockDim.x + threadIdx.x, sgi = blockIdx.y;
    if (x >= w) return;
    const int a = k / 2, he = h + k - 1, SL = (he + BLUR_NSEG - 1) / BLUR_NSEG;
    const int j0 = sgi * SL, j1 = min(j0 + SL, he);
    double acc = 0;
    for (int i = 0; i < sgi; i++) acc += seg[(size_t)i * w + x];
    int j = j0;
    for (; j + 3 < j1; j += 4) {
        double r0 = rows[(size_t)reflect101(j - a, h) * w + x], r1 = rows[(size_t)reflect101(j + 1 - a, h) * w + x];
        double r2 = rows[(size_t)reflect101(j + 2 - a, h) * w + x], r3 = rows[(size_t)reflect101(j + 3 - a, h) * w + x];
        acc += r0; pe[(size_t)j * w + x] = acc;
        acc += r1; pe[(size_t)(j + 1) * w + x] = acc;
        acc += r2; pe[(size_t)(j + 2) * w + x] = acc;
        acc += r3; pe[(size_t)(j + 3) * w + x] = acc;
    }
    for (; j < j1; j++) { acc += rows[(size_t)reflect101(j - a, h) * w + x]; pe[(size_t)j * w + x] = acc; }
}
// two columns per thread: 16-byte loads of both prefix rows when the row pitch allows (w even)
__global__ void blur_cols_out_kernel(const double* __restrict__ pe, int w, int h, int k, int logn, float* __restrict__ dst,
                                     const int* __restrict__ flag) {
    if (!blur_is_exact(flag, logn)) return;
    int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x), y = blockIdx.y;
    if (x >= w) return;
    const double scale = 1.0 / ((double)k * (double)k);
    const double* hi = pe + (size_t)(y + k - 1) * w + x;
    const double* lo = pe + (size_t)(y > 0 ? y - 1 : 0) * w + x;
    float* o = dst + (size_t)y * w + x;
    if ((w & 1) == 0) {
        double2 ph = *reinterpret_cast<const double2*>(hi);
        double2 pl = y > 0 ? *reinterpret_cast<const double2*>(lo) : make_double2(0.0, 0.0);
        *reinterpret_cast<float2*>(o) = make_float2((float)((ph.x - pl.x) * scale), (float)((ph.y - pl.y) * scale));
    } else {
        o[0] = (float)((hi[0] - (y > 0 ? lo[0] : 0.0)) * scale);
        if (x + 1 < w) o[1] = (float)((hi[1] - (y > 0 ? lo[1] : 0.0)) * scale);
    }
}
// the sequential kernels, skipped when the order-free path applies
__global__ void blur_rows_seq_kernel(const float* __restrict__ src, int w, int h, int k, int logn, double* __restrict__ rows,
                                     const int* __restrict__ flag) {
    if (blur_is_exact(flag, logn)) return;
    int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= h) return;
    const float* s = src + (size_t)y * w;
    double* o = rows + (size_t)y * w;
    int a = k / 2;
    double acc = 0;
    for (int i = 0; i < k; i++) acc += (double)s[reflect101(i - a, w)];
    o[0] = acc;
    for (int x = 1; x < w; x++) {
        acc += (double)s[reflect101(x + k - 1 - a, w)] - (double)s[reflect101(x - 1 - a, w)];
        o[x] = acc;
    }
}
__global__ void blur_cols_seq_kernel(const double* __restrict__ rows, int w, int h, int k, int logn, float* __restrict__ dst,
                                     const int* __restrict__ flag) {
    if (blur_is_exact(flag, logn)) return;
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= w) return;
    int a = k / 2;
    double scale = 1.0 / ((double)k * (double)k);
    double sum = 0;
    for (int i = 0; i < k - 1; i++) sum += rows[(size_t)reflect101(i - a, h) * w + x];
    for (int y = 0; y < h; y++) {
        double s0 = sum + rows[(size_t)reflect101(y - a + k - 1, h) * w + x];
        dst[(size_t)y * w + x] = (float)(s0 * scale);
        sum = s0 - rows[(size_t)reflect101(y - a, h) * w + x];
    }
}

int recon_box_blur(sr_ctx* ctx, const float* d_src, float* d_dst, int w, int h, int k, int known_u8) {
    double* rows = (double*)sr_scratch(ctx, 1, (size_t)w * h * sizeof(double));
    if (!rows) return SR_E_CUDA;
    static const char* slow_env = getenv("SR_BLUR_SEQUENTIAL");   // test hook: force the recurrence kernels
    const bool force_seq = slow_env && atoi(slow_env) != 0;
    const size_t smem = (size_t)(w + k - 1) * sizeof(double);
    if (force_seq || smem > 200 * 1024 || k < 1) {
        blur_rows_kernel<<<cdiv(h, 64), 64, 0, ctx->stream>>>(d_src, w, h, k, rows);
        SR_LAUNCHED(ctx);
        blur_cols_kernel<<<cdiv(w, 64), 64, 0, ctx->stream>>>(rows, w, h, k, d_dst);
        SR_LAUNCHED(ctx);
        return SR_OK;
    }
    const int he = h + k - 1;
    // scratch: prefix array (he x w) + segment sums (NSEG x w) + 4-int flag
    char* buf = (char*)sr_scratch(ctx, 15, ((size_t)he * w + (size_t)BLUR_NSEG * w) * sizeof(double) + 64);
    if (!buf) return SR_E_CUDA;
    double* pe = (double*)buf; double* seg = pe + (size_t)he * w; int* flag = (int*)(seg + (size_t)BLUR_NSEG * w);
    static const int flag_init[4] = {-1000, 1000, 0, 0};
    static const int flag_u8[4] = {8, 0, 0, 0};              // integers 0..255: below 2^8, multiples of 2^0
    SR_CUDA(ctx, cudaMemcpyAsync(flag, known_u8 ? flag_u8 : flag_init, sizeof flag_init, cudaMemcpyHostToDevice, ctx->stream));
    size_t n = (size_t)w * h;
    const int logn = blur_logn(w, h, k);
    if (!known_u8) {   // the 8-bit hand-off (quantize_kernel's output) needs no look at the data
        const size_t nvec = ((uintptr_t)d_src & 15) ? 0 : n / 4;
        blur_exact_check_kernel<<<ctx->sms * 8, 256, 0, ctx->stream>>>(d_src, n, nvec, flag);
        SR_LAUNCHED(ctx);
    }
    static DeviceOnce configured;
    if (configured.need(ctx->device)) {
        SR_CUDA(ctx, cudaFuncSetAttribute(blur_rows_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        configured.done(ctx->device);
    }
    blur_rows_fast_kernel<<<h, BLUR_T, smem, ctx->stream>>>(d_src, w, h, k, logn, rows, flag);
    SR_LAUNCHED(ctx);
    dim3 gs(cdiv(w, 128), BLUR_NSEG);
    blur_cols_seg_kernel<<<gs, 128, 0, ctx->stream>>>(rows, w, h, k, logn, seg, flag);
    SR_LAUNCHED(ctx);
    blur_cols_prefix_kernel<<<gs, 128, 0, ctx->stream>>>(rows, w, h, k, logn, seg, pe, flag);
    SR_LAUNCHED(ctx);
    blur_cols_out_kernel<<<dim3(cdiv(w, 256), h), 128, 0, ctx->stream>>>(pe, w, h, k, logn, d_dst, flag);
    SR_LAUNCHED(ctx);
    // inputs outside the proof (huge dynamic range, inf/NaN): the recurrences, exactly as OpenCV runs them
    blur_rows_seq_kernel<<<cdiv(h, 64), 64, 0, ctx->stream>>>(d_src, w, h, k, logn, rows, flag);
    SR_LAUNCHED(ctx);
    blur_cols_seq_kernel<<<cdiv(w, 64), 64, 0, ctx->stream>>>(rows, w, h, k, logn, d_dst, flag);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// reconstruction.cpp:11-17.  Four pixels per thread (16-byte loads of both maps when vec != 0).
__device__ __forceinline__ float outlier1(float d, float b, float thr_front, float thr_back) {
    return (d > thr_front * b || d < thr_back * b) ? b : d;
}
__global__ void outlier_replace_kernel(float* __restrict__ disp, const float* __restrict__ blurred, size_t n,
                                       float thr_front, float thr_back, int vec) {
    size_t i = 4 * ((size_t)blockIdx.x * blockDim.x + threadIdx.x);
    if (i >= n) return;
    if (vec && i + 3 < n) {
        float4 d = *reinterpret_cast<const float4*>(disp + i), b = *reinterpret_cast<const float4*>(blurred + i);
        d.x = outlier1(d.x, b.x, thr_front, thr_back); d.y = outlier1(d.y, b.y, thr_front, thr_back);
        d.z = outlier1(d.z, b.z, thr_front, thr_back); d.w = outlier1(d.w, b.w, thr_front, thr_back);
        *reinterpret_cast<float4*>(disp + i) = d;
    } else {
        for (size_t e = min(i + 4, n); i < e; i++) disp[i] = outlier1(disp[i], blurred[i], thr_front, thr_back);
    }
}

int recon_remove_outliers(sr_ctx* ctx, float* d_disp, int w, int h, int k, float thr_front, float thr_back, int known_u8) {
    float* blurred = (float*)sr_scratch(ctx, 2, (size_t)w * h * sizeof(float));
    if (!blurred) return SR_E_CUDA;
    int rc = recon_box_blur(ctx, d_disp, blurred, w, h, k, known_u8);
    if (rc != SR_OK) return rc;
    size_t n = (size_t)w * h;
    const int vec = (((uintptr_t)d_disp | (uintptr_t)blurred) & 15) == 0;
    outlier_replace_kernel<<<(unsigned)((n + 1023) / 1024), 256, 0, ctx->stream>>>(d_disp, blurred, n, thr_front, thr_back, vec);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// reconstruction.cpp:21-27
__global__ void scale_kernel(float* __restrict__ disp, size_t n, float f) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) disp[i] = f * disp[i];
}
int recon_scale(sr_ctx* ctx, float* d_disp, size_t n, float f) {
    scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(d_disp, n, f);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// Fused disparity -> depth -> back-projected vertex (reconstruction.cpp:30-43 and :169-196).
// depth = d == 0 ? -inf : (focal*baseline)/d;  x = (w*depth - cX*depth)/fX (two rounded products,
// a rounded subtract, a rounded divide).  Either output may be NULL.  bgr: tight 3*w pitch.
__global__ void depth_backproject_kernel(const float* __restrict__ disp, const float* __restrict__ depth_in,
                                         const uint8_t* __restrict__ bgr, size_t bgr_pitch, int w, int h, float focal,
                                         float baseline, float fx, float fy, float cx, float cy,
                                         float* __restrict__ depth_out, sr_vertex* __restrict__ vert) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y;
    if (x >= w) return;
    size_t i = (size_t)y * w + x;
    float depth;
    if (disp) {
        float d = disp[i];
        depth = d == 0.0f ? -INFINITY : focal * baseline / d;
        if (depth_out) depth_out[i] = depth;
    } else {
        depth = depth_in[i];
    }
    if (!vert) return;
    float4 pos;
    uint32_t rgba;
    if (depth == -INFINITY) {
        pos = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        rgba = 0;
    } else {
        float xc = ((float)x * depth - cx * depth) / fx;
        float yc = ((float)y * depth - cy * depth) / fy;
        pos = make_float4(xc, yc, depth, 1.0f);
        const uint8_t* p = bgr + (size_t)y * bgr_pitch + 3 * (size_t)x;
        rgba = (uint32_t)p[2] | ((uint32_t)p[1] << 8) | ((uint32_t)p[0] << 16) | 0xFF000000u;
    }
    // 32-byte vertex: two 16-byte stores (position; colour + zeroed padding)
    float4* o = reinterpret_cast<float4*>(vert + i);
    o[0] = pos;
    o[1] = make_float4(__uint_as_float(rgba), 0.f, 0.f, 0.f);
}

int recon_depth_backproject(sr_ctx* ctx, const float* d_disp, const float* d_depth_in, const uint8_t* d_bgr, size_t bgr_pitch,
                            int w, int h, float focal, float baseline, float fx, float fy, float cx, float cy,
                            float* d_depth_out, sr_vertex* d_vert) {
    dim3 grid(cdiv(w, 128), h);
    depth_backproject_kernel<<<grid, 128, 0, ctx->stream>>>(d_disp, d_depth_in, d_bgr, bgr_pitch, w, h, focal, baseline, fx, fy,
                                                            cx, cy, d_depth_out, d_vert);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// ---- WriteMesh face list (reconstruction.cpp:46-69, :92-113) ----------------------------------------
// Grid cells in row-major order, per cell the upper-left triangle (a,b,c) then the lower-right
// (b,d,c).  Validity is embarrassingly parallel; the reference's face ORDER is kept with an
// ordered stream compaction: per-block counts -> exclusive scan of block totals -> scatter.
// The reference compares sqrtf(s) > thr with s = dx^2 + dy^2 + dz^2 (:53-65).  sqrtf is correctly rounded and monotone,
// so there is one float S(thr) = max{ s : sqrtf(s) <= thr } and the test is s > S(thr), exactly -- no square root per
// edge (three per pixel in the fused kernel, which is issue-bound: 71 % issue-active at 43 % of the DRAM peak).  S(thr)
// is found on the host (sqrt_threshold below) from thr * thr by stepping neighbouring floats.
__device__ __forceinline__ float edge_sq(const float4& a, const float4& b) {
    float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z;
    return dx * dx + dy * dy + dz * dz;  // powf(v,2) == v*v; sums left to right; no FMA
}
__device__ __forceinline__ bool tri_valid(const float4& a, const float4& b, const float4& c, float thr_sq) {
    if (a.x == -INFINITY || b.x == -INFINITY || c.x == -INFINITY) return false;
    float l1 = edge_sq(a, b), l2 = edge_sq(a, c), l3 = edge_sq(b, c);
    return !(l1 > thr_sq || l2 > thr_sq || l3 > thr_sq);
}
static float sqrt_threshold(float thr) {
    if (thr != thr) return INFINITY;                   // l > NaN: never
    if (thr < 0.0f) return -INFINITY;                  // l >= 0 > thr: always (s is NaN for no edge of an accepted triangle)
    if (thr == INFINITY) return INFINITY;
    float c = thr * thr;
    while (sqrtf(c) > thr) c = nextafterf(c, -INFINITY);
    for (;;) {
        const float n = nextafterf(c, INFINITY);
        if (n == c || !(sqrtf(n) <= thr)) break;
        c = n;
    }
    return c;
}

#define FACE_THREADS 256
// Positions are read from the Vertex array (one 32-byte sector per vertex, 16 bytes of it used).  Recomputing them
// from the depth map inside the fused pipeline was measured and dropped: 33 MB instead of 271 MB of DRAM traffic per
// 4K map, but eight IEEE divisions per cell made the kernel ALU-bound (104 us against 64 us).
struct PosFromVertices {
    const sr_vertex* v;
    __device__ __forceinline__ float4 operator()(unsigned x, unsigned y, int w) const {
        return *reinterpret_cast<const float4*>(v + (size_t)y * w + x);
    }
};
// flags of cell `cell`: bit0 = upper-left triangle valid, bit1 = lower-right valid
template <class Pos>
__device__ __forceinline__ uint32_t cell_flags(const Pos& pos, int w, unsigned cell, float thr_sq) {
    unsigned cw = (unsigned)(w - 1);
    unsigned y = cell / cw, x = cell - y * cw;
    const float4 pa = pos(x, y, w), pb = pos(x, y + 1, w), pc = pos(x + 1, y, w), pd = pos(x + 1, y + 1, w);
    return (tri_valid(pa, pb, pc, thr_sq) ? 1u : 0u) | (tri_valid(pb, pd, pc, thr_sq) ? 2u : 0u);
}

template <class Pos>
__global__ void __launch_bounds__(FACE_THREADS) face_count_kernel(Pos pos, int w, float thr_sq, unsigned ncells,
                                                                  uint8_t* __restrict__ flags, uint32_t* __restrict__ block_total) {
    __shared__ uint32_t wsum[FACE_THREADS / 32];
    unsigned cell = blockIdx.x * FACE_THREADS + threadIdx.x;
    uint32_t f = cell < ncells ? cell_flags(pos, w, cell, thr_sq) : 0u;
    if (cell < ncells) flags[cell] = (uint8_t)f;
    uint32_t n = __popc(f);
    for (int o = 16; o; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int i = 0; i < FACE_THREADS / 32; i++) t += wsum[i];
        block_total[blockIdx.x] = t;
    }
}

// ---- fused: disparity -> depth -> vertex -> face validity (main.cpp:56-65 in one pass) --------------------
// depth_backproject_kernel plus the validity flags of WriteMesh, so that the 32-byte vertices are written once and
// never read back (face_count_kernel reads 271 MB of them per 4K map).  A CTA owns 127 columns x 16 rows of
// vertices and marches down its rows: every thread keeps its previous position in registers, the right-hand
// neighbour's comes through shared memory (double-buffered: one barrier per row); the 128th column and the 17th row
// are halo (computed, not stored).  The next row's disparity and colour are loaded before the barrier of the current
// one.  Positions are the values the plain kernel stores (same expressions, no FMA), so the flags equal
// cell_flags() on the stored vertices.
#define DBF_T 128
#define DBF_ROWS 16
// where the disparity of a pixel comes from: a float map, or -- the 8-bit hand-off with removeDisparityOutliers -- the
// u8 map and the int32 column prefixes of its row window sums (recon_u8_pipeline below): the blurred value and the
// outlier rule (reconstruction.cpp:8-17) are evaluated on the fly, the filtered map is stored only when asked for
struct DispFromMap {
    const float* disp;
    __device__ __forceinline__ float operator()(int x, int y, int w) const { return disp[(size_t)y * w + x]; }
};
struct DispFromU8Blur {
    const uint8_t* q8; const int* pe; int k; double scale; float thr_front, thr_back; float* disp_out;
    __device__ __forceinline__ float operator()(int x, int y, int w) const {
        const size_t i = (size_t)y * w + x;
        const int hi = pe[(size_t)(y + k - 1) * w + x], lo = y > 0 ? pe[(size_t)(y - 1) * w + x] : 0;
        const float b = (float)((double)(hi - lo) * scale);   // ColumnSum<double,float>: (float)(sum * 1/(k*k)), sum exact
        const float d = (float)q8[i];
        const float v = (d > thr_front * b || d < thr_back * b) ? b : d;
        if (disp_out) disp_out[i] = v;
        return v;
    }
};
template <class Disp>
__global__ void __launch_bounds__(DBF_T) depth_backproject_faces_kernel(
    Disp disp, const uint8_t* __restrict__ bgr, size_t bgr_pitch, int w, int h, float focal,
    float baseline, float fx, float fy, float cx, float cy, float thr /* S(thrMesh): squared-length threshold */, float* __restrict__ depth_out,
    sr_vertex* __restrict__ vert, uint8_t* __restrict__ flags) {
    __shared__ float4 spos[2][DBF_T];
    __shared__ float svert[2][DBF_T];
    const int t = threadIdx.x;
    const int x = blockIdx.x * (DBF_T - 1) + t;
    const int y0 = blockIdx.y * DBF_ROWS;
    const bool incol = x < w;
    const bool own = incol && t < DBF_T - 1;
    float dn = 0.f;
    uint32_t cn = 0;
    if (incol) {
        dn = disp(x, y0, w);
        const uint8_t* p = bgr + (size_t)y0 * bgr_pitch + 3 * (size_t)x;
        cn = (uint32_t)p[2] | ((uint32_t)p[1] << 8) | ((uint32_t)p[0] << 16) | 0xFF000000u;
    }
    float4 prev = make_float4(0.f, 0.f, 0.f, 0.f), prev_right = prev;
    float top = 0.f;                                         // |pos(x,y-1) - pos(x+1,y-1)|^2 (all edge values are squared lengths)
    for (int r = 0; r <= DBF_ROWS; r++) {
        const int y = y0 + r;
        if (y >= h) break;                                   // uniform over the CTA
        const float d = dn;
        const uint32_t colour = cn;
        if (incol && r < DBF_ROWS && y + 1 < h) {            // prefetch the next row
            dn = disp(x, y + 1, w);
            const uint8_t* p = bgr + (size_t)(y + 1) * bgr_pitch + 3 * (size_t)x;
            cn = (uint32_t)p[2] | ((uint32_t)p[1] << 8) | ((uint32_t)p[0] << 16) | 0xFF000000u;
        }
        float4 cur = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        if (incol) {
            const float depth = d == 0.0f ? -INFINITY : focal * baseline / d;
            uint32_t rgba = 0;
            if (depth != -INFINITY) {
                float xc = ((float)x * depth - cx * depth) / fx;
                float yc = ((float)y * depth - cy * depth) / fy;
                cur = make_float4(xc, yc, depth, 1.0f);
                rgba = colour;
            }
            if (own && r < DBF_ROWS) {
                const size_t i = (size_t)y * w + x;
                depth_out[i] = depth;
                float4* o = reinterpret_cast<float4*>(vert + i);
                o[0] = cur;
                o[1] = make_float4(__uint_as_float(rgba), 0.f, 0.f, 0.f);
            }
        }
        // Edge lengths are symmetric bit for bit ((a-b)^2 == (b-a)^2), so each is computed once: the vertical edge
        // of this column here (the left-hand neighbour reads it as its cell's right side), the horizontal edge of
        // this row (top side of the cell below, bottom side of the cell above) and the diagonal.
        const float vert_edge = edge_sq(prev, cur);
        spos[r & 1][t] = cur;
        svert[r & 1][t] = vert_edge;
        __syncthreads();
        const float4 cur_right = t < DBF_T - 1 ? spos[r & 1][t + 1] : cur;
        const float vert_right = t < DBF_T - 1 ? svert[r & 1][t + 1] : 0.f;
        const float bottom = edge_sq(cur, cur_right);
        if (r > 0 && own && x < w - 1) {
            // cell (x, y - 1): a = prev, b = cur, c = prev_right, d = cur_right (reconstruction.cpp:92-113);
            // triangle (a,b,c): edges ab, ac, bc; triangle (b,d,c): edges bd, bc, dc (:46-69)
            const float diag = edge_sq(cur, prev_right);
            const bool t1 = !(prev.x == -INFINITY || cur.x == -INFINITY || prev_right.x == -INFINITY) &&
                            !(vert_edge > thr || top > thr || diag > thr);
            const bool t2 = !(cur.x == -INFINITY || cur_right.x == -INFINITY || prev_right.x == -INFINITY) &&
                            !(bottom > thr || diag > thr || vert_right > thr);
            flags[(size_t)(y - 1) * (w - 1) + x] = (uint8_t)((t1 ? 1u : 0u) | (t2 ? 2u : 0u));
        }
        prev = cur;
        prev_right = cur_right;
        top = bottom;
    }
}
// per-block face totals from flags that the fused kernel wrote (what face_count_kernel produces beside its flags):
// 16 flag bytes per thread (one 16-byte load), 16 consecutive lanes make up one block of FACE_THREADS cells
__global__ void __launch_bounds__(256) face_blocktotal_kernel(const uint8_t* __restrict__ flags, unsigned ncells, unsigned nblocks,
                                                              uint32_t* __restrict__ block_total) {
    static_assert(FACE_THREADS == 256, "16 lanes x 16 cells per block total");
    const unsigned g = blockIdx.x * 256 + threadIdx.x;      // group of 16 cells
    const size_t c0 = (size_t)g * 16;
    uint32_t n = 0;
    if (c0 < ncells) {
        uint4 v = *reinterpret_cast<const uint4*>(flags + c0);   // the buffer is padded to a multiple of 16 bytes
        if (c0 + 16 > ncells) {                                  // ragged tail: only the first ncells - c0 bytes are flags
            uint32_t wv[4] = {v.x, v.y, v.z, v.w};
            const int keep = (int)(ncells - c0);
            for (int i = 0; i < 4; i++) {
                int kb = keep - 4 * i;
                wv[i] &= kb >= 4 ? 0xFFFFFFFFu : (kb <= 0 ? 0u : ((1u << (8 * kb)) - 1u));
            }
            v = make_uint4(wv[0], wv[1], wv[2], wv[3]);
        }
        n = __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);   // flag bytes are 0..3: one bit per face
    }
    for (int o = 8; o; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o, 16);
    const unsigned blk = g >> 4;
    if ((threadIdx.x & 15) == 0 && blk < nblocks) block_total[blk] = n;
}

// single-CTA exclusive scan of the block totals (in place), four per thread and 4096 per sweep (16-byte loads: the
// array is 16-byte aligned and padded, see recon_mesh_faces); total -> block_total[nblocks]
__global__ void __launch_bounds__(1024) face_scan_kernel(uint32_t* __restrict__ block_total, unsigned nblocks) {
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (unsigned base = 0; base < nblocks; base += 4096) {
        unsigned i = base + 4 * threadIdx.x;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (i < nblocks) {
            v = *reinterpret_cast<const uint4*>(block_total + i);
            if (i + 1 >= nblocks) v.y = 0;
            if (i + 2 >= nblocks) v.z = 0;
            if (i + 3 >= nblocks) v.w = 0;
        }
        uint32_t tot = v.x + v.y + v.z + v.w, inc = tot;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if ((threadIdx.x & 31) >= o) inc += t;
        }
        if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = inc;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t s = wsum[threadIdx.x], si = s;
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(0xffffffffu, si, o);
                if (threadIdx.x >= o) si += t;
            }
            wsum[threadIdx.x] = si - s;  // exclusive warp offsets
        }
        __syncthreads();
        uint32_t excl = carry + wsum[threadIdx.x >> 5] + inc - tot;
        if (i + 3 < nblocks) {
            *reinterpret_cast<uint4*>(block_total + i) = make_uint4(excl, excl + v.x, excl + v.x + v.y, excl + v.x + v.y + v.z);
        } else if (i < nblocks) {
            block_total[i] = excl;
            if (i + 1 < nblocks) block_total[i + 1] = excl + v.x;
            if (i + 2 < nblocks) block_total[i + 2] = excl + v.x + v.y;
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + tot;
        __syncthreads();
    }
    if (threadIdx.x == 0) block_total[nblocks] = carry;
}

// The CTA's faces are contiguous in the list: staged in shared memory, then stored with coalesced 4-byte words.
__global__ void __launch_bounds__(FACE_THREADS) face_scatter_kernel(int w, unsigned ncells, const uint8_t* __restrict__ flags,
                                                                    const uint32_t* __restrict__ block_off,
                                                                    uint32_t* __restrict__ tri) {
    __shared__ uint32_t wsum[FACE_THREADS / 32];
    __shared__ uint32_t stage[FACE_THREADS * 6];
    unsigned cell = blockIdx.x * FACE_THREADS + threadIdx.x;
    uint32_t f = cell < ncells ? flags[cell] : 0u;
    uint32_t n = __popc(f), inc = n;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if ((threadIdx.x & 31) >= o) inc += t;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = inc;
    __syncthreads();
    uint32_t off = inc - n, total = 0;                       // offset inside the CTA, faces of the CTA
    for (int i = 0; i < FACE_THREADS / 32; i++) {
        if (i < (int)(threadIdx.x >> 5)) off += wsum[i];
        total += wsum[i];
    }
    if (f) {
        unsigned cw = (unsigned)(w - 1);
        unsigned y = cell / cw, x = cell - y * cw;
        uint32_t a = y * w + x, b = (y + 1) * w + x, c = y * w + x + 1, d = (y + 1) * w + x + 1;
        uint32_t* st = stage + 3 * off;
        if (f & 1u) { st[0] = a; st[1] = b; st[2] = c; st += 3; }
        if (f & 2u) { st[0] = b; st[1] = d; st[2] = c; }
    }
    __syncthreads();
    uint32_t* o = tri + 3 * (size_t)block_off[blockIdx.x];
    for (unsigned i = threadIdx.x; i < 3 * total; i += FACE_THREADS) o[i] = stage[i];
}

// flags (one byte per cell) + block totals (+ grand total) in scratch slot 3
static uint8_t* face_buffers(sr_ctx* ctx, unsigned ncells, unsigned nblocks, uint32_t** block_total) {
    uint8_t* flags = (uint8_t*)sr_scratch(ctx, 3, (size_t)ncells + ((size_t)nblocks + 1) * 4 + 32);
    if (flags) *block_total = (uint32_t*)(flags + (((size_t)ncells + 15) & ~(size_t)15));
    return flags;
}
// d_vert != NULL: validity from the vertices; NULL: the flags are already in place (recon_depth_backproject_faces)
int recon_mesh_faces(sr_ctx* ctx, const sr_vertex* d_vert, int w, int h, float thr, uint32_t* d_tri, uint32_t* n_faces) {
    *n_faces = 0;
    if (w < 2 || h < 2) return SR_OK;
    unsigned ncells = (unsigned)(w - 1) * (unsigned)(h - 1);
    unsigned nblocks = (ncells + FACE_THREADS - 1) / FACE_THREADS;
    uint32_t* block_total = nullptr;
    uint8_t* flags = face_buffers(ctx, ncells, nblocks, &block_total);
    if (!flags) return SR_E_CUDA;
    if (d_vert) face_count_kernel<PosFromVertices><<<nblocks, FACE_THREADS, 0, ctx->stream>>>(PosFromVertices{d_vert}, w, sqrt_threshold(thr), ncells, flags, block_total);
    else face_blocktotal_kernel<<<cdiv((int)nblocks, 16), 256, 0, ctx->stream>>>(flags, ncells, nblocks, block_total);
    SR_LAUNCHED(ctx);
    face_scan_kernel<<<1, 1024, 0, ctx->stream>>>(block_total, nblocks);
    SR_LAUNCHED(ctx);
    if (d_tri) {
        face_scatter_kernel<<<nblocks, FACE_THREADS, 0, ctx->stream>>>(w, ncells, flags, block_total, d_tri);
        SR_LAUNCHED(ctx);
    }
    SR_CUDA(ctx, cudaMemcpyAsync(n_faces, block_total + nblocks, 4, cudaMemcpyDeviceToHost, ctx->stream));
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SR_OK;
}
int recon_depth_backproject_faces(sr_ctx* ctx, const float* d_disp, const uint8_t* d_bgr, size_t bgr_pitch, int w, int h,
                                  float focal, float baseline, float fx, float fy, float cx, float cy, float thr,
                                  float* d_depth_out, sr_vertex* d_vert) {
    if (w < 2 || h < 2)
        return recon_depth_backproject(ctx, d_disp, nullptr, d_bgr, bgr_pitch, w, h, focal, baseline, fx, fy, cx, cy, d_depth_out, d_vert);
    unsigned ncells = (unsigned)(w - 1) * (unsigned)(h - 1);
    unsigned nblocks = (ncells + FACE_THREADS - 1) / FACE_THREADS;
    uint32_t* block_total = nullptr;
    uint8_t* flags = face_buffers(ctx, ncells, nblocks, &block_total);
    if (!flags) return SR_E_CUDA;
    dim3 grid(cdiv(w, DBF_T - 1), cdiv(h, DBF_ROWS));
    depth_backproject_faces_kernel<DispFromMap><<<grid, DBF_T, 0, ctx->stream>>>(DispFromMap{d_disp}, d_bgr, bgr_pitch, w, h, focal, baseline,
                                                                                 fx, fy, cx, cy, sqrt_threshold(thr), d_depth_out, d_vert, flags);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// ---- main.cpp:41-65 for the 8-bit hand-off, integer-exact ----------------------------------------------------------
// After the PNG hand-off every disparity is an integer 0..255, so every sum cv::blur forms is an integer far below 2^31
// (a window holds at most 255 k^2, a column prefix at most 255 k (h + k - 1)): the whole box filter runs in int32 --
// half the bytes of the fp64 prefix path, exact by construction (no proof kernel) -- and the last step,
// (float)((double)sum * (1/(k*k))), is the one OpenCV performs.  Four launches instead of nine:
//   u8_rowsum_kernel     CV_64F map -> u8 map (rint half-even, saturate) + row window sums (one CTA per row, prefix in smem)
//   blur_cols_seg/prefix column prefixes of the row sums (int32)
//   depth_backproject_faces_kernel<DispFromU8Blur>   blurred value + outlier rule + depth + vertex + face flags
// ~75 B/pixel of traffic against ~125 through the general kernels.
#define U8R_T 512
__global__ void __launch_bounds__(U8R_T) u8_rowsum_kernel(const double* __restrict__ src, int w, int h, int k, uint8_t* __restrict__ q8,
                                                          int* __restrict__ rows) {
    extern __shared__ int pi[];                  // w + k - 1 prefix values of the reflect-101 extended row
    __shared__ int wsum[U8R_T / 32];
    const int y = blockIdx.x, t = threadIdx.x, a = k / 2, we = w + k - 1;
    const double* sr = src + (size_t)y * w;
    // quantise the row once (main.cpp:41-50), staged at its place inside the extended row
    for (int x = t; x < w; x += U8R_T) {
        const double v = sr[x];
        long long q = (v != v) ? 0 : __double2ll_rn(fmax(-1e9, fmin(1e9, v)));
        q = q < 0 ? 0 : (q > 255 ? 255 : q);
        pi[x + a] = (int)q;
        q8[(size_t)y * w + x] = (uint8_t)q;
    }
    __syncthreads();
    for (int j = t; j < we; j += U8R_T) {        // the reflected borders (reads of the staged centre part only)
        const int xs = j - a;
        if (xs < 0 || xs >= w) pi[j] = pi[reflect101(xs, w) + a];
    }
    __syncthreads();
    const int C = (we + U8R_T - 1) / U8R_T, j0 = min(t * C, we), j1 = min(j0 + C, we);
    int acc = 0;
    for (int j = j0; j < j1; j++) { acc += pi[j]; pi[j] = acc; }
    int v = acc;
    for (int off = 1; off < 32; off <<= 1) { int u = __shfl_up_sync(0xffffffffu, v, off); if ((t & 31) >= off) v += u; }
    if ((t & 31) == 31) wsum[t >> 5] = v;
    __syncthreads();
    int base = v - acc;
    for (int i = 0; i < (t >> 5); i++) base += wsum[i];
    for (int j = j0; j < j1; j++) pi[j] += base;
    __syncthreads();
    int* o = rows + (size_t)y * w;
    for (int x = t; x < w; x += U8R_T) o[x] = pi[x + k - 1] - (x > 0 ? pi[x - 1] : 0);
}
#define U8_NSEG 64
// Latency-bound (each thread walks ~40 rows of one column): eight independent loads in flight per thread in the
// segment sums (17.1 -> 13.6 us on the 4K map; the same in the prefix kernel below measured slower, 19.5 against 17.3 us).
__global__ void u8_cols_seg_kernel(const int* __restrict__ rows, int w, int h, int k, int* __restrict__ seg) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, sgi = blockIdx.y;
    if (x >= w) return;
    const int a = k / 2, he = h + k - 1, SL = (he + U8_NSEG - 1) / U8_NSEG;
    const int j0 = sgi * SL, j1 = min(j0 + SL, he);
    int acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, j = j0;
    for (; j + 7 < j1; j += 8) {
#pragma unroll
        for (int u = 0; u < 8; u++) acc[u] += rows[(size_t)reflect101(j + u - a, h) * w + x];
    }
    for (; j < j1; j++) acc[0] += rows[(size_t)reflect101(j - a, h) * w + x];
    seg[(size_t)sgi * w + x] = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((acc[4] + acc[5]) + (acc[6] + acc[7]));
}
__global__ void u8_cols_prefix_kernel(const int* __restrict__ rows, int w, int h, int k, const int* __restrict__ seg, int* __restrict__ pe) {
    int x = blockIdx.x * blockDim.x + threadIdx.x, sgi = blockIdx.y;
    if (x >= w) return;
    const int a = k / 2, he = h + k - 1, SL = (he + U8_NSEG - 1) / U8_NSEG;
    const int j0 = sgi * SL, j1 = min(j0 + SL, he);
    int acc = 0;
    for (int i = 0; i < sgi; i++) acc += seg[(size_t)i * w + x];
    int j = j0;
    for (; j + 3 < j1; j += 4) {
        int r0 = rows[(size_t)reflect101(j - a, h) * w + x], r1 = rows[(size_t)reflect101(j + 1 - a, h) * w + x];
        int r2 = rows[(size_t)reflect101(j + 2 - a, h) * w + x], r3 = rows[(size_t)reflect101(j + 3 - a, h) * w + x];
        acc += r0; pe[(size_t)j * w + x] = acc;
        acc += r1; pe[(size_t)(j + 1) * w + x] = acc;
        acc += r2; pe[(size_t)(j + 2) * w + x] = acc;
        acc += r3; pe[(size_t)(j + 3) * w + x] = acc;
    }
    for (; j < j1; j++) { acc += rows[(size_t)reflect101(j - a, h) * w + x]; pe[(size_t)j * w + x] = acc; }
}

// Applies when the sums fit int32 and a row fits shared memory; the caller falls back to the general kernels otherwise.
bool recon_u8_pipeline_applies(int w, int h, int k) {
    if (w < 2 || h < 2 || k < 1) return false;
    if ((size_t)(w + k - 1) * sizeof(int) > 200 * 1024) return false;
    return 255.0 * k * ((double)h + k - 1) < 2147483647.0 && 255.0 * ((double)w + k - 1) < 2147483647.0;
}
int recon_u8_pipeline(sr_ctx* ctx, const double* d_disp64, const uint8_t* d_bgr, size_t bgr_pitch, int w, int h, int k, float thr_front,
                      float thr_back, float focal, float baseline, float fx, float fy, float cx, float cy, float thr,
                      float* d_disp_out, float* d_depth_out, sr_vertex* d_vert) {
    const int he = h + k - 1;
    const size_t n = (size_t)w * h;
    // scratch: row sums (slot 1), u8 map (slot 2), prefix array + segment sums (slot 15)
    int* rows = (int*)sr_scratch(ctx, 1, n * sizeof(int));
    uint8_t* q8 = (uint8_t*)sr_scratch(ctx, 2, n + 16);
    int* pe = (int*)sr_scratch(ctx, 15, ((size_t)he * w + (size_t)U8_NSEG * w) * sizeof(int));
    if (!rows || !q8 || !pe) return SR_E_CUDA;
    int* seg = pe + (size_t)he * w;
    unsigned ncells = (unsigned)(w - 1) * (unsigned)(h - 1);
    unsigned nblocks = (ncells + FACE_THREADS - 1) / FACE_THREADS;
    uint32_t* block_total = nullptr;
    uint8_t* flags = face_buffers(ctx, ncells, nblocks, &block_total);
    if (!flags) return SR_E_CUDA;
    static DeviceOnce configured;
    if (configured.need(ctx->device)) {
        SR_CUDA(ctx, cudaFuncSetAttribute(u8_rowsum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        configured.done(ctx->device);
    }
    u8_rowsum_kernel<<<h, U8R_T, (size_t)(w + k - 1) * sizeof(int), ctx->stream>>>(d_disp64, w, h, k, q8, rows);
    SR_LAUNCHED(ctx);
    dim3 gs(cdiv(w, 128), U8_NSEG);
    u8_cols_seg_kernel<<<gs, 128, 0, ctx->stream>>>(rows, w, h, k, seg);
    SR_LAUNCHED(ctx);
    u8_cols_prefix_kernel<<<gs, 128, 0, ctx->stream>>>(rows, w, h, k, seg, pe);
    SR_LAUNCHED(ctx);
    DispFromU8Blur src{q8, pe, k, 1.0 / ((double)k * (double)k), thr_front, thr_back, d_disp_out};
    dim3 grid(cdiv(w, DBF_T - 1), cdiv(h, DBF_ROWS));
    depth_backproject_faces_kernel<DispFromU8Blur><<<grid, DBF_T, 0, ctx->stream>>>(src, d_bgr, bgr_pitch, w, h, focal, baseline, fx, fy, cx,
                                                                                    cy, sqrt_threshold(thr), d_depth_out, d_vert, flags);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// main.cpp:41-50: CV_64F -> 8-bit PNG (rint half-even, saturate) -> float32; or a plain cast.
__device__ __forceinline__ float quantize1(double v, int quantize_u8) {
    if (!quantize_u8) return (float)v;
    long long q = (v != v) ? 0 : __double2ll_rn(fmax(-1e9, fmin(1e9, v)));
    q = q < 0 ? 0 : (q > 255 ? 255 : q);
    return (float)q;
}
// four pixels per thread: two 16-byte loads in flight, one 16-byte store (vec != 0: both arrays 16-byte aligned)
__global__ void quantize_kernel(const double* __restrict__ disp, float* __restrict__ out, size_t n, int quantize_u8, int vec) {
    size_t i = 4 * ((size_t)blockIdx.x * blockDim.x + threadIdx.x);
    if (i >= n) return;
    if (vec && i + 3 < n) {
        double2 a = *reinterpret_cast<const double2*>(disp + i), b = *reinterpret_cast<const double2*>(disp + i + 2);
        *reinterpret_cast<float4*>(out + i) = make_float4(quantize1(a.x, quantize_u8), quantize1(a.y, quantize_u8),
                                                           quantize1(b.x, quantize_u8), quantize1(b.y, quantize_u8));
    } else {
        for (size_t e = min(i + 4, n); i < e; i++) out[i] = quantize1(disp[i], quantize_u8);
    }
}
int recon_quantize(sr_ctx* ctx, const double* d_disp, float* d_out, size_t n, int quantize_u8) {
    const int vec = (((uintptr_t)d_disp | (uintptr_t)d_out) & 15) == 0;
    quantize_kernel<<<(unsigned)((n + 1023) / 1024), 256, 0, ctx->stream>>>(d_disp, d_out, n, quantize_u8, vec);
    SR_LAUNCHED(ctx);
    return SR_OK;
}
