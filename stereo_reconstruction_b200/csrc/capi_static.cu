// capi_static.cu -- sr_box_blur_static: cv::blur(32F, k x k) on device buffers with NO allocation of any kind.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of
// the sequential blur pair of recon.cu (blur_rows_kernel / blur_cols_kernel, reconstruction.cpp:8) whose
// intermediate -- the fp64 row sums, w x h doubles -- lives in a file-scope __device__ array instead of the
// context's grow-only cudaMalloc scratch.  Every GPU that loads the library holds the array (8 MiB) until exit.
//
// The array's name is a host-side stand-in in host code, not a device address: the device address is taken with
// cudaGetSymbolAddress once per device, with that device current, and that pointer is what the kernels receive.
#include <mutex>
#include "sr_internal.cuh"

static_assert(SR_BLUR_STATIC_MAX_PIXELS == (1u << 20), "keep the array below in step with include/sr_capi.h");
static __device__ __align__(256) double g_blur_rows[SR_BLUR_STATIC_MAX_PIXELS];

namespace {

constexpr int MAXDEV = 64;
std::mutex g_setup_mutex;
DeviceOnce g_setup;
double* g_rows_addr[MAXDEV];   // device address of g_blur_rows on each device, valid once g_setup says so

// cv::borderInterpolate(p, len, BORDER_REFLECT_101)
__device__ __forceinline__ int reflect101s(int p, int len) {
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        if (p < 0) p = -p; else p = 2 * (len - 1) - p;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

// RowSum<float,double>: first window summed left to right, then s += new - old (one thread per row: the
// recurrence is sequential in x and kept as it is, so the sums equal OpenCV's for any float input).
__global__ void static_blur_rows_kernel(const float* __restrict__ src, int w, int h, int k, double* __restrict__ rows) {
    const int y = blockIdx.x * blockDim.x + threadIdx.x;
    if (y >= h) return;
    const float* s = src + (size_t)y * w;
    double* o = rows + (size_t)y * w;
    const int a = k / 2;
    double acc = 0;
    for (int i = 0; i < k; i++) acc += (double)s[reflect101s(i - a, w)];
    o[0] = acc;
    for (int x = 1; x < w; x++) {
        acc += (double)s[reflect101s(x + k - 1 - a, w)] - (double)s[reflect101s(x - 1 - a, w)];
        o[x] = acc;
    }
}

// ColumnSum<double,float>: running column sum over k rows, out = (float)(sum * 1/(k*k)); one thread per column.
__global__ void static_blur_cols_kernel(const double* __restrict__ rows, int w, int h, int k, float* __restrict__ dst) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= w) return;
    const int a = k / 2;
    const double scale = 1.0 / ((double)k * (double)k);
    double sum = 0;
    for (int i = 0; i < k - 1; i++) sum += rows[(size_t)reflect101s(i - a, h) * w + x];
    for (int y = 0; y < h; y++) {
        const double s0 = sum + rows[(size_t)reflect101s(y - a + k - 1, h) * w + x];
        dst[(size_t)y * w + x] = (float)(s0 * scale);
        sum = s0 - rows[(size_t)reflect101s(y - a, h) * w + x];
    }
}

// Set-up, once per device, with that device current: the device address of the static array.
int static_scratch(sr_ctx* ctx, double** out) {
    const int dev = ctx->device;
    if (dev < 0 || dev >= MAXDEV) return sr_fail(ctx, SR_E_BADARG, "sr_box_blur_static: device %d out of range", dev);
    if (g_setup.need(dev)) {
        std::lock_guard<std::mutex> lock(g_setup_mutex);
        if (g_setup.need(dev)) {
            int cur = -1;
            SR_CUDA(ctx, cudaGetDevice(&cur));
            if (cur != dev) return sr_fail(ctx, SR_E_CUDA, "sr_box_blur_static: device %d is not current (%d is)", dev, cur);
            void* p = nullptr;
            SR_CUDA(ctx, cudaGetSymbolAddress(&p, g_blur_rows));
            if (!p) return sr_fail(ctx, SR_E_CUDA, "sr_box_blur_static: cudaGetSymbolAddress returned no address");
            g_rows_addr[dev] = (double*)p;
            g_setup.done(dev);
        }
    }
    *out = g_rows_addr[dev];
    return SR_OK;
}

bool device_readable(const void* p, int dev) {
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    if (at.type == cudaMemoryTypeManaged) return true;
    return at.type == cudaMemoryTypeDevice && at.device == dev;
}

}  // namespace

extern "C" int sr_box_blur_static(sr_ctx* ctx, const float* d_src, float* d_dst, int w, int h, int ksize, void* stream) {
    if (!ctx) return SR_E_BADARG;
    if (!d_src || !d_dst || w <= 0 || h <= 0 || ksize <= 0)
        return sr_fail(ctx, SR_E_BADARG, "sr_box_blur_static: bad argument");
    if ((uint64_t)w * (uint64_t)h > (uint64_t)SR_BLUR_STATIC_MAX_PIXELS)
        return sr_fail(ctx, SR_E_SHAPE, "sr_box_blur_static: %d x %d = %llu pixels exceed the static scratch of %u pixels "
                       "(SR_BLUR_STATIC_MAX_PIXELS); use sr_box_blur", w, h, (unsigned long long)w * h,
                       (unsigned)SR_BLUR_STATIC_MAX_PIXELS);
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!device_readable(d_src, ctx->device) || !device_readable(d_dst, ctx->device))
        return sr_fail(ctx, SR_E_BADARG, "sr_box_blur_static: d_src / d_dst must be device memory of GPU %d", ctx->device);
    double* rows = nullptr;
    int rc = static_scratch(ctx, &rows);
    if (rc != SR_OK) return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : ctx->stream;
    // Nothing in the scratch has to start at zero: the row pass overwrites all w*h sums the column pass reads, and
    // whatever an earlier (larger) call left behind that range is never read.
    static_blur_rows_kernel<<<cdiv(h, 32), 32, 0, st>>>(d_src, w, h, ksize, rows);
    SR_LAUNCHED(ctx);
    static_blur_cols_kernel<<<cdiv(w, 64), 64, 0, st>>>(rows, w, h, ksize, d_dst);
    SR_LAUNCHED(ctx);
    return SR_OK;
}
