// sr_measure.cu -- measurement helpers used by bench.py: L2 flush, per-kernel stopwatch and a
// live integer-issue peak (the roofline denominator of the search kernel).
#include "sr_internal.cuh"

extern "C" {

// Writes a buffer larger than the 126 MB L2 on the context stream.
int sr_l2_flush(sr_ctx* ctx) {
    if (!ctx) return SR_E_BADARG;
    const size_t bytes = 384ull << 20;
    void* p = sr_scratch(ctx, 7, bytes);
    if (!p) return SR_E_CUDA;
    SR_CUDA(ctx, cudaMemsetAsync(p, 0x5A, bytes, ctx->stream));
    return SR_OK;
}

// Device time of the dominant search kernel of the last sr_search_run (events recorded around
// that launch on the context stream).  Synchronises.
int sr_search_kernel_ms(sr_ctx* ctx, float* ms) {
    if (!ctx || !ms) return SR_E_BADARG;
    if (!ctx->kev_valid) return sr_fail(ctx, SR_E_BADARG, "no search kernel timed yet");
    SR_CUDA(ctx, cudaEventSynchronize(ctx->kev1));
    SR_CUDA(ctx, cudaEventElapsedTime(ms, ctx->kev0, ctx->kev1));
    return SR_OK;
}

}  // extern "C"

// 8 independent chains per thread; MODE 0: VABSDIFF4 + IDP.4A pairs (ALU pipe + FMA pipe, the
// search kernel's pointwise cost), MODE 1: VABSDIFF4 only (one pipe).
template <int MODE>
__global__ void __launch_bounds__(1024) int_peak_kernel(uint32_t* out, uint32_t seed, int iters) {
    uint32_t a[8], b = seed + threadIdx.x, c = seed * 3 + 1;
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed + i * 17 + threadIdx.x;
#pragma unroll 4
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            asm volatile("vabsdiff4.u32.u32.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(0));
            if (MODE == 0) asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[(i + 1) & 7]) : "r"(b), "r"(c));
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

extern "C" {

// Measured integer issue peak in lane-ops/s: dual = VABSDIFF4+IDP.4A interleaved (both pipes),
// single = VABSDIFF4 alone (INT/ALU pipe only).  Runs ~1 ms each on the context stream.
int sr_int_peak(sr_ctx* ctx, double* dual_ops_per_s, double* single_ops_per_s) {
    if (!ctx || !dual_ops_per_s || !single_ops_per_s) return SR_E_BADARG;
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    int sms = 0;
    SR_CUDA(ctx, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
    const int blocks = sms * 2, threads = 1024, iters = 4096;
    uint32_t* out = (uint32_t*)sr_scratch(ctx, 6, (size_t)blocks * threads * 4);
    if (!out) return SR_E_CUDA;
    cudaEvent_t e0, e1;
    SR_CUDA(ctx, cudaEventCreate(&e0));
    SR_CUDA(ctx, cudaEventCreate(&e1));
    float ms[2] = {0, 0};
    for (int mode = 0; mode < 2; mode++) {
        float bestms = 1e30f;
        for (int rep = 0; rep < 4; rep++) {  // first rep warms up
            SR_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
            if (mode == 0) int_peak_kernel<0><<<blocks, threads, 0, ctx->stream>>>(out, 12345u, iters);
            else int_peak_kernel<1><<<blocks, threads, 0, ctx->stream>>>(out, 12345u, iters);
            SR_LAUNCHED(ctx);
            SR_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
            SR_CUDA(ctx, cudaEventSynchronize(e1));
            float t;
            SR_CUDA(ctx, cudaEventElapsedTime(&t, e0, e1));
            if (rep > 0 && t < bestms) bestms = t;
        }
        ms[mode] = bestms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    double lanes = (double)blocks * threads * iters * 8;
    *dual_ops_per_s = lanes * 2 / (ms[0] * 1e-3);
    *single_ops_per_s = lanes / (ms[1] * 1e-3);
    return SR_OK;
}

}  // extern "C"
