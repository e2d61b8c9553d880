// ubench_pipes.cu -- measures the per-SM issue throughput of the integer instructions the search
// kernel is made of (the INT roofline denominator of SURVEY.md 8d must be measured, not assumed).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_pipes ubench_pipes.cu
// Prints one JSON line per instruction: lane-ops / clk / SM at full occupancy.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048
#define NACC 8

template <int OP>
__global__ void __launch_bounds__(1024) k(uint32_t* out, uint32_t seed, long long* cycles) {
    uint32_t a[NACC], b = seed + threadIdx.x, c = seed * 3 + 1;
#pragma unroll
    for (int i = 0; i < NACC; i++) a[i] = seed + i * 17 + threadIdx.x;
    long long t0 = clock64();
#pragma unroll 8
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NACC; i++) {
            if (OP == 0) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
            if (OP == 1) asm volatile("vabsdiff4.u32.u32.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(0));
            if (OP == 2) asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 3) asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 4) asm volatile("min.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b + i));
            if (OP == 5) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 6) { uint32_t t; asm volatile("min.u32 %0, %1, %2;" : "=r"(t) : "r"(a[i]), "r"(b)); asm volatile("min.u32 %0, %1, %2;" : "=r"(a[i]) : "r"(t), "r"(c + i)); }
            if (OP == 7) asm volatile("shfl.sync.down.b32 %0, %0, 1, 0x1f, 0xffffffff;" : "+r"(a[i]));
            if (OP == 8) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 9) { asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b)); asm volatile("sub.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(c)); }
            if (OP == 10) { asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[i]) : "r"(b), "r"(c)); asm volatile("add.u32 %0, %0, %1;" : "+r"(a[(i + 1) % NACC]) : "r"(b)); }
            if (OP == 11) { asm volatile("vabsdiff4.u32.u32.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(0)); asm volatile("dp4a.u32.u32 %0, %1, %2, %0;" : "+r"(a[(i + 1) % NACC]) : "r"(b), "r"(c)); }
            if (OP == 12) { asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c)); asm volatile("add.u32 %0, %0, %1;" : "+r"(a[(i + 1) % NACC]) : "r"(b)); }
            if (OP == 13) asm volatile("vadd2.u32.u32.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(0));
            if (OP == 14) { uint32_t r; asm volatile("redux.sync.min.u32 %0, %1, 0xffffffff;" : "=r"(r) : "r"(a[i])); a[i] += r; }
        }
    }
    long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int per_iter_ops) {
    int dev; cudaGetDevice(&dev);
    cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
    int blocks = p.multiProcessorCount * 2, threads = 1024;
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, (size_t)blocks * threads * 4); cudaMalloc(&cyc, blocks * 8);
    k<OP><<<blocks, threads>>>(out, 12345, cyc);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<OP><<<blocks, threads>>>(out, 12345, cyc);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[8]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    double ops_per_block = (double)threads * ITERS * NACC * per_iter_ops;
    // 2 resident CTAs of 1024 threads per SM run concurrently for h[0] cycles
    double per_clk_sm = 2.0 * ops_per_block / (double)h[0];
    double tops = (double)blocks * ops_per_block / (ms * 1e-3) / 1e12;
    printf("{\"op\": \"%s\", \"lane_ops_per_clk_per_sm\": %.1f, \"chip_Tops\": %.2f, \"ms\": %.3f, \"cycles\": %lld}\n", name, per_clk_sm, tops, ms, h[0]);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    cudaError_t e = cudaSetDevice(0);
    if (e != cudaSuccess) { printf("no device: %s\n", cudaGetErrorString(e)); return 1; }
    run<0>("IADD (add.u32)", 1);
    run<1>("VABSDIFF4", 1);
    run<2>("VABSDIFF4.ACC", 1);
    run<3>("IDP.4A", 1);
    run<4>("VIMNMX (min.u32)", 1);
    run<5>("IMAD (mad.lo)", 1);
    run<6>("2x min (VIMNMX3 if fused)", 2);
    run<7>("SHFL.DOWN", 1);
    run<8>("LOP3", 1);
    run<9>("add+sub (IADD3 if fused)", 2);
    run<10>("IDP.4A + IADD mix", 2);
    run<11>("VABSDIFF4 + IDP.4A mix", 2);
    run<12>("IMAD + IADD mix", 2);
    run<13>("VIADD.16x2 (vadd2)", 1);
    run<14>("REDUX.MIN + add", 2);
    return 0;
}
