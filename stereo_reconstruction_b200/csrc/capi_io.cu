// capi_io.cu -- host-side neighbours of the path: the Middlebury PFM reader and the calib.txt
// parser that DataLoader needs to feed configs 1-2 (data_loader.cpp:110-125,141-164), computeAverageDisparity
// (utils.cpp:8-20) and the row-band split of the multi-GPU path.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include "sr_internal.cuh"

// computeAverageDisparity (utils.cpp:8-20) is ONE float accumulator walked over the map in raster order: every add is
// rounded, so the order is part of the result.  One thread follows it exactly (a dependent FADD chain, ~2 ns per
// pixel); the other lanes of the warp only prefetch.  A parallel sum would be faster and different in the last bits.
__global__ void average_disparity_kernel(const float* __restrict__ d, size_t n, float* __restrict__ out_sum, int* __restrict__ out_cnt) {
    float sum = 0.f;
    int cnt = 0;
    const int lane = threadIdx.x;
    for (size_t base = 0; base < n; base += 32 * 8) {
        float v[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {   // coalesced: lane l holds elements base + 32*k + l
            size_t i = base + 32 * k + lane;
            v[k] = i < n ? d[i] : INFINITY;
        }
#pragma unroll
        for (int k = 0; k < 8; k++)
            for (int l = 0; l < 32; l++) {
                float x = __shfl_sync(0xffffffffu, v[k], l);
                if (lane == 0 && !isinf(x)) { sum = __fadd_rn(sum, x); cnt++; }
            }
    }
    if (lane == 0) { *out_sum = sum; *out_cnt = cnt; }
}

extern "C" {

int sr_band_rows(int h, int rank, int nranks, int* row0, int* row1) {
    if (h < 0 || nranks < 1 || rank < 0 || rank >= nranks || !row0 || !row1) return SR_E_BADARG;
    *row0 = (int)((long long)h * rank / nranks);
    *row1 = (int)((long long)h * (rank + 1) / nranks);
    return SR_OK;
}

// Middlebury PFM as the SDK's ReadImageVerb / ReadFilePFM reads it (data_loader.cpp:110-125): "Pf", width height,
// scale (its sign gives the byte order, its magnitude is NOT applied), float32 rows stored bottom to top -> returned
// top to bottom; unknown pixels are +inf.
int sr_pfm_info(const char* path, int* w, int* h) {
    if (!path || !w || !h) return SR_E_BADARG;
    FILE* f = fopen(path, "rb");
    if (!f) return SR_E_IO;
    char magic[3] = {0, 0, 0};
    float scale = 0;
    int ok = fscanf(f, "%2s %d %d %f", magic, w, h, &scale) == 4 && !strcmp(magic, "Pf") && *w > 0 && *h > 0;
    fclose(f);
    return ok ? SR_OK : SR_E_IO;
}

int sr_read_pfm(const char* path, float* out, int w, int h) {
    if (!path || !out || w <= 0 || h <= 0) return SR_E_BADARG;
    FILE* f = fopen(path, "rb");
    if (!f) return SR_E_IO;
    char magic[3] = {0, 0, 0};
    int fw = 0, fh = 0;
    float scale = 0;
    if (fscanf(f, "%2s %d %d %f", magic, &fw, &fh, &scale) != 4 || strcmp(magic, "Pf") || fw != w || fh != h) { fclose(f); return SR_E_IO; }
    if (fgetc(f) == '\r') fgetc(f);   // the single whitespace byte after the header
    const bool big_endian = scale > 0;
    for (int y = h - 1; y >= 0; y--) {
        float* row = out + (size_t)y * w;
        if (fread(row, sizeof(float), (size_t)w, f) != (size_t)w) { fclose(f); return SR_E_IO; }
        if (big_endian)
            for (int x = 0; x < w; x++) {
                uint32_t u; memcpy(&u, row + x, 4);
                u = (u >> 24) | ((u >> 8) & 0xFF00u) | ((u << 8) & 0xFF0000u) | (u << 24);
                memcpy(row + x, &u, 4);
            }
    }
    fclose(f);
    return SR_OK;
}

// calib.txt of a Middlebury v3 scene.  cam0 / cam1: the first two lines, parsed the way
// DataLoader::readCameraMatrices does (data_loader.cpp:141-164: strip "camN=[" and the last character, ';' -> ' ', nine
// floats, row-major).  The key=value lines that follow (the reference ignores them; configs 1-2 need ndisp) go to
// `calib` when it is not NULL.
int sr_read_calib(const char* path, float cam0[9], float cam1[9], sr_calib* calib) {
    if (!path || !cam0 || !cam1) return SR_E_BADARG;
    FILE* f = fopen(path, "r");
    if (!f) return SR_E_IO;
    char buf[512];
    int rc = SR_OK;
    for (int i = 0; i < 2 && rc == SR_OK; i++) {
        if (!fgets(buf, sizeof buf, f)) { rc = SR_E_IO; break; }
        std::string line(buf);
        while (!line.empty() && (line.back() == '\n' || line.back() == '\r')) line.pop_back();
        if (line.size() < 8) { rc = SR_E_IO; break; }
        line.erase(0, 6);          // "cam0=["
        line.pop_back();           // "]"
        for (char& c : line) if (c == ';') c = ' ';
        float* m = i == 0 ? cam0 : cam1;
        if (sscanf(line.c_str(), "%f %f %f %f %f %f %f %f %f", m, m + 1, m + 2, m + 3, m + 4, m + 5, m + 6, m + 7, m + 8) != 9) rc = SR_E_IO;
    }
    if (calib && rc == SR_OK) {
        memset(calib, 0, sizeof *calib);
        while (fgets(buf, sizeof buf, f)) {
            char key[64]; double v;
            if (sscanf(buf, "%63[^=]=%lf", key, &v) != 2) continue;
            if (!strcmp(key, "doffs")) calib->doffs = v; else if (!strcmp(key, "baseline")) calib->baseline = v;
            else if (!strcmp(key, "width")) calib->width = (int)v; else if (!strcmp(key, "height")) calib->height = (int)v;
            else if (!strcmp(key, "ndisp")) calib->ndisp = (int)v; else if (!strcmp(key, "isint")) calib->isint = (int)v;
            else if (!strcmp(key, "vmin")) calib->vmin = (int)v; else if (!strcmp(key, "vmax")) calib->vmax = (int)v;
            else if (!strcmp(key, "dyavg")) calib->dyavg = v; else if (!strcmp(key, "dymax")) calib->dymax = v;
        }
    }
    fclose(f);
    return rc;
}

// computeAverageDisparity(disparityMap), utils.cpp:8-20: float sum of the non-infinite pixels in raster order / count.
int sr_average_disparity(sr_ctx* ctx, const float* disp, int w, int h, size_t pitch, float* avg, int* valid) {
    if (!ctx || !disp || !avg || w <= 0 || h <= 0 || pitch < (size_t)w * 4) return sr_fail(ctx, SR_E_BADARG, "sr_average_disparity: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)w * h;
    float* d = (float*)sr_scratch(ctx, 4, n * 4 + 16);
    if (!d) return SR_E_CUDA;
    float* dsum = d + n;
    int* dcnt = (int*)(dsum + 1);
    SR_CUDA(ctx, cudaMemcpy2DAsync(d, (size_t)w * 4, disp, pitch, (size_t)w * 4, h, cudaMemcpyHostToDevice, ctx->stream));
    average_disparity_kernel<<<1, 32, 0, ctx->stream>>>(d, n, dsum, dcnt);
    SR_LAUNCHED(ctx);
    float s; int c;
    SR_CUDA(ctx, cudaMemcpyAsync(&s, dsum, 4, cudaMemcpyDeviceToHost, ctx->stream));
    SR_CUDA(ctx, cudaMemcpyAsync(&c, dcnt, 4, cudaMemcpyDeviceToHost, ctx->stream));
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *avg = s / (float)c;   // utils.cpp:19 (0/0 = NaN like the reference)
    if (valid) *valid = c;
    return SR_OK;
}

}  // extern "C"
