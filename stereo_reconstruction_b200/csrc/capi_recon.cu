// capi_recon.cu -- Reconstruction entry points of the C-ABI: host-pointer wrappers around recon.cu.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "recon.cuh"

enum { SL_DISP = 4, SL_DEPTH = 5, SL_BGR = 8, SL_VERT = 9, SL_TRI = 10, SL_A = 11, SL_B = 12, SL_C = 13 };

static int up2d(sr_ctx* ctx, void* dst, size_t dpitch, const void* src, size_t spitch, size_t rowbytes, int rows) {
    SR_CUDA(ctx, cudaMemcpy2DAsync(dst, dpitch, src, spitch, rowbytes, rows, cudaMemcpyHostToDevice, ctx->stream));
    return SR_OK;
}
static int down2d(sr_ctx* ctx, void* dst, size_t dpitch, const void* src, size_t spitch, size_t rowbytes, int rows) {
    SR_CUDA(ctx, cudaMemcpy2DAsync(dst, dpitch, src, spitch, rowbytes, rows, cudaMemcpyDeviceToHost, ctx->stream));
    return SR_OK;
}
#define TRY(x) do { int rc__ = (x); if (rc__ != SR_OK) return rc__; } while (0)
#define NEED(ctx, cond, what) do { if (!(cond)) return sr_fail(ctx, SR_E_BADARG, what); } while (0)

extern "C" {

int sr_box_blur(sr_ctx* ctx, const float* src, float* dst, int w, int h, int ksize) {
    NEED(ctx, ctx && src && dst && w > 0 && h > 0 && ksize > 0, "sr_box_blur: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)w * h;
    float* d = (float*)sr_scratch(ctx, SL_DISP, n * 4);
    float* o = (float*)sr_scratch(ctx, SL_DEPTH, n * 4);
    if (!d || !o) return SR_E_CUDA;
    TRY(up2d(ctx, d, (size_t)w * 4, src, (size_t)w * 4, (size_t)w * 4, h));
    TRY(recon_box_blur(ctx, d, o, w, h, ksize));
    TRY(down2d(ctx, dst, (size_t)w * 4, o, (size_t)w * 4, (size_t)w * 4, h));
    return sr_ctx_sync(ctx);
}

int sr_remove_outliers(sr_ctx* ctx, float* disp, int w, int h, size_t pitch, int ksize, float thr_front, float thr_back) {
    NEED(ctx, ctx && disp && w > 0 && h > 0 && ksize > 0 && pitch >= (size_t)w * 4, "sr_remove_outliers: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    float* d = (float*)sr_scratch(ctx, SL_DISP, (size_t)w * h * 4);
    if (!d) return SR_E_CUDA;
    TRY(up2d(ctx, d, (size_t)w * 4, disp, pitch, (size_t)w * 4, h));
    TRY(recon_remove_outliers(ctx, d, w, h, ksize, thr_front, thr_back));
    TRY(down2d(ctx, disp, pitch, d, (size_t)w * 4, (size_t)w * 4, h));
    return sr_ctx_sync(ctx);
}

int sr_scale(sr_ctx* ctx, float* disp, int w, int h, size_t pitch, float f) {
    NEED(ctx, ctx && disp && w > 0 && h > 0 && pitch >= (size_t)w * 4, "sr_scale: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    float* d = (float*)sr_scratch(ctx, SL_DISP, (size_t)w * h * 4);
    if (!d) return SR_E_CUDA;
    TRY(up2d(ctx, d, (size_t)w * 4, disp, pitch, (size_t)w * 4, h));
    TRY(recon_scale(ctx, d, (size_t)w * h, f));
    TRY(down2d(ctx, disp, pitch, d, (size_t)w * 4, (size_t)w * 4, h));
    return sr_ctx_sync(ctx);
}

int sr_disparity_to_depth(sr_ctx* ctx, const float* disp, float* depth, int w, int h, float focal, float baseline) {
    NEED(ctx, ctx && disp && depth && w > 0 && h > 0, "sr_disparity_to_depth: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)w * h;
    float* d = (float*)sr_scratch(ctx, SL_DISP, n * 4);
    float* o = (float*)sr_scratch(ctx, SL_DEPTH, n * 4);
    if (!d || !o) return SR_E_CUDA;
    TRY(up2d(ctx, d, (size_t)w * 4, disp, (size_t)w * 4, (size_t)w * 4, h));
    TRY(recon_depth_backproject(ctx, d, nullptr, nullptr, 0, w, h, focal, baseline, 1.f, 1.f, 0.f, 0.f, o, nullptr));
    TRY(down2d(ctx, depth, (size_t)w * 4, o, (size_t)w * 4, (size_t)w * 4, h));
    return sr_ctx_sync(ctx);
}

int sr_backproject(sr_ctx* ctx, const uint8_t* bgr, size_t pitch, const float* depth, int w, int h, float fx, float fy,
                   float cx, float cy, sr_vertex* out) {
    NEED(ctx, ctx && bgr && depth && out && w > 0 && h > 0 && pitch >= (size_t)w * 3, "sr_backproject: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)w * h;
    float* d = (float*)sr_scratch(ctx, SL_DEPTH, n * 4);
    uint8_t* b = (uint8_t*)sr_scratch(ctx, SL_BGR, n * 3);
    sr_vertex* v = (sr_vertex*)sr_scratch(ctx, SL_VERT, n * sizeof(sr_vertex));
    if (!d || !b || !v) return SR_E_CUDA;
    TRY(up2d(ctx, d, (size_t)w * 4, depth, (size_t)w * 4, (size_t)w * 4, h));
    TRY(up2d(ctx, b, (size_t)w * 3, bgr, pitch, (size_t)w * 3, h));
    TRY(recon_depth_backproject(ctx, nullptr, d, b, (size_t)w * 3, w, h, 0.f, 0.f, fx, fy, cx, cy, nullptr, v));
    SR_CUDA(ctx, cudaMemcpyAsync(out, v, n * sizeof(sr_vertex), cudaMemcpyDeviceToHost, ctx->stream));
    return sr_ctx_sync(ctx);
}

int sr_mesh_faces(sr_ctx* ctx, const sr_vertex* v, int w, int h, float thr, uint32_t* tri, uint32_t* n_faces) {
    NEED(ctx, ctx && v && n_faces && w > 0 && h > 0, "sr_mesh_faces: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)w * h;
    size_t cells = (size_t)(w > 1 ? w - 1 : 0) * (h > 1 ? h - 1 : 0);
    sr_vertex* dv = (sr_vertex*)sr_scratch(ctx, SL_VERT, n * sizeof(sr_vertex));
    uint32_t* dt = tri ? (uint32_t*)sr_scratch(ctx, SL_TRI, (cells ? cells : 1) * 6 * 4) : nullptr;
    if (!dv || (tri && !dt)) return SR_E_CUDA;
    SR_CUDA(ctx, cudaMemcpyAsync(dv, v, n * sizeof(sr_vertex), cudaMemcpyHostToDevice, ctx->stream));
    TRY(recon_mesh_faces(ctx, dv, w, h, thr, dt, n_faces));
    if (tri && *n_faces) SR_CUDA(ctx, cudaMemcpyAsync(tri, dt, (size_t)*n_faces * 12, cudaMemcpyDeviceToHost, ctx->stream));
    return sr_ctx_sync(ctx);
}

// sr_pair_reconstruct(_warp) with the 8-bit hand-off, outlier removal and a face count / list wanted: the integer-exact
// pipeline of recon.cu (four launches, the float map is only formed when the caller wants it back)
static int reconstruct_device_u8(sr_ctx* ctx, const double* d_disp64, const uint8_t* d_bgr, int w, int h, int ksize, float thr_front,
                                 float thr_back, float focal, float baseline, float fx, float fy, float cx, float cy, float thr_mesh,
                                 float* disp_filtered, float* depth, sr_vertex* vertices, uint32_t* tri, uint32_t* n_faces) {
    size_t n = (size_t)w * h;
    size_t cells = (size_t)(w - 1) * (h - 1);
    float* d_disp = disp_filtered ? (float*)sr_scratch(ctx, SL_DISP, n * 4) : nullptr;
    float* d_depth = (float*)sr_scratch(ctx, SL_DEPTH, n * 4);
    sr_vertex* d_v = (sr_vertex*)sr_scratch(ctx, SL_VERT, n * sizeof(sr_vertex));
    uint32_t* d_t = tri ? (uint32_t*)sr_scratch(ctx, SL_TRI, (cells ? cells : 1) * 6 * 4) : nullptr;
    if ((disp_filtered && !d_disp) || !d_depth || !d_v || (tri && !d_t)) return SR_E_CUDA;
    TRY(recon_u8_pipeline(ctx, d_disp64, d_bgr, (size_t)w * 3, w, h, ksize, thr_front, thr_back, focal, baseline, fx, fy, cx, cy, thr_mesh,
                          d_disp, d_depth, d_v));
    if (disp_filtered) TRY(down2d(ctx, disp_filtered, (size_t)w * 4, d_disp, (size_t)w * 4, (size_t)w * 4, h));
    if (depth) TRY(down2d(ctx, depth, (size_t)w * 4, d_depth, (size_t)w * 4, (size_t)w * 4, h));
    if (vertices) SR_CUDA(ctx, cudaMemcpyAsync(vertices, d_v, n * sizeof(sr_vertex), cudaMemcpyDeviceToHost, ctx->stream));
    TRY(recon_mesh_faces(ctx, nullptr, w, h, thr_mesh, d_t, n_faces));
    if (tri && *n_faces) SR_CUDA(ctx, cudaMemcpyAsync(tri, d_t, (size_t)*n_faces * 12, cudaMemcpyDeviceToHost, ctx->stream));
    return sr_ctx_sync(ctx);
}
static bool use_u8_pipeline(int quantize_u8, int ksize, const uint32_t* n_faces, int w, int h) {
    const char* off = getenv("SR_RECON_GENERAL");   // test hook: force the general kernels (read per call)
    return quantize_u8 && ksize > 0 && n_faces && !(off && atoi(off)) && recon_u8_pipeline_applies(w, h, ksize);
}

// shared tail of sr_reconstruct / sr_pair_reconstruct: d_disp (device, tight) -> outputs
static int reconstruct_device(sr_ctx* ctx, float* d_disp, const uint8_t* d_bgr, int w, int h, int ksize, float thr_front,
                              float thr_back, float focal, float baseline, float fx, float fy, float cx, float cy,
                              float thr_mesh, float* disp_filtered, float* depth, sr_vertex* vertices, uint32_t* tri,
                              uint32_t* n_faces, int disp_is_u8 = 0) {
    size_t n = (size_t)w * h;
    size_t cells = (size_t)(w > 1 ? w - 1 : 0) * (h > 1 ? h - 1 : 0);
    float* d_depth = (float*)sr_scratch(ctx, SL_DEPTH, n * 4);
    sr_vertex* d_v = (sr_vertex*)sr_scratch(ctx, SL_VERT, n * sizeof(sr_vertex));
    uint32_t* d_t = tri ? (uint32_t*)sr_scratch(ctx, SL_TRI, (cells ? cells : 1) * 6 * 4) : nullptr;
    if (!d_depth || !d_v || (tri && !d_t)) return SR_E_CUDA;
    if (ksize > 0) TRY(recon_remove_outliers(ctx, d_disp, w, h, ksize, thr_front, thr_back, disp_is_u8));   // main.cpp:53
    if (disp_filtered) TRY(down2d(ctx, disp_filtered, (size_t)w * 4, d_disp, (size_t)w * 4, (size_t)w * 4, h));
    // main.cpp:56-65; with a face list wanted, the validity flags come out of the same pass
    if (n_faces) TRY(recon_depth_backproject_faces(ctx, d_disp, d_bgr, (size_t)w * 3, w, h, focal, baseline, fx, fy, cx, cy, thr_mesh, d_depth, d_v));
    else TRY(recon_depth_backproject(ctx, d_disp, nullptr, d_bgr, (size_t)w * 3, w, h, focal, baseline, fx, fy, cx, cy, d_depth, d_v));
    if (depth) TRY(down2d(ctx, depth, (size_t)w * 4, d_depth, (size_t)w * 4, (size_t)w * 4, h));
    if (vertices) SR_CUDA(ctx, cudaMemcpyAsync(vertices, d_v, n * sizeof(sr_vertex), cudaMemcpyDeviceToHost, ctx->stream));
    if (n_faces) {
        TRY(recon_mesh_faces(ctx, (w < 2 || h < 2) ? d_v : nullptr, w, h, thr_mesh, d_t, n_faces));
        if (tri && *n_faces) SR_CUDA(ctx, cudaMemcpyAsync(tri, d_t, (size_t)*n_faces * 12, cudaMemcpyDeviceToHost, ctx->stream));
    }
    return sr_ctx_sync(ctx);
}

int sr_reconstruct(sr_ctx* ctx, const uint8_t* bgr, size_t pitch, const float* disp, int w, int h, int ksize, float thr_front,
                   float thr_back, float focal, float baseline, float fx, float fy, float cx, float cy, float thr_mesh,
                   float* disp_filtered, float* depth, sr_vertex* vertices, uint32_t* tri, uint32_t* n_faces) {
    NEED(ctx, ctx && bgr && disp && w > 0 && h > 0 && pitch >= (size_t)w * 3, "sr_reconstruct: bad argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)w * h;
    float* d = (float*)sr_scratch(ctx, SL_DISP, n * 4);
    uint8_t* b = (uint8_t*)sr_scratch(ctx, SL_BGR, n * 3);
    if (!d || !b) return SR_E_CUDA;
    TRY(up2d(ctx, d, (size_t)w * 4, disp, (size_t)w * 4, (size_t)w * 4, h));
    TRY(up2d(ctx, b, (size_t)w * 3, bgr, pitch, (size_t)w * 3, h));
    return reconstruct_device(ctx, d, b, w, h, ksize, thr_front, thr_back, focal, baseline, fx, fy, cx, cy, thr_mesh,
                              disp_filtered, depth, vertices, tri, n_faces);
}

int sr_pair_reconstruct(sr_ctx* ctx, sr_pair* pair, int quantize_u8, int ksize, float thr_front, float thr_back, float focal,
                        float baseline, float fx, float fy, float cx, float cy, float thr_mesh, float* disp_filtered,
                        float* depth, sr_vertex* vertices, uint32_t* tri, uint32_t* n_faces) {
    NEED(ctx, ctx && pair && pair->ow > 0 && pair->oh > 0, "sr_pair_reconstruct: no search result on this pair");
    NEED(ctx, pair->ow == pair->w2 && pair->oh == pair->h2, "sr_pair_reconstruct: needs a right-referenced result (main.cpp:61 uses imageRight)");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    int w = pair->ow, h = pair->oh;
    size_t n = (size_t)w * h;
    if (use_u8_pipeline(quantize_u8, ksize, n_faces, w, h))
        return reconstruct_device_u8(ctx, pair->out, pair->rawR, w, h, ksize, thr_front, thr_back, focal, baseline, fx, fy, cx, cy, thr_mesh,
                                     disp_filtered, depth, vertices, tri, n_faces);
    float* d = (float*)sr_scratch(ctx, SL_DISP, n * 4);
    if (!d) return SR_E_CUDA;
    TRY(recon_quantize(ctx, pair->out, d, n, quantize_u8));   // main.cpp:41-50 (8-bit hand-off) or plain cast
    return reconstruct_device(ctx, d, pair->rawR, w, h, ksize, thr_front, thr_back, focal, baseline, fx, fy, cx, cy, thr_mesh,
                              disp_filtered, depth, vertices, tri, n_faces, quantize_u8 != 0);
}

int sr_pair_reconstruct_warp(sr_ctx* ctx, sr_pair* pair, const double M[9], int dw, int dh, const sr_image* color, int quantize_u8,
                             int ksize, float thr_front, float thr_back, float focal, float baseline, float fx, float fy, float cx,
                             float cy, float thr_mesh, double* disp_warped, float* disp_filtered, float* depth,
                             sr_vertex* vertices, uint32_t* tri, uint32_t* n_faces) {
    NEED(ctx, ctx && pair && M && pair->ow > 0 && pair->oh > 0 && dw > 0 && dh > 0, "sr_pair_reconstruct_warp: bad argument / no search result");
    NEED(ctx, pair->ow == pair->w2 && pair->oh == pair->h2, "sr_pair_reconstruct_warp: needs a right-referenced result");
    NEED(ctx, color ? (color->bgr && color->w == dw && color->h == dh && color->pitch >= (size_t)dw * 3) : (dw == pair->w2 && dh == pair->h2),
         "sr_pair_reconstruct_warp: the colour image must have the warped map's size");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)dw * dh;
    double* dwarp = (double*)sr_scratch(ctx, SL_B, n * 8);
    float* d = (float*)sr_scratch(ctx, SL_DISP, n * 4);
    if (!dwarp || !d) return SR_E_CUDA;
    const uint8_t* d_bgr = pair->rawR;
    if (color) {
        uint8_t* b = (uint8_t*)sr_scratch(ctx, SL_BGR, n * 3);
        if (!b) return SR_E_CUDA;
        TRY(up2d(ctx, b, (size_t)dw * 3, color->bgr, color->pitch, (size_t)dw * 3, dh));
        d_bgr = b;
    }
    TRY(sr_launch_warp_nearest(ctx, pair->out, pair->ow, pair->oh, dwarp, dw, dh, M));   // rectification.cpp:82-87
    if (disp_warped) SR_CUDA(ctx, cudaMemcpyAsync(disp_warped, dwarp, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (use_u8_pipeline(quantize_u8, ksize, n_faces, dw, dh))
        return reconstruct_device_u8(ctx, dwarp, d_bgr, dw, dh, ksize, thr_front, thr_back, focal, baseline, fx, fy, cx, cy, thr_mesh,
                                     disp_filtered, depth, vertices, tri, n_faces);
    TRY(recon_quantize(ctx, dwarp, d, n, quantize_u8));                                   // main.cpp:41-50
    return reconstruct_device(ctx, d, d_bgr, dw, dh, ksize, thr_front, thr_back, focal, baseline, fx, fy, cx, cy, thr_mesh,
                              disp_filtered, depth, vertices, tri, n_faces, quantize_u8 != 0);
}

// WriteMesh (reconstruction.cpp:72-149): validity + face order on the GPU, COFF text on the host.
int sr_write_mesh(sr_ctx* ctx, const sr_vertex* v, int w, int h, const char* filename, float thr) {
    NEED(ctx, ctx && v && filename && w > 0 && h > 0, "sr_write_mesh: bad argument");
    size_t cells = (size_t)(w > 1 ? w - 1 : 0) * (h > 1 ? h - 1 : 0);
    std::vector<uint32_t> tri((cells ? cells : 1) * 6);
    uint32_t nf = 0;
    TRY(sr_mesh_faces(ctx, v, w, h, thr, tri.data(), &nf));
    FILE* f = fopen(filename, "w");
    if (!f) return sr_fail(ctx, SR_E_IO, "cannot open %s", filename);   // :117 returns false
    std::vector<char> buf(1 << 20);
    setvbuf(f, buf.data(), _IOFBF, buf.size());
    fprintf(f, "COFF\n%u %u 0\n", (unsigned)((size_t)w * h), nf);      // :120-121
    for (size_t i = 0; i < (size_t)w * h; i++) {                        // :124-137, ostream << float == %g
        const sr_vertex& p = v[i];
        if (p.pos[0] == -INFINITY) fprintf(f, "0 0 0 %u %u %u %u\n", p.rgba[0], p.rgba[1], p.rgba[2], p.rgba[3]);
        else fprintf(f, "%g %g %g %u %u %u %u\n", (double)p.pos[0], (double)p.pos[1], (double)p.pos[2], p.rgba[0], p.rgba[1],
                     p.rgba[2], p.rgba[3]);
    }
    for (uint32_t i = 0; i < nf; i++) fprintf(f, "3 %u %u %u\n", tri[3 * i], tri[3 * i + 1], tri[3 * i + 2]);  // :140-143
    bool ok = fclose(f) == 0;
    return ok ? SR_OK : sr_fail(ctx, SR_E_IO, "write to %s failed", filename);
}

int sr_write_pfm(const char* path, const float* data, int w, int h) {
    if (!path || !data || w <= 0 || h <= 0) return SR_E_BADARG;
    FILE* f = fopen(path, "wb");
    if (!f) return SR_E_IO;
    fprintf(f, "Pf\n%d %d\n-1.0\n", w, h);   // Middlebury layout: little endian, rows bottom to top
    for (int y = h - 1; y >= 0; y--) fwrite(data + (size_t)y * w, sizeof(float), (size_t)w, f);
    return fclose(f) == 0 ? SR_OK : SR_E_IO;
}

int sr_write_ply(const char* path, const sr_vertex* v, int w, int h) {
    if (!path || !v || w <= 0 || h <= 0) return SR_E_BADARG;
    size_t n = 0, tot = (size_t)w * h;
    for (size_t i = 0; i < tot; i++) if (v[i].pos[0] != -INFINITY) n++;
    FILE* f = fopen(path, "w");
    if (!f) return SR_E_IO;
    fprintf(f, "ply\nformat ascii 1.0\nelement vertex %zu\nproperty float x\nproperty float y\nproperty float z\n"
               "property uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n", n);
    for (size_t i = 0; i < tot; i++)
        if (v[i].pos[0] != -INFINITY)
            fprintf(f, "%.9g %.9g %.9g %u %u %u\n", (double)v[i].pos[0], (double)v[i].pos[1], (double)v[i].pos[2], v[i].rgba[0],
                    v[i].rgba[1], v[i].rgba[2]);
    return fclose(f) == 0 ? SR_OK : SR_E_IO;
}

}  // extern "C"
