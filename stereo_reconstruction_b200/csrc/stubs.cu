// stubs.cu -- entry points not implemented yet (fail loudly, never fall back).
#include "sr_internal.cuh"
#define NOTYET(ctx, name) return sr_fail(ctx, SR_E_BADARG, name ": not implemented yet")
int sr_launch_search_left(sr_ctx* ctx, sr_pair* pair, int block, int max_disp, double smooth) { NOTYET(ctx, "left variant"); }
int sr_launch_search_linear(sr_ctx* ctx, sr_pair* pair, double smooth) { NOTYET(ctx, "linear variant"); }
int sr_search_run_ext(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p, int row0, int row1) { NOTYET(ctx, "extension costs"); }
extern "C" {
int sr_band_rows(int h, int rank, int nranks, int* row0, int* row1) {
    if (h < 0 || nranks < 1 || rank < 0 || rank >= nranks || !row0 || !row1) return SR_E_BADARG;
    *row0 = (int)((long long)h * rank / nranks);
    *row1 = (int)((long long)h * (rank + 1) / nranks);
    return SR_OK;
}
int sr_block_search_batch(sr_ctx* ctx, int n, const sr_image* left, const sr_image* right, const sr_search_params* p, double* const* disp_out, size_t out_pitch) { NOTYET(ctx, "sr_block_search_batch"); }
int sr_remove_outliers(sr_ctx* ctx, float* disp, int w, int h, size_t pitch, int ksize, float thr_front, float thr_back) { NOTYET(ctx, "sr_remove_outliers"); }
int sr_box_blur(sr_ctx* ctx, const float* src, float* dst, int w, int h, int ksize) { NOTYET(ctx, "sr_box_blur"); }
int sr_scale(sr_ctx* ctx, float* disp, int w, int h, size_t pitch, float f) { NOTYET(ctx, "sr_scale"); }
int sr_disparity_to_depth(sr_ctx* ctx, const float* disp, float* depth, int w, int h, float focal, float baseline) { NOTYET(ctx, "sr_disparity_to_depth"); }
int sr_backproject(sr_ctx* ctx, const uint8_t* bgr, size_t pitch, const float* depth, int w, int h, float fx, float fy, float cx, float cy, sr_vertex* out) { NOTYET(ctx, "sr_backproject"); }
int sr_mesh_faces(sr_ctx* ctx, const sr_vertex* v, int w, int h, float thr, uint32_t* tri, uint32_t* n_faces) { NOTYET(ctx, "sr_mesh_faces"); }
int sr_reconstruct(sr_ctx* ctx, const uint8_t* bgr, size_t pitch, const float* disp, int w, int h, int ksize, float thr_front, float thr_back, float focal, float baseline, float fx, float fy, float cx, float cy, float thr_mesh, float* disp_filtered, float* depth, sr_vertex* vertices, uint32_t* tri, uint32_t* n_faces) { NOTYET(ctx, "sr_reconstruct"); }
int sr_pair_reconstruct(sr_ctx* ctx, sr_pair* pair, int quantize_u8, int ksize, float thr_front, float thr_back, float focal, float baseline, float fx, float fy, float cx, float cy, float thr_mesh, float* disp_filtered, float* depth, sr_vertex* vertices, uint32_t* tri, uint32_t* n_faces) { NOTYET(ctx, "sr_pair_reconstruct"); }
int sr_write_mesh(sr_ctx* ctx, const sr_vertex* v, int w, int h, const char* filename, float thr) { NOTYET(ctx, "sr_write_mesh"); }
int sr_warp_nearest_f64(sr_ctx* ctx, const double* src, int sw, int sh, double* dst, int dw, int dh, const double M[9]) { NOTYET(ctx, "sr_warp_nearest_f64"); }
int sr_evaldisp(sr_ctx* ctx, const float* disp, const float* gt, const uint8_t* mask, int w, int h, float badthresh, float maxdisp, int rounddisp, sr_eval* out) { NOTYET(ctx, "sr_evaldisp"); }
int sr_write_pfm(const char* path, const float* data, int w, int h) { return SR_E_IO; }
int sr_write_ply(const char* path, const sr_vertex* v, int w, int h) { return SR_E_IO; }
}
