// recon.cuh -- device-buffer launchers of the Reconstruction kernels (recon.cu).
#pragma once
#include "sr_internal.cuh"
// known_u8 != 0: the caller guarantees integer values 0..255 (the 8-bit hand-off), the exactness check is skipped
int recon_box_blur(sr_ctx* ctx, const float* d_src, float* d_dst, int w, int h, int k, int known_u8 = 0);
int recon_remove_outliers(sr_ctx* ctx, float* d_disp, int w, int h, int k, float thr_front, float thr_back, int known_u8 = 0);
int recon_scale(sr_ctx* ctx, float* d_disp, size_t n, float f);
int recon_depth_backproject(sr_ctx* ctx, const float* d_disp, const float* d_depth_in, const uint8_t* d_bgr, size_t bgr_pitch,
                            int w, int h, float focal, float baseline, float fx, float fy, float cx, float cy,
                            float* d_depth_out, sr_vertex* d_vert);
// depth + vertices + the face-validity flags of WriteMesh in one pass (the vertices are not read back); follow with
// recon_mesh_faces(ctx, NULL, ...) for the face list
int recon_depth_backproject_faces(sr_ctx* ctx, const float* d_disp, const uint8_t* d_bgr, size_t bgr_pitch, int w, int h,
                                  float focal, float baseline, float fx, float fy, float cx, float cy, float thr,
                                  float* d_depth_out, sr_vertex* d_vert);
// face list in the reference's order; d_tri holds up to 6*(w-1)*(h-1) indices; *n_faces on host (synchronises).
// d_vert == NULL: the validity flags were left by recon_depth_backproject_faces on the same w x h map.
int recon_mesh_faces(sr_ctx* ctx, const sr_vertex* d_vert, int w, int h, float thr, uint32_t* d_tri, uint32_t* n_faces);
int recon_quantize(sr_ctx* ctx, const double* d_disp, float* d_out, size_t n, int quantize_u8);
// main.cpp:41-65 on a CV_64F map with the 8-bit hand-off and removeDisparityOutliers(k), integer-exact (recon.cu):
// quantise + row sums, int32 column prefixes, then blurred value + outlier rule + depth + vertex + face flags in one pass.
// d_disp_out (optional) receives the filtered float map.  Follow with recon_mesh_faces(ctx, NULL, ...).
bool recon_u8_pipeline_applies(int w, int h, int k);
int recon_u8_pipeline(sr_ctx* ctx, const double* d_disp64, const uint8_t* d_bgr, size_t bgr_pitch, int w, int h, int k, float thr_front,
                      float thr_back, float focal, float baseline, float fx, float fy, float cx, float cy, float thr,
                      float* d_disp_out, float* d_depth_out, sr_vertex* d_vert);
