/*
 * sr_capi.h -- C-ABI of libsr_b200.so: the B200 (sm_100a) implementation of the dense stereo
 * hot path of Bizilizi/stereo_reconstruction (src/WindowSearch + src/Reconstruction).
 *
 * This is the drop-in boundary.  The reference has no FFI: its "operator API" for this path
 * is the C++ header surface
 *     src/WindowSearch/BlockSearch.h:9-46      (class BlockSearch)
 *     src/WindowSearch/LinearSearch.h:9-20     (class LinearSearch)
 *     src/Reconstruction/reconstruction.h:13-34 (struct Vertex + free functions)
 * The header-compatible C++ shims under stereo_reconstruction_b200/shim/ forward those classes
 * and functions to the entry points below; each entry point cites the reference interface
 * it replaces.  Plain pointers and sizes only.  All paths run on the GPU; there is no CPU
 * fallback: every call fails with SR_E_CUDA when no device / kernel image is available.
 *
 * Threading: a sr_ctx owns one CUDA stream and its device/pinned buffers; it is not
 * thread-safe, distinct contexts are independent.  Host pointers may be pageable or pinned
 * (sr_host_alloc); pinned ones are DMA'd directly.
 */
#ifndef SR_CAPI_H
#define SR_CAPI_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define SR_OK 0
#define SR_E_BADARG (-1)
#define SR_E_SHAPE (-2)  /* the reference would throw cv::Exception / read out of bounds */
#define SR_E_CUDA (-3)
#define SR_E_NCCL (-4)
#define SR_E_IO (-5)

typedef struct sr_ctx sr_ctx;

/* CV_8UC3 interleaved BGR image, row pitch in bytes (cv::Mat::data / ::step). */
typedef struct {
    const uint8_t* bgr;
    int w, h;
    size_t pitch;
} sr_image;

enum { SR_VARIANT_RIGHT = 0, SR_VARIANT_LEFT = 1, SR_VARIANT_LINEAR = 2 };
enum { SR_COST_SSD_L2 = 0, SR_COST_SAD = 1, SR_COST_NCC = 2 };

/* Parameters of one search.  For the reference-exact variants (cost == SR_COST_SSD_L2,
 * subpixel == 0, lr_check == 0) the fields map 1:1 on the reference's arguments:
 *   block, min_disp, max_disp  BlockSearch::BlockSearch(..., blockSize, minDisparity, maxDisparity)
 *                              BlockSearch.cpp:6-13
 *   smooth, var_block, thres   computeDisparityMapRight(smoothFactor, varBlock, thres) :88
 *                              computeDisparityMapLeft(smoothFactor) :24
 *                              LinearSearch::computeDisparityMap(smoothFactor) LinearSearch.cpp:10
 * Extensions (no reference counterpart): cost = SAD / NCC, subpixel, lr_check. */
typedef struct {
    int variant;
    int block;
    int min_disp, max_disp;
    double smooth;
    int var_block;
    double thres;
    int cost;
    int subpixel;
    int lr_check;
} sr_search_params;

/* struct Vertex, reconstruction.h:13-22: Eigen Vector4f (16 B) + Vector4uc (4 B) + 12 B pad. */
typedef struct {
    float pos[4];
    uint8_t rgba[4];
    uint8_t pad_[12];
} sr_vertex;

/* ---- context ---------------------------------------------------------------------------- */
int sr_ctx_create(int device, sr_ctx** out);
void sr_ctx_destroy(sr_ctx* ctx);
/* Message of the last failing call on ctx (or of sr_ctx_create when ctx == NULL). */
const char* sr_last_error(const sr_ctx* ctx);
const char* sr_version(void);
/* Pinned host memory for callers that want direct DMA (e.g. a cv::MatAllocator). */
int sr_host_alloc(void** p, size_t bytes);
int sr_host_free(void* p);
/* Pins an existing host buffer (cv::Mat data, std::vector storage, ...) in place so that uploads / downloads
 * are direct DMA and sr_block_search can overlap them with the search; undo with sr_host_unregister. */
int sr_host_register(void* p, size_t bytes);
int sr_host_unregister(void* p);
/* The context's CUDA stream (cudaStream_t) and a CUDA-event stopwatch on it. */
void* sr_ctx_stream(sr_ctx* ctx);
int sr_ctx_sync(sr_ctx* ctx);
int sr_timer_start(sr_ctx* ctx);
int sr_timer_stop(sr_ctx* ctx, float* ms);          /* records, synchronises, returns elapsed ms */
/* Measurement helpers (bench.py): write a buffer larger than L2 on the context stream; device time
 * of the dominant search kernel of the last sr_search_run; live integer-issue peak in lane-ops/s
 * (dual = VABSDIFF4 + IDP.4A on both pipes, single = INT pipe alone). */
int sr_l2_flush(sr_ctx* ctx);
int sr_search_kernel_ms(sr_ctx* ctx, float* ms);
int sr_int_peak(sr_ctx* ctx, double* dual_ops_per_s, double* single_ops_per_s);
/* Number of kernels this library launched on ctx since creation. */
uint64_t sr_kernel_launches(const sr_ctx* ctx);
/* Name of the search kernel family the last sr_search_* call dispatched to
 * ("tile", "direct", "linear", ...), for tests and profiles. */
const char* sr_last_kernel(const sr_ctx* ctx);

/* ---- WindowSearch: one-shot host-pointer calls -------------------------------------------- */

/* Replaces BlockSearch(left,right,block,minD,maxD).computeDisparityMapRight / Left
 * (BlockSearch.cpp:88-179 / :24-86) and LinearSearch(left,right).computeDisparityMap
 * (LinearSearch.cpp:10-59), selected by p->variant.
 *   disp_out   CV_64F map, h2 x w2 (RIGHT, LINEAR) or h1 x w1 (LEFT), row pitch in bytes
 *   max_block  (optional) the "max block size" the reference prints (BlockSearch.cpp:177)
 * With p->subpixel the map holds d* + parabola offset; with p->lr_check inconsistent pixels
 * are set to 0.  Errors: SR_E_SHAPE where the reference throws or reads out of bounds. */
int sr_block_search(sr_ctx* ctx, const sr_image* left, const sr_image* right,
                    const sr_search_params* p, double* disp_out, size_t out_pitch, int* max_block);

/* n independent pairs of identical shape (config 5: batch sharded by pair); pairs are
 * pipelined H2D / compute / D2H on internal streams.  disp_out[i] as in sr_block_search. */
int sr_block_search_batch(sr_ctx* ctx, int n, const sr_image* left, const sr_image* right,
                          const sr_search_params* p, double* const* disp_out, size_t out_pitch);

/* ---- WindowSearch: device-resident form (what the one-shot call is made of) ------------- */
typedef struct sr_pair sr_pair;
int sr_pair_create(sr_ctx* ctx, int w1, int h1, int w2, int h2, sr_pair** out);
void sr_pair_destroy(sr_ctx* ctx, sr_pair* pair);
int sr_pair_upload(sr_ctx* ctx, sr_pair* pair, const sr_image* left, const sr_image* right);
/* The same with the rectifying warp on the device: the un-rectified images go up, cv::warpPerspective(image, H) of
 * rectification.cpp:486-493 (sr_warp_linear_bgr) fills the pair, whose sizes are the rectified sizes. */
int sr_pair_upload_warped(sr_ctx* ctx, sr_pair* pair, const sr_image* left, const double H[9], const sr_image* right, const double Hp[9]);
/* Kernels only, asynchronous on the context stream; result stays in HBM.
 * rows [row0,row1) of the output are computed (pass 0,h for the full image; used for
 * row-band sharding -- a band needs smooth == 1). */
int sr_search_run(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p, int row0, int row1);
int sr_search_download(sr_ctx* ctx, sr_pair* pair, double* disp_out, size_t out_pitch, int row0, int row1);
/* The same for n uploaded pairs (config 5: batches of small pairs): pairs of one shape share ONE launch of the fused
 * cost + WTA kernel in groups of up to 16 (a 1280x720 / 128-disparity pair alone gives 148 SMs too few work items),
 * where the request allows it (Right variant, SSD / SAD, plain or sub-pixel result); results as sr_search_run per pair. */
int sr_search_run_batch(sr_ctx* ctx, sr_pair* const* pairs, int n, const sr_search_params* p);
/* One rank's share of a row-band sharded pair, from host images (north_star: "by row bands with halo
 * rows for single very large pairs"): uploads only the image rows that output rows [row0,row1) and their
 * window halo read, searches the band and -- when disp_out is not NULL -- downloads it to rows
 * [row0,row1) of disp_out; with pinned host buffers upload / search / download overlap by row chunks.
 * disp_out == NULL leaves the band in HBM (sr_search_result_dev) for the device-side gather to rank 0.
 * Blocking.  The band's result equals rows [row0,row1) of sr_block_search on the whole pair. */
int sr_pair_search_band(sr_ctx* ctx, sr_pair* pair, const sr_image* left, const sr_image* right,
                        const sr_search_params* p, int row0, int row1, double* disp_out, size_t out_pitch);
/* Device pointer (double*, tightly packed rows of the output width) of the last result. */
void* sr_search_result_dev(sr_ctx* ctx, sr_pair* pair);
/* int16 wire format of integer disparity maps for the device-side band gather ("16.6 MB as u16
 * if the integer-disparity wire format is used and widened on rank 0"; signed because the reference stores
 * -x where no candidate survives, BlockSearch.cpp:111,174).  sr_search_pack16 narrows rows [row0,row1) of the
 * CV_64F result into the pair's int16 buffer (SR_E_BADARG if a value is not an integer in int16 range, e.g. a
 * sub-pixel map), sr_search_unpack16 widens rows back into the CV_64F result; both asynchronous on the context
 * stream.  sr_search_result16_dev returns the int16 buffer (w2 x h2, tight). */
int sr_search_pack16(sr_ctx* ctx, sr_pair* pair, int row0, int row1);
int sr_search_unpack16(sr_ctx* ctx, sr_pair* pair, int row0, int row1);
void* sr_search_result16_dev(sr_ctx* ctx, sr_pair* pair);
/* Band gather over peer memory (NVLink / NVSwitch), the gather of SURVEY.md 8e without a collective library on
 * the data path.  Rank 0 (the root) exports a gather window of its pair as a CUDA IPC handle (64 bytes; carry
 * them to the other ranks by any means, e.g. torch.distributed.broadcast_object_list); the other ranks attach
 * to it.  Per step `seq` (1, 2, 3, ... -- the same number on every rank): rank r > 0 calls
 * sr_search_push_band after its band search -- one kernel narrows rows [row0,row1) to the int16 wire format,
 * stores them into the root's HBM and publishes the step in the root's flag word of rank r; the root calls
 * sr_search_collect_bands after its own band search -- one kernel widens every other rank's rows (bands of
 * sr_band_rows(h, r, nranks)) into its CV_64F result as their flags arrive and acknowledges the step, which
 * lets the pushers overwrite the window.  All asynchronous on the context stream; a band that does not arrive in
 * time sets the pair's sticky status (see sr_search_run_push) instead of hanging the GPU.
 * sr_pair_peer_attach_local attaches to a root pair of the same process (single-GPU tests of the protocol). */
#define SR_PEER_HANDLE_BYTES 64
int sr_pair_peer_export(sr_ctx* ctx, sr_pair* pair, void* handle);
int sr_pair_peer_attach(sr_ctx* ctx, sr_pair* pair, const void* handle);
int sr_pair_peer_attach_local(sr_ctx* ctx, sr_pair* pair, sr_pair* root);
int sr_pair_peer_detach(sr_ctx* ctx, sr_pair* pair);
int sr_search_push_band(sr_ctx* ctx, sr_pair* pair, int row0, int row1, int rank, unsigned seq);
/* The band search of a rank > 0 with the gather fused into it: sr_search_run on rows [row0,row1) whose resolve pass
 * writes the int16 wire format straight into the root's window and publishes step `seq` -- the CV_64F band is never
 * formed and no push kernel runs (falls back to sr_search_run + sr_search_push_band where the fused pass does not
 * apply).  Every export / attach starts a new generation of the protocol (steps restart at 1).  A band or an
 * acknowledgement that does not arrive within SR_PEER_TIMEOUT_MS (default 10000) is never published / acknowledged;
 * the failure is sticky and reported by every later sr_search_run_push / sr_search_push_band /
 * sr_search_collect_bands / sr_search_download / sr_ctx_sync / sr_pair_peer_status on that rank. */
int sr_search_run_push(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p, int row0, int row1, int rank, unsigned seq);
int sr_pair_peer_status(sr_ctx* ctx, sr_pair* pair);
int sr_search_collect_bands(sr_ctx* ctx, sr_pair* pair, int nranks, unsigned seq);
int sr_search_max_block(sr_ctx* ctx, sr_pair* pair, int* max_block);
/* The left-referenced integer disparity map (w1 x h1, int32) that the last run with lr_check = 1
 * computed for the consistency test (extension); SR_E_BADARG when there is none. */
int sr_search_left_disparity(sr_ctx* ctx, sr_pair* pair, int32_t* disp_left, size_t out_pitch);
/* Effective (reference-semantics) number of (pixel, candidate) evaluations of the last run
 * restricted to its rows: sum over non-black pixels of the candidates before the break
 * (BlockSearch.cpp:147-149). */
int sr_search_effective_evals(sr_ctx* ctx, sr_pair* pair, uint64_t* evals);

/* ---- Reconstruction ------------------------------------------------------------------------ */
/* removeDisparityOutliers(disp, k, thrFront, thrBack), reconstruction.cpp:5-18.  In place,
 * CV_32FC1 h x w, row pitch in bytes. */
int sr_remove_outliers(sr_ctx* ctx, float* disp, int w, int h, size_t pitch, int ksize,
                       float thr_front, float thr_back);
/* cv::blur(disp, Size(k,k)) alone (reconstruction.cpp:8), tightly packed output. */
int sr_box_blur(sr_ctx* ctx, const float* src, float* dst, int w, int h, int ksize);
/* The same blur for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO counterpart in the reference.  d_src and d_dst are
 * DEVICE pointers (tightly packed h x w floats, d_dst != d_src); the only intermediate, the fp64 row sums, lives in a
 * file-scope __device__ array of SR_BLUR_STATIC_MAX_PIXELS doubles (8 MiB, held by every GPU that loads the library
 * until exit).  The call makes no cudaMalloc / cudaFree / allocating library call, enqueues two kernels on `stream`
 * (a cudaStream_t; NULL = the context's stream) and returns without synchronising; the context must have been
 * created beforehand, and the first call per device also looks up the array's device address (cudaGetSymbolAddress),
 * so make that call outside a capture.  w * h > SR_BLUR_STATIC_MAX_PIXELS -> SR_E_SHAPE (use sr_box_blur).  All calls
 * share the one array: it keeps the last call's contents (nothing in it needs clearing, every value read is written
 * first), and two calls that overlap on the GPU -- different streams, or contexts, of one device -- race.  Results
 * are bit-identical to sr_box_blur. */
#define SR_BLUR_STATIC_MAX_PIXELS (1u << 20)
int sr_box_blur_static(sr_ctx* ctx, const float* d_src, float* d_dst, int w, int h, int ksize, void* stream);
/* scaleDisparityMap(disp, f), reconstruction.cpp:21-27. */
int sr_scale(sr_ctx* ctx, float* disp, int w, int h, size_t pitch, float f);
/* convertDisparityToDepth(disp, focal, baseline), reconstruction.cpp:30-43. */
int sr_disparity_to_depth(sr_ctx* ctx, const float* disp, float* depth, int w, int h,
                          float focal, float baseline);
/* Back-projection loop of reconstruction(), reconstruction.cpp:169-196.  fx,fy,cx,cy =
 * K(0,0), K(1,1), K(0,2), K(1,2) of the (column-major) Eigen Matrix3f. */
int sr_backproject(sr_ctx* ctx, const uint8_t* bgr, size_t pitch, const float* depth, int w, int h,
                   float fx, float fy, float cx, float cy, sr_vertex* out);
/* Face list of WriteMesh in the reference's order, reconstruction.cpp:92-113
 * (CheckTriangularValidity :46-69).  tri holds up to 6*(w-1)*(h-1) indices. */
int sr_mesh_faces(sr_ctx* ctx, const sr_vertex* v, int w, int h, float thr, uint32_t* tri, uint32_t* n_faces);
/* Fused device pipeline of main.cpp:53-65 on a float disparity map: [outlier removal ->]
 * depth -> back-projection -> face list, one upload, no intermediate host round trip.
 * ksize <= 0 skips removeDisparityOutliers.  Any output pointer may be NULL. */
int sr_reconstruct(sr_ctx* ctx, const uint8_t* bgr, size_t pitch, const float* disp, int w, int h,
                   int ksize, float thr_front, float thr_back, float focal, float baseline,
                   float fx, float fy, float cx, float cy, float thr_mesh,
                   float* disp_filtered, float* depth, sr_vertex* vertices, uint32_t* tri, uint32_t* n_faces);
/* Search result (already in HBM) -> [u8 quantisation of main.cpp:41-50] -> the pipeline above,
 * with no host round trip of the disparity map.  quantize_u8 != 0 reproduces the 8-bit PNG
 * hand-off (rint half-even, saturate). */
int sr_pair_reconstruct(sr_ctx* ctx, sr_pair* pair, int quantize_u8,
                        int ksize, float thr_front, float thr_back, float focal, float baseline,
                        float fx, float fy, float cx, float cy, float thr_mesh,
                        float* disp_filtered, float* depth, sr_vertex* vertices, uint32_t* tri, uint32_t* n_faces);
/* The same with the inverse warp of ImageRectifier::computeDisparityMapRight in between (rectification.cpp:82-87 ->
 * main.cpp:41-65): search result (rectified frame, in HBM) -> cv::warpPerspective(disp, M, Size(dw, dh),
 * INTER_NEAREST) with M = H_.inv() as the reference passes it -> [u8 hand-off] -> outliers -> depth -> vertices ->
 * faces, all in HBM.  `color` is the un-rectified right image (dw x dh BGR, main.cpp:61 imageRight) whose pixels
 * colour the vertices; NULL uses the pair's own right image (then dw x dh must be its size).  disp_warped (optional,
 * CV_64F dw x dh, tight) receives the warped map, i.e. what computeDisparityMapRight returns. */
int sr_pair_reconstruct_warp(sr_ctx* ctx, sr_pair* pair, const double M[9], int dw, int dh, const sr_image* color,
                             int quantize_u8, int ksize, float thr_front, float thr_back, float focal, float baseline,
                             float fx, float fy, float cx, float cy, float thr_mesh, double* disp_warped,
                             float* disp_filtered, float* depth, sr_vertex* vertices, uint32_t* tri, uint32_t* n_faces);
/* WriteMesh(vertices, w, h, filename, thr), reconstruction.cpp:72-149: validity on the GPU,
 * COFF text on the host.  Returns SR_E_IO where the reference returns false. */
int sr_write_mesh(sr_ctx* ctx, const sr_vertex* v, int w, int h, const char* filename, float thr);

/* ---- neighbours of the path ------------------------------------------------- */
/* cv::warpPerspective(src, dst, M, size, INTER_NEAREST) on CV_64F, rectification.cpp:82-87. */
int sr_warp_nearest_f64(sr_ctx* ctx, const double* src, int sw, int sh, double* dst, int dw, int dh,
                        const double M[9]);
/* cv::warpPerspective(image, dst, M, size) with the default INTER_LINEAR / BORDER_CONSTANT(0) on CV_8UC3: the rectifying
 * warp of the two input images, rectification.cpp:486-493 (M = H_ / Hp_ as passed; inverted inside like OpenCV does).
 * OpenCV's fixed-point bilinear path (1/32-pixel positions, 15-bit weights), bit-identical to cv2.warpPerspective.
 * Row pitches in bytes.  SR_E_SHAPE for sources of 32768 pixels or more per side (cv::remap's short coordinates). */
int sr_warp_linear_bgr(sr_ctx* ctx, const uint8_t* src, int sw, int sh, size_t src_pitch, uint8_t* dst, int dw, int dh,
                       size_t dst_pitch, const double M[9]);
/* evaldisp, utils.cpp:123-168.  Counts exact; serr is a double-precision sum. */
typedef struct { int n, bad, invalid; double serr; float bad_pct, invalid_pct, total_bad_pct, avg_err; } sr_eval;
int sr_evaldisp(sr_ctx* ctx, const float* disp, const float* gt, const uint8_t* mask, int w, int h,
                float badthresh, float maxdisp, int rounddisp, sr_eval* out);
/* computeAverageDisparity(disparityMap), utils.cpp:8-20: one float accumulator over the non-infinite pixels in raster
 * order, divided by their count -- the order of the additions is reproduced, so the result is bit-identical.
 * valid (optional) receives the count. */
int sr_average_disparity(sr_ctx* ctx, const float* disp, int w, int h, size_t pitch, float* avg, int* valid);
/* Middlebury PFM / ASCII PLY writers (host). */
int sr_write_pfm(const char* path, const float* data, int w, int h);
int sr_write_ply(const char* path, const sr_vertex* v, int w, int h);
/* Middlebury PFM reader with the semantics of the SDK call in DataLoader::loadDisparityMatrices
 * (data_loader.cpp:110-125): rows returned top to bottom, the scale's sign selects the byte order and its magnitude is
 * not applied, unknown pixels stay +inf.  sr_pfm_info gives the size to allocate. */
int sr_pfm_info(const char* path, int* w, int* h);
int sr_read_pfm(const char* path, float* out, int w, int h);
/* calib.txt: cam0 / cam1 (row-major 3x3) parsed like DataLoader::readCameraMatrices (data_loader.cpp:141-164); the
 * key=value lines (ndisp is the maxDisparity of BASELINE.json configs 1-2) into `calib` when not NULL. */
typedef struct { double doffs, baseline, dyavg, dymax; int width, height, ndisp, isint, vmin, vmax; } sr_calib;
int sr_read_calib(const char* path, float cam0[9], float cam1[9], sr_calib* calib);

/* ---- multi-GPU (config 4) ------------------------------------------------------------------ */
/* Balanced band of output rows for `rank` of `nranks`.  Ranks run sr_search_run on their band and
 * the caller gathers the bands to rank 0 over NCCL (stereo_reconstruction_b200/sharding.py). */
int sr_band_rows(int h, int rank, int nranks, int* row0, int* row1);

#ifdef __cplusplus
}
#endif
#endif
