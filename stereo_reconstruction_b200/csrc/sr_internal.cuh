// sr_internal.cuh -- shared declarations of libsr_b200.so (sm_100a only).
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/sr_capi.h"

#define SR_NONE64 0xFFFFFFFFFFFFFFFFull

// Left padding (pixels) of the packed images: tiles start at column -16 (see search_tile.cu).
#define SR_PADL 32

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
};

// cached schedule of a tile-kernel launch (search_tile.cu: tile_work_list)
struct TileWork {
    static constexpr int NKEY = 12;
    int key[NKEY];
    int4* dev = nullptr;
    int n = 0;
    uint64_t stamp = 0;
};

struct sr_stage_pool;   // host threads for pageable <-> pinned staging copies (host_stage.cu)

// cudaFuncSetAttribute applies to the current device only: one flag per device for each kernel instantiation that
// raises its dynamic shared-memory limit (a process may hold contexts on several GPUs; setting twice is harmless).
struct DeviceOnce {
    std::atomic<uint64_t> mask{0};
    bool need(int dev) const { return ((mask.load(std::memory_order_acquire) >> (dev & 63)) & 1ull) == 0; }
    void done(int dev) { mask.fetch_or(1ull << (dev & 63), std::memory_order_release); }
};

struct sr_ctx {
    int device = 0;
    int sms = 148;                 // cudaDevAttrMultiProcessorCount of `device` (B200: 148), set by sr_ctx_create
    cudaStream_t stream = nullptr;
    cudaStream_t aux[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // 0/1: copy streams of the pipelines, 2/3: extra compute streams, 4: pack kernels (high priority)
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t kev0 = nullptr, kev1 = nullptr;  // around the dominant search kernel of the last run
    bool kev_valid = false;
    std::string err;
    uint64_t launches = 0;
    const char* last_kernel = "";
    DevBuf scratch[16];        // grow-only device scratch, reused across calls
    void* pinned[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t pinned_cap[4] = {0, 0, 0, 0};
    std::vector<sr_pair*> batch_pairs;
    std::vector<sr_pair*> peer_pairs;  // pairs with a gather window: sr_ctx_sync reports their protocol status
    std::vector<cudaEvent_t> pipe_ev;  // events of the row-chunk pipeline of sr_block_search
    sr_stage_pool* stage_pool = nullptr;
    std::vector<TileWork> tile_work;   // schedules of the tile kernel, by geometry
    uint64_t tile_work_clock = 0;
    bool more_chunks_follow = false;   // set by the row-chunk pipeline around all launches but the last chunk's
    bool chunked_smoothing = false;    // ... and while it runs the smoothing rule chunk by chunk (rows above are resolved)
    cudaEvent_t raster_wait = nullptr; // the raster pass of the current chunk waits for this event (previous chunk, other stream)
    void* nccl_comm = nullptr;
    int rank = 0, nranks = 1;
};

// Packed image: one uint32 per pixel (B | G<<8 | R<<16, top byte 0), SR_PADL pixels of left
// padding and right padding so that every tile / candidate load stays inside the row.
struct PackedImage {
    uint32_t* base = nullptr;  // allocation
    uint32_t* px = nullptr;    // pointer to pixel (0,0)
    int w = 0, h = 0;
    int pitch = 0;             // pixels per row (multiple of 32)
};

struct SearchGeom {
    int w1, h1, w2, h2, height;  // height = min(h1,h2)
    int block, half;             // blockSize_ and (block-1)/2
    int dlo, dhi;                // candidate range [dlo, dhi)
    int row0, row1;              // output rows to compute
};

struct sr_pair {
    int w1 = 0, h1 = 0, w2 = 0, h2 = 0;
    uint8_t *rawL = nullptr, *rawR = nullptr;  // BGR as uploaded, tight pitch 3*w
    PackedImage L, R;
    PackedImage Rs;            // right image packed with a column shift (Left variant through the tile kernel, search_tile.cu)
    PackedImage Lm, Rm;        // mirror-packed copies (stored[s] = image[w1-2-s]) for the left-referenced tile search
    int pad_disp = 0;          // right padding of L covers this many disparities
    // per-pixel records of the reference-side (output) image
    int ow = 0, oh = 0;        // output size (RIGHT/LINEAR: w2 x h2; LEFT: w1 x h1)
    uint32_t* planes = nullptr;           // tile kernel: per-pass key planes (ctx scratch), cost*64 + d_local
    int npass = 0, plane_dbase = 0;       // planes held (1 when the passes were folded into one plane), first candidate
    int key_shift = 6;                    // key = cost << key_shift | candidate index within the plane's range
    unsigned long long* best = nullptr;   // (cost << 32 | d) first minimum over the searched range
    unsigned long long* c0 = nullptr;     // cost at d == 0 (SR_NONE64 when not a candidate)
    bool c0_from_tile = false;            // the last tile search left the raw d == 0 costs in c0 as a u32 plane
    unsigned long long* cr = nullptr;     // wide-cost first minimum (direct kernels)
    int* dr = nullptr;
    int* nwin = nullptr;                  // window pixel count; -1 = skipped (black) pixel
    uint32_t* bits = nullptr;             // bit planes for the smoothing pass
    double* alt = nullptr;
    double* out = nullptr;                // result, ow x oh, tight
    float* outf = nullptr;                // float copy for the reconstruction pipeline
    short* out16 = nullptr;               // int16 wire format of the result (band gather)
    void* peer_win = nullptr;             // gather window over peer memory (capi_peer.cu)
    int peer_mode = 0;                    // 0 none, 1 root (owns the window), 2 attached through CUDA IPC, 3 attached in-process
    unsigned* peer_counter = nullptr;     // last-block counter of the push / collect kernels
    int* peer_status_h = nullptr;         // sticky status of the gather protocol: one word of mapped pinned host memory that the
    int* peer_status_d = nullptr;         // kernels write (time-out, non-integer map) and the host reads without a copy
    int wire_rank = 0;                    // set by sr_search_run_push around the run: the resolve pass writes the int16 wire
    unsigned wire_seq = 0;                // format straight into the root's window (rank, step) instead of the CV_64F map
    bool wire_done = false;               // ... and reports here that it did
    int* maxblock = nullptr;              // device scalar
    unsigned long long* evals = nullptr;  // device scalar
    size_t npx_cap = 0;
    int last_row0 = 0, last_row1 = 0;
    bool evals_valid = false;
    void* ncc_tab = nullptr;              // window-moment tables of the NCC search (search_ncc.cu)
    void* ncc_tab_m = nullptr;            // the same for the mirror-packed pair (left-referenced NCC search)
    const int* disp_left = nullptr;       // left-referenced map of the last lr_check run (= disp_left_own, or NULL)
    int* disp_left_own = nullptr;         // pair-owned buffer: searches on other pairs of the context cannot overwrite it
    SearchGeom last_geom;
};

int sr_fail(sr_ctx* ctx, int code, const char* fmt, ...);
int sr_cuda_check(sr_ctx* ctx, cudaError_t e, const char* what);
void* sr_scratch(sr_ctx* ctx, int slot, size_t bytes);  // nullptr on failure (error set)
void* sr_pinned(sr_ctx* ctx, int slot, size_t bytes);

// ---- host staging pool (host_stage.cu) ----
sr_stage_pool* sr_stage_pool_get(sr_ctx* ctx);   // created on first use; nullptr when no thread can be started
void sr_stage_pool_destroy(sr_stage_pool* p);
int sr_stage_threads(const sr_stage_pool* p);
void sr_stage_run(sr_stage_pool* p, void (*fn)(int t, int nthreads, void* arg), void* arg);   // returns at once
void sr_stage_wait(sr_stage_pool* p);
// thread t's share of rows [row0, row1) (row indices apply to both sides)
void sr_stage_stripe(char* dst, size_t dpitch, const char* src, size_t spitch, size_t rowbytes, int row0, int row1,
                     int t, int nthreads);
// blocking host copy of `rows` rows, striped over the pool when it is large enough to pay for the wake-up
int sr_stage_copy_rows(sr_ctx* ctx, void* dst, size_t dpitch, const void* src, size_t spitch, size_t rowbytes, int rows);

#define SR_CUDA(ctx, call)                                              \
    do {                                                                \
        cudaError_t e__ = (call);                                       \
        if (e__ != cudaSuccess) return sr_cuda_check(ctx, e__, #call);  \
    } while (0)
#define SR_LAUNCHED(ctx)                                                        \
    do {                                                                        \
        (ctx)->launches++;                                                      \
        cudaError_t e__ = cudaGetLastError();                                   \
        if (e__ != cudaSuccess) return sr_cuda_check(ctx, e__, "kernel launch"); \
    } while (0)

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// ---- launchers implemented in the .cu files (all asynchronous on ctx->stream) ----
int sr_launch_pack(sr_ctx* ctx, const uint8_t* raw, size_t raw_pitch, int w, int h, PackedImage* img);
// rows [y0, y1) only (raw points at row 0 of the device copy)
// mirror packing about the left image's width: dst[s] = src[wref-2-s] (0 outside the source image)
int sr_launch_pack_mirror(sr_ctx* ctx, const uint8_t* raw, size_t raw_pitch, int w, int h, int wref, PackedImage* img);
int sr_alloc_packed(sr_ctx* ctx, PackedImage* img, int w, int h, int padr);
int sr_launch_pack_rows(sr_ctx* ctx, const uint8_t* raw, size_t raw_pitch, int w, int y0, int y1, PackedImage* img);
// the same for two images in one launch (rows [ya0,ya1) of a, [yb0,yb1) of b)
int sr_launch_pack_rows2(sr_ctx* ctx, const uint8_t* rawa, size_t pitcha, int wa, int ya0, int ya1, PackedImage* imga,
                         const uint8_t* rawb, size_t pitchb, int wb, int yb0, int yb1, PackedImage* imgb);

// Fast path: fixed window, SSD/SAD, first minimum over [dlo,dhi) -> pair->best.
// Returns 1 when it handled the request, 0 when the shape is not supported (caller falls back).
int sr_launch_search_tile(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, int cost, int left_ref, int want_c0 = 0);
// Left variant's unbiased winners through the tile kernel (odd blocks 3..17): key planes over the index i = *off - d
int sr_launch_search_tile_leftvariant(sr_ctx* ctx, sr_pair* pair, int block, int max_disp, int* off);
// The same in ONE launch over n <= 16 pairs of one shape (batches of small pairs: more work items per band).
int sr_launch_search_tile_multi(sr_ctx* ctx, sr_pair* const* pairs, int n, const SearchGeom& g, int cost);
// Right-variant search of n pairs of one shape with one tile launch (plain result, or sub-pixel without LR check);
// 0 when the request needs the per-pair path (capi_search.cu)
int sr_search_run_multi(sr_ctx* ctx, sr_pair* const* pairs, int n, const sr_search_params* p);
// Universal path: per-pixel windows (var_block), any size; wide costs -> c0/cr/dr/nwin.
int sr_launch_search_direct(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, int cost, int var_block,
                            double thres, int want_c0);
// cost at d == 0 for the fixed window -> pair->c0 as a u32 plane (resolve: have_c0 == 2)
int sr_launch_cost0(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, int cost);
// best/c0 (tile path) or c0/cr/dr/nwin (direct path) -> pair->out, reference rules.
int sr_launch_resolve_right(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, double smooth, int from_direct,
                            int have_c0);
int sr_launch_search_left(sr_ctx* ctx, sr_pair* pair, int block, int max_disp, double smooth);
int sr_launch_search_linear(sr_ctx* ctx, sr_pair* pair, double smooth);
// cv::warpPerspective(..., INTER_NEAREST) on device buffers (misc.cu); M as passed to warpPerspective
int sr_launch_warp_nearest(sr_ctx* ctx, const double* d_src, int sw, int sh, double* d_dst, int dw, int dh, const double M[9]);
// cv::warpPerspective(..., INTER_LINEAR, BORDER_CONSTANT 0) on CV_8UC3 device buffers (misc.cu), rectification.cpp:486-493
int sr_launch_warp_linear_bgr(sr_ctx* ctx, const uint8_t* d_src, int sw, int sh, size_t spitch, uint8_t* d_dst, int dw, int dh,
                              size_t dpitch, const double M[9]);
// band gather over peer memory (capi_peer.cu)
int sr_peer_check(sr_ctx* ctx, sr_pair* pair);                 // SR_OK, or the sticky protocol error of this pair
int sr_peer_wait_ack(sr_ctx* ctx, sr_pair* pair, unsigned seq); // stream-ordered: the root has consumed step seq-1
short* sr_peer_window16(sr_pair* pair);                        // int16 map of the window this pair is attached to / owns
volatile unsigned* sr_peer_flags(sr_pair* pair);
int sr_launch_count_evals(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g);
// NCC through sliding sums: interior pixels -> per-pass (score, d) planes (search_ncc.cu)
int sr_launch_search_ncc(sr_ctx* ctx, sr_pair* pair, const SearchGeom& g, float** pscore, int** pdisp, void** tab);
bool sr_ext_subpix_out_applies(const sr_search_params* p);
int sr_launch_ext_subpix_out(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p, int row0, int row1);
int sr_launch_ext_subpix_out_multi(sr_ctx* ctx, sr_pair* const* pairs, int n, const sr_search_params* p, int row0, int row1);
// NCC / sub-pixel / LR-check extension path (search_ext.cu)
int sr_search_run_ext(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p, int row0, int row1);
