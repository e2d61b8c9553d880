// capi_search.cu -- WindowSearch entry points of the C-ABI (dispatch, download, one-shot calls).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <thread>
#include <vector>
#include "sr_internal.cuh"

static int rows_fit(int h1, int h2, int half) {
    // BlockSearch.cpp:151-154: the left ROI [y-up, y+down) must lie inside the left image,
    // otherwise the reference throws cv::Exception.
    int height = h1 < h2 ? h1 : h2;
    for (int y = (height - half - 2 > 0 ? height - half - 2 : 0); y < height; y++) {
        int down = h2 - y - 1 < half ? h2 - y - 1 : half;
        if (y + down > h1) return 0;
    }
    return 1;
}

static int validate(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p) {
    if (!ctx || !pair || !p) return sr_fail(ctx, SR_E_BADARG, "NULL argument");
    if (p->block < 1) return sr_fail(ctx, SR_E_BADARG, "block size %d < 1", p->block);
    if (p->cost != SR_COST_SSD_L2 && p->cost != SR_COST_SAD && p->cost != SR_COST_NCC)
        return sr_fail(ctx, SR_E_BADARG, "unknown cost %d", p->cost);
    if (p->variant != SR_VARIANT_RIGHT && p->variant != SR_VARIANT_LEFT && p->variant != SR_VARIANT_LINEAR)
        return sr_fail(ctx, SR_E_BADARG, "unknown variant %d", p->variant);
    return SR_OK;
}

// ---- row-chunk pipeline of the one-shot call ------------------------------------------------------
// H2D + pack of chunk c+1 (copy stream 0), search + resolve of chunk c (compute stream) and D2H of
// chunk c-1 (copy stream 1) overlap; only the first upload and the last download are exposed.
// Applies where output rows are independent (Right variant, tile kernel, no raster smoothing rule)
// Pageable host buffers pass through pinned staging, filled and drained chunk by chunk by the staging threads.
// ---- int16 wire format of integer disparity maps (band gather) ------------------------------------
__global__ void pack16_kernel(const double* __restrict__ src, short* __restrict__ dst, size_t n, int* __restrict__ status) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = src[i];
    int q = (int)v;
    if ((double)q != v || q < -32768 || q > 32767) { *status = SR_E_BADARG; q = 0; }
    dst[i] = (short)q;
}
__global__ void unpack16_kernel(const short* __restrict__ src, double* __restrict__ dst, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = (double)src[i];
}

static bool host_pinned(const void* p) {
    cudaPointerAttributes attr;
    bool pinned = cudaPointerGetAttributes(&attr, p) == cudaSuccess &&
                  (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    return pinned;
}

static int pipeline_chunks(const sr_pair* pair, const sr_search_params* p, int rows = -1) {
    if (rows < 0) rows = pair->h2;
    const char* env = getenv("SR_PIPE_CHUNKS");   // tuning / test override
    int c = env ? atoi(env) : 0;
    if (c <= 0) {
        // large pairs: ~0.8 M right pixels per chunk (4K pair, three compute streams: 6 chunks 2.46 ms, 8 -> 2.36,
        // 10 -> 2.32, 12 -> 2.31).  Medium and small pairs pay for overlap too -- measured from and to pinned buffers
        // (tools/trace_e2e_c1.py), 1 / 2 / 3 / 4 chunks: 1436x992 (145, block 17) 0.82 / 0.67 / 0.59 / 0.61 ms;
        // 1920x1080 (256) 0.79 / 0.79 / 0.67 / 0.72; 1280x720 0.42 / 0.36 / 0.32 / 0.32; 900x750 0.52 / 0.43 / 0.43 / 0.45;
        // 640x480 0.20 / 0.17 / 0.17 / 0.19 -- so three chunks from ~0.8 M pixels, two below.
        double px = (double)pair->w2 * rows;
        c = (int)(px / 0.8e6);
        if (c < 3) c = px >= 0.8e6 ? 3 : (px >= 0.15e6 ? 2 : 1);
        // smoothing rule (chunks resolve in row order, see block_search_pipelined): more, shorter chunks keep less of the
        // serial raster pass exposed -- 1436x992, smooth 0.9: 1 / 2 / 3 / 4 / 6 chunks 0.98 / 0.77 / 0.70 / 0.71 / 0.66 ms
        if (p->smooth != 1.0 && p->min_disp == 0 && px >= 0.8e6 && c < 6) c = 6;
        if (c > 12) c = 12;
    }
    int maxc = rows / (4 * p->block + 16);
    if (c > maxc) c = maxc;
    return c < 1 ? 1 : c;
}

// First row of chunk c of n (c == n: the end) when `rows` rows are cut into n weighted chunks: a short first
// chunk starts the search early; the last two chunks shrink (their downloads queue behind each other once the
// search is over, so little data should be left by then); the chunks in between are equal.
static int chunk_start(int rows, int n, int c, int half) {
    if (c <= 0) return 0;
    if (c >= n) return rows;
    static const char* e0 = getenv("SR_PIPE_EDGE");
    static const char* e1 = getenv("SR_PIPE_EDGE_LAST");
    static const char* e2 = getenv("SR_PIPE_EDGE_PEN");
    double w_first = e0 ? atof(e0) : 0.45, w_last = e1 ? atof(e1) : 0.3, w_pen = e2 ? atof(e2) : 0.6;
    if (n < 4) { w_first = w_last = w_pen = 1.0; }
    auto weight = [&](int i) { return i == 0 ? w_first : i == n - 1 ? w_last : i == n - 2 ? w_pen : 1.0; };
    double total = 0, upto = 0;
    for (int i = 0; i < n; i++) { total += weight(i); if (i < c) upto += weight(i); }
    int r = (int)(rows * upto / total + 0.5);
    int minrows = 2 * half + 2;
    if (r < c * minrows) r = c * minrows;       // every chunk keeps at least a window's worth of rows
    if (r > rows - (n - c) * minrows) r = rows - (n - c) * minrows;
    return r;
}

static bool can_pipeline(sr_ctx* ctx, sr_pair* pair, const sr_image* left, const sr_image* right,
                         const sr_search_params* p, const double* disp_out) {
    if (p->variant != SR_VARIANT_RIGHT || p->var_block || p->cost == SR_COST_NCC || p->subpixel || p->lr_check) return false;
    if (p->min_disp < 0 || p->block < 3 || (p->block - 1) / 2 > 8) return false;
    // (the smoothing rule is a raster dependency between rows: its chunks resolve in order, see block_search_pipelined)
    if (pair->w2 == 1 || pair->h2 == 1) return false;
    if (!rows_fit(pair->h1, pair->h2, (p->block - 1) / 2)) return false;
    if (pipeline_chunks(pair, p) < 2) return false;
    return true;   // pageable buffers go through the staging threads (host_stage.cu)
}

// rows [band0, band1) of the output only (row-band sharding: a rank uploads just the image rows its band
// and the window halo need); disp_out == nullptr leaves the band in HBM (sr_search_result_dev) for a
// device-side gather.

// Host side of the pipeline when a buffer is pageable (a cv::Mat): the staging threads copy the image rows
// of chunk c into pinned staging before the main thread queues that chunk's H2D, and copy finished output
// chunks from pinned staging to the caller's buffer as their D2H events complete.
namespace {
struct PipeChunk { int r0, r1, l0, l1, q0, q1; };   // output rows; left / right image rows this chunk adds
struct PipeStage {
    std::vector<PipeChunk> ch;
    const sr_image *left, *right;
    char *sL, *sR, *sO;                  // pinned staging (nullptr: that side is pinned already)
    char* out; size_t out_pitch, rowL, rowR, rowb;
    std::vector<std::atomic<int>> in_done, out_ready;
    std::atomic<int> abort{0};
    explicit PipeStage(int n) : ch(n), in_done(n), out_ready(n) {
        for (int i = 0; i < n; i++) { in_done[i].store(0); out_ready[i].store(0); }
    }
};
inline void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
}
bool wait_count(const std::atomic<int>& f, int want, const std::atomic<int>& abort) {
    for (int spins = 0; f.load(std::memory_order_acquire) < want; spins++) {
        if (abort.load(std::memory_order_relaxed)) return false;
        if (spins < 4000) cpu_relax(); else std::this_thread::yield();
    }
    return true;
}
void pipe_stage_worker(int t, int nthreads, void* arg) {
    PipeStage* j = (PipeStage*)arg;
    const int n = (int)j->ch.size();
    for (int c = 0; c < n; c++) {
        const PipeChunk& k = j->ch[c];
        if (j->sR) sr_stage_stripe(j->sR, j->rowR, (const char*)j->right->bgr, j->right->pitch, j->rowR, k.q0, k.q1, t, nthreads);
        if (j->sL) sr_stage_stripe(j->sL, j->rowL, (const char*)j->left->bgr, j->left->pitch, j->rowL, k.l0, k.l1, t, nthreads);
        j->in_done[c].fetch_add(1, std::memory_order_release);
    }
    if (!j->sO) return;
    for (int c = 0; c < n; c++) {
        if (!wait_count(j->out_ready[c], 1, j->abort)) return;
        sr_stage_stripe(j->out, j->out_pitch, j->sO, j->rowb, j->rowb, j->ch[c].r0, j->ch[c].r1, t, nthreads);
    }
}
}  // namespace

static int block_search_pipelined(sr_ctx* ctx, sr_pair* pair, const sr_image* left, const sr_image* right,
                                  const sr_search_params* p, double* disp_out, size_t out_pitch, int band0, int band1) {
    const int nch = pipeline_chunks(pair, p, band1 - band0);
    const int half = (p->block - 1) / 2;
    const bool smoothing = p->smooth != 1.0 && p->min_disp == 0;
    if (smoothing && band0 != 0) return sr_fail(ctx, SR_E_BADARG, "row bands need smooth == 1 (the smoothing rule is a raster dependency)");
    const int oh = band1 - band0, ow = pair->w2;
    const size_t rowb = (size_t)ow * sizeof(double);
    if (disp_out && out_pitch < rowb) return sr_fail(ctx, SR_E_BADARG, "output pitch %zu smaller than a row (%zu)", out_pitch, rowb);
    if (left->pitch < (size_t)left->w * 3 || right->pitch < (size_t)right->w * 3)
        return sr_fail(ctx, SR_E_BADARG, "image pitch smaller than a row");
    while ((int)ctx->pipe_ev.size() < 4 * nch + 1) {
        cudaEvent_t e;
        SR_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ctx->pipe_ev.push_back(e);
    }
    const size_t rowL = (size_t)left->w * 3, rowR = (size_t)right->w * 3;
    // chunk table.  Uneven chunks: a short first chunk starts the search early, short last chunks leave a
    // short exposed download; the chunks in between are larger (fewer partial waves in the search kernel)
    PipeStage job(nch);
    {
        int up0 = band0 - half > 0 ? band0 - half : 0;   // first image row the band's windows read
        int upL = up0 < pair->h1 ? up0 : pair->h1, upR = up0 < pair->h2 ? up0 : pair->h2;
        for (int c = 0; c < nch; c++) {
            PipeChunk& k = job.ch[c];
            k.r0 = band0 + chunk_start(oh, nch, c, half);
            k.r1 = band0 + chunk_start(oh, nch, c + 1, half);
            // input rows this chunk needs: window rows [r0 - half, r1 + half) of both images
            int needR = k.r1 + half + 1 < pair->h2 ? k.r1 + half + 1 : pair->h2;
            int needL = k.r1 + half + 1 < pair->h1 ? k.r1 + half + 1 : pair->h1;
            if (c == nch - 1 && band1 >= pair->h2) { needR = pair->h2; needL = pair->h1; }
            k.q0 = upR; k.q1 = needR > upR ? needR : upR; upR = k.q1;
            k.l0 = upL; k.l1 = needL > upL ? needL : upL; upL = k.l1;
        }
    }
    // pageable sides go through pinned staging, filled / drained by the staging threads
    const bool stL = !host_pinned(left->bgr), stR = !host_pinned(right->bgr), stO = disp_out && !host_pinned(disp_out);
    sr_stage_pool* pool = nullptr;
    job.left = left; job.right = right;
    job.sL = job.sR = job.sO = nullptr;
    job.out = (char*)disp_out; job.out_pitch = out_pitch; job.rowL = rowL; job.rowR = rowR; job.rowb = rowb;
    if (stL || stR || stO) {
        if (!(pool = sr_stage_pool_get(ctx))) return SR_E_CUDA;
        if (stL && !(job.sL = (char*)sr_pinned(ctx, 0, rowL * pair->h1))) return SR_E_CUDA;
        if (stR && !(job.sR = (char*)sr_pinned(ctx, 1, rowR * pair->h2))) return SR_E_CUDA;
        if (stO && !(job.sO = (char*)sr_pinned(ctx, 2, rowb * pair->h2))) return SR_E_CUDA;
        sr_stage_run(pool, pipe_stage_worker, &job);
    }
    const int nst = pool ? sr_stage_threads(pool) : 0;
    const char* srcL = stL ? job.sL : (const char*)left->bgr;
    const char* srcR = stR ? job.sR : (const char*)right->bgr;
    const size_t pitL = stL ? rowL : left->pitch, pitR = stR ? rowR : right->pitch;
    char* dstO = stO ? job.sO : (char*)disp_out;
    const size_t pitO = stO ? rowb : out_pitch;

    cudaStream_t compute = ctx->stream, s_up = ctx->aux[0], s_down = ctx->aux[1], s_pack = ctx->aux[4];
    int rc = SR_OK;
    auto cuda_ok = [&](cudaError_t e, const char* what) { if (e != cudaSuccess && rc == SR_OK) rc = sr_cuda_check(ctx, e, what); return rc == SR_OK; };
    // the copy streams start after whatever the compute stream still has queued (previous call)
    cudaEvent_t ev_start = ctx->pipe_ev[4 * nch];
    cuda_ok(cudaEventRecord(ev_start, compute), "cudaEventRecord");
    for (cudaStream_t s : {s_up, s_down, s_pack, ctx->aux[2], ctx->aux[3]}) cuda_ok(cudaStreamWaitEvent(s, ev_start, 0), "cudaStreamWaitEvent");
    // compute streams the chunks rotate over (4K/512, 10 chunks: 2 -> 2.40 ms, 3 -> 2.32 ms)
    static const int ncs = getenv("SR_PIPE_STREAMS") ? atoi(getenv("SR_PIPE_STREAMS")) : 3;
    // SR_PIPE_TRACE=1: per-chunk timeline (ms since the start of the call) on stderr, development aid
    static const bool trace = getenv("SR_PIPE_TRACE") && atoi(getenv("SR_PIPE_TRACE")) != 0;
    std::vector<cudaEvent_t> tev;
    if (trace) {
        tev.resize(4 * nch + 1);
        for (auto& e : tev) cudaEventCreate(&e);
        cudaEventRecord(tev[4 * nch], compute);
    }
    for (int c = 0; c < nch && rc == SR_OK; c++) {
        const PipeChunk& k = job.ch[c];
        if ((stL || stR) && !wait_count(job.in_done[c], nst, job.abort)) break;
        // the copy stream carries nothing but copies, so the link never waits for a pack kernel to find an SM;
        // one pack launch per chunk (both images) on a high-priority stream of its own
        if (k.q1 > k.q0)
            cuda_ok(cudaMemcpy2DAsync(pair->rawR + (size_t)k.q0 * rowR, rowR, srcR + (size_t)k.q0 * pitR, pitR, rowR, k.q1 - k.q0,
                                      cudaMemcpyHostToDevice, s_up), "cudaMemcpy2DAsync(H2D right rows)");
        if (rc == SR_OK && k.l1 > k.l0)
            cuda_ok(cudaMemcpy2DAsync(pair->rawL + (size_t)k.l0 * rowL, rowL, srcL + (size_t)k.l0 * pitL, pitL, rowL, k.l1 - k.l0,
                                      cudaMemcpyHostToDevice, s_up), "cudaMemcpy2DAsync(H2D left rows)");
        if (rc != SR_OK) break;
        if (!cuda_ok(cudaEventRecord(ctx->pipe_ev[4 * c + 3], s_up), "cudaEventRecord")) break;
        if (!cuda_ok(cudaStreamWaitEvent(s_pack, ctx->pipe_ev[4 * c + 3], 0), "cudaStreamWaitEvent")) break;
        ctx->stream = s_pack;
        rc = sr_launch_pack_rows2(ctx, pair->rawR, rowR, right->w, k.q0, k.q1, &pair->R, pair->rawL, rowL, left->w, k.l0, k.l1, &pair->L);
        ctx->stream = compute;
        if (rc != SR_OK) break;
        // chunks rotate over the compute streams: the first CTAs of the next chunks fill the SMs that
        // the last wave of chunk c leaves idle (one 512-thread CTA per SM, so no other overlap exists)
        const int si = ncs >= 3 ? c % 3 : ncs == 2 ? (c & 1) : 0;
        cudaStream_t cs = si == 0 ? compute : ctx->aux[1 + si];
        if (!cuda_ok(cudaEventRecord(ctx->pipe_ev[4 * c], s_pack), "cudaEventRecord")) break;
        if (trace) cudaEventRecord(tev[4 * c], s_pack);
        if (!cuda_ok(cudaStreamWaitEvent(cs, ctx->pipe_ev[4 * c], 0), "cudaStreamWaitEvent")) break;
        if (trace) cudaEventRecord(tev[4 * c + 1], cs);
        ctx->stream = cs;
        ctx->more_chunks_follow = c + 1 < nch;
        // smoothing rule: the chunks' raster passes run in row order -- each waits for the previous chunk's (which ran on
        // another compute stream); search, prepare and the copies of different chunks still overlap
        ctx->chunked_smoothing = smoothing;
        ctx->raster_wait = (smoothing && c > 0) ? ctx->pipe_ev[4 * (c - 1) + 1] : nullptr;
        rc = sr_search_run(ctx, pair, p, k.r0, k.r1);
        ctx->more_chunks_follow = false;
        ctx->chunked_smoothing = false;
        ctx->raster_wait = nullptr;
        ctx->stream = compute;
        if (rc != SR_OK) break;
        if (trace) cudaEventRecord(tev[4 * c + 2], cs);
        if (!disp_out && !smoothing) continue;
        if (!cuda_ok(cudaEventRecord(ctx->pipe_ev[4 * c + 1], cs), "cudaEventRecord")) break;
        if (!disp_out) continue;
        if (!cuda_ok(cudaStreamWaitEvent(s_down, ctx->pipe_ev[4 * c + 1], 0), "cudaStreamWaitEvent")) break;
        if (!cuda_ok(cudaMemcpy2DAsync(dstO + (size_t)k.r0 * pitO, pitO, pair->out + (size_t)k.r0 * ow, rowb, rowb, k.r1 - k.r0,
                                       cudaMemcpyDeviceToHost, s_down), "cudaMemcpy2DAsync(D2H rows)")) break;
        if (stO && !cuda_ok(cudaEventRecord(ctx->pipe_ev[4 * c + 2], s_down), "cudaEventRecord")) break;
        if (trace) cudaEventRecord(tev[4 * c + 3], s_down);
    }
    if (pool) {
        // hand finished output chunks to the staging threads in order; on an error release them instead
        for (int c = 0; c < nch && stO && rc == SR_OK; c++) {
            if (cuda_ok(cudaEventSynchronize(ctx->pipe_ev[4 * c + 2]), "cudaEventSynchronize"))
                job.out_ready[c].store(1, std::memory_order_release);
        }
        if (rc != SR_OK) job.abort.store(1);
        sr_stage_wait(pool);
    }
    cudaStreamSynchronize(s_up);
    cudaStreamSynchronize(s_pack);
    cudaStreamSynchronize(compute);
    cudaStreamSynchronize(ctx->aux[2]);
    cudaStreamSynchronize(ctx->aux[3]);
    cudaStreamSynchronize(s_down);
    if (trace) {
        for (int c = 0; c < nch && rc == SR_OK; c++) {
            float t[4] = {0, 0, 0, 0};
            for (int k = 0; k < 4; k++)
                if (k < 3 || disp_out) cudaEventElapsedTime(&t[k], tev[4 * nch], tev[4 * c + k]);
            fprintf(stderr, "[pipe] chunk %d: upload done %.3f  search start %.3f  end %.3f  download done %.3f\n", c, t[0], t[1], t[2], t[3]);
        }
        for (auto& e : tev) cudaEventDestroy(e);
    }
    if (rc != SR_OK) return rc;
    pair->last_row0 = band0; pair->last_row1 = band1;
    pair->last_geom.row0 = band0; pair->last_geom.row1 = band1;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return sr_cuda_check(ctx, e, "row-chunk pipeline");
    return SR_OK;
}

// n pairs of one shape, whole maps, ONE tile-kernel launch (sr_launch_search_tile_multi), then per pair the resolve pass
// or the sub-pixel pass.  Returns 1 when done, 0 when the request needs the per-pair path (nothing was launched), < 0 on
// error.  Same results as sr_search_run on each pair.
int sr_search_run_multi(sr_ctx* ctx, sr_pair* const* pairs, int n, const sr_search_params* p) {
    if (n < 2 || n > 16 || p->variant != SR_VARIANT_RIGHT || p->var_block || p->lr_check || p->min_disp < 0) return 0;
    if (p->cost != SR_COST_SSD_L2 && p->cost != SR_COST_SAD) return 0;
    if (p->subpixel ? (!sr_ext_subpix_out_applies(p) || p->smooth != 1.0) : (p->smooth != 1.0 && p->min_disp == 0)) return 0;
    sr_pair* p0 = pairs[0];
    for (int q = 1; q < n; q++)
        if (pairs[q]->w1 != p0->w1 || pairs[q]->h1 != p0->h1 || pairs[q]->w2 != p0->w2 || pairs[q]->h2 != p0->h2) return 0;
    SearchGeom g;
    g.w1 = p0->w1; g.h1 = p0->h1; g.w2 = p0->w2; g.h2 = p0->h2;
    g.height = g.h1 < g.h2 ? g.h1 : g.h2;
    g.block = p->block; g.half = (p->block - 1) / 2;
    g.dlo = p->min_disp; g.dhi = p->max_disp;
    if (g.dhi > g.w1) g.dhi = g.w1;
    if (g.dhi < g.dlo) g.dhi = g.dlo;
    g.row0 = 0; g.row1 = g.h2;
    if (g.half < 1 || g.half > 8 || g.w2 == 1 || g.h2 == 1 || !rows_fit(g.h1, g.h2, g.half)) return 0;
    for (int q = 0; q < n; q++) {
        sr_pair* pair = pairs[q];
        pair->evals_valid = false;
        pair->disp_left = nullptr;
        int init0[2] = {p->block, SR_OK};
        SR_CUDA(ctx, cudaMemcpyAsync(pair->maxblock, init0, sizeof init0, cudaMemcpyHostToDevice, ctx->stream));
        pair->ow = g.w2; pair->oh = g.h2;
        pair->last_row0 = 0; pair->last_row1 = g.h2;
    }
    int handled = sr_launch_search_tile_multi(ctx, pairs, n, g, p->cost);
    if (handled <= 0) return handled;
    if (p->subpixel) {                                   // one launch for the group
        int rc = sr_launch_ext_subpix_out_multi(ctx, pairs, n, p, 0, g.h2);
        if (rc != SR_OK) return rc;
    }
    for (int q = 0; q < n; q++) {
        if (!p->subpixel) {
            int rc = sr_launch_resolve_right(ctx, pairs[q], g, 1.0, 0, 0);
            if (rc != SR_OK) return rc;
        }
        pairs[q]->last_geom = g;
        pairs[q]->evals_valid = !p->subpixel;
    }
    ctx->last_kernel = p->subpixel ? "tile+ext" : "tile";
    return 1;
}

extern "C" {

int sr_search_run(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p, int row0, int row1) {
    int rc = validate(ctx, pair, p);
    if (rc != SR_OK) return rc;
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    pair->evals_valid = false;
    pair->disp_left = nullptr;
    {
        // "max block size" (BlockSearch.cpp:98) and the device status word.  A small pageable H2D copy (staged by the
        // driver at call time) on purpose: a 1-thread kernel would have to wait for an SM behind the previous chunk's
        // one-CTA-per-SM search grid and measurably delays the row-chunk pipeline.
        int init0[2] = {p->block, SR_OK};
        SR_CUDA(ctx, cudaMemcpyAsync(pair->maxblock, init0, sizeof init0, cudaMemcpyHostToDevice, ctx->stream));
    }
    if (p->variant == SR_VARIANT_LEFT) {
        if (p->cost != SR_COST_SSD_L2 || p->subpixel || p->lr_check)
            return sr_fail(ctx, SR_E_BADARG, "the Left variant is reference-exact only (SSD, no extensions)");
        if ((p->block & 1) == 0) return sr_fail(ctx, SR_E_SHAPE, "Left variant with an even block overruns the ROI (cv::Exception in the reference)");
        pair->ow = pair->w1; pair->oh = pair->h1;
        pair->last_row0 = 0; pair->last_row1 = pair->h1;
        ctx->last_kernel = "left";
        return sr_launch_search_left(ctx, pair, p->block, p->max_disp, p->smooth);
    }
    if (p->variant == SR_VARIANT_LINEAR) {
        if (pair->h1 < pair->h2 || pair->w1 < pair->w2)
            return sr_fail(ctx, SR_E_SHAPE, "LinearSearch reads leftImage(i,j) for every right pixel: needs h1>=h2, w1>=w2");
        pair->ow = pair->w2; pair->oh = pair->h2;
        pair->last_row0 = 0; pair->last_row1 = pair->h2;
        ctx->last_kernel = "linear";
        return sr_launch_search_linear(ctx, pair, p->smooth);
    }
    // ---- Right variant (BlockSearch.cpp:88-179) and its extension costs ----
    if (p->min_disp < 0) return sr_fail(ctx, SR_E_SHAPE, "min_disp < 0 makes the left ROI start at a negative column");
    if (p->cost == SR_COST_NCC || p->subpixel || p->lr_check) return sr_search_run_ext(ctx, pair, p, row0, row1);
    SearchGeom g;
    g.w1 = pair->w1; g.h1 = pair->h1; g.w2 = pair->w2; g.h2 = pair->h2;
    g.height = g.h1 < g.h2 ? g.h1 : g.h2;
    g.block = p->block; g.half = (p->block - 1) / 2;
    g.dlo = p->min_disp; g.dhi = p->max_disp;
    if (g.dhi > g.w1) g.dhi = g.w1;  // no candidate beyond: x + d + r < w1
    if (g.dhi < g.dlo) g.dhi = g.dlo;
    g.row0 = row0 < 0 ? 0 : row0; g.row1 = row1 > g.h2 ? g.h2 : row1;
    if (g.row1 < g.row0) g.row1 = g.row0;
    if (!p->var_block && !rows_fit(g.h1, g.h2, g.half))
        return sr_fail(ctx, SR_E_SHAPE, "left image too short for the window rows (h1=%d, h2=%d): cv::Exception in the reference", g.h1, g.h2);
    bool smoothing = p->smooth != 1.0 && g.dlo == 0;
    if (smoothing && g.row0 != 0 && !ctx->chunked_smoothing)
        return sr_fail(ctx, SR_E_BADARG, "row bands need smooth == 1 (the smoothing rule is a raster dependency)");
    pair->ow = g.w2; pair->oh = g.h2;
    pair->last_row0 = g.row0; pair->last_row1 = g.row1;
    bool degenerate = g.half == 0 || g.w2 == 1 || g.h2 == 1;  // empty windows: every cost is NaN
    int handled = 0;
    if (!p->var_block && !degenerate) {
        SearchGeom gt = g;
        if (smoothing) gt.dlo = 1 > g.dlo ? 1 : g.dlo;
        if (gt.dhi < gt.dlo) gt.dhi = gt.dlo;
        handled = sr_launch_search_tile(ctx, pair, gt, p->cost, 0, smoothing);
        if (handled < 0) return handled;
        if (handled) {
            ctx->last_kernel = "tile";
            // the d == 0 costs come out of the tile kernel (warp 0 of pass 0 holds those sums); the separable box-sum
            // kernels only where it could not (range not starting in the first pass, nothing searched)
            if (smoothing && !(pair->c0_from_tile && pair->npass > 0) && (rc = sr_launch_cost0(ctx, pair, g, p->cost)) != SR_OK) return rc;
            rc = sr_launch_resolve_right(ctx, pair, g, p->smooth, 0, smoothing ? 2 : 0);   // 2: c0 is the u32 plane of sr_launch_cost0
        }
    }
    if (!handled) {
        ctx->last_kernel = "direct";
        if ((rc = sr_launch_search_direct(ctx, pair, g, p->cost, p->var_block, p->thres, 1)) != SR_OK) return rc;
        rc = sr_launch_resolve_right(ctx, pair, g, p->smooth, 1, 1);
    }
    if (rc != SR_OK) return rc;
    pair->last_geom = g;
    pair->evals_valid = true;  // counted on demand (sr_search_effective_evals)
    return SR_OK;
}

// Band search of a rank > 0 of a row-band sharded pair with the gather fused into it: sr_search_run on rows
// [row0,row1) whose resolve pass writes the int16 wire format straight into the root's gather window and publishes step
// `seq` (no CV_64F band, no separate push kernel).  Where the fused pass does not apply (odd width, direct kernels,
// smoothing rule) it is sr_search_run followed by sr_search_push_band -- same protocol, same result on the root.
int sr_search_run_push(sr_ctx* ctx, sr_pair* pair, const sr_search_params* p, int row0, int row1, int rank, unsigned seq) {
    if (!ctx || !pair || !p) return sr_fail(ctx, SR_E_BADARG, "sr_search_run_push: NULL argument");
    if (pair->peer_mode != 2 && pair->peer_mode != 3) return sr_fail(ctx, SR_E_BADARG, "sr_search_run_push: not attached to a root window");
    if (rank < 1 || rank >= 32 || seq == 0) return sr_fail(ctx, SR_E_BADARG, "sr_search_run_push: rank 1..31, step >= 1");
    int st = sr_peer_check(ctx, pair);
    if (st != SR_OK) return st;
    pair->wire_rank = rank; pair->wire_seq = seq; pair->wire_done = false;
    int rc = sr_search_run(ctx, pair, p, row0, row1);
    bool done = pair->wire_done;
    pair->wire_rank = 0; pair->wire_seq = 0; pair->wire_done = false;
    if (rc != SR_OK) return rc;
    return done ? SR_OK : sr_search_push_band(ctx, pair, row0, row1, rank, seq);
}

// Device-resident batch form: kernels only, asynchronous on the context stream, results stay in HBM.  Pairs of one shape
// in groups of up to 16 share ONE tile-kernel launch where the request allows it (Right variant, SSD / SAD, plain or
// sub-pixel result); otherwise each pair runs sr_search_run.
int sr_search_run_batch(sr_ctx* ctx, sr_pair* const* pairs, int n, const sr_search_params* p) {
    if (!ctx || !pairs || !p || n < 0) return sr_fail(ctx, SR_E_BADARG, "sr_search_run_batch: bad argument");
    for (int i = 0; i < n; i++) { int rc = validate(ctx, pairs[i], p); if (rc != SR_OK) return rc; }
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    for (int i0 = 0; i0 < n; i0 += 16) {
        int cnt = n - i0 < 16 ? n - i0 : 16;
        int done = cnt >= 2 ? sr_search_run_multi(ctx, pairs + i0, cnt, p) : 0;
        if (done < 0) return done;
        if (!done)
            for (int j = 0; j < cnt; j++) {
                int oh = p->variant == SR_VARIANT_LEFT ? pairs[i0 + j]->h1 : pairs[i0 + j]->h2;
                int rc = sr_search_run(ctx, pairs[i0 + j], p, 0, oh);
                if (rc != SR_OK) return rc;
            }
    }
    return SR_OK;
}

void* sr_search_result_dev(sr_ctx* ctx, sr_pair* pair) { return pair ? (void*)pair->out : nullptr; }

int sr_search_download(sr_ctx* ctx, sr_pair* pair, double* disp_out, size_t out_pitch, int row0, int row1) {
    if (!ctx || !pair || !disp_out) return sr_fail(ctx, SR_E_BADARG, "sr_search_download: NULL argument");
    if (row0 < 0) row0 = 0;
    if (row1 > pair->oh) row1 = pair->oh;
    if (row1 <= row0) return SR_OK;
    size_t rowb = (size_t)pair->ow * sizeof(double);
    if (out_pitch < rowb) return sr_fail(ctx, SR_E_BADARG, "output pitch %zu smaller than a row (%zu)", out_pitch, rowb);
    int rows = row1 - row0;
    const double* src = pair->out + (size_t)row0 * pair->ow;
    double* dst = (double*)((char*)disp_out + (size_t)row0 * out_pitch);
    cudaPointerAttributes attr;
    bool pinned = cudaPointerGetAttributes(&attr, disp_out) == cudaSuccess &&
                  (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    if (pinned) {
        SR_CUDA(ctx, cudaMemcpy2DAsync(dst, out_pitch, src, rowb, rowb, rows, cudaMemcpyDeviceToHost, ctx->stream));
        SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    } else {
        double* st = (double*)sr_pinned(ctx, 2, rowb * rows);
        if (!st) return SR_E_CUDA;
        SR_CUDA(ctx, cudaMemcpyAsync(st, src, rowb * rows, cudaMemcpyDeviceToHost, ctx->stream));
        SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        sr_stage_copy_rows(ctx, dst, out_pitch, st, rowb, rowb, rows);
    }
    int status[2];
    SR_CUDA(ctx, cudaMemcpy(status, pair->maxblock, sizeof status, cudaMemcpyDeviceToHost));
    if (status[1] == SR_E_BADARG) return sr_fail(ctx, status[1], "sr_search_pack16: the map holds values that are not integers in int16 range");
    { int st = sr_peer_check(ctx, pair); if (st != SR_OK) return st; }   // band gather: sticky protocol status
    if (status[1] != SR_OK) return sr_fail(ctx, status[1], "the grown window runs past the left image (cv::Exception in the reference)");
    return SR_OK;
}

int sr_search_max_block(sr_ctx* ctx, sr_pair* pair, int* max_block) {
    if (!ctx || !pair || !max_block) return SR_E_BADARG;
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    SR_CUDA(ctx, cudaMemcpy(max_block, pair->maxblock, sizeof(int), cudaMemcpyDeviceToHost));
    return SR_OK;
}

static int rows16(sr_ctx* ctx, sr_pair* pair, int& row0, int& row1) {
    if (!ctx || !pair) return sr_fail(ctx, SR_E_BADARG, "NULL argument");
    if (row0 < 0) row0 = 0;
    if (row1 > pair->oh) row1 = pair->oh;
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!pair->out16) SR_CUDA(ctx, cudaMalloc((void**)&pair->out16, pair->npx_cap * sizeof(short)));
    return SR_OK;
}

int sr_search_pack16(sr_ctx* ctx, sr_pair* pair, int row0, int row1) {
    int rc = rows16(ctx, pair, row0, row1);
    if (rc != SR_OK || row1 <= row0) return rc;
    size_t n = (size_t)(row1 - row0) * pair->ow, off = (size_t)row0 * pair->ow;
    pack16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(pair->out + off, pair->out16 + off, n, pair->maxblock + 1);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

int sr_search_unpack16(sr_ctx* ctx, sr_pair* pair, int row0, int row1) {
    int rc = rows16(ctx, pair, row0, row1);
    if (rc != SR_OK || row1 <= row0) return rc;
    size_t n = (size_t)(row1 - row0) * pair->ow, off = (size_t)row0 * pair->ow;
    unpack16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(pair->out16 + off, pair->out + off, n);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

void* sr_search_result16_dev(sr_ctx* ctx, sr_pair* pair) {
    int r0 = 0, r1 = 0;
    if (rows16(ctx, pair, r0, r1) != SR_OK) return nullptr;
    return (void*)pair->out16;
}

int sr_search_left_disparity(sr_ctx* ctx, sr_pair* pair, int32_t* disp_left, size_t out_pitch) {
    if (!ctx || !pair || !disp_left) return sr_fail(ctx, SR_E_BADARG, "sr_search_left_disparity: NULL argument");
    if (!pair->disp_left) return sr_fail(ctx, SR_E_BADARG, "the last run had no left-right check");
    size_t rowb = (size_t)pair->w1 * sizeof(int32_t);
    if (out_pitch < rowb) return sr_fail(ctx, SR_E_BADARG, "output pitch %zu smaller than a row (%zu)", out_pitch, rowb);
    SR_CUDA(ctx, cudaMemcpy2DAsync(disp_left, out_pitch, pair->disp_left, rowb, rowb, pair->h1, cudaMemcpyDeviceToHost, ctx->stream));
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SR_OK;
}

int sr_search_effective_evals(sr_ctx* ctx, sr_pair* pair, uint64_t* evals) {
    if (!ctx || !pair || !evals) return SR_E_BADARG;
    if (!pair->evals_valid) return sr_fail(ctx, SR_E_BADARG, "no evaluation count for the last run");
    int rc = sr_launch_count_evals(ctx, pair, pair->last_geom);
    if (rc != SR_OK) return rc;
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    unsigned long long v = 0;
    SR_CUDA(ctx, cudaMemcpy(&v, pair->evals, 8, cudaMemcpyDeviceToHost));
    *evals = v;
    return SR_OK;
}

int sr_pair_search_band(sr_ctx* ctx, sr_pair* pair, const sr_image* left, const sr_image* right, const sr_search_params* p,
                        int row0, int row1, double* disp_out, size_t out_pitch) {
    if (!ctx || !pair || !left || !right || !p || !left->bgr || !right->bgr) return sr_fail(ctx, SR_E_BADARG, "sr_pair_search_band: NULL argument");
    if (left->w != pair->w1 || left->h != pair->h1 || right->w != pair->w2 || right->h != pair->h2)
        return sr_fail(ctx, SR_E_BADARG, "sr_pair_search_band: image sizes differ from the pair's");
    int rc = validate(ctx, pair, p);
    if (rc != SR_OK) return rc;
    if (p->variant != SR_VARIANT_RIGHT) return sr_fail(ctx, SR_E_BADARG, "row bands are defined for the Right variant");
    if (row0 < 0) row0 = 0;
    if (row1 > pair->h2) row1 = pair->h2;
    if (row1 <= row0) return SR_OK;
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    const double* pin_probe = disp_out ? disp_out : (const double*)left->bgr;
    bool piped = can_pipeline(ctx, pair, left, right, p, pin_probe) && pipeline_chunks(pair, p, row1 - row0) >= 1;
    if (piped) return block_search_pipelined(ctx, pair, left, right, p, disp_out, out_pitch, row0, row1);
    // general form: whole upload, band search, band download
    if ((rc = sr_pair_upload(ctx, pair, left, right)) != SR_OK) return rc;
    if ((rc = sr_search_run(ctx, pair, p, row0, row1)) != SR_OK) return rc;
    if (disp_out) return sr_search_download(ctx, pair, disp_out, out_pitch, row0, row1);
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return SR_OK;
}

int sr_block_search(sr_ctx* ctx, const sr_image* left, const sr_image* right, const sr_search_params* p,
                    double* disp_out, size_t out_pitch, int* max_block) {
    if (!ctx || !left || !right || !p || !disp_out) return sr_fail(ctx, SR_E_BADARG, "sr_block_search: NULL argument");
    if (!left->bgr || !right->bgr || left->w <= 0 || left->h <= 0 || right->w <= 0 || right->h <= 0)
        return sr_fail(ctx, SR_E_BADARG, "sr_block_search: empty image");
    // one cached pair object per context, re-created when the shape changes
    sr_pair* pair = ctx->batch_pairs.empty() ? nullptr : ctx->batch_pairs[0];
    if (pair && (pair->w1 != left->w || pair->h1 != left->h || pair->w2 != right->w || pair->h2 != right->h)) {
        for (sr_pair* q : ctx->batch_pairs) sr_pair_destroy(ctx, q);
        ctx->batch_pairs.clear();
        pair = nullptr;
    }
    int rc;
    if (!pair) {
        if ((rc = sr_pair_create(ctx, left->w, left->h, right->w, right->h, &pair)) != SR_OK) return rc;
        ctx->batch_pairs.push_back(pair);
    }
    if (validate(ctx, pair, p) == SR_OK && can_pipeline(ctx, pair, left, right, p, disp_out)) {
        if ((rc = block_search_pipelined(ctx, pair, left, right, p, disp_out, out_pitch, 0, pair->h2)) != SR_OK) return rc;
        if (max_block) *max_block = p->block;   // fixed window: BlockSearch.cpp:98
        return SR_OK;
    }
    if ((rc = sr_pair_upload(ctx, pair, left, right)) != SR_OK) return rc;
    int oh = p->variant == SR_VARIANT_LEFT ? left->h : right->h;
    if ((rc = sr_search_run(ctx, pair, p, 0, oh)) != SR_OK) return rc;
    if ((rc = sr_search_download(ctx, pair, disp_out, out_pitch, 0, oh)) != SR_OK) return rc;
    if (max_block) return sr_search_max_block(ctx, pair, max_block);
    return SR_OK;
}

}  // extern "C"
