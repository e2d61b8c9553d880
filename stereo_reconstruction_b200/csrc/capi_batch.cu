// capi_batch.cu -- sr_block_search_batch: n independent pairs of one shape (config 5), pipelined
// H2D + pack (copy stream 0) / search kernels (compute stream) / D2H (copy stream 1) over a ring
// of device-resident pair objects.  With pinned host buffers the three stages overlap fully; pageable buffers
// go through pinned staging, copied by the host staging threads (host_stage.cu).
#include <cstdlib>
#include "sr_internal.cuh"

namespace {
constexpr int NB = 3;  // pairs in flight (per-pair pipeline)
constexpr int GP = 8;  // pairs per launch and GR groups in flight (grouped pipeline)
constexpr int GR = 2;
struct StreamSwap {
    sr_ctx* c; cudaStream_t saved;
    StreamSwap(sr_ctx* ctx, cudaStream_t s) : c(ctx), saved(ctx->stream) { ctx->stream = s; }
    ~StreamSwap() { c->stream = saved; }
};
}  // namespace

extern "C" int sr_block_search_batch(sr_ctx* ctx, int n, const sr_image* left, const sr_image* right, const sr_search_params* p,
                                     double* const* disp_out, size_t out_pitch) {
    if (!ctx || !left || !right || !p || !disp_out || n < 0) return sr_fail(ctx, SR_E_BADARG, "sr_block_search_batch: bad argument");
    if (n == 0) return SR_OK;
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    int w1 = left[0].w, h1 = left[0].h, w2 = right[0].w, h2 = right[0].h;
    for (int i = 0; i < n; i++)
        if (left[i].w != w1 || left[i].h != h1 || right[i].w != w2 || right[i].h != h2 || !left[i].bgr || !right[i].bgr || !disp_out[i])
            return sr_fail(ctx, SR_E_BADARG, "sr_block_search_batch: pair %d has a different shape or a NULL buffer", i);
    // Grouped pipeline: batches of small pairs (a 720p pair alone is 6 (tile, pass) columns on 148 SMs) run GP pairs per
    // tile-kernel launch; uploads of group k+1 (copy stream 0), the kernels of group k (compute stream) and the downloads
    // of group k-1 (copy stream 1) overlap.  Pinned host buffers; anything else takes the per-pair pipeline below.
    {
        bool grouped = n >= 2 * GP && p->variant == SR_VARIANT_RIGHT && !getenv("SR_BATCH_PER_PAIR") && (size_t)w2 * h2 <= (size_t)2048 * 1152;
        for (int i = 0; i < n && grouped; i++) {
            cudaPointerAttributes attr;
            grouped = cudaPointerGetAttributes(&attr, disp_out[i]) == cudaSuccess && attr.type == cudaMemoryTypeHost &&
                      cudaPointerGetAttributes(&attr, left[i].bgr) == cudaSuccess && attr.type == cudaMemoryTypeHost &&
                      cudaPointerGetAttributes(&attr, right[i].bgr) == cudaSuccess && attr.type == cudaMemoryTypeHost;
            cudaGetLastError();
        }
        if (grouped) {
            bool have = (int)ctx->batch_pairs.size() >= GP * GR;
            for (sr_pair* q : ctx->batch_pairs) have = have && q->w1 == w1 && q->h1 == h1 && q->w2 == w2 && q->h2 == h2;
            if (!have) {
                SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
                for (sr_pair* q : ctx->batch_pairs) sr_pair_destroy(ctx, q);
                ctx->batch_pairs.clear();
                for (int i = 0; i < GP * GR; i++) {
                    sr_pair* q = nullptr;
                    int rc = sr_pair_create(ctx, w1, h1, w2, h2, &q);
                    if (rc != SR_OK) return rc;
                    ctx->batch_pairs.push_back(q);
                }
            }
            cudaEvent_t up[GR], comp[GR], down[GR];
            for (int i = 0; i < GR; i++) {
                SR_CUDA(ctx, cudaEventCreateWithFlags(&up[i], cudaEventDisableTiming));
                SR_CUDA(ctx, cudaEventCreateWithFlags(&comp[i], cudaEventDisableTiming));
                SR_CUDA(ctx, cudaEventCreateWithFlags(&down[i], cudaEventDisableTiming));
            }
            const size_t rowb = (size_t)w2 * sizeof(double);
            int rc = out_pitch < rowb ? sr_fail(ctx, SR_E_BADARG, "output pitch smaller than a row") : SR_OK;
            cudaStream_t compute = ctx->stream;
            int ngroups = (n + GP - 1) / GP;
            for (int k = 0; k < ngroups && rc == SR_OK; k++) {
                const int s = k % GR, i0 = k * GP, cnt = n - i0 < GP ? n - i0 : GP;
                sr_pair** grp = ctx->batch_pairs.data() + s * GP;
                if (k >= GR) {   // the slot's previous kernels are done with the images, its results are downloaded
                    cudaStreamWaitEvent(ctx->aux[0], comp[s], 0);
                    cudaStreamWaitEvent(compute, down[s], 0);
                }
                {
                    StreamSwap sw(ctx, ctx->aux[0]);
                    for (int j = 0; j < cnt && rc == SR_OK; j++) rc = sr_pair_upload(ctx, grp[j], &left[i0 + j], &right[i0 + j]);
                }
                if (rc != SR_OK) break;
                cudaEventRecord(up[s], ctx->aux[0]);
                cudaStreamWaitEvent(compute, up[s], 0);
                int done = sr_search_run_multi(ctx, grp, cnt, p);
                if (done < 0) { rc = done; break; }
                if (!done)
                    for (int j = 0; j < cnt && rc == SR_OK; j++) rc = sr_search_run(ctx, grp[j], p, 0, h2);
                if (rc != SR_OK) break;
                cudaEventRecord(comp[s], compute);
                cudaStreamWaitEvent(ctx->aux[1], comp[s], 0);
                for (int j = 0; j < cnt; j++) {
                    cudaError_t e = cudaMemcpy2DAsync(disp_out[i0 + j], out_pitch, grp[j]->out, rowb, rowb, h2, cudaMemcpyDeviceToHost, ctx->aux[1]);
                    if (e != cudaSuccess) { rc = sr_cuda_check(ctx, e, "cudaMemcpy2DAsync(batch D2H)"); break; }
                }
                cudaEventRecord(down[s], ctx->aux[1]);
            }
            cudaStreamSynchronize(ctx->aux[0]);
            cudaStreamSynchronize(compute);
            cudaStreamSynchronize(ctx->aux[1]);
            for (int i = 0; i < GR; i++) { cudaEventDestroy(up[i]); cudaEventDestroy(comp[i]); cudaEventDestroy(down[i]); }
            if (rc == SR_OK) {
                cudaError_t e = cudaGetLastError();
                if (e != cudaSuccess) rc = sr_cuda_check(ctx, e, "grouped batch pipeline");
            }
            return rc;
        }
    }
    // ring of pair objects cached on the context
    bool ok = (int)ctx->batch_pairs.size() >= NB;
    for (sr_pair* q : ctx->batch_pairs) ok = ok && q->w1 == w1 && q->h1 == h1 && q->w2 == w2 && q->h2 == h2;
    if (!ok) {
        SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        for (sr_pair* q : ctx->batch_pairs) sr_pair_destroy(ctx, q);
        ctx->batch_pairs.clear();
        for (int i = 0; i < NB; i++) {
            sr_pair* q = nullptr;
            int rc = sr_pair_create(ctx, w1, h1, w2, h2, &q);
            if (rc != SR_OK) return rc;
            ctx->batch_pairs.push_back(q);
        }
    }
    cudaEvent_t up[NB], comp[NB], down[NB];
    for (int i = 0; i < NB; i++) {
        SR_CUDA(ctx, cudaEventCreateWithFlags(&up[i], cudaEventDisableTiming));
        SR_CUDA(ctx, cudaEventCreateWithFlags(&comp[i], cudaEventDisableTiming));
        SR_CUDA(ctx, cudaEventCreateWithFlags(&down[i], cudaEventDisableTiming));
    }
    int oh = p->variant == SR_VARIANT_LEFT ? h1 : h2, ow = p->variant == SR_VARIANT_LEFT ? w1 : w2;
    size_t rowb = (size_t)ow * sizeof(double);
    int rc = SR_OK;
    cudaStream_t compute = ctx->stream;
    // pageable outputs (cv::Mat memory): each slot downloads into its own pinned staging; the striped host copy
    // to the caller's buffer happens when the slot comes round again, while the GPU works on the pairs in between
    bool stage_out = false;
    for (int i = 0; i < n && !stage_out; i++) {
        cudaPointerAttributes attr;
        stage_out = !(cudaPointerGetAttributes(&attr, disp_out[i]) == cudaSuccess &&
                      (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged));
        cudaGetLastError();
    }
    char* stage = nullptr;
    const size_t slot_bytes = rowb * (size_t)oh;
    if (stage_out && !(stage = (char*)sr_pinned(ctx, 3, slot_bytes * NB))) return SR_E_CUDA;
    auto drain = [&](int i) {   // pair i's staged result -> the caller's buffer
        cudaError_t e = cudaEventSynchronize(down[i % NB]);
        if (e != cudaSuccess) return sr_cuda_check(ctx, e, "cudaEventSynchronize(batch D2H)");
        return sr_stage_copy_rows(ctx, disp_out[i], out_pitch, stage + slot_bytes * (i % NB), rowb, rowb, oh);
    };
    for (int i = 0; i < n && rc == SR_OK; i++) {
        int s = i % NB;
        sr_pair* pair = ctx->batch_pairs[s];
        if (i >= NB) {  // the slot's previous search must be done with the images, its result downloaded
            if (stage_out && (rc = drain(i - NB)) != SR_OK) break;
            cudaStreamWaitEvent(ctx->aux[0], comp[s], 0);
            cudaStreamWaitEvent(compute, down[s], 0);
        }
        {
            StreamSwap sw(ctx, ctx->aux[0]);
            rc = sr_pair_upload(ctx, pair, &left[i], &right[i]);
        }
        if (rc != SR_OK) break;
        cudaEventRecord(up[s], ctx->aux[0]);
        cudaStreamWaitEvent(compute, up[s], 0);
        rc = sr_search_run(ctx, pair, p, 0, oh);
        if (rc != SR_OK) break;
        cudaEventRecord(comp[s], compute);
        cudaStreamWaitEvent(ctx->aux[1], comp[s], 0);
        if (out_pitch < rowb) { rc = sr_fail(ctx, SR_E_BADARG, "output pitch smaller than a row"); break; }
        cudaError_t e = stage_out ? cudaMemcpyAsync(stage + slot_bytes * s, pair->out, slot_bytes, cudaMemcpyDeviceToHost, ctx->aux[1])
                                  : cudaMemcpy2DAsync(disp_out[i], out_pitch, pair->out, rowb, rowb, oh, cudaMemcpyDeviceToHost, ctx->aux[1]);
        if (e != cudaSuccess) { rc = sr_cuda_check(ctx, e, "cudaMemcpy2DAsync(batch D2H)"); break; }
        cudaEventRecord(down[s], ctx->aux[1]);
    }
    if (stage_out && rc == SR_OK)
        for (int i = n > NB ? n - NB : 0; i < n && rc == SR_OK; i++) rc = drain(i);
    cudaStreamSynchronize(ctx->aux[0]);
    cudaStreamSynchronize(compute);
    cudaStreamSynchronize(ctx->aux[1]);
    for (int i = 0; i < NB; i++) { cudaEventDestroy(up[i]); cudaEventDestroy(comp[i]); cudaEventDestroy(down[i]); }
    if (rc == SR_OK) {
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) rc = sr_cuda_check(ctx, e, "batch pipeline");
    }
    return rc;
}
