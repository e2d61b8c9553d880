// sr_core.cu -- context, error handling, buffers, image packing, pair objects.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include "sr_internal.cuh"

static thread_local std::string g_create_err;

int sr_fail(sr_ctx* ctx, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf; else g_create_err = buf;
    return code;
}

int sr_cuda_check(sr_ctx* ctx, cudaError_t e, const char* what) {
    if (e == cudaSuccess) return SR_OK;
    return sr_fail(ctx, SR_E_CUDA, "CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
}

void* sr_scratch(sr_ctx* ctx, int slot, size_t bytes) {
    DevBuf& b = ctx->scratch[slot];
    if (b.cap >= bytes && b.p) return b.p;
    if (b.p) { cudaStreamSynchronize(ctx->stream); cudaFree(b.p); b.p = nullptr; b.cap = 0; }
    size_t cap = bytes + (bytes >> 3) + 256;
    cudaError_t e = cudaMalloc(&b.p, cap);
    if (e != cudaSuccess) { sr_cuda_check(ctx, e, "cudaMalloc(scratch)"); b.p = nullptr; return nullptr; }
    b.cap = cap;
    return b.p;
}

void* sr_pinned(sr_ctx* ctx, int slot, size_t bytes) {
    if (ctx->pinned_cap[slot] >= bytes && ctx->pinned[slot]) return ctx->pinned[slot];
    if (ctx->pinned[slot]) { cudaStreamSynchronize(ctx->stream); cudaFreeHost(ctx->pinned[slot]); ctx->pinned[slot] = nullptr; }
    size_t cap = bytes + (bytes >> 3) + 4096;
    cudaError_t e = cudaMallocHost(&ctx->pinned[slot], cap);
    if (e != cudaSuccess) { sr_cuda_check(ctx, e, "cudaMallocHost"); ctx->pinned[slot] = nullptr; ctx->pinned_cap[slot] = 0; return nullptr; }
    ctx->pinned_cap[slot] = cap;
    return ctx->pinned[slot];
}

extern "C" {

const char* sr_version(void) { return "stereo_reconstruction_b200 0.1 (sm_100a)"; }

int sr_ctx_create(int device, sr_ctx** out) {
    if (!out) return sr_fail(nullptr, SR_E_BADARG, "sr_ctx_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return sr_fail(nullptr, SR_E_CUDA, "sr_ctx_create: no CUDA device (%s); this library has no CPU fallback",
                       e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    if (device < 0 || device >= n) return sr_fail(nullptr, SR_E_BADARG, "sr_ctx_create: device %d out of range (%d devices)", device, n);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return sr_cuda_check(nullptr, e, "cudaGetDeviceProperties");
    if (prop.major != 10)
        return sr_fail(nullptr, SR_E_CUDA, "sr_ctx_create: device %d is sm_%d%d; this library carries sm_100a code only", device, prop.major, prop.minor);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return sr_cuda_check(nullptr, e, "cudaSetDevice");
    sr_ctx* c = new sr_ctx();
    c->device = device;
    { int n = 0; if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device) == cudaSuccess && n > 0) c->sms = n; }
    // the copy streams outrank the compute stream: their small pack kernels must not queue behind the
    // search kernel's CTAs (one per SM) when the pipelines overlap upload, search and download
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
    if ((e = cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_lo)) != cudaSuccess ||
        (e = cudaStreamCreateWithPriority(&c->aux[0], cudaStreamNonBlocking, prio_hi)) != cudaSuccess ||
        (e = cudaStreamCreateWithPriority(&c->aux[1], cudaStreamNonBlocking, prio_hi)) != cudaSuccess ||
        (e = cudaStreamCreateWithPriority(&c->aux[2], cudaStreamNonBlocking, prio_lo)) != cudaSuccess ||
        (e = cudaStreamCreateWithPriority(&c->aux[3], cudaStreamNonBlocking, prio_lo)) != cudaSuccess ||
        (e = cudaStreamCreateWithPriority(&c->aux[4], cudaStreamNonBlocking, prio_hi)) != cudaSuccess ||
        (e = cudaEventCreate(&c->ev0)) != cudaSuccess || (e = cudaEventCreate(&c->ev1)) != cudaSuccess ||
        (e = cudaEventCreate(&c->kev0)) != cudaSuccess || (e = cudaEventCreate(&c->kev1)) != cudaSuccess) {
        int rc = sr_cuda_check(nullptr, e, "stream/event creation");
        delete c;
        return rc;
    }
    *out = c;
    return SR_OK;
}

void sr_ctx_destroy(sr_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    sr_stage_pool_destroy(ctx->stage_pool);
    for (sr_pair* p : ctx->batch_pairs) sr_pair_destroy(ctx, p);
    ctx->batch_pairs.clear();
    for (cudaEvent_t e : ctx->pipe_ev) cudaEventDestroy(e);
    for (TileWork& t : ctx->tile_work) cudaFree(t.dev);
    for (auto& b : ctx->scratch) if (b.p) cudaFree(b.p);
    for (auto& p : ctx->pinned) if (p) cudaFreeHost(p);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaEventDestroy(ctx->kev0);
    cudaEventDestroy(ctx->kev1);
    cudaStreamDestroy(ctx->aux[0]);
    cudaStreamDestroy(ctx->aux[1]);
    cudaStreamDestroy(ctx->aux[2]);
    cudaStreamDestroy(ctx->aux[3]);
    cudaStreamDestroy(ctx->aux[4]);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* sr_last_error(const sr_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int sr_host_alloc(void** p, size_t bytes) {
    if (!p) return SR_E_BADARG;
    return cudaMallocHost(p, bytes ? bytes : 1) == cudaSuccess ? SR_OK : SR_E_CUDA;
}
int sr_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? SR_OK : SR_E_CUDA; }
int sr_host_register(void* p, size_t bytes) {
    if (!p || !bytes) return SR_E_BADARG;
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); return SR_OK; }
    return e == cudaSuccess ? SR_OK : SR_E_CUDA;
}
int sr_host_unregister(void* p) { return p && cudaHostUnregister(p) == cudaSuccess ? SR_OK : SR_E_CUDA; }

void* sr_ctx_stream(sr_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int sr_ctx_sync(sr_ctx* ctx) {
    if (!ctx) return SR_E_BADARG;
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    // a failed step of a band gather (time-out, non-integer map) is reported by every rank at its next synchronisation
    for (sr_pair* p : ctx->peer_pairs) { int st = sr_peer_check(ctx, p); if (st != SR_OK) return st; }
    return SR_OK;
}
int sr_timer_start(sr_ctx* ctx) {
    if (!ctx) return SR_E_BADARG;
    SR_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    return SR_OK;
}
int sr_timer_stop(sr_ctx* ctx, float* ms) {
    if (!ctx || !ms) return SR_E_BADARG;
    SR_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    SR_CUDA(ctx, cudaEventSynchronize(ctx->ev1));
    SR_CUDA(ctx, cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
    return SR_OK;
}
uint64_t sr_kernel_launches(const sr_ctx* ctx) { return ctx ? ctx->launches : 0; }
const char* sr_last_kernel(const sr_ctx* ctx) { return ctx ? ctx->last_kernel : ""; }

}  // extern "C"

// ---- packing: BGR u8 (pitch) -> one uint32 per pixel, padded rows ---------------------------
// The padding of a packed image is zeroed once at allocation; packing writes the image columns only.  One thread
// makes four pixels: 12 source bytes from four aligned words (rows of a tight 3*w pitch start at any byte
// offset, hence the funnel shifts), one 16-byte store.  Two images in one launch (blockIdx.z): the pipelines pack
// while search CTAs own every SM, so each launch waits for an SM to come free.
struct PackJob {
    const uint8_t* raw; size_t raw_pitch; size_t raw_bytes; int w; int y0, y1;
    uint32_t* px; int pitch_px;      // pointer to pixel (0,0), row pitch in pixels
};

__global__ void __launch_bounds__(256) pack_bgr_kernel(PackJob ja, PackJob jb) {
    const PackJob& j = blockIdx.z ? jb : ja;
    const int x0 = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int y = j.y0 + blockIdx.y;
    if (x0 >= j.w || y >= j.y1) return;
    const size_t A = (size_t)y * j.raw_pitch + 3 * (size_t)x0;      // first source byte
    const uint32_t* words = (const uint32_t*)j.raw;
    const size_t wd = A >> 2, nwords = (j.raw_bytes + 3) >> 2;
    const uint32_t sh = (uint32_t)(A & 3) * 8;
    uint32_t w0 = __ldg(words + wd), w1 = wd + 1 < nwords ? __ldg(words + wd + 1) : 0u;
    uint32_t w2 = wd + 2 < nwords ? __ldg(words + wd + 2) : 0u, w3 = wd + 3 < nwords ? __ldg(words + wd + 3) : 0u;
    uint32_t t0 = __funnelshift_r(w0, w1, sh), t1 = __funnelshift_r(w1, w2, sh), t2 = __funnelshift_r(w2, w3, sh);
    uint4 o;
    o.x = t0 & 0xFFFFFFu;
    o.y = __funnelshift_r(t0, t1, 24) & 0xFFFFFFu;
    o.z = __funnelshift_r(t1, t2, 16) & 0xFFFFFFu;
    o.w = t2 >> 8;
    if (x0 + 1 >= j.w) o.y = 0;      // columns past the image: padding stays zero
    if (x0 + 2 >= j.w) o.z = 0;
    if (x0 + 3 >= j.w) o.w = 0;
    *(uint4*)(j.px + (size_t)y * j.pitch_px + x0) = o;
}

static PackJob pack_job(const uint8_t* raw, size_t raw_pitch, int w, int h, int y0, int y1, PackedImage* img) {
    PackJob j;
    j.raw = raw; j.raw_pitch = raw_pitch; j.raw_bytes = raw_pitch * (size_t)h; j.w = w; j.y0 = y0; j.y1 = y1;
    j.px = img->px; j.pitch_px = img->pitch;
    return j;
}

int sr_launch_pack_rows2(sr_ctx* ctx, const uint8_t* rawa, size_t pitcha, int wa, int ya0, int ya1, PackedImage* imga,
                         const uint8_t* rawb, size_t pitchb, int wb, int yb0, int yb1, PackedImage* imgb) {
    int rows = ya1 - ya0 > yb1 - yb0 ? ya1 - ya0 : yb1 - yb0;
    if (rows <= 0) return SR_OK;
    int w = wa > wb ? wa : wb;
    PackJob a = pack_job(rawa, pitcha, wa, imga->h, ya0, ya1, imga), b = pack_job(rawb, pitchb, wb, imgb->h, yb0, yb1, imgb);
    pack_bgr_kernel<<<dim3(cdiv(w, 1024), rows, 2), 256, 0, ctx->stream>>>(a, b);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

int sr_launch_pack_rows(sr_ctx* ctx, const uint8_t* raw, size_t raw_pitch, int w, int y0, int y1, PackedImage* img) {
    if (y1 <= y0) return SR_OK;
    PackJob a = pack_job(raw, raw_pitch, w, img->h, y0, y1, img);
    pack_bgr_kernel<<<dim3(cdiv(w, 1024), y1 - y0, 1), 256, 0, ctx->stream>>>(a, a);
    SR_LAUNCHED(ctx);
    return SR_OK;
}
int sr_launch_pack(sr_ctx* ctx, const uint8_t* raw, size_t raw_pitch, int w, int h, PackedImage* img) {
    return sr_launch_pack_rows(ctx, raw, raw_pitch, w, 0, h, img);
}

__global__ void pack_mirror_kernel(const uint8_t* __restrict__ raw, size_t raw_pitch, int w, int wref,
                                   uint32_t* __restrict__ dst, int pitch_px, int padl) {
    int xp = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y;
    if (xp >= pitch_px) return;
    int x = wref - 2 - (xp - padl);   // source column of stored column s = xp - padl
    uint32_t v = 0;
    if (x >= 0 && x < w) {
        const uint8_t* p = raw + (size_t)y * raw_pitch + 3 * (size_t)x;
        v = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16);
    }
    dst[(size_t)y * pitch_px + xp] = v;
}

int sr_launch_pack_mirror(sr_ctx* ctx, const uint8_t* raw, size_t raw_pitch, int w, int h, int wref, PackedImage* img) {
    dim3 grid(cdiv(img->pitch, 256), h);
    pack_mirror_kernel<<<grid, 256, 0, ctx->stream>>>(raw, raw_pitch, w, wref, img->base, img->pitch, SR_PADL);
    SR_LAUNCHED(ctx);
    return SR_OK;
}

int sr_alloc_packed(sr_ctx* ctx, PackedImage* img, int w, int h, int padr);
static int alloc_packed(sr_ctx* ctx, PackedImage* img, int w, int h, int padr) { return sr_alloc_packed(ctx, img, w, h, padr); }
int sr_alloc_packed(sr_ctx* ctx, PackedImage* img, int w, int h, int padr) {
    img->w = w; img->h = h;
    img->pitch = ((SR_PADL + w + padr + 31) / 32) * 32;
    SR_CUDA(ctx, cudaMalloc((void**)&img->base, (size_t)img->pitch * h * sizeof(uint32_t)));
    SR_CUDA(ctx, cudaMemset(img->base, 0, (size_t)img->pitch * h * sizeof(uint32_t)));   // the padding stays zero for good
    img->px = img->base + SR_PADL;
    return SR_OK;
}

extern "C" {

// Right padding of the packed left image: enough for max_disp up to SR_MAX_DISP_PAD plus a tile.
#define SR_TILE_OVERSHOOT 640

int sr_pair_create(sr_ctx* ctx, int w1, int h1, int w2, int h2, sr_pair** out) {
    if (!ctx || !out) return SR_E_BADARG;
    *out = nullptr;
    if (w1 <= 0 || h1 <= 0 || w2 <= 0 || h2 <= 0) return sr_fail(ctx, SR_E_BADARG, "sr_pair_create: empty image");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    sr_pair* p = new sr_pair();
    p->w1 = w1; p->h1 = h1; p->w2 = w2; p->h2 = h2;
    int rc;
    // the left image is read at x + d: pad by the widest candidate range we can meet (the
    // range is clipped to the image width by the break rule) plus one tile of overshoot
    p->pad_disp = w1 + w2;
    if ((rc = alloc_packed(ctx, &p->L, w1, h1, w2 + SR_TILE_OVERSHOOT + 64)) != SR_OK ||
        (rc = alloc_packed(ctx, &p->R, w2, h2, w1 + SR_TILE_OVERSHOOT + 64)) != SR_OK) {
        sr_pair_destroy(ctx, p);
        return rc;
    }
    size_t npx = (size_t)(w1 > w2 ? w1 : w2) * (size_t)(h1 > h2 ? h1 : h2);
    p->npx_cap = npx;
    cudaError_t e;
    if ((e = cudaMalloc((void**)&p->rawL, (size_t)w1 * h1 * 3)) != cudaSuccess ||
        (e = cudaMalloc((void**)&p->rawR, (size_t)w2 * h2 * 3)) != cudaSuccess ||
        (e = cudaMalloc((void**)&p->best, npx * 8)) != cudaSuccess ||
        (e = cudaMalloc((void**)&p->c0, npx * 8)) != cudaSuccess ||
        (e = cudaMalloc((void**)&p->out, npx * 8)) != cudaSuccess ||
        (e = cudaMalloc((void**)&p->maxblock, 16)) != cudaSuccess) {
        int r = sr_cuda_check(ctx, e, "cudaMalloc(pair)");
        sr_pair_destroy(ctx, p);
        return r;
    }
    p->evals = (unsigned long long*)(p->maxblock + 2);
    *out = p;
    return SR_OK;
}

void sr_pair_destroy(sr_ctx* ctx, sr_pair* p) {
    if (!p) return;
    if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
    sr_pair_peer_detach(ctx, p); cudaFree(p->peer_counter);
    if (p->peer_status_h) cudaFreeHost(p->peer_status_h);
    if (ctx) for (size_t i = 0; i < ctx->peer_pairs.size(); i++) if (ctx->peer_pairs[i] == p) { ctx->peer_pairs.erase(ctx->peer_pairs.begin() + i); break; }
    cudaFree(p->L.base); cudaFree(p->R.base); cudaFree(p->Lm.base); cudaFree(p->Rm.base); cudaFree(p->Rs.base); cudaFree(p->ncc_tab); cudaFree(p->ncc_tab_m); cudaFree(p->out16); cudaFree(p->disp_left_own); cudaFree(p->rawL); cudaFree(p->rawR);
    cudaFree(p->best); cudaFree(p->c0); cudaFree(p->cr); cudaFree(p->dr); cudaFree(p->nwin);
    cudaFree(p->bits); cudaFree(p->alt); cudaFree(p->out); cudaFree(p->outf); cudaFree(p->maxblock);
    delete p;
}

static int upload_one(sr_ctx* ctx, const sr_image* im, uint8_t* raw, PackedImage* pk, int slot) {
    size_t row = (size_t)im->w * 3;
    if (im->pitch < row) return sr_fail(ctx, SR_E_BADARG, "image pitch %zu smaller than a row (%zu)", im->pitch, row);
    cudaPointerAttributes attr;
    bool pinned = cudaPointerGetAttributes(&attr, im->bgr) == cudaSuccess &&
                  (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    if (pinned) {
        SR_CUDA(ctx, cudaMemcpy2DAsync(raw, row, im->bgr, im->pitch, row, im->h, cudaMemcpyHostToDevice, ctx->stream));
    } else {
        // pageable source: stage through the context's pinned buffer (striped host copy, one DMA);
        // the buffer is reused, so the previous upload's DMA must have drained
        SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        uint8_t* st = (uint8_t*)sr_pinned(ctx, slot, row * im->h);
        if (!st) return SR_E_CUDA;
        sr_stage_copy_rows(ctx, st, row, im->bgr, im->pitch, row, im->h);
        SR_CUDA(ctx, cudaMemcpyAsync(raw, st, row * im->h, cudaMemcpyHostToDevice, ctx->stream));
    }
    return sr_launch_pack(ctx, raw, row, im->w, im->h, pk);
}

int sr_pair_upload(sr_ctx* ctx, sr_pair* pair, const sr_image* left, const sr_image* right) {
    if (!ctx || !pair || !left || !right || !left->bgr || !right->bgr) return sr_fail(ctx, SR_E_BADARG, "sr_pair_upload: NULL argument");
    if (left->w != pair->w1 || left->h != pair->h1 || right->w != pair->w2 || right->h != pair->h2)
        return sr_fail(ctx, SR_E_BADARG, "sr_pair_upload: image sizes differ from the pair's");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = upload_one(ctx, left, pair->rawL, &pair->L, 0);
    if (rc != SR_OK) return rc;
    return upload_one(ctx, right, pair->rawR, &pair->R, 1);
}

// The un-rectified images go up, the rectified ones are made on the device (rectification.cpp:486-493: bilinear
// warpPerspective by H_ / Hp_) straight into the pair's raw BGR buffers and packed there: the rectified images never
// exist on the host.  The pair's sizes are the rectified sizes.
static int upload_warped_one(sr_ctx* ctx, const sr_image* im, const double M[9], uint8_t* raw, int w, int h, PackedImage* pk, int slot) {
    size_t row = (size_t)im->w * 3;
    if (im->pitch < row) return sr_fail(ctx, SR_E_BADARG, "image pitch %zu smaller than a row (%zu)", im->pitch, row);
    if (im->w >= 32768 || im->h >= 32768) return sr_fail(ctx, SR_E_SHAPE, "source beyond 32767 pixels (cv::remap's short coordinates)");
    uint8_t* d_src = (uint8_t*)sr_scratch(ctx, 11 + slot, row * im->h);
    if (!d_src) return SR_E_CUDA;
    SR_CUDA(ctx, cudaMemcpy2DAsync(d_src, row, im->bgr, im->pitch, row, im->h, cudaMemcpyHostToDevice, ctx->stream));
    int rc = sr_launch_warp_linear_bgr(ctx, d_src, im->w, im->h, row, raw, w, h, (size_t)w * 3, M);
    if (rc != SR_OK) return rc;
    return sr_launch_pack(ctx, raw, (size_t)w * 3, w, h, pk);
}

int sr_pair_upload_warped(sr_ctx* ctx, sr_pair* pair, const sr_image* left, const double H[9], const sr_image* right, const double Hp[9]) {
    if (!ctx || !pair || !left || !right || !left->bgr || !right->bgr || !H || !Hp)
        return sr_fail(ctx, SR_E_BADARG, "sr_pair_upload_warped: NULL argument");
    if (left->w <= 0 || left->h <= 0 || right->w <= 0 || right->h <= 0) return sr_fail(ctx, SR_E_BADARG, "sr_pair_upload_warped: empty image");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = upload_warped_one(ctx, left, H, pair->rawL, pair->w1, pair->h1, &pair->L, 0);
    if (rc != SR_OK) return rc;
    return upload_warped_one(ctx, right, Hp, pair->rawR, pair->w2, pair->h2, &pair->R, 1);
}

}  // extern "C"
