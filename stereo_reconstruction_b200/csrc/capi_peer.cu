// capi_peer.cu -- band gather over peer memory (NVLink / NVSwitch), no collective library on the data path.
//
// Row-band sharding of one very large pair ends with every rank's band travelling to rank 0.
// Here rank 0 ("root") exports a gather window -- synchronisation words followed by an int16 map -- as a CUDA
// IPC handle; the other ranks map it and their push kernel narrows the band to int16 and stores it straight
// into the root's HBM through NVLink, then publishes the step number in the root's flag word of that rank.
// The root's collect kernel widens the rows of a rank into the CV_64F result as soon as that rank's flag
// shows the step, so early bands are converted while late ones still arrive, and acknowledges the step so
// that the next push may overwrite the window.  One kernel per rank and step, no staging copy, no rendezvous
// with the host.
//
// Window layout (one cudaMalloc of the root, >= 4 MiB so that it is an allocation of its own):
//   word 0..31   flag[r]: last step pushed by rank r        (written by rank r through NVLink, polled by the root)
//   word 32      ack: last step the root has collected      (written by the root, polled remotely by the pushers)
//   byte 4096..  int16 map, output width x height, tight
#include "sr_internal.cuh"
#include <stdlib.h>
#include <string.h>

namespace {

constexpr int WIN_HDR = 4096;
constexpr int WIN_ACK = 32;
constexpr int MAX_RANKS = 32;

struct BandTable { int n; int start[MAX_RANKS + 1]; };

// A lost peer must not hang the GPU: every wait gives up after `timeout_ns` (SR_PEER_TIMEOUT_MS, default 10 s -- far
// above the first-step skew between processes: lazy module loading, work-list allocation) and records SR_E_CUDA in the
// pair's sticky status word (mapped host memory: the host sees it without a copy, nothing resets it but a new
// export / attach).
__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ bool spin_until(const volatile unsigned* word, unsigned want, int* status, unsigned long long timeout_ns) {
    unsigned long long t0 = now_ns();
    unsigned ns = 32;
    while ((int)(*word - want) < 0) {
        if (*(volatile int*)status == SR_E_CUDA) return false;      // another block gave up already
        if (now_ns() - t0 > timeout_ns) { *status = SR_E_CUDA; return false; }
        __nanosleep(ns);
        if (ns < 1024) ns *= 2;
    }
    return true;
}

unsigned long long peer_timeout_ns() {
    static const char* e = getenv("SR_PEER_TIMEOUT_MS");
    double ms = e ? atof(e) : 10000.0;
    if (!(ms > 0)) ms = 10000.0;
    return (unsigned long long)(ms * 1e6);
}

// one thread: returns when the root has acknowledged step seq-1 (or the wait timed out: status set)
__global__ void peer_wait_ack_kernel(const volatile unsigned* win, unsigned seq, int* status, unsigned long long timeout_ns) {
    spin_until(win + 32, seq - 1, status, timeout_ns);
}

__device__ __forceinline__ short narrow(double v, int* status) {
    int q = (int)v;
    if ((double)q != v || q < -32768 || q > 32767) { *status = SR_E_BADARG; q = 0; }
    return (short)q;
}

// rows [e0, e1) (element indices of the tight map) of the local CV_64F result -> the root's int16 window
__global__ void __launch_bounds__(256) push16_kernel(const double* __restrict__ src, short* dst, size_t e0, size_t e1,
                                                     volatile unsigned* win, int rank, unsigned seq,
                                                     unsigned* counter, int* status, unsigned long long timeout_ns) {
    __shared__ int ok;
    if (threadIdx.x == 0) ok = spin_until(win + WIN_ACK, seq - 1, status, timeout_ns);   // previous step consumed by the root
    __syncthreads();
    if (ok) {
        // quads of the absolute element index space, so that the 8-byte stores are aligned in the window
        size_t q0 = e0 >> 2, q1 = (e1 + 3) >> 2;
        for (size_t q = q0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < q1; q += (size_t)gridDim.x * blockDim.x) {
            size_t a = q << 2;
            if (a >= e0 && a + 4 <= e1) {
                double2 v0 = *(const double2*)(src + a), v1 = *(const double2*)(src + a + 2);
                short4 o = make_short4(narrow(v0.x, status), narrow(v0.y, status), narrow(v1.x, status), narrow(v1.y, status));
                *(short4*)(dst + a) = o;
            } else {
                for (int k = 0; k < 4; k++)
                    if (a + k >= e0 && a + k < e1) dst[a + k] = narrow(src[a + k], status);
            }
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(counter, 1u) == gridDim.x - 1) {   // last block: every block's stores are fenced
            *counter = 0;
            __threadfence_system();
            if (*(volatile int*)status == SR_OK) win[rank] = seq;   // a failed step is never published
        }
    }
}

// root: one block per row of the bands of ranks 1..n-1; waits for the owner's flag, widens the row
__global__ void __launch_bounds__(256) collect16_kernel(const short* win16, double* __restrict__ out, int ow, BandTable bt,
                                                        volatile unsigned* win, unsigned seq, unsigned* counter, int* status,
                                                        unsigned long long timeout_ns) {
    __shared__ int ok;
    const int y = bt.start[1] + blockIdx.x;
    if (threadIdx.x == 0) {
        int r = 1;
        while (r + 1 < bt.n && y >= bt.start[r + 1]) r++;
        ok = spin_until(win + r, seq, status, timeout_ns);
        __threadfence_system();
    }
    __syncthreads();
    if (ok) {
        size_t e0 = (size_t)y * ow, e1 = e0 + ow;
        size_t q0 = e0 >> 2, q1 = (e1 + 3) >> 2;
        for (size_t q = q0 + threadIdx.x; q < q1; q += blockDim.x) {
            size_t a = q << 2;
            if (a >= e0 && a + 4 <= e1) {
                // .cg: the rows were written by another GPU into this GPU's L2; L1 may hold the previous step's
                uint2 raw = __ldcg((const uint2*)(win16 + a));
                short s0 = (short)(raw.x & 0xFFFF), s1 = (short)(raw.x >> 16), s2 = (short)(raw.y & 0xFFFF), s3 = (short)(raw.y >> 16);
                *(double2*)(out + a) = make_double2((double)s0, (double)s1);
                *(double2*)(out + a + 2) = make_double2((double)s2, (double)s3);
            } else {
                for (int k = 0; k < 4; k++)
                    if (a + k >= e0 && a + k < e1) out[a + k] = (double)__ldcg(win16 + a + k);
            }
        }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(counter, 1u) == gridDim.x - 1) {
            *counter = 0;
            __threadfence_system();
            // the pushers may overwrite the window now -- unless a band is missing: then the step is NOT acknowledged,
            // the pushers' next wait times out as well and every rank reports the failure
            if (*(volatile int*)status == SR_OK) win[WIN_ACK] = seq;
        }
    }
}

size_t window_bytes(const sr_pair* pair) {
    size_t b = WIN_HDR + pair->npx_cap * sizeof(short);
    return b < ((size_t)4 << 20) ? ((size_t)4 << 20) : b;
}

int ensure_counter(sr_ctx* ctx, sr_pair* pair) {
    if (!pair->peer_counter) {
        SR_CUDA(ctx, cudaMalloc((void**)&pair->peer_counter, 256));
        SR_CUDA(ctx, cudaHostAlloc((void**)&pair->peer_status_h, 64, cudaHostAllocMapped));
        SR_CUDA(ctx, cudaHostGetDevicePointer((void**)&pair->peer_status_d, pair->peer_status_h, 0));
        bool known = false;
        for (sr_pair* q : ctx->peer_pairs) known = known || q == pair;
        if (!known) ctx->peer_pairs.push_back(pair);
    }
    // a new export / attach starts a new generation of the protocol: counter and status start clean
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    SR_CUDA(ctx, cudaMemset(pair->peer_counter, 0, 256));
    *(volatile int*)pair->peer_status_h = SR_OK;
    return SR_OK;
}

}  // namespace

int sr_peer_check(sr_ctx* ctx, sr_pair* pair) {
    if (!pair || !pair->peer_status_h) return SR_OK;
    int st = *(volatile int*)pair->peer_status_h;
    if (st == SR_OK) return SR_OK;
    if (st == SR_E_BADARG) return sr_fail(ctx, st, "band gather: the map holds values that are not integers in int16 range (e.g. a sub-pixel map)");
    return sr_fail(ctx, st, "band gather over peer memory: a band or an acknowledgement did not arrive within SR_PEER_TIMEOUT_MS; "
                            "the step was not published (sticky until the window is exported / attached again)");
}
short* sr_peer_window16(sr_pair* pair) { return (short*)((char*)pair->peer_win + WIN_HDR); }
volatile unsigned* sr_peer_flags(sr_pair* pair) { return (volatile unsigned*)pair->peer_win; }
int sr_peer_wait_ack(sr_ctx* ctx, sr_pair* pair, unsigned seq) {
    peer_wait_ack_kernel<<<1, 1, 0, ctx->stream>>>(sr_peer_flags(pair), seq, pair->peer_status_d, peer_timeout_ns());
    SR_LAUNCHED(ctx);
    return SR_OK;
}

extern "C" {

int sr_pair_peer_export(sr_ctx* ctx, sr_pair* pair, void* handle) {
    if (!ctx || !pair) return sr_fail(ctx, SR_E_BADARG, "sr_pair_peer_export: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == SR_PEER_HANDLE_BYTES, "handle size");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    if (pair->peer_mode == 2 || pair->peer_mode == 3) return sr_fail(ctx, SR_E_BADARG, "the pair is attached to another root's window");
    if (!pair->peer_win) {
        SR_CUDA(ctx, cudaMalloc(&pair->peer_win, window_bytes(pair)));
        pair->peer_mode = 1;
    }
    int rc = ensure_counter(ctx, pair);   // synchronises the stream
    if (rc != SR_OK) return rc;
    // Every export starts a new generation: flags and acknowledgement go back to 0, so that a second gather on the same
    // pair (steps restart at 1) cannot mistake the previous generation's step numbers for its own.  The pushers of the
    // previous generation must have detached (they attach again with the handle this call returns).
    SR_CUDA(ctx, cudaMemset(pair->peer_win, 0, WIN_HDR));
    if (handle) {
        cudaIpcMemHandle_t h;
        SR_CUDA(ctx, cudaIpcGetMemHandle(&h, pair->peer_win));
        memcpy(handle, &h, sizeof h);
    }
    return SR_OK;
}

int sr_pair_peer_detach(sr_ctx* ctx, sr_pair* pair) {
    if (!pair) return SR_OK;
    if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
    if (pair->peer_mode == 1) cudaFree(pair->peer_win);
    else if (pair->peer_mode == 2) cudaIpcCloseMemHandle(pair->peer_win);
    pair->peer_win = nullptr;
    pair->peer_mode = 0;
    if (pair->peer_status_h) *(volatile int*)pair->peer_status_h = SR_OK;
    return SR_OK;
}

int sr_pair_peer_attach(sr_ctx* ctx, sr_pair* pair, const void* handle) {
    if (!ctx || !pair || !handle) return sr_fail(ctx, SR_E_BADARG, "sr_pair_peer_attach: NULL argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    sr_pair_peer_detach(ctx, pair);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    void* base = nullptr;
    SR_CUDA(ctx, cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    pair->peer_win = base;
    pair->peer_mode = 2;
    return ensure_counter(ctx, pair);
}

int sr_pair_peer_attach_local(sr_ctx* ctx, sr_pair* pair, sr_pair* root) {
    if (!ctx || !pair || !root || !root->peer_win || root->peer_mode != 1)
        return sr_fail(ctx, SR_E_BADARG, "sr_pair_peer_attach_local: the root pair has no exported window");
    if (pair->w2 != root->w2 || pair->h2 != root->h2) return sr_fail(ctx, SR_E_BADARG, "sr_pair_peer_attach_local: shapes differ");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    sr_pair_peer_detach(ctx, pair);
    pair->peer_win = root->peer_win;
    pair->peer_mode = 3;
    return ensure_counter(ctx, pair);
}

int sr_search_push_band(sr_ctx* ctx, sr_pair* pair, int row0, int row1, int rank, unsigned seq) {
    if (!ctx || !pair) return sr_fail(ctx, SR_E_BADARG, "sr_search_push_band: NULL argument");
    if (pair->peer_mode != 2 && pair->peer_mode != 3) return sr_fail(ctx, SR_E_BADARG, "sr_search_push_band: not attached to a root window");
    if (rank < 1 || rank >= MAX_RANKS || seq == 0) return sr_fail(ctx, SR_E_BADARG, "sr_search_push_band: rank 1..%d, step >= 1", MAX_RANKS - 1);
    if (pair->ow != pair->w2 || pair->oh != pair->h2) return sr_fail(ctx, SR_E_BADARG, "sr_search_push_band: the last run was not right-referenced");
    if (row0 < 0) row0 = 0;
    if (row1 > pair->oh) row1 = pair->oh;
    if (row1 < row0) row1 = row0;
    int st = sr_peer_check(ctx, pair);   // a failure of an earlier step is reported by every later call
    if (st != SR_OK) return st;
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    size_t e0 = (size_t)row0 * pair->ow, e1 = (size_t)row1 * pair->ow;
    size_t quads = ((e1 + 3) >> 2) - (e0 >> 2);
    unsigned grid = (unsigned)((quads + 255) / 256);
    if (grid > (unsigned)ctx->sms * 4) grid = (unsigned)ctx->sms * 4;
    if (grid < 1) grid = 1;
    unsigned* win = (unsigned*)pair->peer_win;
    push16_kernel<<<grid, 256, 0, ctx->stream>>>(pair->out, (short*)((char*)pair->peer_win + WIN_HDR), e0, e1, win, rank, seq,
                                                 pair->peer_counter, pair->peer_status_d, peer_timeout_ns());
    SR_LAUNCHED(ctx);
    return SR_OK;
}

int sr_search_collect_bands(sr_ctx* ctx, sr_pair* pair, int nranks, unsigned seq) {
    if (!ctx || !pair) return sr_fail(ctx, SR_E_BADARG, "sr_search_collect_bands: NULL argument");
    if (pair->peer_mode != 1) return sr_fail(ctx, SR_E_BADARG, "sr_search_collect_bands: the pair has no exported window");
    if (nranks < 1 || nranks > MAX_RANKS || seq == 0) return sr_fail(ctx, SR_E_BADARG, "sr_search_collect_bands: 1..%d ranks, step >= 1", MAX_RANKS);
    if (pair->ow != pair->w2 || pair->oh != pair->h2) return sr_fail(ctx, SR_E_BADARG, "sr_search_collect_bands: the last run was not right-referenced");
    int st = sr_peer_check(ctx, pair);
    if (st != SR_OK) return st;
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    BandTable bt;
    bt.n = nranks;
    for (int r = 0; r < nranks; r++) { int a, b; sr_band_rows(pair->oh, r, nranks, &a, &b); bt.start[r] = a; bt.start[r + 1] = b; }
    int rows = pair->oh - bt.start[1];
    unsigned* win = (unsigned*)pair->peer_win;
    if (nranks == 1 || rows <= 0) return SR_OK;
    collect16_kernel<<<rows, 256, 0, ctx->stream>>>((const short*)((char*)pair->peer_win + WIN_HDR), pair->out, pair->ow, bt, win, seq,
                                                    pair->peer_counter, pair->peer_status_d, peer_timeout_ns());
    SR_LAUNCHED(ctx);
    return SR_OK;
}

// Synchronises the context stream and reports the sticky status of the pair's gather protocol (every rank, any time).
int sr_pair_peer_status(sr_ctx* ctx, sr_pair* pair) {
    if (!ctx || !pair) return sr_fail(ctx, SR_E_BADARG, "sr_pair_peer_status: NULL argument");
    SR_CUDA(ctx, cudaSetDevice(ctx->device));
    SR_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return sr_peer_check(ctx, pair);
}

}  // extern "C"
