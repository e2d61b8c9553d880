// host_stage.cu -- worker threads that move pageable host memory to and from the context's pinned staging.
//
// A cv::Mat (the reference's image and disparity containers, main.cpp:80-101) lives in pageable memory, which
// the copy engines cannot read: every byte has to pass through a pinned buffer first.  One host thread
// copies ~10 GB/s, far below the ~55 GB/s of the PCIe link, so the staging copy is split into row stripes
// over a small pool of threads that is created on first use and kept for the life of the context.  Host code
// only: nothing here touches the device.
#include "sr_internal.cuh"
#include <string.h>
#include <stdlib.h>
#include <condition_variable>
#include <mutex>
#include <thread>
#if defined(__x86_64__)
#include <emmintrin.h>
#endif

struct sr_stage_pool {
    std::vector<std::thread> th;
    std::mutex m;
    std::condition_variable cv_work, cv_done;
    uint64_t gen = 0;
    int running = 0;
    bool quit = false;
    void (*fn)(int, int, void*) = nullptr;
    void* arg = nullptr;
};

static void stage_worker(sr_stage_pool* p, int t, int nthreads) {
    uint64_t seen = 0;
    for (;;) {
        void (*fn)(int, int, void*);
        void* arg;
        {
            std::unique_lock<std::mutex> lk(p->m);
            p->cv_work.wait(lk, [&] { return p->quit || p->gen != seen; });
            if (p->quit) return;
            seen = p->gen;
            fn = p->fn;
            arg = p->arg;
        }
        fn(t, nthreads, arg);
        {
            std::lock_guard<std::mutex> lk(p->m);
            if (--p->running == 0) p->cv_done.notify_all();
        }
    }
}

sr_stage_pool* sr_stage_pool_get(sr_ctx* ctx) {
    if (ctx->stage_pool) return ctx->stage_pool;
    const char* env = getenv("SR_STAGE_THREADS");
    int n = env ? atoi(env) : 0;
    if (n <= 0) {
        // half the cores, at most 8 (4K pair on a 16-core host: 4 threads 3.7 ms, 8 -> 2.8, 12 -> 3.1, 16 -> 3.5)
        int hc = (int)std::thread::hardware_concurrency();
        n = hc / 2;
        if (n > 8) n = 8;
    }
    if (n < 1) n = 1;
    if (n > 64) n = 64;
    sr_stage_pool* p = new sr_stage_pool;
    try {
        for (int t = 0; t < n; t++) p->th.emplace_back(stage_worker, p, t, n);
    } catch (...) {
        // fewer threads than asked for (thread limit of the process): the stripes are cut by the creation-time
        // count, so start over with what could be created
        int got = (int)p->th.size();
        { std::lock_guard<std::mutex> lk(p->m); p->quit = true; }
        p->cv_work.notify_all();
        for (auto& t : p->th) t.join();
        delete p;
        if (got < 1) { sr_fail(ctx, SR_E_CUDA, "cannot start a host staging thread"); return nullptr; }
        p = new sr_stage_pool;
        for (int t = 0; t < got; t++) p->th.emplace_back(stage_worker, p, t, got);
    }
    ctx->stage_pool = p;
    return p;
}

void sr_stage_pool_destroy(sr_stage_pool* p) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(p->m);
        p->quit = true;
    }
    p->cv_work.notify_all();
    for (auto& t : p->th) t.join();
    delete p;
}

int sr_stage_threads(const sr_stage_pool* p) { return (int)p->th.size(); }

void sr_stage_run(sr_stage_pool* p, void (*fn)(int, int, void*), void* arg) {
    {
        std::lock_guard<std::mutex> lk(p->m);
        p->fn = fn;
        p->arg = arg;
        p->running = (int)p->th.size();
        p->gen++;
    }
    p->cv_work.notify_all();
}

void sr_stage_wait(sr_stage_pool* p) {
    std::unique_lock<std::mutex> lk(p->m);
    p->cv_done.wait(lk, [&] { return p->running == 0; });
}

// Large copies with streaming stores: the destination lines are written without being read first and without
// displacing the source from the caches (a third less memory traffic than a write-allocate copy).  The fence
// makes the data globally visible before the caller publishes "staged" to the thread that queues the DMA.
static const bool g_stream_copy = !(getenv("SR_STAGE_STREAM") && atoi(getenv("SR_STAGE_STREAM")) == 0);
static void copy_bytes(char* dst, const char* src, size_t n) {
#if defined(__x86_64__)
    if (g_stream_copy && n >= 16384) {
        size_t head = (16 - ((uintptr_t)dst & 15)) & 15;
        memcpy(dst, src, head);
        dst += head; src += head; n -= head;
        size_t blocks = n / 64;
        for (size_t i = 0; i < blocks; i++) {
            __m128i a = _mm_loadu_si128((const __m128i*)src), b = _mm_loadu_si128((const __m128i*)(src + 16));
            __m128i c = _mm_loadu_si128((const __m128i*)(src + 32)), d = _mm_loadu_si128((const __m128i*)(src + 48));
            _mm_stream_si128((__m128i*)dst, a);
            _mm_stream_si128((__m128i*)(dst + 16), b);
            _mm_stream_si128((__m128i*)(dst + 32), c);
            _mm_stream_si128((__m128i*)(dst + 48), d);
            src += 64; dst += 64;
        }
        _mm_sfence();
        n -= blocks * 64;
    }
#endif
    memcpy(dst, src, n);
}

void sr_stage_stripe(char* dst, size_t dpitch, const char* src, size_t spitch, size_t rowbytes, int row0, int row1,
                     int t, int nthreads) {
    int n = row1 - row0;
    if (n <= 0) return;
    int per = (n + nthreads - 1) / nthreads;
    int a = row0 + t * per, b = a + per < row1 ? a + per : row1;
    if (a >= b) return;
    if (dpitch == rowbytes && spitch == rowbytes) {
        copy_bytes(dst + (size_t)a * dpitch, src + (size_t)a * spitch, (size_t)(b - a) * rowbytes);
        return;
    }
    for (int y = a; y < b; y++) copy_bytes(dst + (size_t)y * dpitch, src + (size_t)y * spitch, rowbytes);
}

namespace {
struct CopyJob {
    char* dst; size_t dpitch; const char* src; size_t spitch; size_t rowbytes; int rows;
};
void copy_job(int t, int nthreads, void* arg) {
    const CopyJob* j = (const CopyJob*)arg;
    sr_stage_stripe(j->dst, j->dpitch, j->src, j->spitch, j->rowbytes, 0, j->rows, t, nthreads);
}
}  // namespace

int sr_stage_copy_rows(sr_ctx* ctx, void* dst, size_t dpitch, const void* src, size_t spitch, size_t rowbytes, int rows) {
    if (rows <= 0 || !rowbytes) return SR_OK;
    sr_stage_pool* p = rowbytes * (size_t)rows >= ((size_t)1 << 20) ? sr_stage_pool_get(ctx) : nullptr;
    if (!p || sr_stage_threads(p) == 1) {
        sr_stage_stripe((char*)dst, dpitch, (const char*)src, spitch, rowbytes, 0, rows, 0, 1);
        return SR_OK;
    }
    CopyJob j{(char*)dst, dpitch, (const char*)src, spitch, rowbytes, rows};
    sr_stage_run(p, copy_job, &j);
    sr_stage_wait(p);
    return SR_OK;
}
