// ofb200 - host-side copy pool for movie ingest / result egress (no GPU code; built with g++ into
// libofb200_host.so, bound with ctypes from opticalflow_b200/hostio.py).
//
// The reference hands float64 arrays back (optical_flow.py:769, :1189-1205); the device computes in float32, so
// every result plane is widened on the host: ~0.8 GB of freshly allocated destination per 16 pairs of 1024^2, every
// ~20 ms.  Doing that from Python (a thread pool of numpy.copyto pieces) costs a future per 4 MB piece and keeps
// handing the GIL back and forth with the thread that launches the kernels.  Here one call describes a whole
// batch of copies; a persistent pool of plain threads (blocked on a condition variable when idle, no spinning)
// works through it in 1 MB pieces while the caller - which has released the GIL - helps and then waits.
#include <atomic>
#include <condition_variable>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>
#include <immintrin.h>

extern "C" {
typedef struct ofb_host_copy {
    const void* src;
    void* dst;
    size_t count;       // elements
    int kind;           // OFB_COPY_*
    int reserved;
} ofb_host_copy;
}

enum { OFB_COPY_BYTES = 0, OFB_COPY_F32_F64 = 1, OFB_COPY_U8_F64 = 2, OFB_COPY_U16_F64 = 3, OFB_COPY_U8_F32 = 4, OFB_COPY_U16_F32 = 5,
       OFB_COPY_F64_F32 = 6 };

namespace {

bool g_nt = true;        // non-temporal stores for the widening loops

__attribute__((target("avx2"))) void widen_f32_f64_avx2(const float* src, double* dst, size_t n, bool nt) {
    size_t k = 0;
    while (k < n && ((uintptr_t)(dst + k) & 31)) { dst[k] = (double)src[k]; ++k; }
    if (nt) {
        for (; k + 8 <= n; k += 8) {
            const __m256 f = _mm256_loadu_ps(src + k);
            _mm256_stream_pd(dst + k, _mm256_cvtps_pd(_mm256_castps256_ps128(f)));
            _mm256_stream_pd(dst + k + 4, _mm256_cvtps_pd(_mm256_extractf128_ps(f, 1)));
        }
        _mm_sfence();
    } else {
        for (; k + 8 <= n; k += 8) {
            const __m256 f = _mm256_loadu_ps(src + k);
            _mm256_store_pd(dst + k, _mm256_cvtps_pd(_mm256_castps256_ps128(f)));
            _mm256_store_pd(dst + k + 4, _mm256_cvtps_pd(_mm256_extractf128_ps(f, 1)));
        }
    }
    for (; k < n; ++k) dst[k] = (double)src[k];
}

template <typename S, typename D> void convert(const void* src, void* dst, size_t n) {
    const S* s = (const S*)src;
    D* d = (D*)dst;
    for (size_t k = 0; k < n; ++k) d[k] = (D)s[k];
}

size_t elem_sizes(int kind, size_t* dst_size) {
    switch (kind) {
        case OFB_COPY_F32_F64: *dst_size = 8; return 4;
        case OFB_COPY_U8_F64: *dst_size = 8; return 1;
        case OFB_COPY_U16_F64: *dst_size = 8; return 2;
        case OFB_COPY_U8_F32: *dst_size = 4; return 1;
        case OFB_COPY_U16_F32: *dst_size = 4; return 2;
        case OFB_COPY_F64_F32: *dst_size = 4; return 8;
        default: *dst_size = 1; return 1;
    }
}

void run_piece(int kind, const void* src, void* dst, size_t n) {
    static const bool avx2 = __builtin_cpu_supports("avx2");
    switch (kind) {
        case OFB_COPY_F32_F64:
            if (avx2) widen_f32_f64_avx2((const float*)src, (double*)dst, n, g_nt);
            else convert<float, double>(src, dst, n);
            break;
        case OFB_COPY_U8_F64: convert<uint8_t, double>(src, dst, n); break;
        case OFB_COPY_U16_F64: convert<uint16_t, double>(src, dst, n); break;
        case OFB_COPY_U8_F32: convert<uint8_t, float>(src, dst, n); break;
        case OFB_COPY_U16_F32: convert<uint16_t, float>(src, dst, n); break;
        case OFB_COPY_F64_F32: convert<double, float>(src, dst, n); break;
        default: memcpy(dst, src, n); break;
    }
}

struct Piece { int kind; const void* src; void* dst; size_t n; };

// One job at a time (jobs are serialised by job_mutex); workers sleep between jobs.
struct Pool {
    std::vector<std::thread> workers;
    std::mutex m, job_mutex;
    std::condition_variable wake, done_cv;
    const std::vector<Piece>* pieces = nullptr;
    std::atomic<size_t> next{0};
    size_t finished = 0;         // pieces completed of the current job (guarded by m)
    int inside = 0;              // workers currently holding a pointer to the job's piece list (guarded by m)
    uint64_t generation = 0;     // bumped per job (guarded by m)
    bool stop = false;

    void work(const std::vector<Piece>& ps) {
        size_t mine = 0;
        for (;;) {
            const size_t k = next.fetch_add(1);
            if (k >= ps.size()) break;
            run_piece(ps[k].kind, ps[k].src, ps[k].dst, ps[k].n);
            ++mine;
        }
        if (mine) {
            std::lock_guard<std::mutex> lk(m);
            finished += mine;
            if (finished >= ps.size()) done_cv.notify_all();
        }
    }
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            const std::vector<Piece>* ps;
            {
                std::unique_lock<std::mutex> lk(m);
                wake.wait(lk, [&] { return stop || generation != seen; });
                if (stop) return;
                seen = generation;
                ps = pieces;
                if (ps) ++inside;
            }
            if (ps) {
                work(*ps);
                std::lock_guard<std::mutex> lk(m);
                if (--inside == 0) done_cv.notify_all();
            }
        }
    }
    void start(int n) {
        for (int k = 0; k < n; ++k) workers.emplace_back([this] { loop(); });
    }
    void run(const std::vector<Piece>& ps) {
        std::lock_guard<std::mutex> job(job_mutex);
        {
            std::lock_guard<std::mutex> lk(m);
            pieces = &ps;
            next.store(0);
            finished = 0;
            ++generation;
        }
        wake.notify_all();
        work(ps);                                   // the caller helps
        std::unique_lock<std::mutex> lk(m);
        pieces = nullptr;                           // late wakers find nothing; the list outlives everyone who holds it
        done_cv.wait(lk, [&] { return finished >= ps.size() && inside == 0; });
    }
};

Pool* g_pool = nullptr;
std::mutex g_pool_mutex;

}  // namespace

extern "C" {

int ofb_host_version(void) { return 1; }

// Start the pool (once; later calls are ignored).  Returns the number of worker threads.
int ofb_host_pool_start(int n_threads, int non_temporal) {
    std::lock_guard<std::mutex> lk(g_pool_mutex);
    if (g_pool == nullptr) {
        g_nt = non_temporal != 0;
        g_pool = new Pool;
        g_pool->start(n_threads > 0 ? n_threads : 1);
    }
    return (int)g_pool->workers.size();
}

// Perform all copies of `descs` (regions must not overlap); returns when every byte has been written.
int ofb_host_copy_many(const ofb_host_copy* descs, int n, size_t piece_bytes) {
    if (n <= 0) return 0;
    if (g_pool == nullptr) ofb_host_pool_start(4, 1);
    if (piece_bytes < 4096) piece_bytes = 1 << 20;
    std::vector<Piece> pieces;
    for (int d = 0; d < n; ++d) {
        size_t dsz = 1;
        const size_t ssz = elem_sizes(descs[d].kind, &dsz);
        const size_t step = piece_bytes / dsz > 0 ? piece_bytes / dsz : 1;
        for (size_t a = 0; a < descs[d].count; a += step) {
            const size_t cnt = descs[d].count - a < step ? descs[d].count - a : step;
            pieces.push_back({descs[d].kind, (const char*)descs[d].src + a * ssz, (char*)descs[d].dst + a * dsz, cnt});
        }
    }
    g_pool->run(pieces);
    return 0;
}

}  // extern "C"
