/* b2d_hostcopy.cpp — the one host-side copy of the streaming paths: caller's (pageable) bytes -> page-locked staging,
 * and page-locked staging -> caller's buffers.  Host code only (no CUDA).
 *
 * The reference copies every block once as well (`data = bytes(b)`, gzip_ng_threaded.py:315), on the thread that
 * holds the GIL.  Here the copy is the slowest stage of the Python-facing writer once the kernels run at tens of GB/s,
 * so it gets (a) a persistent pool of worker threads instead of a thread spawn per call and (b) non-temporal stores:
 * the destination is read next by the GPU's DMA engine (or is a buffer far larger than the caches), so there is no
 * point in reading its cache lines in before overwriting them — that saves a third of the copy's DRAM traffic.
 */
#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

#include "b2d_hostcopy.h"

namespace {

#if defined(__x86_64__)
__attribute__((target("avx2"))) void copy_stream_avx2(uint8_t *d, const uint8_t *s, size_t n) {
    size_t head = (32 - ((uintptr_t)d & 31)) & 31;          /* up to 32-byte alignment of the destination */
    if (head > n) head = n;
    memcpy(d, s, head);
    d += head; s += head; n -= head;
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256((const __m256i *)(s + i)), b = _mm256_loadu_si256((const __m256i *)(s + i + 32));
        const __m256i c = _mm256_loadu_si256((const __m256i *)(s + i + 64)), e = _mm256_loadu_si256((const __m256i *)(s + i + 96));
        _mm256_stream_si256((__m256i *)(d + i), a);
        _mm256_stream_si256((__m256i *)(d + i + 32), b);
        _mm256_stream_si256((__m256i *)(d + i + 64), c);
        _mm256_stream_si256((__m256i *)(d + i + 96), e);
    }
    _mm_sfence();
    memcpy(d + i, s + i, n - i);
}
const bool g_avx2 = __builtin_cpu_supports("avx2");
#endif
/* B2D_COPY_NT=0 switches the non-temporal path off (A/B measurements) */
const bool g_nt = !(getenv("B2D_COPY_NT") && getenv("B2D_COPY_NT")[0] == '0');

inline void copy_piece(uint8_t *d, const uint8_t *s, size_t n) {
#if defined(__x86_64__)
    if (g_nt && g_avx2 && n >= 65536) { copy_stream_avx2(d, s, n); return; }
#endif
    memcpy(d, s, n);
}

/* persistent helper threads: a task is a closure; the submitter keeps whatever the closure points at alive until every
 * copy of it has run (b2d_par_memcpy waits for that) */
struct pool {
    std::mutex mu;
    std::condition_variable cv;
    std::deque<std::function<void()>> q;
    unsigned threads = 0, busy = 0;
    void worker() {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv.wait(lk, [&] { return !q.empty(); });
            std::function<void()> t = std::move(q.front());
            q.pop_front();
            ++busy;
            lk.unlock();
            t();
            lk.lock();
            --busy;
        }
    }
    void run(unsigned copies, const std::function<void()> &f) {
        if (!copies) return;
        {
            std::lock_guard<std::mutex> g(mu);
            for (unsigned i = 0; i < copies; ++i) q.push_back(f);
            /* a thread for everything queued or running (several callers may be copying at once), at most 48 */
            while (threads < q.size() + busy && threads < 48) {
                std::thread([this] { worker(); }).detach();     /* the pool lives as long as the process */
                ++threads;
            }
        }
        cv.notify_all();
    }
};
/* on the heap and never destroyed: its threads are still waiting on it when the process runs its static destructors */
pool &g_pool = *new pool;

}  // namespace

void b2d_par_memcpy(void *dst, const void *src, size_t n, unsigned nt) {
    const size_t piece = 2ull << 20;
    if (nt > 32) nt = 32;
    if (nt < 2 || n < 2 * piece) { copy_piece((uint8_t *)dst, (const uint8_t *)src, n); return; }
    /* fork / join over the pool; this frame holds the job until the last helper has left it */
    struct shared {
        uint8_t *dst; const uint8_t *src; size_t n, piece, pieces;
        std::atomic<size_t> next{0};
        std::atomic<unsigned> left{0};    /* helpers that have not finished yet */
        std::mutex mu; std::condition_variable cv;
    } J;
    J.dst = (uint8_t *)dst; J.src = (const uint8_t *)src; J.n = n; J.piece = piece; J.pieces = (n + piece - 1) / piece;
    unsigned helpers = nt - 1;
    if ((size_t)helpers + 1 > J.pieces) helpers = (unsigned)(J.pieces - 1);
    J.left.store(helpers);
    shared *j = &J;
    auto body = [j] {
        for (;;) {
            const size_t k = j->next.fetch_add(1);
            if (k >= j->pieces) return;
            const size_t o = k * j->piece, m = j->n - o < j->piece ? j->n - o : j->piece;
            copy_piece(j->dst + o, j->src + o, m);
        }
    };
    g_pool.run(helpers, [j, body] {
        body();
        std::lock_guard<std::mutex> g(j->mu);       /* under the lock: the owner cannot leave between the store and the notify */
        j->left.fetch_sub(1);
        j->cv.notify_all();
    });
    body();                                          /* the calling thread works too */
    std::unique_lock<std::mutex> lk(J.mu);
    J.cv.wait(lk, [&] { return J.left.load() == 0; });
}
