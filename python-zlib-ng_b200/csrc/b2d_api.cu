/* b2d_api.cu — host side of libb2d.so: the C ABI declared in include/b2d.h.
 *
 * A context owns one CUDA stream, the per-batch work buffers of the pipeline
 * (b2d_kernels.cu), a pinned staging pair for host-buffer calls and the
 * checksum workspace (b2d_checksum.cu).  Big inputs are cut into batches of
 * `max_batch` bytes; batch j > 0 is primed with the 32 KiB of input before it,
 * so the result does not depend on the batch size (blocks are
 * independent given the raw input).
 */
#include <cuda_runtime.h>
#include <sys/mman.h>
#include <condition_variable>
#include <map>
#include <mutex>
#include <thread>
#include <vector>
#include <chrono>
#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/b2d.h"
#include "b2d_kernels.h"
#include "b2d_inflate.h"
#include "b2d_hostcopy.h"

static thread_local char g_err[512];
static int fail_cuda(cudaError_t e, const char *where) {
    snprintf(g_err, sizeof g_err, "CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), where);
    cudaGetLastError();
    return e == cudaErrorMemoryAllocation ? B2D_MEM_ERROR : B2D_CUDA_ERROR;
}
static int fail(int code, const char *msg) { snprintf(g_err, sizeof g_err, "%s", msg); return code; }
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return fail_cuda(e_, #x); } while (0)

extern "C" const char *b2d_last_error(void) { return g_err; }
extern "C" int b2d_device_count(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; } return n; }

static inline uint64_t align_up(uint64_t x, uint64_t a) { return (x + a - 1) / a * a; }

extern "C" uint64_t b2d_deflate_bound(uint64_t n) {
    const uint64_t nsub = (n + B2D_SUB - 1) / B2D_SUB, nblk = (n + B2D_BLOCK - 1) / B2D_BLOCK;
    /* stored sub-block: <= 1 byte of header+pad, 4 bytes LEN/NLEN; block marker: <= 5 bytes; final 03 00 */
    return n + 5 * nsub + 5 * (nblk ? nblk : 1) + 2 + 8;
}

/* ---- page-locked host memory ------------------------------------------------------------------------------
 * cudaHostAlloc pins fresh pages at ~2.7 GB/s on the test box (it faults and pins them 4 KiB at a time on one thread):
 * 1.6 GiB of staging per GPU would cost the first call of a process 0.6 s.  Memory that is mapped with transparent huge
 * pages, touched by several threads and then registered is ready 5-6 times faster (tools/pin_probe.cu: 15 GB/s), and
 * DMA treats it like any other page-locked memory.  Falls back to cudaHostAlloc where registering is refused. */
static std::mutex g_pin_mu;
static std::map<void *, size_t> g_pin_mapped;          /* blocks that came from mmap + cudaHostRegister */

static cudaError_t pinned_alloc(void **p, size_t bytes) {
    if (bytes == 0) bytes = 1;
    const size_t huge = 2ull << 20;
    const size_t len = (bytes + huge - 1) / huge * huge;
    if (len >= (8ull << 20)) {
        void *m = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (m != MAP_FAILED) {
            madvise(m, len, MADV_HUGEPAGE);
            {   /* first touch on several threads */
                unsigned nt = (unsigned)(len / (32ull << 20));
                if (nt < 1) nt = 1;
                if (nt > 8) nt = 8;
                std::thread th[8];
                const size_t per = (len / nt + huge - 1) / huge * huge;
                for (unsigned i = 0; i < nt; ++i)
                    th[i] = std::thread([=] {
                        const size_t o = (size_t)i * per;
                        const size_t k = o >= len ? 0 : (len - o < per ? len - o : per);
                        for (size_t q = 0; q < k; q += 4096) ((volatile char *)m)[o + q] = 0;
                    });
                for (unsigned i = 0; i < nt; ++i) th[i].join();
            }
            if (cudaHostRegister(m, len, cudaHostRegisterPortable) == cudaSuccess) {
                std::lock_guard<std::mutex> g(g_pin_mu);
                g_pin_mapped[m] = len;
                *p = m;
                return cudaSuccess;
            }
            cudaGetLastError();
            munmap(m, len);
        }
    }
    return cudaHostAlloc(p, bytes, cudaHostAllocPortable);
}
static cudaError_t pinned_free(void *p) {
    if (!p) return cudaSuccess;
    size_t len = 0;
    {
        std::lock_guard<std::mutex> g(g_pin_mu);
        auto it = g_pin_mapped.find(p);
        if (it != g_pin_mapped.end()) { len = it->second; g_pin_mapped.erase(it); }
    }
    if (len) {
        cudaError_t e = cudaHostUnregister(p);
        munmap(p, len);
        return e;
    }
    return cudaFreeHost(p);
}

#define PROF_BATCHES 64
/* decode streams of a host inflate call: neighbouring chunks decode side by side.  (Eight streams — every chunk of a 512 MiB
 * call on its own — measured slower than four: 29 ms against 25 ms per call; a sub-block takes a warp 7-10 ms whatever
 * shares the SM with it, so what matters is how early the first chunks' warps start, tools/reader_trace.py.) */
#define B2D_INF_DEC_STREAMS 4
#ifndef B2D_NLANE
#define B2D_NLANE 4     /* batches of one call in flight on the device, each on its own stream with its own intermediates.  The
                         * chain stage runs one warp per 128 KiB block, so a batch of 64 MiB fills a quarter of the device's warp
                         * slots: four lanes of 64 MiB do what two of 256 MiB do (profiles/ab_lanes_r2.log) with a quarter of the
                         * memory and of the pipeline's fill and drain time */
#endif
#define B2D_NSLOT (B2D_NLANE + 1)   /* input/output slots of the host pipeline: one more than compute lanes */
/* per-batch intermediates of the kernel pipeline.  Two of them ("lanes") let consecutive batches of one
 * call run on two streams, so one batch's kernels fill the SMs the other's tail waves leave idle. */
struct b2d_work {
    uint16_t *prevd; uint32_t *match, *tok, *hist;
    b2d_subinfo *info; b2d_plan *plan;
    uint64_t *block_off;
};
static void work_free(b2d_work *w) {
    cudaFree(w->prevd); cudaFree(w->match); cudaFree(w->tok); cudaFree(w->hist); cudaFree(w->info); cudaFree(w->plan);
    cudaFree(w->block_off);
    memset(w, 0, sizeof(*w));
}
static cudaError_t work_alloc(b2d_work *w, uint64_t max_batch) {
    const uint64_t nsub = max_batch / B2D_SUB, nblk = max_batch / B2D_BLOCK;
    cudaError_t e;
    if ((e = cudaMalloc(&w->prevd, (nblk * (uint64_t)B2D_LINKS + 64) * sizeof(uint16_t))) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w->match, max_batch * sizeof(uint32_t))) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w->tok, max_batch * sizeof(uint32_t))) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w->hist, nsub * B2D_HIST * sizeof(uint32_t))) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w->info, nsub * sizeof(b2d_subinfo))) != cudaSuccess) return e;
    if ((e = cudaMalloc(&w->plan, nsub * sizeof(b2d_plan))) != cudaSuccess) return e;
    return cudaMalloc(&w->block_off, (nblk + 1) * sizeof(uint64_t));
}

struct b2d_ctx {
    int device;
    uint64_t max_batch;
    cudaStream_t st;
    cudaEvent_t ev0, ev1;
    std::recursive_mutex mu;       /* recursive: the compound entry points (compress + index, host inflate) hold it across their parts */
    /* work buffers for one batch */
    uint8_t *d_in;  uint64_t d_in_cap;      /* WINDOW + max_batch + pad (host-buffer calls)   */
    uint8_t *d_out; uint64_t d_out_cap;     /* bound(max_batch) rounded (host-buffer calls)   */
    b2d_work wk[B2D_NLANE];                 /* lanes 1.. are allocated on first use (ensure_second_lane) */
    cudaStream_t stl[B2D_NLANE];            /* lane streams; stl[0] == st                      */
    cudaEvent_t ev_scan[B2D_NLANE], ev_fork, ev_join[B2D_NLANE];
    uint64_t *d_total;                      /* running output byte offset of a call           */
    uint64_t *h_total;                      /* pinned                                         */
    uint8_t *h_stage; uint64_t h_stage_cap, h_stage_max; /* pinned staging for pageable host buffers (one-batch calls) */
    b2d_cksum_ws *ck;
    /* second slot + copy streams: lets b2d_deflate_host overlap H2D(j+1), kernels(j) and D2H(j-1) */
    uint8_t *d_inx[B2D_NSLOT - 1], *d_outx[B2D_NSLOT - 1];      /* slots 1.. (slot 0 = d_in / d_out / d_total / ck) */
    uint64_t *d_totalx[B2D_NSLOT - 1];
    b2d_cksum_ws *ckx[B2D_NSLOT - 1];
    cudaStream_t st_in, st_out;
    uint8_t *h_in_s[B2D_NSLOT], *h_out_s[B2D_NSLOT];   /* pinned staging per slot, for pageable callers (lazy) */
    cudaEvent_t ev_in[B2D_NSLOT], ev_comp[B2D_NSLOT], ev_out[B2D_NSLOT];
    uint64_t launches;
    /* compressed start offset of every device block of the last b2d_deflate_device call (b2d_block_index) */
    uint64_t *d_index; uint64_t index_cap, index_n;
    uint32_t *d_subidx;                      /* bit offset of every sub-block inside its block, B2D_NSUB entries per block (same capacity) */
    uint32_t *d_inf_sub; uint64_t inf_sub_cap;     /* inflate: the caller's sub-block bit offsets on the device */
    /* host-buffer calls keep offsets relative to each batch's output: index_base[j] is where batch j's output starts */
    std::vector<uint64_t> index_base; uint64_t index_total;
    uint64_t index_bpb, index_local;        /* blocks per batch of that call; entries this context holds in d_index */
    /* inflate scratch: 16-bit symbols, block index, error word */
    uint16_t *d_inf16; uint64_t inf16_cap;
    uint64_t *d_inf_index; uint64_t inf_index_cap;
    uint32_t *d_inf_err;
    uint8_t *d_inf_in, *d_inf_out; uint64_t inf_in_cap, inf_out_cap;     /* device copies for b2d_inflate_host */
    uint32_t *d_inf_flags; uint64_t inf_flags_cap;                       /* one ready flag per block */
    cudaStream_t st_inf; cudaEvent_t ev_inf_fork, ev_inf_join;           /* the serial pass runs beside the decode */
    cudaStream_t st_inf_in, st_inf_out;                                  /* host calls: copies of chunk c+1 / c-1 beside chunk c's kernels */
    cudaStream_t st_inf_dec[B2D_INF_DEC_STREAMS];                                          /* ... and the decodes of neighbouring chunks beside each other */
    std::vector<cudaEvent_t> ev_inf_h2d, ev_inf_done;                    /* per chunk */
    /* b2d_inflate_host_decode has queued the decode of a call; b2d_inflate_host_resolve finishes it */
    struct { bool active; uint64_t in_n, nblocks, raw_n, nchunks, dict_n; const uint32_t *d_sub; int mode; } inf_pend;
    /* optional per-stage timing of device-resident calls */
    int profiling;
    cudaEvent_t pev[PROF_BATCHES][B2D_NSTAGE + 1];
    double stage_ms[B2D_NSTAGE];
    uint64_t stage_launches[B2D_NSTAGE];
};

static void ctx_free(b2d_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaFree(c->d_in); cudaFree(c->d_out); for (int i = 0; i < B2D_NLANE; ++i) work_free(&c->wk[i]);
    for (int i = 0; i < B2D_NLANE; ++i) { if (c->ev_scan[i]) cudaEventDestroy(c->ev_scan[i]); if (c->ev_join[i]) cudaEventDestroy(c->ev_join[i]); }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    for (int i = 1; i < B2D_NLANE; ++i) if (c->stl[i]) cudaStreamDestroy(c->stl[i]);
    cudaFree(c->d_total);
    cudaFree(c->d_index); cudaFree(c->d_subidx); cudaFree(c->d_inf_sub); cudaFree(c->d_inf16); cudaFree(c->d_inf_index); cudaFree(c->d_inf_err); cudaFree(c->d_inf_in); cudaFree(c->d_inf_out); cudaFree(c->d_inf_flags);
    if (c->st_inf) cudaStreamDestroy(c->st_inf);
    if (c->st_inf_in) cudaStreamDestroy(c->st_inf_in);
    if (c->st_inf_out) cudaStreamDestroy(c->st_inf_out);
    for (int i = 0; i < B2D_INF_DEC_STREAMS; ++i) if (c->st_inf_dec[i]) cudaStreamDestroy(c->st_inf_dec[i]);
    for (cudaEvent_t e : c->ev_inf_h2d) cudaEventDestroy(e);
    for (cudaEvent_t e : c->ev_inf_done) cudaEventDestroy(e);
    if (c->ev_inf_fork) cudaEventDestroy(c->ev_inf_fork);
    if (c->ev_inf_join) cudaEventDestroy(c->ev_inf_join);
    for (int i = 0; i < B2D_NSLOT - 1; ++i) { cudaFree(c->d_inx[i]); cudaFree(c->d_outx[i]); cudaFree(c->d_totalx[i]); b2d_cksum_ws_destroy(c->ckx[i]); }
    for (int i = 0; i < B2D_NSLOT; ++i) {
        if (c->ev_in[i]) cudaEventDestroy(c->ev_in[i]);
        if (c->ev_comp[i]) cudaEventDestroy(c->ev_comp[i]);
        if (c->ev_out[i]) cudaEventDestroy(c->ev_out[i]);
    }
    for (int i = 0; i < B2D_NSLOT; ++i) { pinned_free(c->h_in_s[i]); pinned_free(c->h_out_s[i]); }
    if (c->st_in) cudaStreamDestroy(c->st_in);
    if (c->st_out) cudaStreamDestroy(c->st_out);
    if (c->h_total) cudaFreeHost(c->h_total);
    pinned_free(c->h_stage);
    b2d_cksum_ws_destroy(c->ck);
    for (int i = 0; i < PROF_BATCHES; ++i) for (int j = 0; j <= B2D_NSTAGE; ++j) if (c->pev[i][j]) cudaEventDestroy(c->pev[i][j]);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->st) cudaStreamDestroy(c->st);
    delete c;
}

/* B2D_TRACE=1: where the time of creating a context goes, on stderr (cold-start work) */
struct init_trace {
    bool on; std::chrono::steady_clock::time_point t;
    init_trace() : on(getenv("B2D_TRACE") && getenv("B2D_TRACE")[0] == '1'), t(std::chrono::steady_clock::now()) {}
    void mark(const char *what, int device) {
        if (!on) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[b2d] device %d: %-28s %7.1f ms\n", device, what, std::chrono::duration<double, std::milli>(now - t).count());
        t = now;
    }
};

extern "C" int b2d_ctx_create(int device, uint64_t max_batch, b2d_ctx **out) {
    if (!out) return fail(B2D_STREAM_ERROR, "null out pointer");
    *out = nullptr;
    init_trace tr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail_cuda(e == cudaSuccess ? cudaErrorNoDevice : e, "b2d_ctx_create: no usable CUDA device (there is no CPU fallback)");
    if (device < 0 || device >= ndev) return fail(B2D_STREAM_ERROR, "bad device index");
    if (max_batch == 0) max_batch = 256ull << 20;
    max_batch = align_up(max_batch, B2D_BLOCK);
    if (max_batch > (2048ull << 20)) return fail(B2D_STREAM_ERROR, "max_batch above 2 GiB");
    tr.mark("cudaGetDeviceCount (cuInit)", device);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail_cuda(e, "cudaSetDevice");
    if ((e = cudaFree(nullptr)) != cudaSuccess) return fail_cuda(e, "primary context");
    tr.mark("primary context", device);
    b2d_ctx *c = new b2d_ctx();
    c->device = device; c->max_batch = max_batch; c->st = nullptr; c->ev0 = c->ev1 = nullptr;
#define CC(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { int rc_ = fail_cuda(e_, #x); ctx_free(c); return rc_; } } while (0)
    CC(cudaSetDevice(device));
    CC(cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking));
    CC(cudaEventCreate(&c->ev0)); CC(cudaEventCreate(&c->ev1));
    CC(b2d_kernels_init());
    tr.mark("stream, events, kernel attrs", device);
    c->d_in_cap = B2D_WINDOW + max_batch + 1024;
    c->d_out_cap = align_up(b2d_deflate_bound(max_batch), 256);
    CC(cudaMalloc(&c->d_in, c->d_in_cap));
    CC(cudaMalloc(&c->d_out, c->d_out_cap));
    CC(work_alloc(&c->wk[0], max_batch));
    CC(cudaMalloc(&c->d_total, sizeof(uint64_t)));
    CC(cudaMallocHost(&c->h_total, (B2D_NSLOT + 1) * sizeof(uint64_t)));   /* a few words: not worth a mapping */
    c->h_stage_max = (B2D_WINDOW + max_batch) > c->d_out_cap ? (B2D_WINDOW + max_batch) : c->d_out_cap;
    if (c->h_stage_max > (64ull << 20)) c->h_stage_max = 64ull << 20;   /* staged in pieces */
    /* h_stage (pageable callers of the one-batch path) starts small and grows with the calls: ensure_stage */
    c->h_stage_cap = c->h_stage_max < (1ull << 20) ? c->h_stage_max : (1ull << 20);
    CC(pinned_alloc((void **)&c->h_stage, c->h_stage_cap));
    tr.mark("device + stage buffers", device);
    CC(b2d_cksum_ws_create(&c->ck));
    tr.mark("checksum workspace", device);
    /* the second slot is allocated on first use (ensure_second_slot) */
#undef CC
    *out = c;
    return B2D_OK;
}

extern "C" void b2d_ctx_destroy(b2d_ctx *c) { ctx_free(c); }
extern "C" uint64_t b2d_launch_count(b2d_ctx *c) { return c ? c->launches : 0; }

/* ---- memory helpers -------------------------------------------------------- */
extern "C" int b2d_device_alloc(b2d_ctx *c, uint64_t bytes, void **p) { CU(cudaSetDevice(c->device)); CU(cudaMalloc(p, bytes)); return B2D_OK; }
extern "C" int b2d_device_free(b2d_ctx *c, void *p) { CU(cudaSetDevice(c->device)); CU(cudaFree(p)); return B2D_OK; }
extern "C" int b2d_host_alloc(b2d_ctx *c, uint64_t bytes, void **p) { CU(cudaSetDevice(c->device)); CU(pinned_alloc(p, bytes)); return B2D_OK; }
extern "C" int b2d_host_free(b2d_ctx *c, void *p) { CU(cudaSetDevice(c->device)); CU(pinned_free(p)); return B2D_OK; }
extern "C" int b2d_copy_h2d(b2d_ctx *c, void *d, const void *h, uint64_t n) {
    CU(cudaSetDevice(c->device)); CU(cudaMemcpyAsync(d, h, n, cudaMemcpyHostToDevice, c->st)); CU(cudaStreamSynchronize(c->st)); return B2D_OK;
}
extern "C" int b2d_copy_d2h(b2d_ctx *c, void *h, const void *d, uint64_t n) {
    CU(cudaSetDevice(c->device)); CU(cudaMemcpyAsync(h, d, n, cudaMemcpyDeviceToHost, c->st)); CU(cudaStreamSynchronize(c->st)); return B2D_OK;
}
/* page-locked host memory without a context (visible to every device: unified addressing), and the test the writers
 * use to hand such memory to the GPU without staging it */
extern "C" int b2d_pinned_alloc(uint64_t bytes, void **p) {
    if (!p) return fail(B2D_STREAM_ERROR, "null argument");
    CU(pinned_alloc(p, bytes));
    return B2D_OK;
}
extern "C" int b2d_pinned_free(void *p) { CU(pinned_free(p)); return B2D_OK; }
extern "C" int b2d_host_is_pinned(const void *p) {
    cudaPointerAttributes at;
    const bool pinned = p && cudaPointerGetAttributes(&at, p) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    return pinned ? 1 : 0;
}
extern "C" int b2d_device_sync(b2d_ctx *c) { CU(cudaSetDevice(c->device)); CU(cudaDeviceSynchronize()); return B2D_OK; }

/* ---- checksums -------------------------------------------------------------- */
extern "C" uint32_t b2d_crc32_combine(uint32_t crc1, uint32_t crc2, int64_t len2) {
    return b2d_host_crc32_combine(crc1, crc2, (uint64_t)len2);
}
extern "C" uint32_t b2d_adler32_combine(uint32_t a1, uint32_t a2, int64_t len2) {
    if (len2 < 0) return 0xffffffffu;
    return b2d_host_adler32_combine(a1, a2, (uint64_t)len2);
}

static uint32_t crc_launches(const uint8_t *d, uint64_t n) {
    uint64_t head = (16 - ((uintptr_t)d & 15)) & 15; if (head > n) head = n;
    return (n - head) / 16384 ? 2u : 1u;
}

extern "C" int b2d_crc32_device(b2d_ctx *c, uint32_t start, const uint8_t *d, uint64_t n, uint32_t *out, float *ms) {
    if (!c || !out) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> g(c->mu);
    CU(cudaSetDevice(c->device));
    if (ms) CU(cudaEventRecord(c->ev0, c->st));
    CU(b2d_crc32_dev(c->ck, d, n, start, out, c->st));
    if (n) c->launches += crc_launches(d, n);
    if (ms) { CU(cudaEventRecord(c->ev1, c->st)); CU(cudaEventSynchronize(c->ev1)); CU(cudaEventElapsedTime(ms, c->ev0, c->ev1)); }
    return B2D_OK;
}
extern "C" int b2d_adler32_device(b2d_ctx *c, uint32_t start, const uint8_t *d, uint64_t n, uint32_t *out, float *ms) {
    if (!c || !out) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> g(c->mu);
    CU(cudaSetDevice(c->device));
    if (ms) CU(cudaEventRecord(c->ev0, c->st));
    CU(b2d_adler32_dev(c->ck, d, n, start, out, c->st));
    if (n) c->launches += 2;
    if (ms) { CU(cudaEventRecord(c->ev1, c->st)); CU(cudaEventSynchronize(c->ev1)); CU(cudaEventElapsedTime(ms, c->ev0, c->ev1)); }
    return B2D_OK;
}

extern "C" uint64_t b2d_noalloc_max_bytes(void) { return B2D_NOALLOC_MAX_BYTES; }
extern "C" int b2d_adler32_device_noalloc(int device, uint32_t start, const uint8_t *d_data, uint64_t n, uint32_t *out) {
    if (!out || (!d_data && n)) return fail(B2D_STREAM_ERROR, "null argument");
    if (n > B2D_NOALLOC_MAX_BYTES) {
        snprintf(g_err, sizeof g_err, "b2d_adler32_device_noalloc: %llu bytes exceed the maximum of %llu (use b2d_adler32_device)",
                 (unsigned long long)n, (unsigned long long)B2D_NOALLOC_MAX_BYTES);
        return B2D_STREAM_ERROR;
    }
    const char *why = "";
    cudaError_t e = b2d_adler32_noalloc(device, d_data, n, start, out, &why);
    if (e != cudaSuccess) return fail_cuda(e, why);
    return B2D_OK;
}

static int checksum_host_pipelined(b2d_ctx *c, const uint8_t *data, uint64_t n, uint32_t start, bool adler, uint32_t *out);

/* the one-batch paths stage pageable memory through h_stage: 1 MiB to begin with (small calls are latency-bound and
 * must not pay for pinning 64 MiB), grown to what a call moves, at most h_stage_max */
static int ensure_stage(b2d_ctx *c, uint64_t want) {
    if (want > c->h_stage_max) want = c->h_stage_max;
    if (want <= c->h_stage_cap) return B2D_OK;
    uint64_t cap = c->h_stage_cap;
    while (cap < want) cap *= 2;
    if (cap > c->h_stage_max) cap = c->h_stage_max;
    CU(cudaStreamSynchronize(c->st));
    pinned_free(c->h_stage); c->h_stage = nullptr; c->h_stage_cap = 0;
    CU(pinned_alloc((void **)&c->h_stage, cap));
    c->h_stage_cap = cap;
    return B2D_OK;
}

/* host bytes -> device in pieces through the pinned stage (or directly when the
 * caller's buffer is itself pinned), calling `fn` on each resident piece */
template <typename F>
static int for_each_staged(b2d_ctx *c, const uint8_t *data, uint64_t n, F fn) {
    cudaPointerAttributes at;
    bool pinned = cudaPointerGetAttributes(&at, data) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    const uint64_t piece_cap = c->max_batch;
    for (uint64_t off = 0; off < n;) {
        const uint64_t m = n - off < piece_cap ? n - off : piece_cap;
        if (pinned) { CU(cudaMemcpyAsync(c->d_in, data + off, m, cudaMemcpyHostToDevice, c->st)); }
        else {
            { int rc = ensure_stage(c, m); if (rc) return rc; }
            for (uint64_t o = 0; o < m; o += c->h_stage_cap) {
                const uint64_t k = m - o < c->h_stage_cap ? m - o : c->h_stage_cap;
                memcpy(c->h_stage, data + off + o, k);
                CU(cudaMemcpyAsync(c->d_in + o, c->h_stage, k, cudaMemcpyHostToDevice, c->st));
                CU(cudaStreamSynchronize(c->st));
            }
        }
        int rc = fn(c->d_in, m);
        if (rc) return rc;
        off += m;
    }
    return B2D_OK;
}

extern "C" int b2d_crc32(b2d_ctx *c, uint32_t start, const uint8_t *data, uint64_t n, uint32_t *out) {
    if (!c || !out || (!data && n)) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> g(c->mu);
    CU(cudaSetDevice(c->device));
    if (n > c->max_batch) return checksum_host_pipelined(c, data, n, start, false, out);
    uint32_t crc = start;
    int rc = for_each_staged(c, data, n, [&](const uint8_t *d, uint64_t m) -> int {
        CU(b2d_crc32_dev(c->ck, d, m, crc, &crc, c->st));
        c->launches += crc_launches(d, m);
        return B2D_OK;
    });
    if (rc) return rc;
    *out = crc;
    return B2D_OK;
}
extern "C" int b2d_adler32(b2d_ctx *c, uint32_t start, const uint8_t *data, uint64_t n, uint32_t *out) {
    if (!c || !out || (!data && n)) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> g(c->mu);
    CU(cudaSetDevice(c->device));
    if (n > c->max_batch) return checksum_host_pipelined(c, data, n, start, true, out);
    uint32_t a = start;
    int rc = for_each_staged(c, data, n, [&](const uint8_t *d, uint64_t m) -> int {
        CU(b2d_adler32_dev(c->ck, d, m, a, &a, c->st));
        c->launches += 2;
        return B2D_OK;
    });
    if (rc) return rc;
    *out = a;
    return B2D_OK;
}

/* ---- deflate ---------------------------------------------------------------- */
static void fill_job(b2d_ctx *c, b2d_job *J, const uint8_t *buf, uint64_t D, uint64_t n, uint64_t readable, int level, uint8_t *out,
                     uint32_t max_dist = B2D_WINDOW, int lane = 0) {
    const b2d_work *w = &c->wk[lane];
    J->max_dist = max_dist;
    J->buf = buf; J->D = D; J->n = n; J->buf_readable = readable; J->level = level;
    J->prevd = w->prevd; J->match = w->match; J->tok = w->tok; J->info = w->info; J->hist = w->hist;
    J->plan = w->plan; J->block_off = w->block_off; J->total = c->d_total; J->out = out;
    J->scan_wait = nullptr; J->scan_done = nullptr;
}

/* second set of intermediates + its stream: only multi-batch calls pay for them */
static int ensure_lanes(b2d_ctx *c, int want) {
    if (!c->ev_fork) {
        c->stl[0] = c->st;
        for (int i = 0; i < B2D_NLANE; ++i) {
            CU(cudaEventCreateWithFlags(&c->ev_scan[i], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&c->ev_join[i], cudaEventDisableTiming));
        }
        CU(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    }
    for (int i = 1; i < want && i < B2D_NLANE; ++i)
        if (!c->stl[i]) { CU(work_alloc(&c->wk[i], c->max_batch)); CU(cudaStreamCreateWithFlags(&c->stl[i], cudaStreamNonBlocking)); }
    return B2D_OK;
}
static int ensure_second_lane(b2d_ctx *c) { return ensure_lanes(c, B2D_NLANE); }
/* lanes a device-resident call uses: two batches of 128 MiB or more fill the device (k_chains: one warp per 128 KiB
 * block, 1776 warp slots), so more lanes would only hold memory (2.7 GiB each at 256 MiB); smaller batches use all four */
static int device_lanes(const b2d_ctx *c) {
    static const int forced = [] { const char *e = getenv("B2D_DEVICE_LANES"); const int v = e ? atoi(e) : 0; return v >= 1 && v <= B2D_NLANE ? v : 0; }();
    if (forced) return forced;          /* measurements only */
    return c->max_batch >= (128ull << 20) ? 2 : B2D_NLANE;
}

/* kernels one batch launches at this level / strategy */
static uint32_t pipeline_launches(int level) {
    const b2d_policy pol = b2d_level_policy(level);
    return 4u + (pol.search == B2D_SEARCH_CHAINS ? 1u : 0u) + (pol.search != B2D_SEARCH_NONE ? 1u : 0u);
}

/* fold the event pairs of `nb` profiled batches into the per-stage totals */
static int prof_drain(b2d_ctx *c, int nb, int level) {
    const b2d_policy pol = b2d_level_policy(level);
    CU(cudaStreamSynchronize(c->st));
    for (int i = 1; i < B2D_NLANE; ++i) if (c->stl[i]) CU(cudaStreamSynchronize(c->stl[i]));
    for (int i = 0; i < nb; ++i)
        for (int j = 0; j < B2D_NSTAGE; ++j) {
            float t = 0;
            CU(cudaEventElapsedTime(&t, c->pev[i][j], c->pev[i][j + 1]));
            c->stage_ms[j] += t;
            if (j >= 2 || (j == 0 && pol.search == B2D_SEARCH_CHAINS) || (j == 1 && pol.search != B2D_SEARCH_NONE)) c->stage_launches[j] += 1;
        }
    return B2D_OK;
}

extern "C" int b2d_profile(b2d_ctx *c, int enable) {
    if (!c) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> g(c->mu);
    CU(cudaSetDevice(c->device));
    if (enable && !c->pev[0][0])
        for (int i = 0; i < PROF_BATCHES; ++i) for (int j = 0; j <= B2D_NSTAGE; ++j) CU(cudaEventCreate(&c->pev[i][j]));
    c->profiling = enable == 2 ? 2 : (enable ? 1 : 0);
    return B2D_OK;
}
extern "C" int b2d_profile_read(b2d_ctx *c, double *stage_ms, uint64_t *stage_launches, int reset) {
    if (!c) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> g(c->mu);
    for (int j = 0; j < B2D_NSTAGE; ++j) {
        if (stage_ms) stage_ms[j] = c->stage_ms[j];
        if (stage_launches) stage_launches[j] = c->stage_launches[j];
        if (reset) { c->stage_ms[j] = 0; c->stage_launches[j] = 0; }
    }
    return B2D_OK;
}

/* room for the block index of a call over n input bytes */
static int ensure_index(b2d_ctx *c, uint64_t n) {
    const uint64_t need = (n + B2D_BLOCK - 1) / B2D_BLOCK + 2;
    if (need > c->index_cap) {
        cudaFree(c->d_index); c->d_index = nullptr; c->index_cap = 0;
        cudaFree(c->d_subidx); c->d_subidx = nullptr;
        CU(cudaMalloc(&c->d_index, need * sizeof(uint64_t)));
        CU(cudaMalloc(&c->d_subidx, need * B2D_NSUB * sizeof(uint32_t)));
        c->index_cap = need;
    }
    CU(cudaMemsetAsync(c->d_subidx, 0, need * B2D_NSUB * sizeof(uint32_t), c->st));     /* sub-blocks a short last block does not have */
    CU(cudaStreamSynchronize(c->st));                                                   /* (the lanes write it from other streams) */
    c->index_n = 0;
    c->index_base.clear();
    c->index_total = 0;
    return B2D_OK;
}

/* keep the bit offset of every sub-block of a batch (b2d_subinfo.bit_off, written by k_scan) next to the block index */
static cudaError_t save_sub_bits(b2d_ctx *c, uint64_t first_block, const b2d_job &J, uint64_t m, cudaStream_t st) {
    const uint64_t nsub = (m + B2D_SUB - 1) / B2D_SUB;
    return cudaMemcpy2DAsync(c->d_subidx + first_block * B2D_NSUB, sizeof(uint32_t),
                             reinterpret_cast<const uint8_t *>(J.info) + offsetof(b2d_subinfo, bit_off), sizeof(b2d_subinfo),
                             sizeof(uint32_t), nsub, cudaMemcpyDeviceToDevice, st);
}

extern "C" int b2d_deflate_device(b2d_ctx *c, int level, const uint8_t *d_buf, uint64_t dict_n, uint64_t n, uint8_t *d_out,
                                  uint64_t out_cap, uint64_t *out_n, uint32_t *crc, float *ms) {
    if (!c || !out_n || (!d_buf && n) || !d_out) return fail(B2D_STREAM_ERROR, "null argument");
    if (!b2d_level_valid(level)) return fail(B2D_STREAM_ERROR, "Bad compression level");
    if (dict_n > B2D_WINDOW) return fail(B2D_STREAM_ERROR, "device dictionary longer than 32768 bytes");
    if (((uintptr_t)d_buf & 15) || ((uintptr_t)d_out & 3)) return fail(B2D_STREAM_ERROR, "misaligned device pointer");
    if (out_cap < b2d_deflate_bound(n)) return fail(B2D_BUF_ERROR, "out_cap below b2d_deflate_bound(n)");
    std::lock_guard<std::recursive_mutex> g(c->mu);
    CU(cudaSetDevice(c->device));
    if (ms) CU(cudaEventRecord(c->ev0, c->st));
    CU(cudaMemsetAsync(d_out, 0, align_up(b2d_deflate_bound(n), 4) <= out_cap ? align_up(b2d_deflate_bound(n), 4) : out_cap, c->st));
    CU(cudaMemsetAsync(c->d_total, 0, sizeof(uint64_t), c->st));
    const uint64_t readable = dict_n + n + B2D_DEVICE_PAD;
    int nprof = 0;
    { int rc = ensure_index(c, n); if (rc) return rc; }
    /* several batches alternate between two lanes (streams + intermediates); b2d_profile(ctx, 2) keeps
     * one lane so that the stage times are those of kernels running alone */
    const bool two = n > c->max_batch && c->profiling != 2;
    const int NL = device_lanes(c);
    if (two) {
        int rc = ensure_lanes(c, NL); if (rc) return rc;
        CU(cudaEventRecord(c->ev_fork, c->st));
        for (int i = 1; i < NL; ++i) CU(cudaStreamWaitEvent(c->stl[i], c->ev_fork, 0));
    }
    uint64_t jb = 0;
    for (uint64_t off = 0; off < n; off += c->max_batch, ++jb) {
        const uint64_t m = n - off < c->max_batch ? n - off : c->max_batch;
        const int lane = two ? (int)(jb % NL) : 0;
        /* re-base so the work buffers are indexed from (at most) 32 KiB + 15 bytes before the batch */
        const uint64_t abs0 = dict_n + off;
        const uint64_t base = abs0 > B2D_WINDOW ? (abs0 - B2D_WINDOW) & ~(uint64_t)15 : 0;
        b2d_job J; fill_job(c, &J, d_buf + base, abs0 - base, m, readable - base, level, d_out, B2D_WINDOW, lane);
        if (two) { J.scan_wait = jb ? c->ev_scan[(lane + NL - 1) % NL] : nullptr; J.scan_done = c->ev_scan[lane]; }
        cudaEvent_t *pe = nullptr;
        if (c->profiling) {
            if (nprof == PROF_BATCHES) { int rc = prof_drain(c, nprof, level); if (rc) return rc; nprof = 0; }
            pe = c->pev[nprof++];
        }
        CU(b2d_launch_pipeline(&J, lane ? c->stl[lane] : c->st, pe));
        c->launches += pipeline_launches(level);
        /* k_scan left the absolute output offset of each block of this batch (and the running total) in the lane's
         * block_off: keep them, the inflate side needs them */
        const uint64_t nb = (m + B2D_BLOCK - 1) / B2D_BLOCK;
        CU(cudaMemcpyAsync(c->d_index + off / B2D_BLOCK, J.block_off, (nb + 1) * sizeof(uint64_t), cudaMemcpyDeviceToDevice,
                           lane ? c->stl[lane] : c->st));
        CU(save_sub_bits(c, off / B2D_BLOCK, J, m, lane ? c->stl[lane] : c->st));
    }
    if (two) for (int i = 1; i < NL; ++i) { CU(cudaEventRecord(c->ev_join[i], c->stl[i])); CU(cudaStreamWaitEvent(c->st, c->ev_join[i], 0)); }
    if (crc && n) { CU(b2d_crc32_dev_launch(c->ck, d_buf + dict_n, n, c->st)); c->launches += crc_launches(d_buf + dict_n, n); }
    CU(cudaMemcpyAsync(c->h_total, c->d_total, sizeof(uint64_t), cudaMemcpyDeviceToHost, c->st));
    if (ms) CU(cudaEventRecord(c->ev1, c->st));
    CU(cudaStreamSynchronize(c->st));
    if (ms) CU(cudaEventElapsedTime(ms, c->ev0, c->ev1));
    if (nprof) { int rc = prof_drain(c, nprof, level); if (rc) return rc; }
    if (crc) *crc = n ? b2d_crc32_dev_finish(c->ck, n, 0) : 0;
    *out_n = *c->h_total;
    c->index_n = n ? (n + B2D_BLOCK - 1) / B2D_BLOCK + 1 : 0;
    if (n == 0) {   /* the empty job is just the sync marker: 00 00 00 ff ff */
        static const uint8_t marker[5] = {0, 0, 0, 0xff, 0xff};
        CU(cudaMemcpyAsync(d_out, marker, 5, cudaMemcpyHostToDevice, c->st));
        CU(cudaStreamSynchronize(c->st));
        *out_n = 5;
    }
    return B2D_OK;
}

/* second input/output slot, copy streams and events: only callers that stream several batches from
 * pinned memory pay for them */
static int ensure_second_slot(b2d_ctx *c) {
    if (c->st_in) return B2D_OK;
    for (int i = 0; i < B2D_NSLOT - 1; ++i) {
        CU(cudaMalloc(&c->d_inx[i], c->d_in_cap));
        CU(cudaMalloc(&c->d_outx[i], c->d_out_cap));
        CU(cudaMalloc(&c->d_totalx[i], sizeof(uint64_t)));
        CU(b2d_cksum_ws_create(&c->ckx[i]));
    }
    CU(cudaStreamCreateWithFlags(&c->st_in, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c->st_out, cudaStreamNonBlocking));
    for (int i = 0; i < B2D_NSLOT; ++i) {
        CU(cudaEventCreateWithFlags(&c->ev_in[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_comp[i], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_out[i], cudaEventDisableTiming));
    }
    return B2D_OK;
}

/* host copies run on a persistent pool with non-temporal stores (b2d_hostcopy.cpp): one thread moves ~14 GB/s */
static void par_memcpy_n(void *dst, const void *src, uint64_t n, unsigned nt) { b2d_par_memcpy(dst, src, n, nt); }
/* threads that stage one batch of a pageable caller: 16 (B2D_STAGE_THREADS) — measured with zlib_ng.compress of 1 GiB:
 * 14.2 / 19.6 / 24.2 GB/s at 4 / 8 / 16 — shared between the device threads of a multi-device call */
static thread_local unsigned g_stage_share = 1;
static unsigned stage_threads() {
    static const unsigned n = [] { const char *e = getenv("B2D_STAGE_THREADS"); int v = e ? atoi(e) : 16; return (unsigned)(v < 1 ? 1 : v > 32 ? 32 : v); }();
    const unsigned t = n / g_stage_share;
    return t < 4 ? 4 : t;
}
static void par_memcpy(void *dst, const void *src, uint64_t n) { par_memcpy_n(dst, src, n, stage_threads()); }

extern "C" int b2d_host_copy(void *dst, const void *src, uint64_t bytes, int threads) {
    if ((!dst || !src) && bytes) return B2D_STREAM_ERROR;
    par_memcpy_n(dst, src, bytes, threads < 1 ? 1u : (unsigned)threads);
    return B2D_OK;
}

static int ensure_staging(b2d_ctx *c, bool in, bool out) {
    for (int i = 0; i < B2D_NSLOT; ++i) {
        if (in && !c->h_in_s[i]) CU(pinned_alloc((void **)&c->h_in_s[i], c->d_in_cap));
        if (out && !c->h_out_s[i]) CU(pinned_alloc((void **)&c->h_out_s[i], c->d_out_cap));
    }
    return B2D_OK;
}

/* What the batches of one call share: where each batch's output starts.  With one context the batches finish in
 * order anyway; with several (b2d_deflate_host_multi: batch j runs on device j % k) a context learns the offset of
 * its batch j from the context that ran batch j - 1, as soon as that batch's kernels are done — its size is known
 * then, its device-to-host copy need not be. */
struct b2d_ledger {
    std::mutex mu;
    std::condition_variable cv;
    std::vector<uint64_t> off;            /* off[j] = output bytes before batch j; valid for j <= known */
    std::vector<uint32_t> crc, adler;     /* per batch, from the neutral start values (0 / 1) */
    uint64_t known = 0;
    int rc = B2D_OK;                      /* first failure of any context; the others stop waiting */
    char err[512] = {0};
    explicit b2d_ledger(uint64_t nb) : off(nb + 1, 0), crc(nb, 0), adler(nb, 1) {}
    void fail_with(int code, const char *msg) {
        std::lock_guard<std::mutex> g(mu);
        if (rc == B2D_OK) { rc = code; snprintf(err, sizeof err, "%s", msg); }
        cv.notify_all();
    }
};

/* Host input and output, more than one batch: three-stage software pipeline over the batches
 * j = first, first + stride, ... of the call (one context takes them all: first 0, stride 1).
 *   st_in : H2D of a batch into slot i % B2D_NSLOT (waits until the kernels of the slot's previous batch released it)
 *   lanes : kernels + CRC of the batch on lane i % B2D_NLANE (waits for its H2D and for the D2H of the slot's previous batch);
 *           consecutive batches overlap, which fills the tail waves of each other's kernels
 *   st_out: D2H of an earlier batch (issued once the host knows its size and where it goes)
 * The host only blocks on "kernels of the batch two back are done", by which time the next two are already queued.
 * Pageable input / output goes through pinned staging slots (copied by a few host threads), so those callers
 * get the same overlap. */
static int deflate_host_pipelined(b2d_ctx *c, int level, uint32_t max_dist, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                  uint64_t dict_n, uint8_t *out, uint64_t out_cap, uint64_t reserve, bool want_crc, bool want_adler,
                                  bool in_pinned, bool out_pinned, uint64_t batch, uint64_t first, uint64_t stride, b2d_ledger *G) {
    int rc = ensure_second_slot(c);
    if (rc) return rc;
    if ((rc = ensure_second_lane(c))) return rc;
    if ((rc = ensure_staging(c, !in_pinned, !out_pinned))) return rc;
    cudaStream_t *cs = c->stl;                  /* local batch i computes on lane i % B2D_NLANE */
    uint8_t *d_in[B2D_NSLOT] = {c->d_in}, *d_out[B2D_NSLOT] = {c->d_out};
    uint64_t *d_total[B2D_NSLOT] = {c->d_total};
    b2d_cksum_ws *ck[B2D_NSLOT] = {c->ck};
    for (int i = 1; i < B2D_NSLOT; ++i) { d_in[i] = c->d_inx[i - 1]; d_out[i] = c->d_outx[i - 1]; d_total[i] = c->d_totalx[i - 1]; ck[i] = c->ckx[i - 1]; }
    const uint64_t nb = (n + batch - 1) / batch;                                  /* batches of the whole call */
    const uint64_t nloc = first < nb ? (nb - first + stride - 1) / stride : 0;    /* ... of this context */
    const uint64_t bpb = batch / B2D_BLOCK;
    if ((rc = ensure_index(c, nloc * batch))) return rc;
    c->index_base.assign(nloc, 0);
    c->index_bpb = bpb;
    uint64_t pend_off = 0, pend_n = 0; int pend_s = -1;
    static const bool trace = getenv("B2D_TRACE") && getenv("B2D_TRACE")[0] == '1';
    double t_comp = 0, t_ledger = 0, t_issue = 0;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_begin = now();
    auto batch_len = [&](uint64_t j) { return (j + 1 == nb) ? n - j * batch : batch; };
    auto drain = [&]() -> int {          /* staged output: copy the batch whose D2H was issued last */
        if (pend_s < 0) return B2D_OK;
        CU(cudaEventSynchronize(c->ev_out[pend_s]));
        par_memcpy(out + pend_off, c->h_out_s[pend_s], pend_n);
        pend_s = -1;
        return B2D_OK;
    };
    auto finish = [&](uint64_t i) -> int {          /* local batch i: wait for its kernels, note its checksums, start its D2H */
        const int s = (int)(i % B2D_NSLOT);
        const uint64_t j = first + i * stride, m = batch_len(j);
        const double t0 = now();
        CU(cudaEventSynchronize(c->ev_comp[s]));
        const double t1 = now();
        t_comp += t1 - t0;
        if (want_crc) G->crc[j] = b2d_crc32_dev_finish(ck[s], m, 0);
        if (want_adler) G->adler[j] = b2d_adler32_dev_finish(ck[s], m, 1);
        const uint64_t got = c->h_total[s];
        uint64_t at;
        {
            std::unique_lock<std::mutex> lk(G->mu);
            G->cv.wait(lk, [&] { return G->known >= j || G->rc != B2D_OK; });
            if (G->rc != B2D_OK) return G->rc;            /* another context failed: its message is in the ledger */
            at = G->off[j];
            if (at + got + reserve > out_cap) {
                lk.unlock();
                G->fail_with(B2D_BUF_ERROR, "Compressed output exceeds buffer size");
                return fail(B2D_BUF_ERROR, "Compressed output exceeds buffer size");
            }
            G->off[j + 1] = at + got;
            G->known = j + 1;
        }
        G->cv.notify_all();
        t_ledger += now() - t1;
        c->index_base[i] = at;
        if (out_pinned) {
            CU(cudaMemcpyAsync(out + at, d_out[s], got, cudaMemcpyDeviceToHost, c->st_out));
            CU(cudaEventRecord(c->ev_out[s], c->st_out));
        } else {
            int rc2 = drain(); if (rc2) return rc2;     /* the previous batch's bytes, while this one's kernels ran */
            CU(cudaMemcpyAsync(c->h_out_s[s], d_out[s], got, cudaMemcpyDeviceToHost, c->st_out));
            CU(cudaEventRecord(c->ev_out[s], c->st_out));
            pend_s = s; pend_off = at; pend_n = got;
        }
        return B2D_OK;
    };
    for (uint64_t i = 0; i < nloc; ++i) {
        const double ti = now();
        const int s = (int)(i % B2D_NSLOT);
        const int ln = (int)(i % B2D_NLANE);  /* compute lane */
        const uint64_t j = first + i * stride;
        const uint64_t off = j * batch;
        const uint64_t m = batch_len(j);
        uint64_t D;
        if (i >= B2D_NSLOT) CU(cudaStreamWaitEvent(c->st_in, c->ev_comp[s], 0));      /* slot's input no longer read */
        if (j == 0) {
            D = dict_n;
            if (D) { CU(cudaMemcpyAsync(d_in[s], dict, D, cudaMemcpyHostToDevice, c->st_in)); }
            if (in_pinned) { CU(cudaMemcpyAsync(d_in[s] + D, data, m, cudaMemcpyHostToDevice, c->st_in)); }
            else {
                par_memcpy(c->h_in_s[s], data, m);
                CU(cudaMemcpyAsync(d_in[s] + D, c->h_in_s[s], m, cudaMemcpyHostToDevice, c->st_in));
            }
        } else {
            D = B2D_WINDOW;
            if (in_pinned) { CU(cudaMemcpyAsync(d_in[s], data + off - B2D_WINDOW, B2D_WINDOW + m, cudaMemcpyHostToDevice, c->st_in)); }
            else {
                /* the slot's previous H2D (three batches back) finished before that batch's kernels, which finish() waited for */
                par_memcpy(c->h_in_s[s], data + off - B2D_WINDOW, B2D_WINDOW + m);
                CU(cudaMemcpyAsync(d_in[s], c->h_in_s[s], B2D_WINDOW + m, cudaMemcpyHostToDevice, c->st_in));
            }
        }
        CU(cudaEventRecord(c->ev_in[s], c->st_in));
        CU(cudaStreamWaitEvent(cs[ln], c->ev_in[s], 0));
        if (i >= B2D_NSLOT) CU(cudaStreamWaitEvent(cs[ln], c->ev_out[s], 0));         /* slot's output already copied out */
        CU(cudaMemsetAsync(d_out[s], 0, align_up(b2d_deflate_bound(m), 4), cs[ln]));
        CU(cudaMemsetAsync(d_total[s], 0, sizeof(uint64_t), cs[ln]));
        b2d_job J; fill_job(c, &J, d_in[s], D, m, c->d_in_cap, level, d_out[s], max_dist, ln);
        J.total = d_total[s];
        CU(b2d_launch_pipeline(&J, cs[ln], nullptr));
        c->launches += pipeline_launches(level);
        CU(cudaMemcpyAsync(&c->h_total[s], d_total[s], sizeof(uint64_t), cudaMemcpyDeviceToHost, cs[ln]));
        CU(cudaMemcpyAsync(c->d_index + i * bpb, J.block_off, ((m + B2D_BLOCK - 1) / B2D_BLOCK) * sizeof(uint64_t),
                           cudaMemcpyDeviceToDevice, cs[ln]));
        CU(save_sub_bits(c, i * bpb, J, m, cs[ln]));
        if (want_crc) { CU(b2d_crc32_dev_launch(ck[s], d_in[s] + D, m, cs[ln])); c->launches += crc_launches(d_in[s] + D, m); }
        if (want_adler) { CU(b2d_adler32_dev_launch(ck[s], d_in[s] + D, m, cs[ln])); c->launches += 2; }
        CU(cudaEventRecord(c->ev_comp[s], cs[ln]));
        t_issue += now() - ti;
        /* the host trails by one batch per lane: the lanes stay busy while it waits for the oldest batch */
        if (i >= B2D_NLANE && (rc = finish(i - B2D_NLANE))) { cudaDeviceSynchronize(); return rc; }
    }
    for (uint64_t k = nloc > B2D_NLANE ? nloc - B2D_NLANE : 0; k < nloc; ++k)
        if ((rc = finish(k))) { cudaDeviceSynchronize(); return rc; }
    if ((rc = drain())) { cudaDeviceSynchronize(); return rc; }
    const double t_tail0 = now();
    CU(cudaStreamSynchronize(c->st_out));
    if (trace)
        fprintf(stderr, "[b2d] device %d: %llu batches of %llu MiB: total %.1f ms = issue %.1f + wait kernels %.1f + ledger/D2H issue %.1f + last D2H %.1f\n",
                c->device, (unsigned long long)nloc, (unsigned long long)(batch >> 20), now() - t_begin, t_issue, t_comp, t_ledger, now() - t_tail0);
    /* the index of this context's blocks (b2d_block_index, collect_index): relative offsets stay on the device */
    c->index_n = 0;
    c->index_local = nloc ? (nloc - 1) * bpb + (batch_len(first + (nloc - 1) * stride) + B2D_BLOCK - 1) / B2D_BLOCK : 0;
    return B2D_OK;
}

/* fold the per-batch checksums of a finished call in stream order */
static void ledger_fold(const b2d_ledger &G, uint64_t n, uint64_t batch, uint32_t *crc_acc, uint32_t *adler_acc) {
    const uint64_t nb = (n + batch - 1) / batch;
    for (uint64_t j = 0; j < nb; ++j) {
        const uint64_t m = (j + 1 == nb) ? n - j * batch : batch;
        if (crc_acc) *crc_acc = b2d_host_crc32_combine(*crc_acc, G.crc[j], m);
        if (adler_acc) *adler_acc = b2d_host_adler32_combine(*adler_acc, G.adler[j], m);
    }
}

/* crc32 / adler32 of a host buffer larger than one batch: copy-in (through pinned staging when the caller's
 * memory is pageable), H2D and the checksum kernels of neighbouring pieces overlap; the host folds the
 * pieces in order (the kernels compute each piece from a neutral start, see b2d_checksum.cu) */
static int checksum_host_pipelined(b2d_ctx *c, const uint8_t *data, uint64_t n, uint32_t start, bool adler, uint32_t *out) {
    cudaPointerAttributes at;
    const bool pinned = cudaPointerGetAttributes(&at, data) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    int rc = ensure_second_slot(c);
    if (rc) return rc;
    if ((rc = ensure_staging(c, !pinned, false))) return rc;
    uint8_t *d_in[B2D_NSLOT] = {c->d_in};
    b2d_cksum_ws *ck[B2D_NSLOT] = {c->ck};
    for (int i = 1; i < B2D_NSLOT; ++i) { d_in[i] = c->d_inx[i - 1]; ck[i] = c->ckx[i - 1]; }
    const uint64_t nb = (n + c->max_batch - 1) / c->max_batch;
    uint32_t acc = start;
    auto piece = [&](uint64_t j) { return (j + 1 == nb) ? n - j * c->max_batch : c->max_batch; };
    auto finish = [&](uint64_t j) -> int {
        const int s = (int)(j % B2D_NSLOT);
        CU(cudaEventSynchronize(c->ev_comp[s]));
        acc = adler ? b2d_adler32_dev_finish(ck[s], piece(j), acc) : b2d_crc32_dev_finish(ck[s], piece(j), acc);
        return B2D_OK;
    };
    for (uint64_t j = 0; j < nb; ++j) {
        const int s = (int)(j % B2D_NSLOT);
        const uint64_t off = j * c->max_batch, m = piece(j);
        if (j >= B2D_NSLOT) CU(cudaStreamWaitEvent(c->st_in, c->ev_comp[s], 0));
        if (pinned) { CU(cudaMemcpyAsync(d_in[s], data + off, m, cudaMemcpyHostToDevice, c->st_in)); }
        else {
            par_memcpy(c->h_in_s[s], data + off, m);
            CU(cudaMemcpyAsync(d_in[s], c->h_in_s[s], m, cudaMemcpyHostToDevice, c->st_in));
        }
        CU(cudaEventRecord(c->ev_in[s], c->st_in));
        CU(cudaStreamWaitEvent(c->st, c->ev_in[s], 0));
        if (adler) { CU(b2d_adler32_dev_launch(ck[s], d_in[s], m, c->st)); c->launches += 2; }
        else { CU(b2d_crc32_dev_launch(ck[s], d_in[s], m, c->st)); c->launches += crc_launches(d_in[s], m); }
        CU(cudaEventRecord(c->ev_comp[s], c->st));
        if (j >= 2 && (rc = finish(j - 2))) { cudaDeviceSynchronize(); return rc; }
    }
    if (nb >= 2 && (rc = finish(nb - 2))) { cudaDeviceSynchronize(); return rc; }
    if ((rc = finish(nb - 1))) { cudaDeviceSynchronize(); return rc; }
    *out = acc;
    return B2D_OK;
}

extern "C" int b2d_deflate_host(b2d_ctx *c, int level, int window_bits, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                uint64_t dict_n, int final, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc, uint32_t *adler) {
    if (!c || !out_n || (!data && n) || (!out && out_cap) || (!dict && dict_n)) return fail(B2D_STREAM_ERROR, "null argument");
    if (!b2d_level_valid(level)) return fail(B2D_STREAM_ERROR, "Bad compression level");
    if (window_bits < 9 || window_bits > 15) return fail(B2D_STREAM_ERROR, "window_bits outside 9..15");
    const uint32_t max_dist = 1u << window_bits;
    std::lock_guard<std::recursive_mutex> g(c->mu);
    CU(cudaSetDevice(c->device));
    if (dict_n > max_dist) { dict += dict_n - max_dist; dict_n = max_dist; }
    if (dict_n > B2D_WINDOW) { dict += dict_n - B2D_WINDOW; dict_n = B2D_WINDOW; }
    uint64_t produced = 0;
    uint32_t crc_acc = crc ? *crc : 0, adler_acc = adler ? *adler : 1;
    cudaPointerAttributes at;
    const bool pinned = n && cudaPointerGetAttributes(&at, data) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    {
        cudaPointerAttributes ot;
        const bool opinned = out && cudaPointerGetAttributes(&ot, out) == cudaSuccess && ot.type == cudaMemoryTypeHost;
        cudaGetLastError();
        if (n > c->max_batch) {
            b2d_ledger G((n + c->max_batch - 1) / c->max_batch);
            int rc = deflate_host_pipelined(c, level, max_dist, data, n, dict, dict_n, out, out_cap, final ? 2 : 0, crc != nullptr,
                                            adler != nullptr, pinned, opinned, c->max_batch, 0, 1, &G);
            if (rc) return rc;
            ledger_fold(G, n, c->max_batch, crc ? &crc_acc : nullptr, adler ? &adler_acc : nullptr);
            produced = G.off[G.off.size() - 1];
            c->index_total = produced;
            c->index_n = (n + B2D_BLOCK - 1) / B2D_BLOCK + 1;
            if (final) { out[produced++] = 0x03; out[produced++] = 0x00; }
            *out_n = produced;
            if (crc) *crc = crc_acc;
            if (adler) *adler = adler_acc;
            return B2D_OK;
        }
    }
    auto h2d = [&](uint8_t *dst, const uint8_t *src, uint64_t bytes, bool src_pinned) -> int {
        if (src_pinned) { CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->st)); return B2D_OK; }
        { int rc = ensure_stage(c, bytes); if (rc) return rc; }
        for (uint64_t o = 0; o < bytes; o += c->h_stage_cap) {
            const uint64_t k = bytes - o < c->h_stage_cap ? bytes - o : c->h_stage_cap;
            CU(cudaStreamSynchronize(c->st));            /* the stage may still be in flight */
            memcpy(c->h_stage, src + o, k);
            CU(cudaMemcpyAsync(dst + o, c->h_stage, k, cudaMemcpyHostToDevice, c->st));
        }
        CU(cudaStreamSynchronize(c->st));
        return B2D_OK;
    };
    { int rc = ensure_index(c, n); if (rc) return rc; }
    for (uint64_t off = 0; off < n; off += c->max_batch) {      /* one trip: longer inputs took the pipelined path */
        const uint64_t m = n - off < c->max_batch ? n - off : c->max_batch;
        uint64_t D;
        int rc;
        if (off == 0) {
            D = dict_n;
            if (D && (rc = h2d(c->d_in, dict, D, false))) return rc;
            if ((rc = h2d(c->d_in + D, data, m, pinned))) return rc;
        } else {
            D = B2D_WINDOW;
            if ((rc = h2d(c->d_in, data + off - B2D_WINDOW, B2D_WINDOW + m, pinned))) return rc;
        }
        const uint64_t bound = align_up(b2d_deflate_bound(m), 4);
        CU(cudaMemsetAsync(c->d_out, 0, bound, c->st));
        CU(cudaMemsetAsync(c->d_total, 0, sizeof(uint64_t), c->st));
        b2d_job J; fill_job(c, &J, c->d_in, D, m, c->d_in_cap, level, c->d_out, max_dist);
        CU(b2d_launch_pipeline(&J, c->st, nullptr));
        c->launches += pipeline_launches(level);
        CU(cudaMemcpyAsync(c->h_total, c->d_total, sizeof(uint64_t), cudaMemcpyDeviceToHost, c->st));
        CU(cudaMemcpyAsync(c->d_index, J.block_off, ((m + B2D_BLOCK - 1) / B2D_BLOCK + 1) * sizeof(uint64_t), cudaMemcpyDeviceToDevice, c->st));
        CU(save_sub_bits(c, 0, J, m, c->st));
        c->index_n = (m + B2D_BLOCK - 1) / B2D_BLOCK + 1;
        if (crc) { CU(b2d_crc32_dev_launch(c->ck, c->d_in + D, m, c->st)); c->launches += crc_launches(c->d_in + D, m); }
        CU(cudaStreamSynchronize(c->st));
        if (crc) crc_acc = b2d_crc32_dev_finish(c->ck, m, crc_acc);
        if (adler) { CU(b2d_adler32_dev(c->ck, c->d_in + D, m, adler_acc, &adler_acc, c->st)); c->launches += 2; }
        const uint64_t got = *c->h_total;
        if (produced + got + (final ? 2 : 0) > out_cap) return fail(B2D_BUF_ERROR, "Compressed output exceeds buffer size");
        /* device -> caller: direct when pinned, else through the stage */
        cudaPointerAttributes ot;
        const bool opinned = cudaPointerGetAttributes(&ot, out) == cudaSuccess && ot.type == cudaMemoryTypeHost;
        cudaGetLastError();
        if (opinned) { CU(cudaMemcpyAsync(out + produced, c->d_out, got, cudaMemcpyDeviceToHost, c->st)); CU(cudaStreamSynchronize(c->st)); }
        else if ((rc = ensure_stage(c, got))) return rc;
        else for (uint64_t o = 0; o < got; o += c->h_stage_cap) {
            const uint64_t k = got - o < c->h_stage_cap ? got - o : c->h_stage_cap;
            CU(cudaMemcpyAsync(c->h_stage, c->d_out + o, k, cudaMemcpyDeviceToHost, c->st));
            CU(cudaStreamSynchronize(c->st));
            memcpy(out + produced + o, c->h_stage, k);
        }
        produced += got;
    }
    if (n == 0) {
        if (5 + (final ? 2u : 0u) > out_cap) return fail(B2D_BUF_ERROR, "Compressed output exceeds buffer size");
        static const uint8_t marker[5] = {0, 0, 0, 0xff, 0xff};
        memcpy(out, marker, 5); produced = 5;
    }
    if (final) { out[produced++] = 0x03; out[produced++] = 0x00; }
    *out_n = produced;
    if (crc) *crc = crc_acc;
    if (adler) *adler = adler_acc;
    return B2D_OK;
}

extern "C" int b2d_compress_and_crc(b2d_ctx *c, int level, const uint8_t *data, uint64_t n, const uint8_t *dict, uint64_t dict_n,
                                    uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc) {
    if (!crc) return fail(B2D_STREAM_ERROR, "null argument");
    *crc = 0;
    return b2d_deflate_host(c, level, 15, data, n, dict, dict_n, 0, out, out_cap, out_n, crc, nullptr);
}


extern "C" int b2d_compress_and_crc_indexed(b2d_ctx *c, int level, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                            uint64_t dict_n, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                                            uint64_t *offsets, uint64_t cap, uint64_t *count) {
    if (!c || !count) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> g(c->mu);       /* no other call on this context between the two parts */
    int rc = b2d_compress_and_crc(c, level, data, n, dict, dict_n, out, out_cap, out_n, crc);
    if (rc) return rc;
    return b2d_block_index(c, offsets, cap, count);
}

/* ------------------------------------------------------------------------- */
/* one stream over several devices: what the reference does with N worker threads on one stream —
 * round-robin dispatch (gzip_ng_threaded.py:316-321), in-order output and the crc32_combine fold of its writer
 * thread (:382-398) — with one GPU context per worker.  No device talks to another: batch j goes to device j % k
 * primed with the 32 KiB of host input before it, its output lands behind batch j - 1's in the caller's buffer, the
 * host folds the checksums. */
/* ------------------------------------------------------------------------- */
#define B2D_MAX_DEVICES 16
struct b2d_mctx {
    int n;
    b2d_ctx *ctx[B2D_MAX_DEVICES];
    uint64_t max_batch;
    std::mutex mu;                 /* one call at a time */
    std::vector<uint64_t> sizes;   /* compressed size of every device block of the last call */
};

extern "C" int b2d_mctx_create(const int *devices, int n_devices, uint64_t max_batch, b2d_mctx **out) {
    if (!out || !devices) return fail(B2D_STREAM_ERROR, "null argument");
    *out = nullptr;
    if (n_devices < 1 || n_devices > B2D_MAX_DEVICES) return fail(B2D_STREAM_ERROR, "between 1 and 16 devices");
    b2d_mctx *m = new b2d_mctx();
    m->n = n_devices;
    for (int i = 0; i < B2D_MAX_DEVICES; ++i) m->ctx[i] = nullptr;
    /* contexts come up side by side: the first CUDA call on a device pays for its primary context */
    int rcs[B2D_MAX_DEVICES];
    char errs[B2D_MAX_DEVICES][512];
    std::thread th[B2D_MAX_DEVICES];
    for (int i = 0; i < n_devices; ++i)
        th[i] = std::thread([&, i] {
            rcs[i] = b2d_ctx_create(devices[i], max_batch, &m->ctx[i]);
            snprintf(errs[i], sizeof errs[i], "%s", rcs[i] ? g_err : "");
        });
    for (int i = 0; i < n_devices; ++i) th[i].join();
    for (int i = 0; i < n_devices; ++i)
        if (rcs[i]) {
            const int rc = rcs[i];
            snprintf(g_err, sizeof g_err, "device %d: %s", devices[i], errs[i]);
            for (int k = 0; k < n_devices; ++k) ctx_free(m->ctx[k]);
            delete m;
            return rc;
        }
    m->max_batch = m->ctx[0]->max_batch;
    *out = m;
    return B2D_OK;
}
extern "C" void b2d_mctx_destroy(b2d_mctx *m) {
    if (!m) return;
    for (int i = 0; i < m->n; ++i) ctx_free(m->ctx[i]);
    delete m;
}
extern "C" int b2d_mctx_devices(b2d_mctx *m) { return m ? m->n : 0; }
extern "C" uint64_t b2d_mctx_launch_count(b2d_mctx *m) {
    uint64_t t = 0;
    if (m) for (int i = 0; i < m->n; ++i) t += m->ctx[i]->launches;
    return t;
}

/* batch size of a multi-device call over n bytes: small enough that every device gets two batches (its two lanes
 * overlap), at least 8 MiB (below that a batch does not fill a device), at most the contexts' max_batch */
static uint64_t multi_batch(const b2d_mctx *m, uint64_t n) {
    uint64_t b = align_up((n + 2ull * m->n - 1) / (2ull * m->n), B2D_BLOCK);
    const uint64_t lo = 8ull << 20;
    if (b < lo) b = lo;
    if (b > m->max_batch) b = m->max_batch;
    return b;
}

static int deflate_host_multi_impl(b2d_mctx *m, int level, int window_bits, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                   uint64_t dict_n, int final, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                                   uint32_t *adler, uint64_t *block_sizes, uint32_t *sub_bits, uint64_t cap, uint64_t *count);

extern "C" int b2d_deflate_host_multi(b2d_mctx *m, int level, int window_bits, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                      uint64_t dict_n, int final, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                                      uint32_t *adler, uint64_t *block_sizes, uint64_t cap, uint64_t *count) {
    return deflate_host_multi_impl(m, level, window_bits, data, n, dict, dict_n, final, out, out_cap, out_n, crc, adler, block_sizes, nullptr, cap, count);
}

extern "C" int b2d_compress_and_crc_multi_sub(b2d_mctx *m, int level, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                              uint64_t dict_n, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                                              uint64_t *block_sizes, uint32_t *sub_bits, uint64_t cap, uint64_t *count) {
    if (!crc) return fail(B2D_STREAM_ERROR, "null argument");
    if (sub_bits && !block_sizes) return fail(B2D_STREAM_ERROR, "sub_bits needs block_sizes");
    *crc = 0;
    return deflate_host_multi_impl(m, level, 15, data, n, dict, dict_n, 0, out, out_cap, out_n, crc, nullptr, block_sizes, sub_bits, cap, count);
}

static int deflate_host_multi_impl(b2d_mctx *m, int level, int window_bits, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                   uint64_t dict_n, int final, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                                   uint32_t *adler, uint64_t *block_sizes, uint32_t *sub_bits, uint64_t cap, uint64_t *count) {
    if (!m || !out_n || (!data && n) || (!out && out_cap) || (!dict && dict_n)) return fail(B2D_STREAM_ERROR, "null argument");
    if (!b2d_level_valid(level)) return fail(B2D_STREAM_ERROR, "Bad compression level");
    if (window_bits < 9 || window_bits > 15) return fail(B2D_STREAM_ERROR, "window_bits outside 9..15");
    std::lock_guard<std::mutex> whole(m->mu);
    const uint64_t nblocks = (n + B2D_BLOCK - 1) / B2D_BLOCK;
    if (count) *count = nblocks;
    if (block_sizes && cap < nblocks) return fail(B2D_BUF_ERROR, "b2d_deflate_host_multi: capacity below the number of blocks");
    const uint64_t batch = multi_batch(m, n);
    const uint64_t nb = (n + batch - 1) / batch;
    if (nb < 2) {
        /* one batch: one device, the single-context path (same bytes: the output does not depend on how batches are cut) */
        b2d_ctx *c = m->ctx[0];
        std::lock_guard<std::recursive_mutex> g(c->mu);
        int rc = b2d_deflate_host(c, level, window_bits, data, n, dict, dict_n, final, out, out_cap, out_n, crc, adler);
        if (rc || !block_sizes || !nblocks) return rc;
        std::vector<uint64_t> offs(nblocks + 1);
        uint64_t cnt = 0;
        if ((rc = b2d_block_index(c, offs.data(), nblocks + 1, &cnt))) return rc;
        if (cnt != nblocks + 1) return fail(B2D_STREAM_ERROR, "block index does not match the input size");
        for (uint64_t i = 0; i < nblocks; ++i) block_sizes[i] = offs[i + 1] - offs[i];
        if (sub_bits && (rc = b2d_subblock_bits(c, sub_bits, nblocks * B2D_NSUB, &cnt))) return rc;
        return B2D_OK;
    }
    const uint32_t max_dist = 1u << window_bits;
    if (dict_n > max_dist) { dict += dict_n - max_dist; dict_n = max_dist; }
    if (dict_n > B2D_WINDOW) { dict += dict_n - B2D_WINDOW; dict_n = B2D_WINDOW; }
    const int k = (uint64_t)m->n < nb ? m->n : (int)nb;      /* contexts that get work */
    b2d_ledger G(nb);
    std::thread th[B2D_MAX_DEVICES];
    std::vector<uint64_t> rel[B2D_MAX_DEVICES];               /* per context: its blocks' offsets relative to their batch */
    std::vector<uint32_t> rsub[B2D_MAX_DEVICES];              /* ... and the bit offsets of their sub-blocks */
    for (int d = 0; d < k; ++d)
        th[d] = std::thread([&, d] {
            b2d_ctx *c = m->ctx[d];
            std::lock_guard<std::recursive_mutex> g(c->mu);
            g_stage_share = (unsigned)k;
            int rc = B2D_OK;
            cudaError_t e = cudaSetDevice(c->device);
            if (e != cudaSuccess) rc = fail_cuda(e, "cudaSetDevice");
            if (!rc) {
                cudaPointerAttributes at;
                const bool ip = cudaPointerGetAttributes(&at, data) == cudaSuccess && at.type == cudaMemoryTypeHost;
                cudaGetLastError();
                const bool op = cudaPointerGetAttributes(&at, out) == cudaSuccess && at.type == cudaMemoryTypeHost;
                cudaGetLastError();
                rc = deflate_host_pipelined(c, level, max_dist, data, n, dict, dict_n, out, out_cap, final ? 2 : 0, crc != nullptr,
                                            adler != nullptr, ip, op, batch, (uint64_t)d, (uint64_t)k, &G);
            }
            if (!rc && block_sizes && c->index_local) {
                rel[d].resize(c->index_local);
                e = cudaMemcpyAsync(rel[d].data(), c->d_index, c->index_local * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->st);
                if (e == cudaSuccess && sub_bits) {
                    rsub[d].resize(c->index_local * B2D_NSUB);
                    e = cudaMemcpyAsync(rsub[d].data(), c->d_subidx, c->index_local * B2D_NSUB * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->st);
                }
                if (e == cudaSuccess) e = cudaStreamSynchronize(c->st);
                if (e != cudaSuccess) rc = fail_cuda(e, "block index copy");
            }
            if (rc) G.fail_with(rc, g_err);       /* wakes the contexts that wait for this one's offsets */
        });
    for (int d = 0; d < k; ++d) th[d].join();
    if (G.rc != B2D_OK) { snprintf(g_err, sizeof g_err, "%s", G.err); return G.rc; }
    uint64_t produced = G.off[nb];
    if (block_sizes) {
        /* absolute start of every block = its batch's place in the output + its offset inside the batch's output */
        const uint64_t bpb = batch / B2D_BLOCK;
        uint64_t prev = 0;
        for (uint64_t b = 0; b < nblocks; ++b) {
            const uint64_t j = b / bpb, d = j % (uint64_t)k, i = j / (uint64_t)k;
            const uint64_t start = G.off[j] + rel[d][i * bpb + b % bpb];
            if (b) block_sizes[b - 1] = start - prev;
            prev = start;
            if (sub_bits) memcpy(sub_bits + b * B2D_NSUB, &rsub[d][(i * bpb + b % bpb) * B2D_NSUB], B2D_NSUB * sizeof(uint32_t));
        }
        block_sizes[nblocks - 1] = produced - prev;
    }
    uint32_t crc_acc = crc ? *crc : 0, adler_acc = adler ? *adler : 1;
    ledger_fold(G, n, batch, crc ? &crc_acc : nullptr, adler ? &adler_acc : nullptr);
    if (final) { out[produced++] = 0x03; out[produced++] = 0x00; }
    *out_n = produced;
    if (crc) *crc = crc_acc;
    if (adler) *adler = adler_acc;
    return B2D_OK;
}

extern "C" int b2d_compress_and_crc_multi(b2d_mctx *m, int level, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                          uint64_t dict_n, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                                          uint64_t *block_sizes, uint64_t cap, uint64_t *count) {
    if (!crc) return fail(B2D_STREAM_ERROR, "null argument");
    *crc = 0;
    return b2d_deflate_host_multi(m, level, 15, data, n, dict, dict_n, 0, out, out_cap, out_n, crc, nullptr, block_sizes, cap, count);
}

/* ------------------------------------------------------------------------- */
/* inflate side (first version): streams of this library + their block index */
/* ------------------------------------------------------------------------- */
extern "C" int b2d_block_index(b2d_ctx *c, uint64_t *offsets, uint64_t cap, uint64_t *count) {
    if (!c || !count) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> g(c->mu);
    *count = c->index_n;
    if (!offsets || c->index_n == 0) return B2D_OK;
    if (cap < c->index_n) return fail(B2D_BUF_ERROR, "b2d_block_index: capacity below the number of entries");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(offsets, c->d_index, c->index_n * sizeof(uint64_t), cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    if (!c->index_base.empty()) {       /* multi-batch host call: offsets were relative to each batch's output */
        const uint64_t bpb = c->index_bpb;
        for (uint64_t i = 0; i + 1 < c->index_n; ++i) offsets[i] += c->index_base[i / bpb];
        offsets[c->index_n - 1] = c->index_total;
    }
    return B2D_OK;
}

/* scratch of the inflate side, grown on demand: 16-bit symbols (streams of this library only), block index, ready
 * flags, error words, the second stream */
static int inflate_prepare(b2d_ctx *c, uint64_t nblocks, uint64_t raw_n, bool need16) {
    if (!need16) raw_n = 0;
    if (raw_n > c->inf16_cap) {
        cudaFree(c->d_inf16); c->d_inf16 = nullptr; c->inf16_cap = 0;
        cudaError_t e = cudaMalloc(&c->d_inf16, raw_n * sizeof(uint16_t));
        if (e != cudaSuccess) { cudaGetLastError(); return fail(B2D_MEM_ERROR, "inflate scratch (2 bytes per output byte) does not fit"); }
        c->inf16_cap = raw_n;
    }
    if (nblocks > c->inf_index_cap) {
        cudaFree(c->d_inf_index); c->d_inf_index = nullptr; c->inf_index_cap = 0;
        CU(cudaMalloc(&c->d_inf_index, nblocks * sizeof(uint64_t)));
        c->inf_index_cap = nblocks;
    }
    if (!c->d_inf_err) CU(cudaMalloc(&c->d_inf_err, 2 * sizeof(uint32_t)));
    if (nblocks > c->inf_flags_cap) {
        cudaFree(c->d_inf_flags); c->d_inf_flags = nullptr; c->inf_flags_cap = 0;
        CU(cudaMalloc(&c->d_inf_flags, nblocks * sizeof(uint32_t)));
        c->inf_flags_cap = nblocks;
    }
    if (!c->st_inf) {
        int lo = 0, hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CU(cudaStreamCreateWithPriority(&c->st_inf, cudaStreamNonBlocking, hi));     /* the serial pass is the critical path */
        CU(cudaEventCreateWithFlags(&c->ev_inf_fork, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->ev_inf_join, cudaEventDisableTiming));
    }
    return B2D_OK;
}

/* report what the decode left in the error words (streams synchronised by the caller) */
static int inflate_verdict(const uint32_t err[2]) {
    if (!err[0]) return B2D_OK;
    static const char *why[] = {"", "invalid block header", "output does not match the block layout", "invalid code", "", "", "", ""};
    snprintf(g_err, sizeof g_err, "inflate: %s in device block %u", why[(err[0] >> 28) & 7], err[0] & 0x0fffffffu);
    return -3;   /* Z_DATA_ERROR */
}

/* how the three inflate kernels are issued: side by side where the device can do that, on one stream where it
 * cannot (or where a tool serialises launches anyway) */
static int inflate_mode(b2d_ctx *c) {
    static const bool forced = (getenv("B2D_INF_SERIAL") && getenv("B2D_INF_SERIAL")[0] == '1') ||
                               (getenv("CUDA_LAUNCH_BLOCKING") && getenv("CUDA_LAUNCH_BLOCKING")[0] == '1');
    if (forced) return B2D_INF_SERIAL;
    int conc = 1;
    if (cudaDeviceGetAttribute(&conc, cudaDevAttrConcurrentKernels, c->device) != cudaSuccess) { cudaGetLastError(); conc = 1; }
    return conc ? B2D_INF_CONCURRENT : B2D_INF_SERIAL;
}

/* the caller's sub-block bit offsets (B2D_NSUB per block, host array) on the device, or null */
static int upload_sub_bits(b2d_ctx *c, const uint32_t *sub_bits, uint64_t nblocks, const uint32_t **d_out) {
    *d_out = nullptr;
    if (!sub_bits) return B2D_OK;
    for (uint64_t i = 0; i < nblocks * B2D_NSUB; ++i)
        if (sub_bits[i] >= 8u * (B2D_BLOCK + 1024u) || ((i % B2D_NSUB) && sub_bits[i] && sub_bits[i] <= sub_bits[i - 1]))
            return fail(B2D_STREAM_ERROR, "bad sub-block index");
    if (nblocks * B2D_NSUB > c->inf_sub_cap) {
        cudaFree(c->d_inf_sub); c->d_inf_sub = nullptr; c->inf_sub_cap = 0;
        CU(cudaMalloc(&c->d_inf_sub, nblocks * B2D_NSUB * sizeof(uint32_t)));
        c->inf_sub_cap = nblocks * B2D_NSUB;
    }
    CU(cudaMemcpyAsync(c->d_inf_sub, sub_bits, nblocks * B2D_NSUB * sizeof(uint32_t), cudaMemcpyHostToDevice, c->st));
    *d_out = c->d_inf_sub;
    return B2D_OK;
}

extern "C" int b2d_subblock_bits(b2d_ctx *c, uint32_t *bits, uint64_t cap, uint64_t *count) {
    if (!c || !count) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> g(c->mu);
    const uint64_t nblocks = c->index_n ? c->index_n - 1 : 0;
    *count = nblocks * B2D_NSUB;
    if (!bits || !nblocks) return B2D_OK;
    if (cap < *count) return fail(B2D_BUF_ERROR, "b2d_subblock_bits: capacity below the number of entries");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(bits, c->d_subidx, *count * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    return B2D_OK;
}

extern "C" int b2d_inflate_device(b2d_ctx *c, const uint8_t *d_in, uint64_t in_n, const uint64_t *block_offsets,
                                  uint64_t nblocks, const uint8_t *d_dict, uint64_t dict_n, uint8_t *d_out, uint64_t raw_n,
                                  float *ms) {
    return b2d_inflate_device_sub(c, d_in, in_n, block_offsets, nullptr, nblocks, d_dict, dict_n, d_out, raw_n, ms);
}

extern "C" int b2d_inflate_device_sub(b2d_ctx *c, const uint8_t *d_in, uint64_t in_n, const uint64_t *block_offsets,
                                      const uint32_t *sub_bits, uint64_t nblocks, const uint8_t *d_dict, uint64_t dict_n,
                                      uint8_t *d_out, uint64_t raw_n, float *ms) {
    if (!c || (!d_in && in_n) || (!d_out && raw_n) || (!block_offsets && nblocks) || (!d_dict && dict_n))
        return fail(B2D_STREAM_ERROR, "null argument");
    if (nblocks != (raw_n + B2D_BLOCK - 1) / B2D_BLOCK) return fail(B2D_STREAM_ERROR, "block index does not match the output size");
    if (nblocks > 0x0fffffffull) return fail(B2D_STREAM_ERROR, "too many blocks");
    if (dict_n > B2D_WINDOW) return fail(B2D_STREAM_ERROR, "device dictionary longer than 32768 bytes");
    if ((uintptr_t)d_out & 15) return fail(B2D_STREAM_ERROR, "misaligned device pointer");
    for (uint64_t b = 0; b < nblocks; ++b)
        if (block_offsets[b] >= in_n || (b && block_offsets[b] <= block_offsets[b - 1]))
            return fail(B2D_STREAM_ERROR, "bad block index");
    if (ms) *ms = 0.f;
    if (raw_n == 0) return B2D_OK;
    std::lock_guard<std::recursive_mutex> g(c->mu);
    CU(cudaSetDevice(c->device));
    { int rc = inflate_prepare(c, nblocks, raw_n, true); if (rc) return rc; }
    CU(cudaMemcpyAsync(c->d_inf_index, block_offsets, nblocks * sizeof(uint64_t), cudaMemcpyHostToDevice, c->st));
    const uint32_t *d_sub = nullptr;
    { int rc = upload_sub_bits(c, sub_bits, nblocks, &d_sub); if (rc) return rc; }
    CU(cudaMemsetAsync(c->d_inf_err, 0, 2 * sizeof(uint32_t), c->st));
    if (ms) CU(cudaEventRecord(c->ev0, c->st));
    CU(cudaMemsetAsync(c->d_inf_flags, 0, nblocks * sizeof(uint32_t), c->st));
    const int mode = inflate_mode(c);
    if (mode == B2D_INF_CONCURRENT) {
        CU(cudaEventRecord(c->ev_inf_fork, c->st));                    /* flags and error words are cleared */
        CU(cudaStreamWaitEvent(c->st_inf, c->ev_inf_fork, 0));
    }
    CU(b2d_launch_inflate(d_in, in_n, c->d_inf_index, d_sub, 0, (uint32_t)nblocks, (uint32_t)nblocks, d_dict, (uint32_t)dict_n, c->d_inf16, d_out, raw_n,
                          c->d_inf_flags, c->d_inf_err, c->st, mode == B2D_INF_CONCURRENT ? c->st_inf : c->st, mode));
    if (mode == B2D_INF_CONCURRENT) {
        CU(cudaEventRecord(c->ev_inf_join, c->st_inf));
        CU(cudaStreamWaitEvent(c->st, c->ev_inf_join, 0));
    }
    c->launches += B2D_INF_LAUNCHES;
    if (ms) CU(cudaEventRecord(c->ev1, c->st));
    uint32_t err[2] = {0, 0};
    CU(cudaMemcpyAsync(err, c->d_inf_err, sizeof err, cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    if (err[1] && !err[0]) {
        /* the pass over the tails gave up waiting for the decode (it could not run beside it and the wait ran out).
         * That says nothing about the data: every block is decoded by now, so resolve again on one stream. */
        CU(cudaMemsetAsync(c->d_inf_err + 1, 0, sizeof(uint32_t), c->st));
        CU(b2d_launch_inflate(d_in, in_n, c->d_inf_index, d_sub, 0, (uint32_t)nblocks, (uint32_t)nblocks, d_dict, (uint32_t)dict_n, c->d_inf16, d_out, raw_n,
                              c->d_inf_flags, c->d_inf_err, c->st, c->st, B2D_INF_RESOLVE_ONLY));
        c->launches += 2;
        if (ms) CU(cudaEventRecord(c->ev1, c->st));
        CU(cudaMemcpyAsync(err, c->d_inf_err, sizeof err, cudaMemcpyDeviceToHost, c->st));
        CU(cudaStreamSynchronize(c->st));
        if (err[1] && !err[0]) return fail(B2D_CUDA_ERROR, "inflate: the resolving pass did not complete");
    }
    if (ms) CU(cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return inflate_verdict(err);
}

#ifndef B2D_INF_CHUNK_BLOCKS
#define B2D_INF_CHUNK_BLOCKS 512u    /* 64 MiB of output per chunk of a host call */
#endif

/* device copies of a host call's input (32 KiB of room for the dictionary in front of it) and output: kept by the
 * context, grown on demand */
static int inflate_host_buffers(b2d_ctx *c, uint64_t in_n, uint64_t out_n) {
    const uint64_t need_in = B2D_WINDOW + in_n + 64, need_out = out_n + 64;
    if (need_in > c->inf_in_cap) {
        cudaFree(c->d_inf_in); c->d_inf_in = nullptr; c->inf_in_cap = 0;
        if (cudaMalloc(&c->d_inf_in, need_in) != cudaSuccess) { cudaGetLastError(); return fail(B2D_MEM_ERROR, "inflate input buffer does not fit"); }
        c->inf_in_cap = need_in;
    }
    if (need_out > c->inf_out_cap) {
        cudaFree(c->d_inf_out); c->d_inf_out = nullptr; c->inf_out_cap = 0;
        if (cudaMalloc(&c->d_inf_out, need_out) != cudaSuccess) { cudaGetLastError(); return fail(B2D_MEM_ERROR, "inflate output buffer does not fit"); }
        c->inf_out_cap = need_out;
    }
    return B2D_OK;
}

extern "C" int b2d_inflate_host(b2d_ctx *c, const uint8_t *in, uint64_t in_n, const uint64_t *block_offsets, uint64_t nblocks,
                                const uint8_t *dict, uint64_t dict_n, uint8_t *out, uint64_t raw_n, uint32_t *crc) {
    return b2d_inflate_host_sub(c, in, in_n, block_offsets, nullptr, nblocks, dict, dict_n, out, raw_n, crc);
}

/* One host call in two halves, so that a reader can queue the decode of its next window before it waits for the window
 * before (a 32 KiB sub-block takes a warp 6-7 ms whatever shares the SM: a single 512 MiB call leaves the device half
 * empty while its last warps finish and its bytes travel back).  The decode needs no dictionary — what it cannot know
 * it leaves as markers — so b2d_inflate_host_decode uploads the compressed bytes and queues k_inf_decode for every chunk
 * without waiting for anything; b2d_inflate_host_resolve, given the dictionary (the 32 KiB the window before ended
 * with), queues the pass over the tails and the finishing pass per chunk, brings the bytes back and checks the
 * verdict.  One decode may be outstanding per context. */
extern "C" int b2d_inflate_host_decode(b2d_ctx *c, const uint8_t *in, uint64_t in_n, const uint64_t *block_offsets, const uint32_t *sub_bits,
                                       uint64_t nblocks, uint64_t dict_n, uint64_t raw_n) {
    if (!c || (!in && in_n) || (!block_offsets && nblocks)) return fail(B2D_STREAM_ERROR, "null argument");
    if (raw_n == 0 || nblocks != (raw_n + B2D_BLOCK - 1) / B2D_BLOCK) return fail(B2D_STREAM_ERROR, "block index does not match the output size");
    if (nblocks > 0x0fffffffull) return fail(B2D_STREAM_ERROR, "too many blocks");
    for (uint64_t b = 0; b < nblocks; ++b)
        if (block_offsets[b] >= in_n || (b && block_offsets[b] <= block_offsets[b - 1])) return fail(B2D_STREAM_ERROR, "bad block index");
    if (dict_n > B2D_WINDOW) dict_n = B2D_WINDOW;
    std::lock_guard<std::recursive_mutex> whole(c->mu);
    if (c->inf_pend.active) return fail(B2D_STREAM_ERROR, "a decode is outstanding on this context: b2d_inflate_host_resolve comes first");
    CU(cudaSetDevice(c->device));
    { int rc = inflate_host_buffers(c, in_n, raw_n); if (rc) return rc; }
    { int rc = inflate_prepare(c, nblocks, raw_n, true); if (rc) return rc; }
    if (!c->st_inf_in) {
        CU(cudaStreamCreateWithFlags(&c->st_inf_in, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&c->st_inf_out, cudaStreamNonBlocking));
        for (int i = 0; i < B2D_INF_DEC_STREAMS; ++i) CU(cudaStreamCreateWithFlags(&c->st_inf_dec[i], cudaStreamNonBlocking));
    }
    const uint64_t nchunks = (nblocks + B2D_INF_CHUNK_BLOCKS - 1) / B2D_INF_CHUNK_BLOCKS;
    while (c->ev_inf_h2d.size() < nchunks) {
        cudaEvent_t e1, e2;
        CU(cudaEventCreateWithFlags(&e1, cudaEventDisableTiming));
        c->ev_inf_h2d.push_back(e1);
        CU(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
        c->ev_inf_done.push_back(e2);
    }
    const int mode = inflate_mode(c);
    const bool conc = mode == B2D_INF_CONCURRENT;
    uint8_t *d_stream = c->d_inf_in + B2D_WINDOW;
    /* Chunks of 512 blocks flow through three stages: the compressed bytes of chunk k + 1 go up (st_inf_in) while chunk
     * k is decoded (one of B2D_INF_DEC_STREAMS decode streams: a block takes a warp milliseconds, so neighbouring chunks must decode
     * side by side to fill the device) and, in the second half of the call, resolved (st_inf, which follows the decodes
     * block by block) while the bytes of chunk k - 1 come back (st_inf_out). */
    CU(cudaMemcpyAsync(c->d_inf_index, block_offsets, nblocks * sizeof(uint64_t), cudaMemcpyHostToDevice, c->st));
    const uint32_t *d_sub = nullptr;
    { int rc = upload_sub_bits(c, sub_bits, nblocks, &d_sub); if (rc) return rc; }
    CU(cudaMemsetAsync(c->d_inf_err, 0, 2 * sizeof(uint32_t), c->st));
    CU(cudaMemsetAsync(c->d_inf_flags, 0, nblocks * sizeof(uint32_t), c->st));
    if (conc) {
        CU(cudaEventRecord(c->ev_inf_fork, c->st));                    /* index, flags, error words are in place */
        CU(cudaStreamWaitEvent(c->st_inf, c->ev_inf_fork, 0));
        for (int i = 0; i < B2D_INF_DEC_STREAMS; ++i) CU(cudaStreamWaitEvent(c->st_inf_dec[i], c->ev_inf_fork, 0));
    }
    for (uint64_t k = 0; k < nchunks; ++k) {
        const uint64_t b0 = k * B2D_INF_CHUNK_BLOCKS, b1 = b0 + B2D_INF_CHUNK_BLOCKS < nblocks ? b0 + B2D_INF_CHUNK_BLOCKS : nblocks;
        const uint64_t lo = block_offsets[b0], hi = b1 < nblocks ? block_offsets[b1] : in_n;
        cudaStream_t st_dec = conc ? c->st_inf_dec[k % B2D_INF_DEC_STREAMS] : c->st;
        CU(cudaMemcpyAsync(d_stream + lo, in + lo, hi - lo, cudaMemcpyHostToDevice, c->st_inf_in));
        CU(cudaEventRecord(c->ev_inf_h2d[k], c->st_inf_in));
        CU(cudaStreamWaitEvent(st_dec, c->ev_inf_h2d[k], 0));
        CU(b2d_launch_inflate(d_stream, in_n, c->d_inf_index, d_sub, (uint32_t)b0, (uint32_t)(b1 - b0), (uint32_t)nblocks, nullptr, (uint32_t)dict_n,
                              c->d_inf16, c->d_inf_out, raw_n, c->d_inf_flags, c->d_inf_err, st_dec, st_dec, B2D_INF_DECODE_ONLY));
        c->launches += 1;
    }
    c->inf_pend.active = true; c->inf_pend.in_n = in_n; c->inf_pend.nblocks = nblocks; c->inf_pend.raw_n = raw_n;
    c->inf_pend.nchunks = nchunks; c->inf_pend.dict_n = dict_n; c->inf_pend.d_sub = d_sub; c->inf_pend.mode = mode;
    return B2D_OK;
}

/* wait for everything a decode has queued (the context's buffers may be reused afterwards) */
static int inflate_settle(b2d_ctx *c) {
    c->inf_pend.active = false;
    CU(cudaStreamSynchronize(c->st_inf_in));
    for (int i = 0; i < B2D_INF_DEC_STREAMS; ++i) CU(cudaStreamSynchronize(c->st_inf_dec[i]));
    CU(cudaStreamSynchronize(c->st_inf));
    CU(cudaStreamSynchronize(c->st_inf_out));
    CU(cudaStreamSynchronize(c->st));
    return B2D_OK;
}

/* out == NULL: give the outstanding decode up (waits for what it queued) */
extern "C" int b2d_inflate_host_resolve(b2d_ctx *c, const uint8_t *dict, uint64_t dict_n, uint8_t *out, uint64_t raw_n, uint32_t *crc) {
    if (!c || (!dict && dict_n)) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> whole(c->mu);
    if (!c->inf_pend.active) return fail(B2D_STREAM_ERROR, "b2d_inflate_host_resolve without b2d_inflate_host_decode");
    CU(cudaSetDevice(c->device));
    if (!out) return inflate_settle(c);
    if (dict_n > B2D_WINDOW) { dict += dict_n - B2D_WINDOW; dict_n = B2D_WINDOW; }
    if (raw_n != c->inf_pend.raw_n || dict_n != c->inf_pend.dict_n) {
        inflate_settle(c);
        return fail(B2D_STREAM_ERROR, "output size or dictionary length differ from what the decode was told");
    }
    const uint64_t in_n = c->inf_pend.in_n, nblocks = c->inf_pend.nblocks, nchunks = c->inf_pend.nchunks;
    const uint32_t *d_sub = c->inf_pend.d_sub;
    const bool conc = c->inf_pend.mode == B2D_INF_CONCURRENT;
    c->inf_pend.active = false;
    uint8_t *d_stream = c->d_inf_in + B2D_WINDOW;
    const uint8_t *d_dict = dict_n ? c->d_inf_in : nullptr;
    cudaStream_t st_res = conc ? c->st_inf : c->st;                    /* where a chunk's bytes become final */
    if (dict_n) CU(cudaMemcpyAsync(c->d_inf_in, dict, dict_n, cudaMemcpyHostToDevice, st_res));
    /* B2D_TRACE=1: when each chunk was resolved and downloaded (device time, ms from the start of this half) */
    static const bool trace = getenv("B2D_TRACE") && getenv("B2D_TRACE")[0] == '1';
    std::vector<cudaEvent_t> tev;
    const auto t_host = std::chrono::steady_clock::now();
    if (trace) {
        tev.resize(2 * nchunks + 1);
        for (auto &e : tev) CU(cudaEventCreate(&e));
        CU(cudaEventRecord(tev[2 * nchunks], st_res));
    }
    for (uint64_t k = 0; k < nchunks; ++k) {
        const uint64_t b0 = k * B2D_INF_CHUNK_BLOCKS, b1 = b0 + B2D_INF_CHUNK_BLOCKS < nblocks ? b0 + B2D_INF_CHUNK_BLOCKS : nblocks;
        /* the pass over the tails follows the decode through the blocks' ready flags (chunk by chunk on one stream: the
         * tails of chunk k build on those of chunk k - 1) */
        CU(b2d_launch_inflate(d_stream, in_n, c->d_inf_index, d_sub, (uint32_t)b0, (uint32_t)(b1 - b0), (uint32_t)nblocks, d_dict, (uint32_t)dict_n,
                              c->d_inf16, c->d_inf_out, raw_n, c->d_inf_flags, c->d_inf_err, st_res, st_res, B2D_INF_RESOLVE_ONLY));
        c->launches += 2;
        CU(cudaEventRecord(c->ev_inf_done[k], st_res));
        if (trace) CU(cudaEventRecord(tev[2 * k], st_res));
        CU(cudaStreamWaitEvent(c->st_inf_out, c->ev_inf_done[k], 0));
        const uint64_t o0 = b0 * (uint64_t)B2D_BLOCK, o1 = b1 * (uint64_t)B2D_BLOCK < raw_n ? b1 * (uint64_t)B2D_BLOCK : raw_n;
        CU(cudaMemcpyAsync(out + o0, c->d_inf_out + o0, o1 - o0, cudaMemcpyDeviceToHost, c->st_inf_out));
        if (trace) CU(cudaEventRecord(tev[2 * k + 1], c->st_inf_out));
        if (k + 1 == nchunks) {
            if (crc) { CU(b2d_crc32_dev_launch(c->ck, c->d_inf_out, raw_n, st_res)); c->launches += crc_launches(c->d_inf_out, raw_n); }
            if (st_res != c->st) { CU(cudaEventRecord(c->ev_inf_join, st_res)); CU(cudaStreamWaitEvent(c->st, c->ev_inf_join, 0)); }
        }
    }
    uint32_t err[2] = {0, 0};
    CU(cudaMemcpyAsync(err, c->d_inf_err, sizeof err, cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    CU(cudaStreamSynchronize(c->st_inf_out));
    if (trace) {
        fprintf(stderr, "b2d_inflate_host_resolve: %llu blocks, host %.2f ms; chunk: resolved / downloaded (ms)\n", (unsigned long long)nblocks,
                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_host).count());
        for (uint64_t k = 0; k < nchunks; ++k) {
            float b = 0, d = 0;
            cudaEventElapsedTime(&b, tev[2 * nchunks], tev[2 * k]);
            cudaEventElapsedTime(&d, tev[2 * nchunks], tev[2 * k + 1]);
            fprintf(stderr, "   %2llu: %6.2f %6.2f\n", (unsigned long long)k, b, d);
        }
        for (auto &e : tev) cudaEventDestroy(e);
    }
    if (err[1] && !err[0]) {
        /* the pass over the tails gave up waiting for a decode (never a statement about the data): once every block is
         * decoded, resolve everything again on one stream and fetch the bytes once more */
        for (int i = 0; i < B2D_INF_DEC_STREAMS; ++i) CU(cudaStreamSynchronize(c->st_inf_dec[i]));
        CU(cudaMemsetAsync(c->d_inf_err + 1, 0, sizeof(uint32_t), c->st));
        CU(b2d_launch_inflate(d_stream, in_n, c->d_inf_index, d_sub, 0, (uint32_t)nblocks, (uint32_t)nblocks, d_dict, (uint32_t)dict_n, c->d_inf16,
                              c->d_inf_out, raw_n, c->d_inf_flags, c->d_inf_err, c->st, c->st, B2D_INF_RESOLVE_ONLY));
        c->launches += 2;
        if (crc) { CU(b2d_crc32_dev_launch(c->ck, c->d_inf_out, raw_n, c->st)); c->launches += crc_launches(c->d_inf_out, raw_n); }
        CU(cudaMemcpyAsync(out, c->d_inf_out, raw_n, cudaMemcpyDeviceToHost, c->st));
        CU(cudaMemcpyAsync(err, c->d_inf_err, sizeof err, cudaMemcpyDeviceToHost, c->st));
        CU(cudaStreamSynchronize(c->st));
        if (err[1] && !err[0]) return fail(B2D_CUDA_ERROR, "inflate: the resolving pass did not complete");
    }
    { int rc = inflate_verdict(err); if (rc) return rc; }
    if (crc) *crc = b2d_crc32_dev_finish(c->ck, raw_n, *crc);
    return B2D_OK;
}

extern "C" int b2d_inflate_host_sub(b2d_ctx *c, const uint8_t *in, uint64_t in_n, const uint64_t *block_offsets, const uint32_t *sub_bits,
                                    uint64_t nblocks, const uint8_t *dict, uint64_t dict_n, uint8_t *out, uint64_t raw_n, uint32_t *crc) {
    if (!c || (!in && in_n) || (!out && raw_n) || (!dict && dict_n) || (!block_offsets && nblocks)) return fail(B2D_STREAM_ERROR, "null argument");
    if (raw_n == 0) return nblocks ? fail(B2D_STREAM_ERROR, "block index does not match the output size") : B2D_OK;
    std::lock_guard<std::recursive_mutex> whole(c->mu);      /* the device copies belong to this call until it returns */
    int rc = b2d_inflate_host_decode(c, in, in_n, block_offsets, sub_bits, nblocks, dict_n, raw_n);
    if (rc) return rc;
    return b2d_inflate_host_resolve(c, dict, dict_n, out, raw_n, crc);
}

/* ------------------------------------------------------------------------- */
/* BGZF (the blocked gzip of htslib / bgzip; the reference ships and reads one: tests/data/test.fastq.bgzip.gz,
 * tests/test_gzip_ng.py:466-473): a series of gzip members of at most 64 KiB, each a complete DEFLATE stream without
 * a dictionary, whose total size sits in a "BC" extra subfield — a block index for free. */
/* ------------------------------------------------------------------------- */
extern "C" int b2d_bgzf_scan(const uint8_t *buf, uint64_t n, uint64_t max_members, uint64_t *payload_off, uint32_t *payload_len,
                             uint32_t *raw_len, uint32_t *crc, uint64_t *count, uint64_t *consumed) {
    if (!count || !consumed || (!buf && n)) return fail(B2D_STREAM_ERROR, "null argument");
    uint64_t pos = 0, k = 0;
    while (k < max_members && pos + 18 <= n) {
        const uint8_t *h = buf + pos;
        if (h[0] != 0x1f || h[1] != 0x8b || h[2] != 8 || h[3] != 4) break;             /* a gzip member with FEXTRA and nothing else */
        const uint32_t xlen = h[10] | (h[11] << 8);
        if (pos + 12 + xlen > n) break;
        uint32_t bsize = 0; bool found = false;
        for (uint32_t o = 0; o + 4 <= xlen;) {                                         /* RFC 1952 2.3.1.1 subfields */
            const uint8_t *f = h + 12 + o;
            const uint32_t slen = f[2] | (f[3] << 8);
            if (o + 4 + slen > xlen) break;
            if (f[0] == 'B' && f[1] == 'C' && slen == 2) { bsize = (f[4] | (f[5] << 8)) + 1u; found = true; }
            o += 4 + slen;
        }
        if (!found || bsize < 12u + xlen + 8u) break;
        if (pos + bsize > n) break;                                                    /* member not complete in this buffer */
        const uint8_t *t = h + bsize - 8;
        const uint32_t isize = t[4] | (t[5] << 8) | (t[6] << 16) | ((uint32_t)t[7] << 24);
        if (isize > 65536u) break;                                                     /* the format's limit */
        if (payload_off) {
            payload_off[k] = pos + 12 + xlen;
            payload_len[k] = bsize - 12u - xlen - 8u;
            crc[k] = t[0] | (t[1] << 8) | (t[2] << 16) | ((uint32_t)t[3] << 24);
            raw_len[k] = isize;
        }
        pos += bsize;
        ++k;
    }
    *count = k;
    *consumed = pos;
    return B2D_OK;
}

/* Inflate `nmembers` independent DEFLATE streams (host buffers): member i is in[payload_off[i], + payload_len[i]) and
 * decodes to raw_len[i] bytes that land back to back in `out`.  One warp per member, bytes written directly.  `crcs`
 * (optional): the CRC-32 every member must have — checked against one device CRC of the whole output, folded on the host
 * with crc32_combine. */
extern "C" int b2d_inflate_members_host(b2d_ctx *c, const uint8_t *in, uint64_t in_n, const uint64_t *payload_off,
                                        const uint32_t *payload_len, const uint32_t *raw_len, const uint32_t *crcs, uint64_t nmembers,
                                        uint8_t *out, uint64_t out_cap, uint64_t *out_n) {
    if (!c || !out_n || (!in && in_n) || (nmembers && (!payload_off || !payload_len || !raw_len))) return fail(B2D_STREAM_ERROR, "null argument");
    if (nmembers > 0x0fffffffull) return fail(B2D_STREAM_ERROR, "too many members");
    std::vector<uint64_t> idx(2 * nmembers + 2);
    uint64_t *in_idx = idx.data(), *out_idx = idx.data() + nmembers;        /* out_idx has nmembers + 1 entries */
    uint64_t total = 0;
    uint32_t want = 0;
    for (uint64_t i = 0; i < nmembers; ++i) {
        if (payload_off[i] + payload_len[i] > in_n || raw_len[i] > 65536u) return fail(B2D_STREAM_ERROR, "bad member index");
        in_idx[i] = payload_off[i];
    }
    for (uint64_t i = 0; i < nmembers; ++i) {
        out_idx[i] = total;
        if (crcs) want = b2d_host_crc32_combine(want, crcs[i], raw_len[i]);
        total += raw_len[i];
    }
    out_idx[nmembers] = total;
    *out_n = total;
    if (total > out_cap) return fail(B2D_BUF_ERROR, "output buffer too small");
    if (total == 0) return B2D_OK;
    if (!out) return fail(B2D_STREAM_ERROR, "null argument");
    std::lock_guard<std::recursive_mutex> whole(c->mu);
    CU(cudaSetDevice(c->device));
    { int rc = inflate_host_buffers(c, in_n, total); if (rc) return rc; }
    { int rc = inflate_prepare(c, 2 * nmembers + 2, 0, false); if (rc) return rc; }
    CU(cudaMemcpyAsync(c->d_inf_index, idx.data(), (2 * nmembers + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, c->st));
    CU(cudaMemsetAsync(c->d_inf_err, 0, 2 * sizeof(uint32_t), c->st));
    CU(cudaMemcpyAsync(c->d_inf_in, in, in_n, cudaMemcpyHostToDevice, c->st));
    CU(b2d_launch_inflate_members(c->d_inf_in, in_n, c->d_inf_index, c->d_inf_index + nmembers, (uint32_t)nmembers, c->d_inf_out, total,
                                  c->d_inf_err, c->st));
    c->launches += 1;
    if (crcs) { CU(b2d_crc32_dev_launch(c->ck, c->d_inf_out, total, c->st)); c->launches += crc_launches(c->d_inf_out, total); }
    CU(cudaMemcpyAsync(out, c->d_inf_out, total, cudaMemcpyDeviceToHost, c->st));
    uint32_t err[2] = {0, 0};
    CU(cudaMemcpyAsync(err, c->d_inf_err, sizeof err, cudaMemcpyDeviceToHost, c->st));
    CU(cudaStreamSynchronize(c->st));
    { int rc = inflate_verdict(err); if (rc) return rc; }
    if (crcs && b2d_crc32_dev_finish(c->ck, total, 0) != want) {
        snprintf(g_err, sizeof g_err, "inflate: CRC check failed");
        return -3;
    }
    return B2D_OK;
}
