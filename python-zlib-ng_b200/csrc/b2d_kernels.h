/* b2d_kernels.h — internal interface between the kernels (b2d_kernels.cu,
 * b2d_checksum.cu) and the C-ABI host layer (b2d_api.cu). Not installed. */
#ifndef B2D_KERNELS_H
#define B2D_KERNELS_H
#include <cuda_runtime.h>
#include <stdint.h>
#include "b2d_common.h"
#include "b2d_huff.h"

typedef struct {
    const uint8_t *buf;        /* device: [dict | data], 16-byte aligned            */
    uint64_t D, n;             /* dictionary bytes (<= 32768 used), data bytes      */
    uint64_t buf_readable;     /* bytes of buf that may be read (>= D + n)          */
    int level;                 /* 0..9                                              */
    uint32_t max_dist;         /* longest match distance allowed (<= 32768)         */
    uint16_t *prevd;           /* [nblocks * B2D_LINKS] per-block link arrays       */
    uint32_t *match;           /* [n]                                               */
    uint32_t *tok;             /* [nsub * B2D_SUB]                                  */
    b2d_subinfo *info;         /* [nsub]                                            */
    uint32_t *hist;            /* [nsub * B2D_HIST]                                 */
    b2d_plan *plan;            /* [nsub]                                            */
    uint64_t *block_off;       /* [nblocks + 1] byte offsets into out               */
    uint64_t *total;           /* device: in = bytes of out already used, out = after this job */
    uint8_t *out;              /* device output, 4-byte aligned, ZEROED by the caller from *total on */
    /* batches of one call may run on two streams; *total is the only state they share, so the scan
     * stage waits for the previous batch's scan and signals its own (both optional) */
    cudaEvent_t scan_wait, scan_done;
} b2d_job;

cudaError_t b2d_kernels_init(void);
#define B2D_NSTAGE 6   /* chains, match, parse, plan, scan, emit */
cudaError_t b2d_launch_pipeline(const b2d_job *J, cudaStream_t st, cudaEvent_t *ev);

/* checksums (b2d_checksum.cu) */
typedef struct b2d_cksum_ws b2d_cksum_ws;
cudaError_t b2d_cksum_ws_create(b2d_cksum_ws **ws);
void b2d_cksum_ws_destroy(b2d_cksum_ws *ws);
/* crc32 / adler32 of n device bytes continuing from `start`; synchronises st */
cudaError_t b2d_crc32_dev(b2d_cksum_ws *ws, const uint8_t *d, uint64_t n, uint32_t start, uint32_t *out, cudaStream_t st);
cudaError_t b2d_adler32_dev(b2d_cksum_ws *ws, const uint8_t *d, uint64_t n, uint32_t start, uint32_t *out, cudaStream_t st);
/* asynchronous halves, for callers that overlap them with the deflate kernels */
cudaError_t b2d_crc32_dev_launch(b2d_cksum_ws *ws, const uint8_t *d, uint64_t n, cudaStream_t st);
uint32_t b2d_crc32_dev_finish(b2d_cksum_ws *ws, uint64_t n, uint32_t start);
cudaError_t b2d_adler32_dev_launch(b2d_cksum_ws *ws, const uint8_t *d, uint64_t n, cudaStream_t st);
uint32_t b2d_adler32_dev_finish(b2d_cksum_ws *ws, uint64_t n, uint32_t start);

/* adler32 of n device bytes without allocating anything: the partial sums live in a file-scope device array
 * (b2d_checksum.cu); synchronous, on the calling thread's per-thread stream */
cudaError_t b2d_adler32_noalloc(int device, const uint8_t *d, uint64_t n, uint32_t start, uint32_t *out, const char **why);

/* host GF(2) helpers (b2d_checksum.cu) */
uint32_t b2d_host_crc32_combine(uint32_t crc1, uint32_t crc2, uint64_t len2);
uint32_t b2d_host_adler32_combine(uint32_t a1, uint32_t a2, uint64_t len2);
#endif
