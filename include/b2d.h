/* b2d.h — C ABI of libb2d.so, the B200-native block-parallel DEFLATE / checksum
 * library.  Plain pointers and sizes only; no Python, torch or CUDA types.
 *
 * Each entry point names the reference interface it stands in for (paths under
 * /root/reference/src/zlib_ng/).  The reference is a CPython extension, so the
 * "binding a maintainer would add" is the ctypes stub in
 * python-zlib-ng_b200/zlib_ng_b200/_native.py.
 *
 * Return value: 0 = ok, negative = error; codes reuse zlib's numbering
 * (zlib-ng.h Z_STREAM_ERROR / Z_MEM_ERROR / Z_BUF_ERROR) so the Python layer can
 * raise exactly what zlib_ngmodule.c raises (zlib_ngmodule.c:68-95,224-225,
 * 1749-1755).  There is no CPU fallback anywhere behind this ABI: if no CUDA
 * device is usable, b2d_ctx_create fails with B2D_CUDA_ERROR.
 */
#ifndef B2D_H
#define B2D_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B2D_OK 0
#define B2D_STREAM_ERROR (-2) /* bad level / bad argument            (Z_STREAM_ERROR) */
#define B2D_MEM_ERROR (-4)    /* host or device allocation failed     (Z_MEM_ERROR)    */
#define B2D_BUF_ERROR (-5)    /* output capacity too small            (Z_BUF_ERROR)    */
#define B2D_CUDA_ERROR (-100) /* CUDA runtime error; see b2d_last_error()              */

#define B2D_DEVICE_PAD 64u    /* bytes that must be readable after device-resident input */

typedef struct b2d_ctx b2d_ctx;

/* One context = one GPU + its work buffers (sized for `max_batch` input bytes per
 * launch sequence; 0 selects 256 MiB; rounded up to 128 KiB).  Calls on one
 * context are serialised internally; use one context per GPU.
 * Stands in for ParallelCompress__new__ / zng_deflateInit2 (zlib_ngmodule.c:1617-1674). */
int b2d_ctx_create(int device, uint64_t max_batch, b2d_ctx **out);
void b2d_ctx_destroy(b2d_ctx *ctx);
/* text of the last error raised on the calling thread */
const char *b2d_last_error(void);
int b2d_device_count(void);

/* Worst-case raw-deflate bytes for n input bytes (stored fallback per 32 KiB
 * sub-block + a sync marker per 128 KiB block + optional final block).
 * Stands in for zng_deflateBound (zlib-ng/deflate.c:661-667). */
uint64_t b2d_deflate_bound(uint64_t n);

/* _ParallelCompress.compress_and_crc(data, zdict) (zlib_ngmodule.c:1696-1782):
 * raw DEFLATE of `data` primed with the last 32 KiB of `dict`, never BFINAL,
 * ends byte-aligned with the sync-flush marker 00 00 ff ff; *crc = crc32(data).
 * B2D_BUF_ERROR if the result does not fit out_cap (-> OverflowError). */
int b2d_compress_and_crc(b2d_ctx *ctx, int level, const uint8_t *data, uint64_t n, const uint8_t *dict,
                         uint64_t dict_n, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc);

/* General host-buffer form used for zlib_ng.compress / compressobj
 * (zlib_compress_impl, zlib_ngmodule.c:199-273; Compress.compress/flush :530-784):
 * same stream as above; `final` != 0 appends the BFINAL empty fixed block 03 00
 * (what gzip_ng_threaded.py:333 appends).  window_bits (9..15) caps match
 * distances at 2^window_bits for the small-wbits forms.  crc / adler may be NULL; adler is the
 * Adler-32 of `data` continued from *adler's incoming value, crc likewise.
 * Every `level` argument of this library is 0..9, optionally with the zlib strategy in bits 8..11
 * (B2D_LEVEL_STRATEGY): 1 Z_FILTERED (matches shorter than 6 are dropped), 2 Z_HUFFMAN_ONLY (no matches,
 * deflate_huff.c), 3 Z_RLE (distance 1 only, deflate_rle.c), 4 Z_FIXED (no dynamic blocks, trees.c:677);
 * what compressobj(strategy=...) passes (zlib_ngmodule.c:376-430). */
#define B2D_LEVEL_STRATEGY(level, strategy) ((level) | ((strategy) << 8))
int b2d_deflate_host(b2d_ctx *ctx, int level, int window_bits, const uint8_t *data, uint64_t n, const uint8_t *dict,
                     uint64_t dict_n, int final, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc, uint32_t *adler);

/* zlib_ng.crc32 / adler32 (zlib_ngmodule.c:1517-1562, 1455-1500) on host bytes */
int b2d_crc32(b2d_ctx *ctx, uint32_t start, const uint8_t *data, uint64_t n, uint32_t *out);
int b2d_adler32(b2d_ctx *ctx, uint32_t start, const uint8_t *data, uint64_t n, uint32_t *out);
/* zlib_ng.crc32_combine (zlib_ngmodule.c:1585-1596 -> crc32_braid_comb.c:17-19); pure host */
uint32_t b2d_crc32_combine(uint32_t crc1, uint32_t crc2, int64_t len2);
/* zng_adler32_combine (zlib-ng/adler32.c:78-100); pure host */
uint32_t b2d_adler32_combine(uint32_t adler1, uint32_t adler2, int64_t len2);

/* ---- device-resident variants (inputs and outputs stay in HBM) --------------
 * d_buf = [dict (dict_n <= 32768 bytes) | data (n bytes)], 16-byte aligned, with
 * B2D_DEVICE_PAD readable bytes after the data.  d_out is 4-byte aligned and
 * out_cap >= b2d_deflate_bound(n).  *kernel_ms (may be NULL) receives the
 * CUDA-event time of the whole launch sequence. */
int b2d_deflate_device(b2d_ctx *ctx, int level, const uint8_t *d_buf, uint64_t dict_n, uint64_t n, uint8_t *d_out,
                       uint64_t out_cap, uint64_t *out_n, uint32_t *crc, float *kernel_ms);
int b2d_crc32_device(b2d_ctx *ctx, uint32_t start, const uint8_t *d_data, uint64_t n, uint32_t *out, float *kernel_ms);
int b2d_adler32_device(b2d_ctx *ctx, uint32_t start, const uint8_t *d_data, uint64_t n, uint32_t *out, float *kernel_ms);

/* Adler-32 of device-resident bytes for callers that may not allocate (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Needs no context and calls no cudaMalloc / cudaFree / allocating routine: the partial sums of the bulk
 * kernel live in a file-scope device array of 24 KiB that every GPU loading the library holds until exit.  At most
 * B2D_NOALLOC_MAX_BYTES (1 GiB) per call: larger sizes are refused with B2D_STREAM_ERROR and a message (b2d_last_error).
 * Synchronous (the calling thread's per-thread stream).  Every call shares the array, which keeps the last call's
 * contents (the entries a call reads are cleared first); calls are serialised inside one process, calls that overlap
 * on the GPU otherwise race.  Same result as b2d_adler32_device, bit for bit. */
#define B2D_NOALLOC_MAX_BYTES (1ull << 30)
uint64_t b2d_noalloc_max_bytes(void);
int b2d_adler32_device_noalloc(int device, uint32_t start, const uint8_t *d_data, uint64_t n, uint32_t *out);

/* ---- memory helpers for callers without their own CUDA binding --------------- */
int b2d_device_alloc(b2d_ctx *ctx, uint64_t bytes, void **ptr);
int b2d_device_free(b2d_ctx *ctx, void *ptr);
int b2d_host_alloc(b2d_ctx *ctx, uint64_t bytes, void **ptr);      /* pinned */
int b2d_host_free(b2d_ctx *ctx, void *ptr);
int b2d_copy_h2d(b2d_ctx *ctx, void *d_dst, const void *h_src, uint64_t bytes);
int b2d_copy_d2h(b2d_ctx *ctx, void *h_dst, const void *d_src, uint64_t bytes);
int b2d_device_sync(b2d_ctx *ctx);
/* page-locked host memory that needs no context (cudaHostAlloc, portable: every device can read and write it), and
 * whether an address lies in page-locked memory — host-buffer calls move such memory by DMA directly, anything else
 * through the context's pinned staging slots */
int b2d_pinned_alloc(uint64_t bytes, void **ptr);
int b2d_pinned_free(void *ptr);
int b2d_host_is_pinned(const void *ptr);
/* memcpy between host buffers split over `threads` native threads: how the writer behind
 * gzip_ng_threaded.open fills its pinned batch from the caller's bytes (the reference copies every block once
 * as well: `data = bytes(b)`, gzip_ng_threaded.py:315).  One thread moves ~14 GB/s on the test box, eight ~70. */
int b2d_host_copy(void *dst, const void *src, uint64_t bytes, int threads);
/* ---- one stream over several GPUs ---------------------------------------------------------
 * The reference runs N compression threads on ONE stream: blocks are dealt round-robin (gzip_ng_threaded.py:316-321),
 * a writer thread emits the results in order and folds their CRCs with crc32_combine (:382-398, :394).  A multi-device
 * context is that with one GPU per worker: the call cuts `data` into batches, batch j runs on device j % n_devices
 * primed with the 32 KiB of input before it (one native thread per device, each with the H2D / kernels / D2H
 * pipeline of b2d_deflate_host), every batch's output lands behind its predecessor's in `out`, and the host folds
 * the checksums.  No device-to-device traffic, no collective.  The bytes produced are those of the single-device
 * call on the same input (blocks are independent given the raw input).
 * `devices` may name a device more than once (several contexts on one GPU): that is how a one-GPU box exercises
 * the sharding. */
typedef struct b2d_mctx b2d_mctx;
int b2d_mctx_create(const int *devices, int n_devices, uint64_t max_batch, b2d_mctx **out);
void b2d_mctx_destroy(b2d_mctx *m);
int b2d_mctx_devices(b2d_mctx *m);
uint64_t b2d_mctx_launch_count(b2d_mctx *m);
/* b2d_deflate_host over all devices of `m`.  block_sizes (optional, `cap` entries) receives the compressed size of
 * every 128 KiB device block, *count their number (what the writer stores as its block index). */
int b2d_deflate_host_multi(b2d_mctx *m, int level, int window_bits, const uint8_t *data, uint64_t n, const uint8_t *dict,
                           uint64_t dict_n, int final, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                           uint32_t *adler, uint64_t *block_sizes, uint64_t cap, uint64_t *count);
/* _ParallelCompress.compress_and_crc (zlib_ngmodule.c:1696-1782) over all devices of `m` */
int b2d_compress_and_crc_multi(b2d_mctx *m, int level, const uint8_t *data, uint64_t n, const uint8_t *dict,
                               uint64_t dict_n, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                               uint64_t *block_sizes, uint64_t cap, uint64_t *count);

/* ... and, with sub_bits (4 entries per block, `cap` counts blocks), the bit offset of every 32 KiB sub-block inside its
 * block (see b2d_subblock_bits): what the writer stores so that readers can decode one warp per sub-block */
int b2d_compress_and_crc_multi_sub(b2d_mctx *m, int level, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                   uint64_t dict_n, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                                   uint64_t *block_sizes, uint32_t *sub_bits, uint64_t cap, uint64_t *count);

/* ---- the other side of the wire: block-parallel inflate of streams written by this library ----
 * (first version.  The reference decompresses with zng_inflate on one thread:
 * zlib_ng.decompress, zlib_ngmodule.c:275-370; _GzipReader.read, :2426-2637; gzip_ng_threaded.py:90-167.)
 *
 * b2d_block_index: where every 128 KiB device block of the LAST deflate call on this context (b2d_deflate_device,
 * b2d_deflate_host, b2d_compress_and_crc) starts
 * in its compressed output: *count = blocks + 1 entries (the last one is the stream length), copied to
 * `offsets` when it is not NULL (B2D_BUF_ERROR if cap < *count).  Blocks start byte-aligned. */
int b2d_block_index(b2d_ctx *ctx, uint64_t *offsets, uint64_t cap, uint64_t *count);
/* b2d_compress_and_crc and b2d_block_index in one call: nothing else runs on the context between the two, so
 * writers that share a context (several files open in one process) get their own block's index */
int b2d_compress_and_crc_indexed(b2d_ctx *ctx, int level, const uint8_t *data, uint64_t n, const uint8_t *dict,
                                 uint64_t dict_n, uint8_t *out, uint64_t out_cap, uint64_t *out_n, uint32_t *crc,
                                 uint64_t *offsets, uint64_t cap, uint64_t *count);
/* Inflate `raw_n` bytes from the raw-DEFLATE stream d_in[0, in_n) given block_offsets[0, nblocks) (a host
 * array; nblocks = ceil(raw_n / 128 KiB)) and the dictionary the stream was primed with (device pointer, at
 * most 32 KiB, may be NULL/0).  Every block is decoded by its own warp; references into the preceding 32 KiB
 * are carried as markers and resolved afterwards (b2d_inflate.cu).  Needs 2 * raw_n bytes of device scratch
 * (owned by the context).  Returns -3 (Z_DATA_ERROR) if a block does not decode to its expected length — and only then:
 * the decode kernel is issued first and waits for nobody, so the result does not depend on kernels running side by side
 * (profilers, CUDA_LAUNCH_BLOCKING). */
int b2d_inflate_device(b2d_ctx *ctx, const uint8_t *d_in, uint64_t in_n, const uint64_t *block_offsets,
                       uint64_t nblocks, const uint8_t *d_dict, uint64_t dict_n, uint8_t *d_out, uint64_t raw_n,
                       float *ms);
/* Four times the parallelism: the writer also knows at which BIT every 32 KiB sub-block (= one DEFLATE block) starts inside
 * its 128 KiB block.  b2d_subblock_bits returns them for the last deflate call on the context (4 entries per block, 0 for
 * the first sub-block and for sub-blocks a short last block does not have); the _sub forms of the inflate calls take them
 * (sub_bits may be NULL: one warp per block as before) and decode one warp per sub-block. */
int b2d_subblock_bits(b2d_ctx *ctx, uint32_t *bits, uint64_t cap, uint64_t *count);
int b2d_inflate_device_sub(b2d_ctx *ctx, const uint8_t *d_in, uint64_t in_n, const uint64_t *block_offsets,
                           const uint32_t *sub_bits, uint64_t nblocks, const uint8_t *d_dict, uint64_t dict_n,
                           uint8_t *d_out, uint64_t raw_n, float *ms);
int b2d_inflate_host_sub(b2d_ctx *ctx, const uint8_t *in, uint64_t in_n, const uint64_t *block_offsets, const uint32_t *sub_bits,
                         uint64_t nblocks, const uint8_t *dict, uint64_t dict_n, uint8_t *out, uint64_t raw_n, uint32_t *crc);
/* b2d_inflate_host_sub in two halves, for readers that work through a file window by window (the reference's reader
 * thread does the same job serially, src/zlib_ng/gzip_ng_threaded.py:98-190): _decode uploads the window's compressed
 * bytes and queues its decode without waiting for anything and WITHOUT the dictionary (dict_n only says how long it will
 * be); _resolve, given the dictionary — the 32 KiB the window before ended with — finishes the window into `out` and
 * returns when the bytes are there.  A reader queues the decode of window w + 1 (on a second context) before it resolves
 * window w, so the device never idles between windows.  One decode may be outstanding per context; out == NULL gives
 * it up. */
int b2d_inflate_host_decode(b2d_ctx *ctx, const uint8_t *in, uint64_t in_n, const uint64_t *block_offsets, const uint32_t *sub_bits,
                            uint64_t nblocks, uint64_t dict_n, uint64_t raw_n);
int b2d_inflate_host_resolve(b2d_ctx *ctx, const uint8_t *dict, uint64_t dict_n, uint8_t *out, uint64_t raw_n, uint32_t *crc);
/* the same with host buffers: chunks of 512 blocks flow through H2D / decode / resolve / D2H side by side.  `crc`
 * (optional) is continued over the output: *crc = crc32(out[0, raw_n), *crc), computed on the device. */
int b2d_inflate_host(b2d_ctx *ctx, const uint8_t *in, uint64_t in_n, const uint64_t *block_offsets, uint64_t nblocks,
                     const uint8_t *dict, uint64_t dict_n, uint8_t *out, uint64_t raw_n, uint32_t *crc);

/* ---- BGZF (blocked gzip: htslib / bgzip; the reference reads such files with its one zng_inflate thread, fixture
 * tests/data/test.fastq.bgzip.gz, tests/test_gzip_ng.py:466-473) ----------------------------------------------------
 * b2d_bgzf_scan walks the members at buf[0, n): gzip members whose "BC" extra subfield holds their total size.  For each
 * of the first *count (<= max_members) complete members it reports where the DEFLATE data lies, the ISIZE and the CRC-32
 * of its trailer; *consumed = offset behind the last complete member (the rest of the buffer is a partial member, or not
 * BGZF).  Pure host code; the arrays may be NULL to count only. */
int b2d_bgzf_scan(const uint8_t *buf, uint64_t n, uint64_t max_members, uint64_t *payload_off, uint32_t *payload_len,
                  uint32_t *raw_len, uint32_t *crc, uint64_t *count, uint64_t *consumed);
/* Inflate independent DEFLATE streams (no dictionary, each at most 64 KiB of output: BGZF members) from a host buffer into
 * `out`, back to back, one warp per member.  crcs (optional): the members' CRC-32s, verified against the output (-3 on
 * mismatch, like a damaged gzip member: zlib_ngmodule.c:2581-2599). */
int b2d_inflate_members_host(b2d_ctx *ctx, const uint8_t *in, uint64_t in_n, const uint64_t *payload_off,
                             const uint32_t *payload_len, const uint32_t *raw_len, const uint32_t *crcs, uint64_t nmembers,
                             uint8_t *out, uint64_t out_cap, uint64_t *out_n);

/* Per-kernel timing of b2d_deflate_device calls (CUDA events on the launching
 * stream, accumulated): stages 0..5 = chains, match, parse, plan, scan, emit.
 * enable = 1 times the kernels as they run in production (consecutive batches of a
 * call overlap on two streams, so a stage's time includes what it shares the SMs
 * with); enable = 2 also keeps every batch on one stream (kernels timed alone). */
#define B2D_NSTAGES 6
int b2d_profile(b2d_ctx *ctx, int enable);
int b2d_profile_read(b2d_ctx *ctx, double *stage_ms, uint64_t *stage_launches, int reset);
/* number of this library's kernel launches since the context was created */
uint64_t b2d_launch_count(b2d_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* B2D_H */
