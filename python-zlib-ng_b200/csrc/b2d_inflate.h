/* b2d_inflate.h — launcher of the block-parallel inflate stages (b2d_inflate.cu). */
#ifndef B2D_INFLATE_H
#define B2D_INFLATE_H
#include <cuda_runtime.h>
#include <stdint.h>

#define B2D_INF_LAUNCHES 3   /* k_inf_decode, k_inf_tails, k_inf_finish */

/* d_index[b] = byte offset of device block b in the compressed stream; d_out16 = 2 * raw_n bytes of scratch;
 * d_flags = nblocks words of scratch; d_err = two words zeroed by the caller: d_err[0] receives
 * (reason << 28) | block of the first block that failed to decode (bad data), d_err[1] becomes non-zero if the
 * pass over the tails gave up waiting for the decode (never bad data: the caller runs B2D_INF_RESOLVE_ONLY once
 * the decode is complete). */
#define B2D_INF_CONCURRENT   0   /* decode first, the tails pass behind it on its own stream */
#define B2D_INF_SERIAL       1   /* decode, tails, finish on one stream */
#define B2D_INF_RESOLVE_ONLY 2   /* tails + finish again over symbols that are already decoded */
#define B2D_INF_DECODE_ONLY  3   /* k_inf_decode alone: the caller resolves later (B2D_INF_RESOLVE_ONLY), e.g. once the dictionary is known */
/* One chunk: blocks [b0, b0 + nb) of a stream of nblocks, chunks in order.  The decode is issued on st_dec, the pass over
 * the tails and the rest of the resolving on st_res.  The caller makes both streams wait until d_flags and d_err are
 * cleared; st_res == st_dec for B2D_INF_SERIAL and B2D_INF_RESOLVE_ONLY. */
/* d_sub_bits (optional): bit offset of every 32 KiB sub-block inside its block, 4 entries per block — one warp per
 * sub-block instead of one per block */
cudaError_t b2d_launch_inflate(const uint8_t *d_in, uint64_t in_n, const uint64_t *d_index, const uint32_t *d_sub_bits, uint32_t b0, uint32_t nb,
                               uint32_t nblocks, const uint8_t *d_dict, uint32_t dict_n, uint16_t *d_out16, uint8_t *d_out,
                               uint64_t raw_n, uint32_t *d_flags, uint32_t *d_err, cudaStream_t st_dec, cudaStream_t st_res, int mode);
/* independent members (BGZF): d_index[b] = where member b's DEFLATE data starts, d_out_index[b] .. d_out_index[b + 1] = its
 * place in the output; bytes are written directly, *d_err as above */
cudaError_t b2d_launch_inflate_members(const uint8_t *d_in, uint64_t in_n, const uint64_t *d_index, const uint64_t *d_out_index,
                                       uint32_t nblocks, uint8_t *d_out, uint64_t raw_n, uint32_t *d_err, cudaStream_t st);
#endif
