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
 * the decode is complete).  st_tails / ev_fork / ev_join: a second stream and two events, so that the serial pass
 * runs beside the decode. */
#define B2D_INF_CONCURRENT   0   /* decode first, the tails pass behind it on st_tails */
#define B2D_INF_SERIAL       1   /* decode, tails, finish on one stream */
#define B2D_INF_RESOLVE_ONLY 2   /* tails + finish again over symbols that are already decoded */
cudaError_t b2d_launch_inflate(const uint8_t *d_in, uint64_t in_n, const uint64_t *d_index, uint32_t nblocks,
                               const uint8_t *d_dict, uint32_t dict_n, uint16_t *d_out16, uint8_t *d_out,
                               uint64_t raw_n, uint32_t *d_flags, uint32_t *d_err, cudaStream_t st, cudaStream_t st_tails,
                               cudaEvent_t ev_fork, cudaEvent_t ev_join, int mode);
#endif
