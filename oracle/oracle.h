/* oracle/oracle.h — TEST INFRASTRUCTURE ONLY (see oracle.c). */
#ifndef B2D_ORACLE_H
#define B2D_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
uint32_t orc_crc32(uint32_t start, const uint8_t *buf, size_t len);
uint32_t orc_crc32_combine(uint32_t crc1, uint32_t crc2, int64_t len2);
uint32_t orc_adler32(uint32_t start, const uint8_t *buf, size_t len);
uint32_t orc_adler32_combine(uint32_t adler1, uint32_t adler2, int64_t len2);
/* Decode one raw DEFLATE stream (RFC 1951), optionally primed with a dictionary.
 * stop_at_final = 0 lets a stream that ends after a sync-flush marker without a
 * BFINAL block count as complete when the input is exhausted on a block
 * boundary.  Returns 0 ok, <0 error; *out_len / *consumed receive the sizes. */
int orc_inflate_raw(const uint8_t *in, size_t in_len, const uint8_t *dict, size_t dict_len,
                    uint8_t *out, size_t out_cap, size_t *out_len, size_t *consumed,
                    int stop_at_final);
#ifdef __cplusplus
}
#endif
#endif
