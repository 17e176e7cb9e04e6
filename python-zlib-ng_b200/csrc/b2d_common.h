/* b2d_common.h — definitions shared by every stage of the B200 block-parallel
 * DEFLATE pipeline.  Pure integer helpers, usable from CUDA device code and
 * (for the CPU twin used ONLY by tests, oracle/b2d_twin.c) from plain C++.
 *
 * Format facts follow RFC 1951 as implemented by the reference's emitter:
 *   block header / types      zlib-ng trees_emit.h:198-207, zutil.h:51-53
 *   length / distance codes   zlib-ng trees_tbl.h:76-132, trees.h:21-25
 *   code-length alphabet      zlib-ng trees.c:412-588, trees.h:26-31
 *   stored / static / dynamic zlib-ng trees.c:633-711
 *   sync-flush marker         zlib-ng deflate.c:951-956, trees.c:593-610
 * The parse / match-finding policy is our own; only the
 * bitstream format is fixed by the reference.
 */
#ifndef B2D_COMMON_H
#define B2D_COMMON_H

#include <stdint.h>

#if defined(__CUDACC__)
#define B2D_HD __host__ __device__ __forceinline__
#define B2D_HDN static __host__ __device__
#else
#define B2D_HD static inline
#define B2D_HDN static
#endif

/* ---- geometry ---------------------------------------------------------- */
#define B2D_BLOCK       131072u   /* device block: input bytes per independent unit      */
#define B2D_WINDOW      32768u    /* dictionary halo = DEFLATE window                     */
#define B2D_SUB         32768u    /* one DEFLATE block per sub-block of input             */
#define B2D_NSUB        (B2D_BLOCK / B2D_SUB)
#define B2D_SPAN        (B2D_BLOCK + B2D_WINDOW)   /* positions a block hashes: itself + its dictionary */
#define B2D_LINKS       (B2D_SPAN + 16u)           /* link entries per block (first up to 7 and last 8+ are alignment slack) */
#define B2D_UNIT        1024u     /* parse unit: no token crosses a multiple of this (so   */
                                  /* every unit is parsed by its own thread); costs <0.2 % */
#define B2D_MIN_MATCH   4u        /* shortest match ever emitted (hash covers 5 bytes)    */
#define B2D_MAX_MATCH   258u
#define B2D_HASH_BITS_MAX 14      /* head table = 2^bits u16 entries; bits is per level (policy) */
#define B2D_HASH_SIZE_MAX (1u << B2D_HASH_BITS_MAX)
#define B2D_LL_CODES    286
#define B2D_D_CODES     30
#define B2D_BL_CODES    19
#define B2D_HDR_WORDS   96        /* packed dynamic-block header, 32-bit words            */
#define B2D_IN_PAD      64u       /* readable slack required after the input bytes        */

/* worst-case compressed bytes of one device block (stored fallback per
 * sub-block: 3 header bits + pad + LEN/NLEN, + sync marker) */
#define B2D_BLOCK_BOUND(len) ((len) + 5u * B2D_NSUB + 8u)
#define B2D_HIST        (B2D_LL_CODES + B2D_D_CODES)   /* 316 counters per sub-block */

/* ---- per-level policy --------------------------------------------------- */
typedef struct {
    int depth;     /* candidates examined per position (hash-chain steps)        */
    int lazy;      /* 0 greedy, 1 one-ahead, 2 two-ahead                         */
    int nice;      /* stop searching once a match this long is found             */
    int hash_bits; /* log2 of the head-table size: a smaller table keeps more blocks   */
                   /* resident per SM in the chain stage, at a little ratio            */
    int level;     /* 0..9                                                        */
    int strategy;  /* Z_DEFAULT_STRATEGY 0, Z_FILTERED 1, Z_HUFFMAN_ONLY 2, Z_RLE 3, Z_FIXED 4 */
    int search;    /* B2D_SEARCH_*: where match candidates come from              */
    int min_take;  /* shortest match the parse accepts (Z_FILTERED: 6, zlib-ng deflate_slow.c "TOO_FAR"/filtered rule) */
    int blocks;    /* B2D_BLOCKS_*: which block types the plan may choose         */
} b2d_policy;
#define B2D_SEARCH_NONE   0   /* literals only: level 0 (stored anyway) and Z_HUFFMAN_ONLY (deflate_huff.c) */
#define B2D_SEARCH_CHAINS 1
#define B2D_SEARCH_RLE    2   /* distance 1 only: Z_RLE (deflate_rle.c)                                      */
#define B2D_BLOCKS_ANY    0
#define B2D_BLOCKS_STORED 1   /* level 0 (deflate_stored.c)                                                  */
#define B2D_BLOCKS_NO_DYN 2   /* Z_FIXED: stored or fixed Huffman only (trees.c:677)                         */
/* the C ABI's `level` argument carries the strategy in bits 8..11 */
#define B2D_LEVEL(level, strategy) ((level) | ((strategy) << 8))
B2D_HD int b2d_level_valid(int x) { return x >= 0 && (x & 0xff) <= 9 && (x >> 8) <= 4; }

#ifndef D_L6
#define D_L5 5
#define D_L6 6
#endif
#ifndef HB_L1
#define HB_L1 12
#define HB_L23 13
#define HB_L456 13
#endif
B2D_HD b2d_policy b2d_level_policy(int level_and_strategy) {
    b2d_policy p;
    const int level = level_and_strategy & 0xff;
    p.level = level;
    p.strategy = (level_and_strategy >> 8) & 0xf;
    switch (level) {
        case 1:  p.depth = 1;  p.lazy = 0; p.nice = 258; p.hash_bits = HB_L1; break;
        case 2:  p.depth = 2;  p.lazy = 0; p.nice = 258; p.hash_bits = HB_L23; break;
        case 3:  p.depth = 3;  p.lazy = 1; p.nice = 258; p.hash_bits = HB_L23; break;
        case 4:  p.depth = 4;  p.lazy = 1; p.nice = 258; p.hash_bits = HB_L456; break;
        case 5:  p.depth = D_L5;  p.lazy = 2; p.nice = 258; p.hash_bits = HB_L456; break;
        case 6:  p.depth = D_L6;  p.lazy = 2; p.nice = 258; p.hash_bits = HB_L456; break;
        case 7:  p.depth = 16; p.lazy = 2; p.nice = 258; p.hash_bits = 14; break;
        case 8:  p.depth = 32; p.lazy = 2; p.nice = 258; p.hash_bits = 14; break;
        default: p.depth = 64; p.lazy = 2; p.nice = 258; p.hash_bits = 14; break; /* 9, and 0 (unused there) */
    }
    p.search = B2D_SEARCH_CHAINS; p.min_take = 4; p.blocks = B2D_BLOCKS_ANY;
    if (p.strategy == 1) p.min_take = 6;
    if (p.strategy == 2) p.search = B2D_SEARCH_NONE;
    if (p.strategy == 3) { p.search = B2D_SEARCH_RLE; p.depth = 1; }
    if (p.strategy == 4) p.blocks = B2D_BLOCKS_NO_DYN;
    if (level == 0) { p.search = B2D_SEARCH_NONE; p.blocks = B2D_BLOCKS_STORED; }
    return p;
}

/* ---- hashing ------------------------------------------------------------ */
/* 5-byte multiplicative hash: w = bytes p..p+3 little-endian, b4 = byte p+4; the top `bits` bits
 * of the product index the head table */
B2D_HD uint32_t b2d_hash5(uint32_t w, uint32_t b4, int bits) {
    uint32_t v = w * 0x9E3779B1u + b4 * 0x85EBCA6Bu;
    v ^= v >> 15;
    v *= 0x2C1B3C6Du;
    return v >> (32 - bits);
}

/* ---- match / token packing ---------------------------------------------- */
/* match record per input position: (len - 3) << 23 | (dist - 1) << 8 | the position's own byte; bits 8.. are 0 where
 * no match was found (len >= 4, so len - 3 is never 0 for a match).  Bits 8-30 are the low bits of the token a taken
 * match becomes, bits 0-7 the token of a literal: the parse reads nothing but these records. */
B2D_HD uint32_t b2d_pack_match(uint32_t len, uint32_t dist, uint32_t lit) { return ((len - 3u) << 23) | ((dist - 1u) << 8) | lit; }
B2D_HD uint32_t b2d_match_len(uint32_t m) { return (m >> 23) ? (m >> 23) + 3u : 0u; }
B2D_HD uint32_t b2d_match_dist(uint32_t m) { return ((m >> 8) & 0x7fffu) + 1u; }
B2D_HD uint32_t b2d_match_lit(uint32_t m) { return m & 0xffu; }

/* token: literal = byte value; match = 1<<31 | (len-3) << 15 | (dist-1) */
B2D_HD uint32_t b2d_tok_match(uint32_t len, uint32_t dist) { return 0x80000000u | ((len - 3u) << 15) | (dist - 1u); }
B2D_HD int      b2d_tok_is_match(uint32_t t) { return (int)(t >> 31); }
B2D_HD uint32_t b2d_tok_len3(uint32_t t) { return (t >> 15) & 0xffu; }   /* len - 3  */
B2D_HD uint32_t b2d_tok_dist1(uint32_t t) { return t & 0x7fffu; }         /* dist - 1 */

B2D_HD int b2d_ilog2(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int)x);
#else
    return 31 - __builtin_clz(x);
#endif
}

/* length symbol (0..28 => codes 257..285), extra-bit count and value for x = len-3 */
B2D_HD void b2d_len_sym(uint32_t x, uint32_t *sym, uint32_t *ebits, uint32_t *eval) {
    if (x < 8u) { *sym = x; *ebits = 0; *eval = 0; }
    else if (x == 255u) { *sym = 28; *ebits = 0; *eval = 0; }
    else {
        uint32_t e = (uint32_t)b2d_ilog2(x) - 2u;
        *sym = 4u + 4u * e + ((x >> e) & 3u);
        *ebits = e; *eval = x & ((1u << e) - 1u);
    }
}
/* distance symbol (0..29), extra-bit count and value for d = dist-1 */
B2D_HD void b2d_dist_sym(uint32_t d, uint32_t *sym, uint32_t *ebits, uint32_t *eval) {
    if (d < 4u) { *sym = d; *ebits = 0; *eval = 0; }
    else {
        uint32_t e = (uint32_t)b2d_ilog2(d) - 1u;
        *sym = 2u * e + 2u + ((d >> e) & 1u);
        *ebits = e; *eval = d & ((1u << e) - 1u);
    }
}
B2D_HD uint32_t b2d_len_extra_bits(uint32_t sym) {   /* sym 0..28 */
    return (sym < 8u || sym == 28u) ? 0u : (sym - 4u) >> 2;
}
B2D_HD uint32_t b2d_dist_extra_bits(uint32_t sym) {  /* sym 0..29 */
    return sym < 4u ? 0u : (sym - 2u) >> 1;
}

/* ---- parse rule ---------------------------------------------------------- */
/* Decide whether the parse, standing at a position whose match length is l0,
 * emits that match (returns l0) or a literal (returns 1). l1/l2 are the match
 * lengths found at the next two positions (0 past the block end). */
B2D_HD uint32_t b2d_parse_step(uint32_t l0, uint32_t l1, uint32_t l2, int lazy) {
    if (l0 == 0u) return 1u;
    if (lazy >= 1 && l1 > l0) return 1u;
    if (lazy >= 2 && l2 > l0 + 1u) return 1u;
    return l0;
}

/* bit reversal of the low `len` bits (canonical codes are sent MSB-first inside
 * an LSB-first stream: zlib-ng trees.c:811-818) */
B2D_HD uint32_t b2d_bitrev(uint32_t code, int len) {
#if defined(__CUDA_ARCH__)
    return len ? __brev(code) >> (32 - len) : 0u;
#else
    uint32_t r = 0;
    for (int i = 0; i < len; i++) { r = (r << 1) | (code & 1u); code >>= 1; }
    return r;
#endif
}

/* ---- per-sub-block plan written by the Huffman stage, read by the emitter -- */
typedef struct {
    uint32_t ll_code[B2D_LL_CODES];   /* (len << 16) | bit-reversed code           */
    uint32_t d_code[B2D_D_CODES];
    uint32_t hdr[B2D_HDR_WORDS];      /* block header bits incl. the 3-bit BTYPE   */
    uint32_t hdr_bits;                /* number of valid bits in hdr               */
    uint32_t type;                    /* 0 stored, 1 fixed, 2 dynamic              */
    uint32_t body_bits;               /* token bits incl. EOB (types 1,2)          */
    uint32_t pad;
} b2d_plan;

/* per-sub-block parse summary written by the parse stage.  Sub-block k of a
 * job covers input bytes [k*B2D_SUB, k*B2D_SUB + in_len) and owns the token
 * slots tok[k*B2D_SUB ...] (a sub-block never has more tokens than bytes). */
typedef struct {
    uint32_t in_len;      /* input bytes covered by this sub-block's tokens      */
    uint32_t tok_count;
    uint32_t bit_off;     /* bit offset of the sub-block inside its device block (scan stage) */
    uint32_t pad;
} b2d_subinfo;

#endif /* B2D_COMMON_H */
