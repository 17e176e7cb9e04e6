/* b2d_kernels.cu — the sm_100a kernels of the block-parallel DEFLATE pipeline.
 *
 * One job = a contiguous device buffer [dict (D bytes) | data (n bytes)].  It is
 * cut into device blocks of B2D_BLOCK input bytes (each ends byte-aligned with a
 * sync-flush marker, so blocks concatenate with plain byte copies, exactly like
 * the outputs of the reference's compress_and_crc, zlib_ngmodule.c:1696-1782)
 * and each block into B2D_NSUB DEFLATE blocks ("sub-blocks") of B2D_SUB bytes.
 *
 * Stages (each a kernel; DESIGN.md §3 has the per-stage byte/roofline figures):
 *   K1 k_chains  one warp per device block (+32 KiB halo): 5-byte hash, 2^14-entry
 *                head table in shared memory, warp match.any to order the 32
 *                positions of a step; writes prevd[p] = distance to the previous
 *                position with the same hash.        (replaces insert_string /
 *                deflateSetDictionary priming, insert_string_tpl.h:50-107)
 *   K2 k_match   one CTA per sub-block: the 32 KiB window + 32 KiB tile and their
 *                chain links staged in shared memory (192 KiB), one thread per
 *                position walks `depth` links, 4-byte compares with funnel shifts.
 *                                                   (replaces longest_match,
 *                match_tpl.h:31-285, compare256.c)
 *   K3 k_parse   one CTA per sub-block: one lane per 1 KiB parse unit chases the
 *                lazy parse through match lengths in shared memory, then all
 *                threads compact the reached positions into tokens and
 *                histograms.                        (replaces deflate_quick /
 *                deflate_medium's decision loop and zng_tr_tally, deflate_p.h:61-86)
 *   K4 k_plan    one thread per sub-block: length-limited Huffman, header,
 *                stored/fixed/dynamic choice.       (trees.c:186-552,633-711)
 *   K5 k_scan    bit offsets of sub-blocks, byte offsets of device blocks.
 *   K6 k_emit    one CTA per sub-block: per-thread token runs, block scan of bit
 *                lengths, 64-bit accumulators -> global words; sync marker.
 *                                                   (compress_block, send_bits,
 *                zng_tr_stored_block: trees.c:593-740, trees_emit.h:42-225)
 */
#include <cuda_runtime.h>
#include <stdint.h>

#include "b2d_common.h"
#include "b2d_huff.h"
#include "b2d_kernels.h"

#define FULL 0xffffffffu

/* ------------------------------------------------------------------------- */
/* K1: hash chains                                                            */
/* ------------------------------------------------------------------------- */
#define K_NOLINK 0xffffu   /* prevd value of a position without a predecessor: beyond any max_dist */
#define K1_BATCH 512u   /* bytes staged per step group (16 steps of 32 positions) */

template <int HB>
__global__ void __launch_bounds__(32) k_chains(const uint8_t *__restrict__ buf, uint64_t D, uint64_t n,
                                               uint64_t buf_readable, uint16_t *__restrict__ prevd) {
    extern __shared__ __align__(16) uint8_t smem1[];
    uint16_t *head = reinterpret_cast<uint16_t *>(smem1);                         /* 2^HB entries */
    /* 2 x 512 B ring + 16 B guard: a copy of the ring's first bytes, so that "the word after" never wraps */
    uint32_t *stage = reinterpret_cast<uint32_t *>(smem1 + (2u << HB));
    const uint32_t lane = threadIdx.x;
    const uint64_t b = blockIdx.x;
    const uint64_t job_end = D + n;
    const uint64_t start = D + b * (uint64_t)B2D_BLOCK;
    const uint64_t end = start + B2D_BLOCK < job_end ? start + B2D_BLOCK : job_end;
    const uint64_t halo = start < B2D_WINDOW ? start : B2D_WINDOW;
    const uint64_t span_start = start - halo;

    {   /* head[] = 0xffff */
        uint4 ff = make_uint4(FULL, FULL, FULL, FULL);
        uint4 *h4 = reinterpret_cast<uint4 *>(head);
        for (uint32_t i = lane; i < (2u << HB) / 16u; i += 32) h4[i] = ff;
    }
    const uint64_t c0 = span_start & ~(uint64_t)(K1_BATCH - 1);
    const uint64_t rd_lim = buf_readable & ~(uint64_t)15;
    auto load16 = [&](uint64_t c) -> uint4 {
        uint64_t off = c + 16u * lane;
        if (off + 16 <= rd_lim) return __ldg(reinterpret_cast<const uint4 *>(buf + off));
        return make_uint4(0, 0, 0, 0);
    };
    uint4 *stage4 = reinterpret_cast<uint4 *>(stage);
    /* the batch at byte offset c lives in ring slot (c / 512) & 1 */
    auto put_batch = [&](uint32_t slot, const uint4 &v) {
        stage4[slot * 32u + lane] = v;
        if (slot == 0 && lane == 0) stage4[2 * K1_BATCH / 16] = v;           /* the guard */
    };
    /* positions are handled as 32-bit offsets r from span_start; a batch that starts before the span
     * (c0 is rounded down) gives wrapped-around offsets that fail every range test */
    const uint32_t span_len = (uint32_t)(end - span_start);
    /* r < act_len: position can be hashed (its 5 bytes end inside the block; nothing after the block is
     * ever looked at, so a block's output does not depend on what follows it) */
    const uint32_t act_len = span_len >= 5u ? span_len - 4u : 0u;
    /* every block keeps the links of its whole span (dictionary included) in its own array: a block's
     * output then depends on nothing but its 32 KiB of history and its data, wherever batches are cut */
    uint16_t *pd = prevd + b * (uint64_t)B2D_LINKS + (span_start & 7u);
    const uint32_t lt_mask = (1u << lane) - 1u;
    /* a lane's positions sit at the same byte of their words in every step (steps are 32 positions apart) */
    const uint32_t sh = (lane & 3u) << 3, b4sel = 0x4440u | (lane & 3u);
    auto hash_of = [&](const uint32_t *w) -> uint32_t {     /* w: the word holding the position's first byte */
        const uint32_t w0 = w[0], w1 = w[1];
        return b2d_hash5(__funnelshift_r(w0, w1, sh), __byte_perm(w1, 0u, b4sel), HB);
    };
    uint32_t slot = (uint32_t)(c0 >> 9) & 1u;
    put_batch(slot, load16(c0));
    uint4 nxt = load16(c0 + K1_BATCH);
    __syncwarp();
    uint32_t h_next = hash_of(stage + slot * (K1_BATCH / 4) + (lane >> 2));
    for (uint64_t c = c0; c < end; c += K1_BATCH, slot ^= 1u) {
        put_batch(slot ^ 1u, nxt);
        nxt = load16(c + 2 * K1_BATCH);
        __syncwarp();
        const uint32_t rb = (uint32_t)(c - span_start) + lane;             /* this lane's offset in the batch's step 0 */
        uint16_t *pb = pd + (int64_t)(int32_t)rb;
        const uint32_t *sw = stage + slot * (K1_BATCH / 4) + (lane >> 2);  /* its word in step 0 */
        const uint32_t *so = stage + (slot ^ 1u) * (K1_BATCH / 4) + (lane >> 2);   /* and in the next batch's step 0 */
#pragma unroll 1
        for (uint32_t half = 0; half < 2; ++half) {
            const uint32_t *sn = half ? so : sw + 64;        /* the step after this half's last one */
            const uint32_t rh = rb + half * (K1_BATCH / 2);
            uint16_t *ph = pb + half * (K1_BATCH / 2);
            const uint32_t *swh = sw + half * 64;
#pragma unroll
            for (uint32_t k = 0; k < K1_BATCH / 64; ++k) {
                /* the next step's hash does not depend on the table: computed here so that its shared loads and
                 * multiplies overlap this step's read / write / read-back chain */
                const uint32_t h = h_next;
                h_next = hash_of(k + 1 < K1_BATCH / 64 ? swh + 8 * (k + 1) : sn);
                const uint32_t r = rh + 32u * k;
                const bool active = r < act_len;
                /* Ordered insert of the step's 32 positions.  Fast path: read the old head, every lane
                 * writes its own position, read back — if every lane still sees its own value no two
                 * lanes share a bucket and `old` is each lane's predecessor.  (Inactive lanes: old = back = r,
                 * i.e. distance 0 = no link, and not lost.) */
                uint32_t old = r, back = r;
                if (active) old = head[h];
                __syncwarp();
                if (active) head[h] = (uint16_t)r;
                __syncwarp();
                if (active) back = head[h];
                const bool lost = ((back ^ r) & 0xffffu) != 0u;
                const uint32_t dd = (r - old) & 0xffffu;
                /* a distance beyond the window is stored as it is: every reader stops at the first link that takes
                 * it past its max_dist (K_NOLINK itself is just the largest such value) */
                uint32_t d = (dd != 0u) & (dd <= r) ? dd : K_NOLINK;
                if (__any_sync(FULL, lost)) {
                    /* Some bucket was hit twice (runs, short periods, small alphabets).  The losers write
                     * again, so the first winner of such a bucket sees a foreign value too: now every lane
                     * of a shared bucket knows it.  match.any costs ~12 cycles per distinct key, so only
                     * those lanes bring their bucket; everyone else shares one dummy key. */
                    const uint32_t r16 = r & 0xffffu;
                    if (lost) head[h] = (uint16_t)r16;
                    __syncwarp();
                    const bool shared = active && (lost || head[h] != r16);
                    const uint32_t m = __match_any_sync(FULL, shared ? h : 0x10000u);
                    if (shared) {
                        const uint32_t lower = m & lt_mask;
                        if (lower) d = lane - (31u - (uint32_t)__clz((int)lower));
                    }
                    __syncwarp();
                    if (shared && (m >> lane) == 1u) head[h] = (uint16_t)r16;   /* highest lane of the bucket is the head */
                    __syncwarp();
                }
                if (r < span_len) ph[32u * k] = (uint16_t)d;
            }
        }
    }
}

/* ------------------------------------------------------------------------- */
/* K2: all-position match search                                              */
/* ------------------------------------------------------------------------- */
#define K2_THREADS 1024
#ifndef K2_TILE
#define K2_TILE B2D_SUB     /* positions per k_match CTA (a divisor of B2D_BLOCK, a multiple of B2D_UNIT) */
#endif
#define K2_IN_BYTES (B2D_WINDOW + B2D_SUB + 16u + 288u)   /* +16 align slack, +258+pad look-ahead (k_match_d1: whole sub-blocks) */
#define K2_IN_BYTES_T (B2D_WINDOW + K2_TILE + 16u + 288u)  /* k_match */

/* shared-memory loads by 32-bit shared address: candidate addresses are `own address - distance`, one subtract and one
 * mask, without the index arithmetic a C++ array expression costs per candidate */
__device__ __forceinline__ uint32_t lds32(uint32_t a) { uint32_t v; asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds32_4(uint32_t a) { uint32_t v; asm("ld.shared.u32 %0, [%1+4];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds32_8(uint32_t a) { uint32_t v; asm("ld.shared.u32 %0, [%1+8];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds32_12(uint32_t a) { uint32_t v; asm("ld.shared.u32 %0, [%1+12];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds32_16(uint32_t a) { uint32_t v; asm("ld.shared.u32 %0, [%1+16];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds16(uint32_t a) { uint32_t v; asm("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds8(uint32_t a) { uint32_t v; asm("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
/* whole equal bytes at the low end of a xor of two little-endian words (4 for x == 0): bit-reverse + count-leading-zeros */
__device__ __forceinline__ uint32_t eq_bytes(uint32_t x) {
    uint32_t r;
    asm("{\n\t.reg .b32 t;\n\tbrev.b32 t, %1;\n\tclz.b32 %0, t;\n\t}" : "=r"(r) : "r"(x));
    return r >> 3;
}

/* 31 - (index of the lowest set bit), as bfind of the reversed word gives it; 0xffffffff for 0 */
__device__ __forceinline__ uint32_t low_bit_from_top(uint32_t x) {
    uint32_t r;
    asm("{\n\t.reg .b32 t;\n\tbrev.b32 t, %1;\n\tbfind.u32 %0, t;\n\t}" : "=r"(r) : "r"(x));
    return r;
}
__device__ __forceinline__ uint32_t mad_u32(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

/* Best match of the position at shared byte address `pa` (= address of s_in + its byte index; `la` = address of its
 * link), at most `maxl` bytes, `depth` candidates (DEPTH > 0: that many, as straight-line code without a step counter).
 * Two codings of the same search, same result:
 *   DENSE = false  quick reject on the byte at index `best` before a candidate is compared; cheapest where hardly any
 *                  candidate is a match (incompressible data: the whole warp skips the compare code);
 *   DENSE = true   the kernel is issue-bound, and on compressible data some lane of a warp needs the compare in nine
 *                  steps of ten: every lane compares the candidate's first eight bytes straight-line, without the
 *                  reject branch.
 * The walk's state is the candidate's byte address `ca` and its link's address `lc` (each link moves both down), and
 * the best match so far is one key, length << 18 | candidate address: a maximum over keys keeps the longest match and,
 * among equals, the nearest one — the one found first, as distances grow along a chain.  Shared addresses are below
 * 2^18.  The shared window starts 4-byte aligned, so a byte address shifted left by 3 is, in its low five bits, the
 * funnel shift that lines the bytes at that address up (shf wraps the amount). */
#define K2_KEY_SHIFT 18
template <bool DENSE, bool FULLWIN, int DEPTH>
__device__ __forceinline__ uint32_t k2_search(uint32_t pa, uint32_t la, uint32_t maxl, int depth, uint32_t max_dist_arg) {
    /* FULLWIN: the 32 KiB window of every caller but the small-wbits forms — a constant in the loop's distance test */
    const uint32_t max_dist = FULLWIN ? B2D_WINDOW : max_dist_arg;
    /* links are distances to the previous position of the same bucket; K_NOLINK (no predecessor)
     * pushes the running distance past max_dist, so one compare ends the walk either way */
    const uint32_t d0 = lds16(la);
    uint32_t key = (3u << K2_KEY_SHIFT) | ((1u << K2_KEY_SHIFT) - 1u);     /* beats every match of three bytes */
    if (maxl >= B2D_MIN_MATCH && d0 <= max_dist) {
        /* the position's first 16 bytes stay in registers: most compares end inside them */
        const uint32_t pw = pa & ~3u, osh = pa << 3;
        uint32_t own[4];
        {
            const uint32_t a0 = lds32(pw), a1 = lds32_4(pw), a2 = lds32_8(pw), a3 = lds32_12(pw), a4 = lds32_16(pw);
            own[0] = __funnelshift_r(a0, a1, osh); own[1] = __funnelshift_r(a1, a2, osh);
            own[2] = __funnelshift_r(a2, a3, osh); own[3] = __funnelshift_r(a3, a4, osh);
        }
        uint32_t own_b = own[0] >> 24;                  /* byte at index `best`: a longer match must repeat it */
        uint32_t boff = 3;                              /* that index (DENSE = false only) */
        uint32_t ca = pa - d0, lc = la - 2u * d0;
        const uint32_t ca_min = pa - max_dist;          /* the window's far end: a shared address, or "negative" (huge) */
        const uint32_t maxl_key = maxl << K2_KEY_SHIFT;
        /* one candidate; true = the walk is over (nothing can beat a match of maxl bytes) */
        auto visit = [&](uint32_t &nd) -> bool {
            nd = lds16(lc);                             /* next link: independent of the compare below */
            if (DENSE || lds8(ca + boff) == own_b) {
                const uint32_t cw = ca & ~3u, csh = ca << 3;
                const uint32_t a0 = lds32(cw), a1 = lds32_4(cw), a2 = lds32_8(cw);
                const uint32_t x0 = __funnelshift_r(a0, a1, csh) ^ own[0];
                const uint32_t x1 = __funnelshift_r(a1, a2, csh) ^ own[1];
                uint32_t x = x0 ? x0 : x1, l = x0 ? 0u : 4u;      /* most compares end in the first 8 bytes */
                if ((x0 | x1) == 0u) {
                    const uint32_t a3 = lds32_12(cw), a4 = lds32_16(cw);
                    if ((x = __funnelshift_r(a2, a3, csh) ^ own[2]) != 0) l = 8;
                    else if ((x = __funnelshift_r(a3, a4, csh) ^ own[3]) != 0) l = 12;
                    else {
                        /* long match: 8 bytes per trip, each aligned word of either side loaded once */
                        l = 16;
                        uint32_t cq = cw + 16u, oq = pw + 16u, cp = a4, op = lds32(oq);
                        while (l < maxl) {
                            const uint32_t c1 = lds32_4(cq), c2 = lds32_8(cq), o1 = lds32_4(oq), o2 = lds32_8(oq);
                            if ((x = __funnelshift_r(cp, c1, csh) ^ __funnelshift_r(op, o1, osh)) != 0) break;
                            l += 4;
                            if ((x = __funnelshift_r(c1, c2, csh) ^ __funnelshift_r(o1, o2, osh)) != 0) break;
                            l += 4; cq += 8u; oq += 8u; cp = c2; op = o2;
                        }
                    }
                }
                /* the key, put together with multiply-adds (the kernel is bound by the ALU pipe, the FMA pipe idles):
                 * 8 * (equal low bytes of x) = ~(index of its lowest set bit, from the top) & 0x18.  x == 0 only when the
                 * long compare ran into maxl, where the extra bytes do not matter.  The length may overshoot maxl (by less
                 * than 12); such a candidate ends the walk, and the length is clamped once, at the end. */
                const uint32_t k = mad_u32(~low_bit_from_top(x) & 0x18u, 1u << (K2_KEY_SHIFT - 3), mad_u32(l, 1u << K2_KEY_SHIFT, ca));
                if (DENSE) {
                    key = k > key ? k : key;
                } else if (k > key) {
                    key = k;
                    boff = k >> K2_KEY_SHIFT;
                    own_b = lds8(pa + boff);    /* >= maxl: a byte past the match, never compared (the walk ends) */
                }
                /* nice_match is 258 at every level (b2d_level_policy; the launcher refuses anything else) and
                 * maxl <= 258: "l >= nice || l >= maxl" is "l >= maxl".  (A second candidate that long is never reached.) */
                return k >= maxl_key;
            }
            return false;
        };
        if (DEPTH > 0) {
#pragma unroll
            for (int i = 0; i < DEPTH; ++i) {
                uint32_t nd;
                const bool over = visit(nd);
                ca -= nd; lc = mad_u32(nd, 0xfffffffeu, lc);      /* lc - 2 * nd, on the FMA pipe */
                /* signed: near the start of the buffer both may be below zero (addresses are < 2^18, distances < 2^17) */
                if (i + 1 < DEPTH && (over | ((int32_t)ca < (int32_t)ca_min))) break;
            }
        } else {
            int step = depth;
            for (;;) {
                uint32_t nd;
                const bool over = visit(nd);
                ca -= nd; lc = mad_u32(nd, 0xfffffffeu, lc);      /* lc - 2 * nd, on the FMA pipe */
                if (over | (--step <= 0) | ((int32_t)ca < (int32_t)ca_min)) break;
            }
        }
    }
    const uint32_t best = key >> K2_KEY_SHIFT;
    /* without the position's own byte (the caller ORs it in) */
    return best > 3u ? b2d_pack_match(best < maxl ? best : maxl, pa - (key & ((1u << K2_KEY_SHIFT) - 1u)), 0u) : 0u;
}

#ifndef K2_DENSE_HITS
#define K2_DENSE_HITS (K2_THREADS / 8)   /* matches among a tile's first 1024 positions that select the dense coding */
#endif

/* DEPTH > 0: the policy's chain depth as a compile-time constant (levels 4-6); 0: the `depth` argument.  Depth 1
 * (level 1, Z_RLE) goes to k_match_d1. */
template <bool FULLWIN, int DEPTH>
__global__ void __launch_bounds__(K2_THREADS, 1)
k_match(const uint8_t *__restrict__ buf, uint64_t D, uint64_t n, uint64_t buf_readable,
        const uint16_t *__restrict__ prevd, uint32_t *__restrict__ match, int depth, uint32_t max_dist) {
    extern __shared__ __align__(16) uint8_t smem2[];
    uint32_t *s_in = reinterpret_cast<uint32_t *>(smem2);                       /* K2_IN_BYTES */
    uint16_t *s_prev = reinterpret_cast<uint16_t *>(smem2 + K2_IN_BYTES_T);       /* 2*(W+SUB)+16 */
    const uint32_t tid = threadIdx.x;
    const uint64_t k = blockIdx.x;
    const uint64_t job_end = D + n;
    const uint64_t tile_start = D + k * (uint64_t)K2_TILE;
    const uint64_t tile_end = tile_start + K2_TILE < job_end ? tile_start + K2_TILE : job_end;
    const uint64_t ws = tile_start >= B2D_WINDOW ? tile_start - B2D_WINDOW : 0;
    const uint64_t ws_al = ws & ~(uint64_t)15;           /* smem byte 0 of s_in   */
    const uint64_t wp_al = ws & ~(uint64_t)7;            /* element 0 of s_prev   */
    const uint64_t rd_lim = buf_readable & ~(uint64_t)15;
    const uint32_t poff = (uint32_t)(wp_al - ws_al);     /* 0 or 8 */
    {
        uint4 *d4 = reinterpret_cast<uint4 *>(s_in);
        const uint32_t n16 = (uint32_t)((tile_end + 272 - ws_al + 15) >> 4);   /* <= K2_IN_BYTES/16 */
        for (uint32_t i = tid; i < n16; i += K2_THREADS) {
            uint64_t off = ws_al + 16ull * i;
            d4[i] = (off + 16 <= rd_lim) ? __ldg(reinterpret_cast<const uint4 *>(buf + off)) : make_uint4(0, 0, 0, 0);
        }
        uint4 *p4 = reinterpret_cast<uint4 *>(s_prev);
        const uint32_t np = (uint32_t)((tile_end - wp_al + 7) >> 3);
        /* links of this tile's block, entry 0 = position (span_start & ~7) */
        const uint64_t blk = k / (B2D_BLOCK / K2_TILE);
        const uint64_t bstart = D + blk * (uint64_t)B2D_BLOCK;
        const uint64_t span_al = (bstart - (bstart < B2D_WINDOW ? bstart : B2D_WINDOW)) & ~(uint64_t)7;
        const uint4 *src = reinterpret_cast<const uint4 *>(prevd + blk * (uint64_t)B2D_LINKS + (wp_al - span_al));
        for (uint32_t i = tid; i < np; i += K2_THREADS) p4[i] = __ldg(src + i);
    }
    __syncthreads();
    const uint32_t tile_len = (uint32_t)(tile_end - tile_start);
    const uint32_t tile_off = (uint32_t)(tile_start - ws_al);
    /* this thread's positions are tid, tid + 1024, ...: K2_THREADS == B2D_UNIT, so all of them sit at the same offset
     * of their parse units and no match crosses a unit (tile starts are unit-aligned relative to the data) */
    static_assert(K2_THREADS == B2D_UNIT, "one thread per offset of a parse unit");
    const uint32_t to_unit = B2D_UNIT - tid < B2D_MAX_MATCH ? B2D_UNIT - tid : B2D_MAX_MATCH;
    /* the loop runs on the position's shared byte address; its link's address and its slot in `match` follow from it */
    const uint32_t pa0 = (uint32_t)__cvta_generic_to_shared(s_in) + tile_off;                     /* the tile's first position */
    const uint32_t la_off = (uint32_t)__cvta_generic_to_shared(s_prev) - 2u * poff
                            - 2u * (uint32_t)__cvta_generic_to_shared(s_in);                      /* link address = 2 * pa + this */
    const uint32_t pa_end = pa0 + tile_len;
    /* the position's slot in `match` is at mb + 4 * pa (one multiply-add per position) */
    const uint64_t mb = (uint64_t)(uintptr_t)(match + (tile_start - D)) - 4ull * pa0;
    auto put = [&](uint32_t p, uint32_t m) { asm volatile("st.global.u32 [%0], %1;" :: "l"(mb + 4ull * p), "r"(m) : "memory"); };
    uint32_t pa = pa0 + tid;
    auto max_len = [&](uint32_t p) -> uint32_t { const uint32_t to_end = pa_end - p; return to_unit < to_end ? to_unit : to_end; };
    /* the tile's first 1024 positions are searched with the quick-reject coding and tell which coding suits the rest */
    uint32_t hit = 0;
    if (pa < pa_end) {
        const uint32_t m = k2_search<false, FULLWIN, DEPTH>(pa, 2u * pa + la_off, max_len(pa), depth, max_dist);
        put(pa, m | lds8(pa));
        hit = m != 0u;
    }
    const bool dense = __syncthreads_count((int)hit) >= (int)K2_DENSE_HITS;
    pa += K2_THREADS;
    if (dense) {
        for (; pa < pa_end; pa += K2_THREADS)
            put(pa, k2_search<true, FULLWIN, DEPTH>(pa, 2u * pa + la_off, max_len(pa), depth, max_dist) | lds8(pa));
    } else {
        for (; pa < pa_end; pa += K2_THREADS)
            put(pa, k2_search<false, FULLWIN, DEPTH>(pa, 2u * pa + la_off, max_len(pa), depth, max_dist) | lds8(pa));
    }
}

/* K2 at depth 1 (level 1, Z_RLE): a position has one candidate, so nothing walks the links and they need not be staged —
 * each thread reads its own links from global memory.  Only the input window sits in shared memory (65 KiB instead of
 * 197), three CTAs share an SM and one tile's loads overlap the other's compares.  A thread takes the four positions of
 * one aligned shared-memory word at a time: their own bytes are three word loads for all four, lined up with constant
 * shifts.  Same results as k_match with depth == 1. */
#define K2D1_THREADS 512
template <bool RLE>
__global__ void __launch_bounds__(K2D1_THREADS, 3)
k_match_d1(const uint8_t *__restrict__ buf, uint64_t D, uint64_t n, uint64_t buf_readable,
           const uint16_t *__restrict__ prevd, uint32_t *__restrict__ match, uint32_t max_dist) {
    extern __shared__ __align__(16) uint8_t smem2[];
    uint32_t *s_in = reinterpret_cast<uint32_t *>(smem2);                       /* K2_IN_BYTES */
    const uint32_t tid = threadIdx.x;
    const uint64_t k = blockIdx.x;
    const uint64_t job_end = D + n;
    const uint64_t tile_start = D + k * (uint64_t)B2D_SUB;
    const uint64_t tile_end = tile_start + B2D_SUB < job_end ? tile_start + B2D_SUB : job_end;
    const uint64_t ws = tile_start >= B2D_WINDOW ? tile_start - B2D_WINDOW : 0;
    const uint64_t ws_al = ws & ~(uint64_t)15;
    const uint64_t rd_lim = buf_readable & ~(uint64_t)15;
    {
        uint4 *d4 = reinterpret_cast<uint4 *>(s_in);
        const uint32_t n16 = (uint32_t)((tile_end + 272 - ws_al + 15) >> 4);
        for (uint32_t i = tid; i < n16; i += K2D1_THREADS) {
            uint64_t off = ws_al + 16ull * i;
            d4[i] = (off + 16 <= rd_lim) ? __ldg(reinterpret_cast<const uint4 *>(buf + off)) : make_uint4(0, 0, 0, 0);
        }
    }
    /* links of this tile's block: entry 0 = position (span_start & ~7) */
    const uint64_t blk = k / B2D_NSUB;
    const uint64_t bstart = D + blk * (uint64_t)B2D_BLOCK;
    const uint64_t span_al = (bstart - (bstart < B2D_WINDOW ? bstart : B2D_WINDOW)) & ~(uint64_t)7;
    const uint16_t *links = prevd + blk * (uint64_t)B2D_LINKS + (tile_start - span_al);
    __syncthreads();
    const uint32_t tile_len = (uint32_t)(tile_end - tile_start);
    const uint32_t tile_off = (uint32_t)(tile_start - ws_al);
    uint32_t *mout = match + (tile_start - D);
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(s_in);
    /* groups = aligned words of the shared window; the first and the last one may reach outside the tile */
    const uint32_t g_end = (tile_off + tile_len + 3u) >> 2;
    /* match slots of a group are 16-byte aligned when the data starts a multiple of 4 bytes after the window's word 0
     * (always, unless a dictionary of odd length comes first) */
    const bool vec_out = (((tile_start - D) - tile_off) & 3u) == 0u;
    for (uint32_t G = (tile_off >> 2) + tid; G < g_end; G += K2D1_THREADS) {
        const uint32_t wa = sbase + 4u * G;
        const uint32_t t0 = 4u * G - tile_off;              /* tile position of the word's byte 0 ("negative" before the tile) */
        const uint16_t *lk = links + (int64_t)(int32_t)t0;
        uint32_t *mo = mout + (int64_t)(int32_t)t0;
        /* all four positions inside the tile (t0 wraps to a huge value before it): their links are one aligned 8-byte
         * load (the window and the link array start at multiples of 4 positions of each other), their results one store */
        const bool inner = t0 + 3u < tile_len && t0 < tile_len;
        uint32_t dist[4], mres[4];
        if (RLE) {
            /* Z_RLE: the only candidate is the previous byte (deflate_rle.c) */
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j) dist[j] = 4u * G + j >= 1u ? 1u : K_NOLINK;
        } else if (inner) {
            const uint2 lv = __ldg(reinterpret_cast<const uint2 *>(lk));
            dist[0] = lv.x & 0xffffu; dist[1] = lv.x >> 16; dist[2] = lv.y & 0xffffu; dist[3] = lv.y >> 16;
        } else {
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j) dist[j] = t0 + j < tile_len ? (uint32_t)__ldg(lk + j) : K_NOLINK;
        }
        const uint32_t w0 = lds32(wa), w1 = lds32_4(wa), w2 = lds32_8(wa);
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j) {
            const uint32_t t = t0 + j;
            mres[j] = (w0 >> (8u * j)) & 0xffu;                /* the position's own byte: what it is as a literal */
            if (t >= tile_len) continue;
            const uint32_t to_unit = B2D_UNIT - (t & (B2D_UNIT - 1u)), to_end = tile_len - t;
            uint32_t maxl = to_unit < to_end ? to_unit : to_end;
            if (maxl > B2D_MAX_MATCH) maxl = B2D_MAX_MATCH;
            uint32_t l = 0;
            if (maxl >= B2D_MIN_MATCH && dist[j] <= max_dist) {
                const uint32_t own0 = j ? __funnelshift_r(w0, w1, 8u * j) : w0, own1 = j ? __funnelshift_r(w1, w2, 8u * j) : w1;
                const uint32_t ca = wa + j - dist[j], cw = ca & ~3u, csh = ca << 3;
                const uint32_t a0 = lds32(cw), a1 = lds32_4(cw), a2 = lds32_8(cw);
                const uint32_t x0 = __funnelshift_r(a0, a1, csh) ^ own0;
                const uint32_t x1 = __funnelshift_r(a1, a2, csh) ^ own1;
                uint32_t x = x0 ? x0 : x1;
                l = x0 ? 0u : 4u;
                if ((x0 | x1) == 0u) {
                    /* longer than 8: 8 bytes per trip, each aligned word of either side loaded once */
                    l = 8;
                    uint32_t cq = cw + 8u, oq = wa + 8u, cp = a2, op = w2;
                    while (l < maxl) {
                        const uint32_t c1 = lds32_4(cq), c2 = lds32_8(cq), p1 = lds32_4(oq), p2 = lds32_8(oq);
                        if ((x = __funnelshift_r(cp, c1, csh) ^ (j ? __funnelshift_r(op, p1, 8u * j) : op)) != 0) break;
                        l += 4;
                        if ((x = __funnelshift_r(c1, c2, csh) ^ (j ? __funnelshift_r(p1, p2, 8u * j) : p1)) != 0) break;
                        l += 4; cq += 8u; oq += 8u; cp = c2; op = p2;
                    }
                }
                l += eq_bytes(x);           /* x == 0: the long compare ran into maxl; four more bytes, clamped next */
                if (l > maxl) l = maxl;
            }
            if (l >= B2D_MIN_MATCH) mres[j] = b2d_pack_match(l, dist[j], mres[j]);
        }
        if (inner && vec_out) *reinterpret_cast<uint4 *>(mo) = make_uint4(mres[0], mres[1], mres[2], mres[3]);
        else {
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j) if (t0 + j < tile_len) mo[j] = mres[j];
        }
    }
}

/* match records without a search: the position's byte, nothing else (four positions per thread) */
__global__ void __launch_bounds__(256) k_literals(const uint8_t *__restrict__ data, uint64_t n, uint32_t *__restrict__ match) {
    const uint64_t i0 = 4ull * (blockIdx.x * 256ull + threadIdx.x);
    for (uint32_t j = 0; j < 4; ++j) if (i0 + j < n) match[i0 + j] = data[i0 + j];
}

/* ------------------------------------------------------------------------- */
/* K3: parse + token compaction + histograms                                  */
/* ------------------------------------------------------------------------- */
#define K3_THREADS 256
#define K3_WORDS (B2D_SUB / 32u)      /* bitmap words per sub-block = 1024 */
#define K3_SEGS 4u                    /* lanes chasing one parse unit speculatively */
#define K3_SEG (B2D_UNIT / K3_SEGS)

__global__ void __launch_bounds__(K3_THREADS, 5)
k_parse(uint64_t n, const uint32_t *__restrict__ match,
        uint32_t *__restrict__ tok, b2d_subinfo *__restrict__ info, uint32_t *__restrict__ hist, int lazy,
        uint32_t min_take) {
    __shared__ __align__(16) uint8_t s_len[B2D_SUB + 16];
    __shared__ uint32_t s_reach[K3_WORDS], s_base[K3_WORDS];
    __shared__ uint32_t s_hist[B2D_HIST];
    __shared__ uint32_t s_wsum[K3_THREADS / 32];
    __shared__ uint32_t s_exit[K3_SEGS * 32];
    __shared__ uint8_t s_lsym[256], s_dsym[512];   /* length symbol of len-3; distance symbol of d-1 (d-1 < 256) or 256 + ((d-1) >> 7) */
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint64_t k = blockIdx.x;
    const uint64_t sub_start = k * (uint64_t)B2D_SUB;
    const uint32_t len = (uint32_t)(n - sub_start < B2D_SUB ? n - sub_start : B2D_SUB);
    const uint32_t *m = match + sub_start;
    const uint32_t take3 = min_take - 3u;                 /* >= 1: "no match" (0) never passes */
    /* match lengths -> one byte each (len-3, 0 = no match); sub_start is a multiple of 4 positions.
     * Whole uint4 groups first, eight loads in flight per thread; then the ragged end and the padding. */
    {
        uint32_t *W = reinterpret_cast<uint32_t *>(s_len);
        const uint4 *m4 = reinterpret_cast<const uint4 *>(m);
        const uint32_t full = len >> 2;                      /* groups of 4 positions entirely inside the sub-block */
        for (uint32_t i0 = 0; i0 < full; i0 += 8u * K3_THREADS) {
            uint4 v[8];
#pragma unroll
            for (uint32_t j = 0; j < 8; ++j) {
                const uint32_t i = i0 + j * K3_THREADS + tid;
                v[j] = i < full ? __ldg(m4 + i) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (uint32_t j = 0; j < 8; ++j) {
                const uint32_t i = i0 + j * K3_THREADS + tid;
                /* the records carry len - 3 in their top bits (0 = none) */
                const uint32_t l0 = v[j].x >> 23, l1 = v[j].y >> 23, l2 = v[j].z >> 23, l3 = v[j].w >> 23;
                if (i < full)      /* matches below min_take (Z_FILTERED) count as none */
                    W[i] = (l0 >= take3 ? l0 : 0u) | ((l1 >= take3 ? l1 : 0u) << 8) |
                           ((l2 >= take3 ? l2 : 0u) << 16) | ((l3 >= take3 ? l3 : 0u) << 24);
            }
        }
        for (uint32_t i = full + tid; i < (B2D_SUB + 16) / 4; i += K3_THREADS) {
            uint32_t pk = 0;
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j) {
                const uint32_t pos = 4 * i + j;
                const uint32_t l = pos < len ? __ldg(m + pos) >> 23 : 0u;
                pk |= (l >= take3 ? l : 0u) << (8 * j);
            }
            W[i] = pk;
        }
    }
    for (uint32_t i = tid; i < B2D_HIST; i += K3_THREADS) s_hist[i] = 0;
    for (uint32_t i = tid; i < K3_WORDS; i += K3_THREADS) s_reach[i] = 0;
    {   /* K3_THREADS == 256: one table row per thread */
        uint32_t sy, eb, ev;
        b2d_len_sym(tid, &sy, &eb, &ev); s_lsym[tid] = (uint8_t)sy;
        b2d_dist_sym(tid, &sy, &eb, &ev); s_dsym[tid] = (uint8_t)sy;
        b2d_dist_sym(tid << 7, &sy, &eb, &ev); s_dsym[256 + tid] = (uint8_t)sy;
    }
    __syncthreads();
    /* lazy rule applied to every position at once (b2d_parse_step): a position keeps its length byte if
     * a match would be taken there, else 0.  Results sit in registers until every thread has read. */
    {
        uint32_t *W = reinterpret_cast<uint32_t *>(s_len);
        uint32_t outw[B2D_SUB / 4 / K3_THREADS];
#pragma unroll
        for (uint32_t r = 0; r < B2D_SUB / 4 / K3_THREADS; ++r) {
            const uint32_t wi = r * K3_THREADS + tid;
            const uint32_t a = W[wi], b = W[wi + 1];
            const uint64_t v = ((uint64_t)b << 32) | a;
            uint32_t o = 0;
#pragma unroll
            for (uint32_t j = 0; j < 4; ++j) {
                const uint32_t l0 = (uint32_t)(v >> (8 * j)) & 255u, l1 = (uint32_t)(v >> (8 * j + 8)) & 255u,
                               l2 = (uint32_t)(v >> (8 * j + 16)) & 255u;
                bool take = l0 != 0;
                if (lazy >= 1 && l1 > l0) take = false;
                if (lazy >= 2 && l2 > l0 + 1u) take = false;
                if (take) o |= l0 << (8 * j);
            }
            outw[r] = o;
        }
        __syncthreads();
#pragma unroll
        for (uint32_t r = 0; r < B2D_SUB / 4 / K3_THREADS; ++r) W[r * K3_THREADS + tid] = outw[r];
    }
    __syncthreads();
    /* Chase.  A unit's parse is a chain i -> i + step(i); K3_SEGS lanes walk it at once, lane (u, g)
     * starting at the g-th quarter of unit u as if that position were reached.  Two chains that ever meet
     * stay together, so a quarter's speculative marks are right from the point where the true chain
     * (entering from the previous quarter) first steps on one of them; the repair below fixes the rest. */
    const uint64_t *L8 = reinterpret_cast<const uint64_t *>(s_len);
    if (warp < K3_SEGS) {
        const uint32_t ustart = lane * B2D_UNIT;
        const uint32_t uend = ustart + B2D_UNIT < len ? ustart + B2D_UNIT : len;
        const uint32_t b = ustart + warp * K3_SEG;
        const uint32_t e = b + K3_SEG < uend ? b + K3_SEG : uend;     /* marks stay below e */
        uint32_t i = b;
        while (i < e) {
            /* the 8-byte group holding i, shifted so that byte 0 is position i: if nothing is left in
             * it, every position up to the group's end is a literal (literal runs go 8 at a time) */
            const uint32_t o = i & 7u;
            const uint64_t g = L8[i >> 3] >> (8u * o);
            const uint32_t v = (uint32_t)g & 255u;
            uint32_t k = 1;                                   /* positions marked reached */
            if (g == 0) { k = 8u - o; if (i + k > e) k = e - i; }
            atomicOr(&s_reach[i >> 5], ((1u << k) - 1u) << (i & 31u));
            i += v ? v + 3u : k;
        }
        s_exit[warp * 32u + lane] = i;                        /* first position at or past e the chain lands on */
    }
    __syncthreads();
    if (warp == 0) {
        /* repair, one lane per unit, quarter by quarter: `at` is where the true chain enters the quarter */
        const uint32_t ustart = lane * B2D_UNIT;
        const uint32_t uend = ustart + B2D_UNIT < len ? ustart + B2D_UNIT : len;
        uint32_t at = s_exit[lane];
        for (uint32_t q = 1; q < K3_SEGS; ++q) {
            const uint32_t b = ustart + q * K3_SEG;
            if (b >= uend) break;
            const uint32_t e = b + K3_SEG < uend ? b + K3_SEG : uend;
            if (at == b) { at = s_exit[q * 32u + lane]; continue; }          /* the guess was right */
            /* where does the true chain meet the speculative one (or leave the quarter)? */
            uint32_t j = at;
            while (j < e && !((s_reach[j >> 5] >> (j & 31u)) & 1u)) { const uint32_t v = s_len[j]; j += v ? v + 3u : 1u; }
            const uint32_t meet = j < e ? j : e;
            /* speculative marks before the meeting point are wrong: clear [b, meet), then mark the true chain */
            for (uint32_t w = b >> 5; w <= (meet - 1u) >> 5 && meet > b; ++w) {
                const uint32_t lo = w == (b >> 5) ? (b & 31u) : 0u;
                const uint32_t hi = w == ((meet - 1u) >> 5) ? ((meet - 1u) & 31u) : 31u;
                s_reach[w] &= ~((0xffffffffu >> (31u - hi)) & (0xffffffffu << lo));
            }
            uint32_t t2 = at;
            while (t2 < meet) { s_reach[t2 >> 5] |= 1u << (t2 & 31u); const uint32_t v = s_len[t2]; t2 += v ? v + 3u : 1u; }
            at = j < e ? s_exit[q * 32u + lane] : j;
        }
    }
    __syncthreads();
    /* exclusive scan of per-word token counts: thread t owns words 4t..4t+3 */
    uint32_t c[4], tsum = 0;
#pragma unroll
    for (uint32_t j = 0; j < 4; ++j) { c[j] = __popc(s_reach[4 * tid + j]); tsum += c[j]; }
    uint32_t incl = tsum;
#pragma unroll
    for (uint32_t o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += v; }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    uint32_t wbase = 0, total = 0;
#pragma unroll
    for (uint32_t w = 0; w < K3_THREADS / 32; ++w) { uint32_t v = s_wsum[w]; if (w < warp) wbase += v; total += v; }
    uint32_t run = wbase + incl - tsum;
#pragma unroll
    for (uint32_t j = 0; j < 4; ++j) { s_base[4 * tid + j] = run; run += c[j]; }
    __syncthreads();
    /* compaction: a warp takes 4 bitmap words per trip (lane = bit) and issues the global loads of
     * all four before using any, so their latency overlaps.  Symbols come from small tables (the
     * zlib-ng _length_code / _dist_code idea, trees_tbl.h:76-132); one store and one literal/length
     * count per token, whatever its kind. */
    const uint32_t tbase = (uint32_t)sub_start;           /* a batch is at most 2 GiB: token indices fit 32 bits */
    const uint32_t lt_mask = (1u << lane) - 1u;
    for (uint32_t w0 = warp * 4u; w0 < K3_WORDS; w0 += (K3_THREADS / 32) * 4u) {
        uint32_t rw[4], val[4];
        bool reached[4], take[4];
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j) {
            rw[j] = s_reach[w0 + j];
            const uint32_t pos = 32u * (w0 + j) + lane;
            reached[j] = (rw[j] >> lane) & 1u;
            take[j] = s_len[pos] != 0;
            val[j] = 0;
            if (reached[j]) val[j] = __ldg(m + pos);     /* the record holds the token either way: match fields and literal */
        }
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j) {
            if (!reached[j]) continue;
            const uint32_t idx = s_base[w0 + j] + __popc(rw[j] & lt_mask);
            uint32_t tv = b2d_match_lit(val[j]), sym = tv;
            if (take[j]) {
                const uint32_t d1 = (val[j] >> 8) & 0x7fffu;
                sym = 257u + s_lsym[val[j] >> 23];
                atomicAdd(&s_hist[B2D_LL_CODES + s_dsym[d1 < 256u ? d1 : 256u + (d1 >> 7)]], 1u);
                tv = 0x80000000u | (val[j] >> 8);         /* b2d_tok_match: (len - 3) << 15 | (dist - 1) */
            }
            atomicAdd(&s_hist[sym], 1u);
            tok[tbase + idx] = tv;
        }
    }
    __syncthreads();
    if (tid == 0) {
        if (total) s_hist[256] = 1;
        info[k].in_len = len; info[k].tok_count = total; info[k].bit_off = 0; info[k].pad = 0;
    }
    __syncthreads();
    uint32_t *H = hist + k * (uint64_t)B2D_HIST;
    for (uint32_t i = tid; i < B2D_HIST; i += K3_THREADS) H[i] = s_hist[i];
}

/* ------------------------------------------------------------------------- */
/* K4: Huffman plan per sub-block                                             */
/* ------------------------------------------------------------------------- */
#define K4_WARPS 7
struct k4_warp_smem {
    b2d_huff_scratch S;
    b2d_plan P;
    uint32_t hist[B2D_HIST];
};

/* ascending sort of S->key[0..m) by the whole warp: rank of each (unique) key = number of smaller keys */
__device__ __forceinline__ void warp_rank_sort(b2d_huff_scratch *S, int m, uint32_t lane) {
    for (int i = (int)lane; i < m; i += 32) {
        const uint32_t k = S->key[i];
        int r = 0;
        for (int j = 0; j < m; ++j) r += S->key[j] < k;
        S->w[r] = k;
    }
    __syncwarp();
    for (int i = (int)lane; i < m; i += 32) S->key[i] = S->w[i];
    __syncwarp();
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
    for (uint32_t o = 16; o; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}

/* b2d_huff_keys by the warp: one key per used symbol (any order: the keys are unique and get sorted) */
__device__ __forceinline__ int warp_huff_keys(const uint32_t *freq, int n, uint8_t *lens, b2d_huff_scratch *S, uint32_t lane) {
    int m = 0;
    for (int base = 0; base < n; base += 32) {
        const int i = base + (int)lane;
        const uint32_t f = i < n ? freq[i] : 0u;
        if (i < n) lens[i] = 0;
        const uint32_t used = __ballot_sync(FULL, f != 0);
        if (f) S->key[m + __popc(used & ((1u << lane) - 1u))] = (f << 9) | (uint32_t)i;
        m += __popc(used);
    }
    __syncwarp();
    if (m < 2 && lane == 0) {          /* at least two symbols always receive a code (b2d_huff_keys) */
        if (m == 0) { S->key[0] = (1u << 9) | 0u; S->key[1] = (1u << 9) | 1u; }
        else { const uint32_t s0 = S->key[0] & 511u; S->key[1] = (1u << 9) | (s0 == 0u ? 1u : 0u); }
    }
    if (m < 2) m = 2;
    __syncwarp();
    return m;
}

/* one code-length set: keys and sort by the warp, the serial tree phases by lane 0 */
__device__ __forceinline__ void warp_huff_lengths(const uint32_t *freq, int n, int maxbits, uint8_t *lens,
                                                  b2d_huff_scratch *S, uint32_t lane) {
    const int m = warp_huff_keys(freq, n, lens, S, lane);
    warp_rank_sort(S, m, lane);
    if (lane == 0) b2d_huff_from_sorted(m, maxbits, lens, S);
    __syncwarp();
}

/* b2d_canonical by the warp: lane b keeps next_code[b]; a symbol's code is next_code[len] plus the number
 * of lower-numbered symbols of the same length */
__device__ __forceinline__ void warp_canonical(const uint8_t *lens, int n, int maxbits, uint32_t *out, uint32_t lane) {
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t cnt = 0;
    for (int base = 0; base < n; base += 32) {
        const int i = base + (int)lane;
        const uint32_t l = i < n ? lens[i] : 0u;
        for (int b = 1; b <= maxbits; ++b) {
            const uint32_t mk = __ballot_sync(FULL, l == (uint32_t)b);
            if (lane == (uint32_t)b) cnt += __popc(mk);
        }
    }
    uint32_t code = 0, next = 0;
    for (int b = 1; b <= maxbits; ++b) {
        const uint32_t below = b >= 2 ? __shfl_sync(FULL, cnt, b - 1) : 0u;
        code = (code + below) << 1;
        if (lane == (uint32_t)b) next = code;
    }
    for (int base = 0; base < n; base += 32) {
        const int i = base + (int)lane;
        const uint32_t l = i < n ? lens[i] : 0u;
        uint32_t mine = 0;
        for (int b = 1; b <= maxbits; ++b) {
            const uint32_t mk = __ballot_sync(FULL, l == (uint32_t)b);
            const uint32_t nb = __shfl_sync(FULL, next, b);
            if (l == (uint32_t)b) mine = nb + __popc(mk & lt);
            if (lane == (uint32_t)b) next += __popc(mk);
        }
        if (i < n) out[i] = l ? ((l << 16) | (__brev(mine) >> (32u - l))) : 0u;
    }
    __syncwarp();
}

__constant__ uint8_t c_bl_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

/* b2d_plan_end by the warp (same results): costs by reduction, code tables and header bits in parallel */
__device__ __forceinline__ void warp_plan_end(const uint32_t *ll_freq, const uint32_t *d_freq, uint32_t in_len,
                                              b2d_plan *P, b2d_huff_scratch *S, int blocks, uint32_t lane) {
    const bool force_stored = blocks == B2D_BLOCKS_STORED, no_dyn = blocks == B2D_BLOCKS_NO_DYN;
    const int hlit = S->hlit, hdist = S->hdist, nseq = S->nseq;
    int hclen = B2D_BL_CODES;
    while (hclen > 4 && S->bl_len[c_bl_order[hclen - 1]] == 0) hclen--;
    uint32_t hdr_part = 0, dyn_part = 0, fix_part = 0;
    for (int i = (int)lane; i < nseq; i += 32) {
        const uint32_t s = S->seq_sym[i];
        hdr_part += S->bl_len[s] + (s == 16u ? 2u : s == 17u ? 3u : s == 18u ? 7u : 0u);
    }
    for (int i = (int)lane; i < B2D_LL_CODES; i += 32) {
        const uint32_t f = ll_freq[i];
        const uint32_t ex = i >= 257 ? b2d_len_extra_bits((uint32_t)i - 257u) : 0u;
        dyn_part += f * (S->ll_len[i] + ex);
        fix_part += f * (b2d_fixed_ll_len((uint32_t)i) + ex);
    }
    if (lane < B2D_D_CODES) {
        const uint32_t f = d_freq[lane], ex = b2d_dist_extra_bits(lane);
        dyn_part += f * (S->d_len[lane] + ex);
        fix_part += f * (5u + ex);
    }
    const uint32_t dyn_hdr = 3u + 5u + 5u + 4u + 3u * (uint32_t)hclen + warp_sum(hdr_part);
    const uint32_t dyn_body = warp_sum(dyn_part), fix_body = warp_sum(fix_part);
    const uint32_t dyn_cost = dyn_hdr + dyn_body, fix_cost = 3u + fix_body;
    const uint32_t sto_cost = 3u + 7u + 32u + 8u * in_len;   /* worst-case alignment pad */
    for (uint32_t i = lane; i < B2D_HDR_WORDS; i += 32) P->hdr[i] = 0;
    __syncwarp();
    if ((force_stored || ((sto_cost <= dyn_cost || no_dyn) && sto_cost <= fix_cost)) && in_len <= 65535u) {
        for (uint32_t i = lane; i < B2D_LL_CODES; i += 32) P->ll_code[i] = 0;
        if (lane < B2D_D_CODES) P->d_code[lane] = 0;
        if (lane == 0) { P->type = 0; P->body_bits = 0; P->pad = 0; P->hdr_bits = 3; }   /* BFINAL=0, BTYPE=00: three zero bits */
    } else if (fix_cost <= dyn_cost || no_dyn) {
        for (uint32_t i = lane; i < B2D_LL_CODES; i += 32) {
            const uint32_t l = b2d_fixed_ll_len(i);
            const uint32_t c = i < 144u ? 0x30u + i : i < 256u ? 0x190u + (i - 144u) : i < 280u ? i - 256u : 0xC0u + (i - 280u);
            P->ll_code[i] = (l << 16) | b2d_bitrev(c, (int)l);
        }
        if (lane < B2D_D_CODES) P->d_code[lane] = (5u << 16) | b2d_bitrev(lane, 5);
        if (lane == 0) { P->type = 1; P->body_bits = fix_body; P->pad = 0; P->hdr[0] = 2u; P->hdr_bits = 3; }   /* BTYPE=01 */
    } else {
        warp_canonical(S->ll_len, B2D_LL_CODES, 15, P->ll_code, lane);
        warp_canonical(S->d_len, B2D_D_CODES, 15, P->d_code, lane);
        warp_canonical(S->bl_len, B2D_BL_CODES, 7, S->bl_code, lane);
        uint32_t nbits = 0;
        if (lane == 0) {
            b2d_put_bits(P->hdr, &nbits, 4u, 3);          /* BFINAL=0, BTYPE=10 */
            b2d_put_bits(P->hdr, &nbits, (uint32_t)(hlit - 257), 5);
            b2d_put_bits(P->hdr, &nbits, (uint32_t)(hdist - 1), 5);
            b2d_put_bits(P->hdr, &nbits, (uint32_t)(hclen - 4), 4);
            for (int i = 0; i < hclen; i++) b2d_put_bits(P->hdr, &nbits, S->bl_len[c_bl_order[i]], 3);
        }
        nbits = __shfl_sync(FULL, nbits, 0);
        __syncwarp();
        /* the run-length coded length sequence: lane i places item i at the prefix sum of the widths */
        for (int base = 0; base < nseq; base += 32) {
            const int i = base + (int)lane;
            uint32_t val = 0, w = 0;
            if (i < nseq) {
                const uint32_t s = S->seq_sym[i], c = S->bl_code[s];
                const uint32_t eb = s == 16u ? 2u : s == 17u ? 3u : s == 18u ? 7u : 0u;
                val = (c & 0xffffu) | ((uint32_t)S->seq_ext[i] << (c >> 16));
                w = (c >> 16) + eb;
            }
            uint32_t incl = w;
#pragma unroll
            for (uint32_t o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += v; }
            if (w) {
                const uint32_t pos = nbits + incl - w, wi = pos >> 5, sh = pos & 31u;
                atomicOr(&P->hdr[wi], val << sh);
                if (sh + w > 32u) atomicOr(&P->hdr[wi + 1], val >> (32u - sh));
            }
            nbits += __shfl_sync(FULL, incl, 31);
        }
        if (lane == 0) { P->type = 2; P->body_bits = dyn_body; P->pad = 0; P->hdr_bits = nbits; }
    }
    __syncwarp();
}

__global__ void __launch_bounds__(K4_WARPS * 32)
k_plan(const uint32_t *__restrict__ hist, const b2d_subinfo *__restrict__ info, b2d_plan *__restrict__ plan,
       uint32_t nsub, int blocks) {
    extern __shared__ __align__(16) uint8_t smem4[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t k = blockIdx.x * K4_WARPS + warp;
    if (k >= nsub) return;
    k4_warp_smem *W = reinterpret_cast<k4_warp_smem *>(smem4) + warp;
    if (!info[k].tok_count) {
        if (lane == 0) { plan[k].hdr_bits = 0; plan[k].body_bits = 0; plan[k].type = 1; }
        return;
    }
    const uint32_t *H = hist + (uint64_t)k * B2D_HIST;
    for (uint32_t i = lane; i < B2D_HIST; i += 32) W->hist[i] = H[i];
    __syncwarp();
    b2d_huff_scratch *S = &W->S;
    warp_huff_lengths(W->hist, B2D_LL_CODES, 15, S->ll_len, S, lane);
    warp_huff_lengths(W->hist + B2D_LL_CODES, B2D_D_CODES, 15, S->d_len, S, lane);
    if (lane == 0) b2d_plan_mid(S);
    __syncwarp();
    warp_huff_lengths(S->bl_freq, B2D_BL_CODES, 7, S->bl_len, S, lane);
    warp_plan_end(W->hist, W->hist + B2D_LL_CODES, info[k].in_len, &W->P, S, blocks, lane);
    const uint32_t *src = reinterpret_cast<const uint32_t *>(&W->P);
    uint32_t *dst = reinterpret_cast<uint32_t *>(&plan[k]);
    for (uint32_t i = lane; i < sizeof(b2d_plan) / 4; i += 32) dst[i] = src[i];
}

/* ------------------------------------------------------------------------- */
/* K5: offsets                                                                */
/* ------------------------------------------------------------------------- */
/* bits the sub-block occupies when it starts at bit offset `at` inside its device block */
__device__ __forceinline__ uint64_t sub_bits(const b2d_plan &P, const b2d_subinfo &I, uint64_t at) {
    if (!I.tok_count) return 0;
    if (P.type == 0) { uint64_t e = (at + P.hdr_bits + 7) & ~7ull; return e + 32 + 8ull * I.in_len - at; }
    return (uint64_t)P.hdr_bits + P.body_bits;
}

__global__ void __launch_bounds__(1024)
k_scan(const b2d_plan *__restrict__ plan, b2d_subinfo *__restrict__ info, uint32_t nsub, uint32_t nblocks,
       uint64_t *__restrict__ block_off, uint64_t *__restrict__ total) {
    __shared__ uint64_t s_w[32];
    __shared__ uint64_t s_carry;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    if (tid == 0) s_carry = *total;      /* bytes already produced by earlier batches of this call */
    __syncthreads();
    for (uint32_t base = 0; base < nblocks; base += 1024) {
        const uint32_t b = base + tid;
        uint64_t bytes = 0;
        if (b < nblocks) {
            uint64_t bits = 0;
            for (uint32_t s = 0; s < B2D_NSUB; ++s) {
                const uint32_t k = b * B2D_NSUB + s;
                if (k >= nsub) break;
                info[k].bit_off = (uint32_t)bits;
                bits += sub_bits(plan[k], info[k], bits);
            }
            bits += 3; bits = (bits + 7) & ~7ull; bits += 32;
            bytes = bits >> 3;
        }
        uint64_t incl = bytes;
#pragma unroll
        for (uint32_t o = 1; o < 32; o <<= 1) { uint64_t v = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += v; }
        if (lane == 31) s_w[warp] = incl;
        __syncthreads();
        uint64_t wb = 0, tot = 0;
        for (uint32_t w = 0; w < 32; ++w) { uint64_t v = s_w[w]; if (w < warp) wb += v; tot += v; }
        const uint64_t carry = s_carry;
        if (b < nblocks) block_off[b] = carry + wb + incl - bytes;
        __syncthreads();
        if (tid == 0) s_carry = carry + tot;
        __syncthreads();
    }
    if (tid == 0) { block_off[nblocks] = s_carry; *total = s_carry; }
}

/* ------------------------------------------------------------------------- */
/* K6: bit emitter                                                            */
/* ------------------------------------------------------------------------- */
#define K6_THREADS 256

struct bitsink {
    uint32_t *out;       /* zero-initialised output words */
    uint64_t acc;        /* pending bits, LSB first        */
    uint32_t fill;       /* valid bits in acc (< 32 after flush) */
    uint64_t word;       /* index of the word acc's bit 0 belongs to */
    bool first;
};
__device__ __forceinline__ void sink_open(bitsink &s, uint32_t *out, uint64_t bitpos) {
    s.out = out; s.word = bitpos >> 5; s.fill = (uint32_t)(bitpos & 31u); s.acc = 0; s.first = true;
}
__device__ __forceinline__ void sink_put(bitsink &s, uint64_t value, uint32_t count) {   /* count <= 32 */
    s.acc |= value << s.fill;
    s.fill += count;
    if (s.fill >= 32u) {
        const uint32_t w = (uint32_t)s.acc;
        if (s.first) { atomicOr(&s.out[s.word], w); s.first = false; } else s.out[s.word] = w;
        s.word++; s.acc >>= 32; s.fill -= 32u;
    }
}
__device__ __forceinline__ void sink_close(bitsink &s) {
    if (s.fill) atomicOr(&s.out[s.word], (uint32_t)s.acc);
}

/* the bits of one token.  lc[len - 3] = the length's code with its extra bits behind it, bit count in the top byte (built
 * once per sub-block); dsym = distance symbol of d-1 (d-1 < 256) or of 256 + ((d-1) >> 7), as in k_parse */
__device__ __forceinline__ void tok_bits(uint32_t x, const uint32_t *ll, const uint32_t *dc, const uint32_t *lc, const uint8_t *dsym,
                                         uint64_t *val, uint32_t *cnt) {
    if (!b2d_tok_is_match(x)) { const uint32_t c = ll[x]; *val = c & 0xffffu; *cnt = c >> 16; return; }
    const uint32_t d1 = b2d_tok_dist1(x);
    const uint32_t c1 = lc[b2d_tok_len3(x)];
    const uint32_t ds = dsym[d1 < 256u ? d1 : 256u + (d1 >> 7)];
    const uint32_t c2 = dc[ds], de = b2d_dist_extra_bits(ds);
    uint64_t v = c1 & 0xffffffu; uint32_t nb = c1 >> 24;
    v |= (uint64_t)(c2 & 0xffffu) << nb; nb += c2 >> 16;
    v |= (uint64_t)(d1 & ((1u << de) - 1u)) << nb; nb += de;
    *val = v; *cnt = nb;                      /* <= 15+5+15+13 = 48 bits */
}

__global__ void __launch_bounds__(K6_THREADS)
k_emit(const uint8_t *__restrict__ data, const uint32_t *__restrict__ tok, const b2d_subinfo *__restrict__ info,
       const b2d_plan *__restrict__ plan, const uint64_t *__restrict__ block_off, uint32_t nsub,
       uint32_t *__restrict__ out) {
    __shared__ uint32_t s_ll[B2D_LL_CODES], s_d[B2D_D_CODES];
    __shared__ uint32_t s_lc[256];
    __shared__ uint8_t s_dsym[512];
    __shared__ uint32_t s_wsum[K6_THREADS / 32];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t k = blockIdx.x;
    const uint32_t b = k / B2D_NSUB;
    const b2d_subinfo I = info[k];
    const b2d_plan *P = &plan[k];
    const uint64_t base = block_off[b] * 8ull + I.bit_off;       /* absolute bit position of this sub-block */
    const bool last_of_block = (k % B2D_NSUB == B2D_NSUB - 1) || (k == nsub - 1);
    uint64_t endbit = base;
    if (I.tok_count) {
        const uint32_t type = P->type, hdr_bits = P->hdr_bits;
        /* header bits: thread j places hdr word j */
        for (uint32_t j = tid; 32u * j < hdr_bits; j += K6_THREADS) {
            const uint32_t nb = hdr_bits - 32u * j < 32u ? hdr_bits - 32u * j : 32u;
            uint32_t v = P->hdr[j];
            if (nb < 32u) v &= (1u << nb) - 1u;
            const uint64_t bp = base + 32ull * j;
            const uint32_t sh = (uint32_t)(bp & 31u);
            atomicOr(&out[bp >> 5], v << sh);
            if (sh && (v >> (32u - sh))) atomicOr(&out[(bp >> 5) + 1], v >> (32u - sh));
        }
        if (type == 0) {
            /* stored: pad to a byte, LEN, NLEN, raw bytes */
            const uint64_t hb = ((base + hdr_bits + 7) & ~7ull) >> 3;          /* byte offset of LEN */
            const uint8_t *src = data + (uint64_t)k * B2D_SUB;
            const uint32_t L = I.in_len;
            const uint64_t total = 4ull + L;                                   /* bytes to place from hb */
            const uint64_t w0 = hb >> 2, w1 = (hb + total + 3) >> 2;
            for (uint64_t w = w0 + tid; w < w1; w += K6_THREADS) {
                uint32_t v = 0;
#pragma unroll
                for (uint32_t j = 0; j < 4; ++j) {
                    const uint64_t byte = 4 * w + j;
                    if (byte < hb || byte >= hb + total) continue;
                    const uint32_t r = (uint32_t)(byte - hb);
                    uint32_t x;
                    if (r == 0) x = L & 0xffu; else if (r == 1) x = (L >> 8) & 0xffu;
                    else if (r == 2) x = (~L) & 0xffu; else if (r == 3) x = ((~L) >> 8) & 0xffu;
                    else x = src[r - 4];
                    v |= x << (8 * j);
                }
                if (w == w0 || w == w1 - 1) atomicOr(&out[w], v); else out[w] = v;
            }
            endbit = (hb + total) * 8ull;
        } else {
            for (uint32_t i = tid; i < B2D_LL_CODES; i += K6_THREADS) s_ll[i] = P->ll_code[i];
            if (tid < B2D_D_CODES) s_d[tid] = P->d_code[tid];
            __syncthreads();
            static_assert(K6_THREADS == 256, "one table row per thread");
            {   /* per length: code | extra bits << code length, total bit count in the top byte (<= 20 bits) */
                uint32_t sy, eb, ev;
                b2d_len_sym(tid, &sy, &eb, &ev);
                const uint32_t c1 = s_ll[257u + sy];
                s_lc[tid] = (((c1 >> 16) + eb) << 24) | (c1 & 0xffffu) | (ev << (c1 >> 16));
                b2d_dist_sym(tid, &sy, &eb, &ev); s_dsym[tid] = (uint8_t)sy;
                b2d_dist_sym(tid << 7, &sy, &eb, &ev); s_dsym[256 + tid] = (uint8_t)sy;
            }
            __syncthreads();
            const uint32_t nt = I.tok_count;
            /* each thread owns a run of tokens whose length is a multiple of 4, read as uint4 with the
             * next group already in flight (the loop is otherwise bound by the load latency) */
            const uint32_t per = ((nt + K6_THREADS - 1) / K6_THREADS + 3u) & ~3u;
            const uint32_t t0 = tid * per < nt ? tid * per : nt;
            const uint32_t t1 = t0 + per < nt ? t0 + per : nt;
            const uint4 *t4 = reinterpret_cast<const uint4 *>(tok + (uint64_t)k * B2D_SUB);
            const uint32_t g0 = t0 >> 2, g1 = (t1 + 3u) >> 2;
            uint32_t mybits = 0;
            if (t1 > t0) {          /* (an empty run has t0 = nt, which need not be a multiple of 4) */
                uint4 q = __ldg(t4 + g0);
                for (uint32_t g = g0; g < g1; ++g) {
                    const uint4 cur = q;
                    if (g + 1 < g1) q = __ldg(t4 + g + 1);
                    const uint32_t x[4] = {cur.x, cur.y, cur.z, cur.w};
#pragma unroll
                    for (uint32_t j = 0; j < 4; ++j) {
                        if (4u * g + j >= t1) break;          /* slots past the last token hold stale words */
                        uint64_t v; uint32_t c; tok_bits(x[j], s_ll, s_d, s_lc, s_dsym, &v, &c);
                        mybits += c;
                    }
                }
            }
            uint32_t incl = mybits;
#pragma unroll
            for (uint32_t o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += v; }
            if (lane == 31) s_wsum[warp] = incl;
            __syncthreads();
            uint32_t wb = 0, tot = 0;
#pragma unroll
            for (uint32_t w = 0; w < K6_THREADS / 32; ++w) { uint32_t v = s_wsum[w]; if (w < warp) wb += v; tot += v; }
            const uint64_t body = base + hdr_bits;
            if (t1 > t0) {
                bitsink s; sink_open(s, out, body + wb + incl - mybits);
                uint4 q = __ldg(t4 + g0);
                for (uint32_t g = g0; g < g1; ++g) {
                    const uint4 cur = q;
                    if (g + 1 < g1) q = __ldg(t4 + g + 1);
                    const uint32_t x[4] = {cur.x, cur.y, cur.z, cur.w};
#pragma unroll
                    for (uint32_t j = 0; j < 4; ++j) {
                        if (4u * g + j >= t1) break;
                        uint64_t v; uint32_t c; tok_bits(x[j], s_ll, s_d, s_lc, s_dsym, &v, &c);
                        if (c > 32u) { sink_put(s, v & 0xffffffffull, 32u); sink_put(s, v >> 32, c - 32u); }
                        else sink_put(s, v, c);
                    }
                }
                sink_close(s);
            }
            if (tid == 0) {   /* end-of-block code */
                const uint32_t c = s_ll[256];
                bitsink s; sink_open(s, out, body + tot); sink_put(s, c & 0xffffu, c >> 16); sink_close(s);
            }
            endbit = body + tot + (s_ll[256] >> 16);
        }
    }
    if (last_of_block && tid == 0) {
        /* sync-flush marker: empty stored block, 000 + pad + 00 00 ff ff (trees.c:593-602) */
        const uint64_t hb = ((endbit + 3 + 7) & ~7ull) >> 3;
        const uint64_t bp = (hb + 2) * 8ull;            /* the two 0xff bytes */
        const uint32_t sh = (uint32_t)(bp & 31u);
        atomicOr(&out[bp >> 5], 0xffffu << sh);
        if (sh > 16u) atomicOr(&out[(bp >> 5) + 1], 0xffffu >> (32u - sh));
    }
}

/* ------------------------------------------------------------------------- */
/* host-side launchers (called from b2d_api.cu)                               */
/* ------------------------------------------------------------------------- */
static size_t k1_smem(int hb) { return (2u << hb) + 2u * K1_BATCH + 16u; }
static size_t k4_smem() { return K4_WARPS * sizeof(k4_warp_smem); }
static size_t k2_smem() { return K2_IN_BYTES_T + 2u * (B2D_WINDOW + K2_TILE) + 32u; }

cudaError_t b2d_kernels_init(void) {
    cudaError_t e = cudaSuccess;
#define K2_ATTR(F, DP) do { e = cudaFuncSetAttribute(k_match<F, DP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k2_smem()); \
                            if (e != cudaSuccess) return e; } while (0)
    K2_ATTR(true, 0); K2_ATTR(true, 4); K2_ATTR(true, 5); K2_ATTR(true, 6); K2_ATTR(false, 0);
#undef K2_ATTR
    e = cudaFuncSetAttribute(k_match_d1<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K2_IN_BYTES);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_match_d1<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K2_IN_BYTES);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_plan, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)k4_smem());
    if (e != cudaSuccess) return e;
    return cudaSuccess;   /* k_chains needs at most 33 KiB: below the default dynamic limit */
}

cudaError_t b2d_launch_pipeline(const b2d_job *J, cudaStream_t st, cudaEvent_t *ev) {
    /* ev (optional): B2D_NSTAGE + 1 events recorded around the stages, for per-kernel timing */
    const uint64_t n = J->n;
    if (n == 0) return cudaSuccess;
    const uint32_t nblocks = (uint32_t)((n + B2D_BLOCK - 1) / B2D_BLOCK);
    const uint32_t nsub = (uint32_t)((n + B2D_SUB - 1) / B2D_SUB);
    const b2d_policy pol = b2d_level_policy(J->level);
    const uint8_t *data = J->buf + J->D;
#define MARK(i) do { if (ev) { cudaError_t e_ = cudaEventRecord(ev[i], st); if (e_ != cudaSuccess) return e_; } } while (0)
    MARK(0);
    if (pol.search == B2D_SEARCH_CHAINS) {
        /* the head table is 2^hash_bits u16: the smaller it is the more blocks an SM keeps in flight */
        switch (pol.hash_bits) {
            case 11: k_chains<11><<<nblocks, 32, k1_smem(11), st>>>(J->buf, J->D, n, J->buf_readable, J->prevd); break;
            case 12: k_chains<12><<<nblocks, 32, k1_smem(12), st>>>(J->buf, J->D, n, J->buf_readable, J->prevd); break;
            case 13: k_chains<13><<<nblocks, 32, k1_smem(13), st>>>(J->buf, J->D, n, J->buf_readable, J->prevd); break;
            default: k_chains<14><<<nblocks, 32, k1_smem(14), st>>>(J->buf, J->D, n, J->buf_readable, J->prevd); break;
        }
    }
    MARK(1);
    if (pol.search != B2D_SEARCH_NONE) {
        if (pol.nice < (int)B2D_MAX_MATCH) return cudaErrorInvalidValue;      /* see k2_search */
        if (pol.depth == 1) {
            if (pol.search == B2D_SEARCH_RLE)
                k_match_d1<true><<<nsub, K2D1_THREADS, K2_IN_BYTES, st>>>(J->buf, J->D, n, J->buf_readable, J->prevd, J->match, J->max_dist);
            else
                k_match_d1<false><<<nsub, K2D1_THREADS, K2_IN_BYTES, st>>>(J->buf, J->D, n, J->buf_readable, J->prevd, J->match, J->max_dist);
        }
        else {
            const uint32_t grid = (uint32_t)((n + K2_TILE - 1) / K2_TILE);
#define K2_GO(F, DP) k_match<F, DP><<<grid, K2_THREADS, k2_smem(), st>>>(J->buf, J->D, n, J->buf_readable, J->prevd, J->match, pol.depth, J->max_dist)
            /* the default window and the depths of levels 4-6 get their own instances (constants instead of loop state) */
            if (J->max_dist != B2D_WINDOW) K2_GO(false, 0);
            else if (pol.depth == 6) K2_GO(true, 6);
            else if (pol.depth == 5) K2_GO(true, 5);
            else if (pol.depth == 4) K2_GO(true, 4);
            else K2_GO(true, 0);
#undef K2_GO
        }
    } else {
        /* no search (level 0, Z_HUFFMAN_ONLY): every record is its position's literal */
        k_literals<<<(uint32_t)((n + 4u * 256u - 1) / (4u * 256u)), 256, 0, st>>>(data, n, J->match);
    }
    MARK(2);
    k_parse<<<nsub, K3_THREADS, 0, st>>>(n, J->match, J->tok, J->info, J->hist, pol.lazy, (uint32_t)pol.min_take);
    MARK(3);
    k_plan<<<(nsub + K4_WARPS - 1) / K4_WARPS, K4_WARPS * 32, k4_smem(), st>>>(J->hist, J->info, J->plan, nsub, pol.blocks);
    MARK(4);
    if (J->scan_wait) { cudaError_t e = cudaStreamWaitEvent(st, J->scan_wait, 0); if (e != cudaSuccess) return e; }
    k_scan<<<1, 1024, 0, st>>>(J->plan, J->info, nsub, nblocks, J->block_off, J->total);
    if (J->scan_done) { cudaError_t e = cudaEventRecord(J->scan_done, st); if (e != cudaSuccess) return e; }
    MARK(5);
    k_emit<<<nsub, K6_THREADS, 0, st>>>(data, J->tok, J->info, J->plan, J->block_off, nsub,
                                        reinterpret_cast<uint32_t *>(J->out));
    MARK(6);
#undef MARK
    return cudaGetLastError();
}
