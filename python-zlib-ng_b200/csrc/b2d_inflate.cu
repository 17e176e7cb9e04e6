/* b2d_inflate.cu — block-parallel inflate of raw-DEFLATE streams written by this library
 * (the step on the other side of the wire; first version).
 *
 * The writer cuts its input into 128 KiB device blocks, each byte-aligned in the output and primed with the
 * previous 32 KiB of *input* (b2d_kernels.cu), and can report where every block starts in the compressed
 * stream (b2d_block_index).  With that index every block can be entropy-decoded on its own; what a block
 * cannot know is the content of its dictionary, i.e. the last 32 KiB its predecessor produces.  So decoding
 * runs in three stages (the two-stage scheme of parallel gzip readers, restated for one warp per block):
 *
 *   k_inf_decode   one warp per device block.  Lane 0 parses the DEFLATE block headers (stored / fixed /
 *                  dynamic: inflate.c's TYPE..CODELENS states, inftrees.c) and decodes symbols with a
 *                  10-bit / 9-bit first-level table in shared memory and the canonical count/offset walk
 *                  for longer codes; the warp executes every copy together.  Output is 16 bits per byte:
 *                  0..255 = the byte, 256 + k = "byte k of the 32 KiB before this block" — copies move these
 *                  markers around like bytes, so a block never waits for its predecessor.
 *   k_inf_tails    one cluster of 8 CTAs walks the blocks in order and resolves the last 32 KiB of each (one
 *                  gather per marker from the predecessor's resolved tail): the only serial part.  It is launched
 *                  right behind the decode, on its own stream, and follows it through one ready flag per block;
 *                  the decode never waits for it, so nothing depends on the two kernels running at the same time.
 *   k_inf_finish   one CTA per block resolves the rest of the block against the now final tails.
 *
 * Reference code this stands in for: zng_inflate (zlib-ng/inflate.c:474-1150), inflate_table
 * (inftrees.c:30-290), inflate_fast (inffast_tpl.h), as driven by zlib_ng.decompress / _GzipReader
 * (zlib_ngmodule.c:275-370, 2426-2637).  Streams from other writers carry no block index and are not
 * accepted (DESIGN.md §0 "Out of scope", §10); BGZF members are (k_inf_decode<INF_MEMBERS>).
 */
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <type_traits>

#include "b2d_common.h"
#include "b2d_inflate.h"

namespace cg = cooperative_groups;

#define FULL 0xffffffffu
#define INF_WARPS 4
#define INF_LL_BITS 10
#define INF_D_BITS 9
#define INF_TOK_EOB 0x7fffffffu
#define INF_TOK_ERR 0x7ffffffeu

struct inf_warp {
    uint16_t ll_tab[1 << INF_LL_BITS];   /* (symbol << 4) | code length, 0 = code longer than the table */
    uint16_t d_tab[1 << INF_D_BITS];
    uint16_t ll_sorted[288];             /* symbols ordered by (length, symbol): canonical order */
    uint16_t d_sorted[32];
    uint16_t ll_count[16];               /* codes per length */
    uint16_t d_count[16];
    uint8_t lens[288 + 32];
};

__constant__ uint16_t c_lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115,
                                     131, 163, 195, 227, 258};
__constant__ uint8_t c_lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
__constant__ uint16_t c_dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537,
                                     2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
__constant__ uint8_t c_dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
__constant__ uint8_t c_clorder[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

/* LSB-first bit reader over global memory (lane 0 only).  After refill() at least 33 bits are buffered: enough
 * for a length code with its extra bits (20) or a distance code with its extra bits (28). */
struct bitrd {
    const uint8_t *in;
    uint64_t pos, lim;       /* next unread byte, end of the stream (bytes beyond read as zero) */
    uint64_t buf;
    uint32_t cnt;
};
__device__ __forceinline__ void br_init(bitrd &r, const uint8_t *in, uint64_t pos, uint64_t lim) {
    r.in = in; r.pos = pos; r.lim = lim; r.buf = 0; r.cnt = 0;
    while ((reinterpret_cast<uintptr_t>(in + r.pos) & 3u) && r.cnt < 24u) {
        const uint32_t b = r.pos < lim ? in[r.pos] : 0u;
        r.buf |= (uint64_t)b << r.cnt; r.cnt += 8; r.pos++;
    }
}
__device__ __forceinline__ void br_refill(bitrd &r) {
    if (r.cnt <= 32u) {
        uint32_t w;
        if (r.pos + 4 <= r.lim) w = *reinterpret_cast<const uint32_t *>(r.in + r.pos);
        else {
            w = 0;
            for (uint32_t i = 0; i < 4; ++i) if (r.pos + i < r.lim) w |= (uint32_t)r.in[r.pos + i] << (8 * i);
        }
        r.buf |= (uint64_t)w << r.cnt; r.cnt += 32; r.pos += 4;
    }
}
__device__ __forceinline__ uint32_t br_take(bitrd &r, uint32_t n) {   /* n <= 16, bits already buffered */
    const uint32_t v = (uint32_t)r.buf & ((1u << n) - 1u);
    r.buf >>= n; r.cnt -= n;
    return v;
}

/* canonical decode of the code at the low end of `bits` (at most maxlen bits): count / sorted as built by
 * inf_sort; returns the symbol and its length, or -1 (puff-style walk; inftrees.c builds tables from the same
 * counts and offsets) */
__device__ __forceinline__ int inf_canon(uint32_t bits, const uint16_t *count, const uint16_t *sorted, int maxlen, uint32_t *len_out) {
    int code = 0, first = 0, index = 0;
    for (int len = 1; len <= maxlen; ++len) {
        code |= (int)(bits & 1u); bits >>= 1;
        const int c = count[len];
        if (code - c < first) { *len_out = (uint32_t)len; return sorted[index + (code - first)]; }
        index += c; first += c; first <<= 1; code <<= 1;
    }
    return -1;
}
/* count[] and sorted[] of n code lengths (lane 0) */
__device__ __forceinline__ void inf_sort(const uint8_t *lens, int n, uint16_t *count, uint16_t *sorted) {
    uint16_t offs[16];
    for (int l = 0; l < 16; ++l) count[l] = 0;
    for (int s = 0; s < n; ++s) count[lens[s]]++;
    count[0] = 0;
    offs[1] = 0;
    for (int l = 1; l < 15; ++l) offs[l + 1] = (uint16_t)(offs[l] + count[l]);
    for (int s = 0; s < n; ++s) if (lens[s]) sorted[offs[lens[s]]++] = (uint16_t)s;
}
/* first-level table by the whole warp: entry i = the symbol whose code is a prefix of i */
__device__ __forceinline__ void inf_fill(uint16_t *tab, int tbits, const uint16_t *count, const uint16_t *sorted, uint32_t lane) {
    for (uint32_t i = lane; i < (1u << tbits); i += 32) {
        uint32_t len = 0;
        const int s = inf_canon(i, count, sorted, tbits, &len);
        tab[i] = s >= 0 ? (uint16_t)(((uint32_t)s << 4) | len) : (uint16_t)0;
    }
}

/* What one warp decodes (MODE):
 *   INF_BLOCKS   a 128 KiB device block of this library's stream: output bytes [b * 128 KiB, ...), may look 32 KiB behind
 *                its own start (markers), writes 16-bit symbols;
 *   INF_MEMBERS  an independent member (BGZF: its own DEFLATE stream, no dictionary): output [out_index[b],
 *                out_index[b + 1]), nothing may reach behind it, bytes are written as bytes, no resolving pass;
 *   INF_SUBS     one 32 KiB sub-block (= one DEFLATE block) of a device block, found through its bit offset inside the
 *                block (sub_bits[4 b + s]): four times the parallelism and a quarter of the latency of INF_BLOCKS.
 *                Markers then mean "byte k of the 32 KiB before this SUB-block"; the warp that finishes a block's last
 *                sub-block rewrites those of sub-blocks 1.. against their predecessors, after which the block looks
 *                exactly as INF_BLOCKS leaves it.
 * A launch covers blocks [b0, b0 + nb). */
#define INF_BLOCKS 0
#define INF_MEMBERS 1
#define INF_SUBS 2
#define INF_READY 0x80000000u        /* flags[b]: the block's symbols are final (low bits: sub-blocks decoded so far) */
template <int MODE>
__global__ void __launch_bounds__(INF_WARPS * 32)
k_inf_decode(const uint8_t *__restrict__ in, uint64_t in_n, const uint64_t *__restrict__ index,
             const uint64_t *__restrict__ out_index, const uint32_t *__restrict__ sub_bits, uint32_t b0, uint32_t nb, uint64_t raw_n,
             uint32_t dict_n, void *outp, uint32_t *flags, uint32_t *err) {
    constexpr bool INDEP = MODE == INF_MEMBERS, SUBS = MODE == INF_SUBS;
    typedef typename std::conditional<INDEP, uint8_t, uint16_t>::type sym_t;
    __shared__ inf_warp s_w[INF_WARPS];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t unit = blockIdx.x * INF_WARPS + warp;
    const uint32_t b = b0 + (SUBS ? unit / B2D_NSUB : unit), sub = SUBS ? unit % B2D_NSUB : 0u;
    if (b >= b0 + nb) return;
    inf_warp *T = &s_w[warp];
    const uint64_t blk_start = INDEP ? out_index[b] : (uint64_t)b * B2D_BLOCK;
    const uint32_t blk_len = INDEP ? (uint32_t)(out_index[b + 1] - blk_start) : (uint32_t)(raw_n - blk_start < B2D_BLOCK ? raw_n - blk_start : B2D_BLOCK);
    const uint32_t nsub_b = (blk_len + B2D_SUB - 1) / B2D_SUB;            /* sub-blocks this block has (SUBS) */
    if (SUBS && sub >= nsub_b) return;
    const uint64_t bs = blk_start + (uint64_t)sub * B2D_SUB;
    const uint32_t expect = SUBS ? (blk_len - sub * B2D_SUB < B2D_SUB ? blk_len - sub * B2D_SUB : B2D_SUB) : blk_len;
    /* how far behind its own start this unit may reach: 32 KiB, or the caller's dictionary at the start of the stream */
    const uint32_t reach = INDEP ? 0u : ((b == 0 && sub == 0) ? dict_n : B2D_WINDOW);
    sym_t *o16 = reinterpret_cast<sym_t *>(outp) + bs;
    bitrd R;
    {
        const uint32_t bit0 = SUBS ? sub_bits[(uint64_t)b * B2D_NSUB + sub] : 0u;
        br_init(R, in, index[b] + (bit0 >> 3), in_n);
        if (SUBS && (bit0 & 7u)) { br_refill(R); br_take(R, bit0 & 7u); }
    }
    uint32_t produced = 0, fail = 0;

    while (produced < expect && !fail) {
        /* ---- block header (lane 0): type, and for dynamic blocks the code lengths into T->lens */
        uint32_t type = 0, st_len = 0, bfinal = 0;
        uint64_t st_off = 0;
        if (lane == 0) {
            br_refill(R);
            bfinal = br_take(R, 1);                         /* BFINAL: our writer never sets it on data blocks; members end with it */
            type = br_take(R, 2);
            if (type == 0) {
                br_take(R, R.cnt & 7u);                     /* to the next byte boundary */
                br_refill(R);
                const uint32_t len = br_take(R, 16), nlen = br_take(R, 16);
                if ((len ^ nlen) != 0xffffu) type = 4;
                st_len = len;
                st_off = R.pos - (R.cnt >> 3);
            } else if (type == 2) {
                br_refill(R);
                const uint32_t hlit = br_take(R, 5) + 257u, hdist = br_take(R, 5) + 1u, hclen = br_take(R, 4) + 4u;
                if (hlit > 286u || hdist > 30u) type = 4;
                else {
                    uint8_t *cl = T->lens + 288;            /* 19 code-length code lengths, parked in the distance area */
                    for (uint32_t i = 0; i < 19; ++i) cl[i] = 0;
                    for (uint32_t i = 0; i < hclen; ++i) { br_refill(R); cl[c_clorder[i]] = (uint8_t)br_take(R, 3); }
                    inf_sort(cl, 19, T->d_count, T->d_sorted);
                    uint32_t idx = 0;
                    const uint32_t total = hlit + hdist;
                    uint8_t *L = T->lens;                   /* lengths of both alphabets back to back (<= 316), split below;
                                                             * the parked 19 bytes may be overwritten: inf_sort has read them */
                    while (idx < total) {
                        br_refill(R);
                        uint32_t l = 0;
                        const int s = inf_canon((uint32_t)R.buf & 0x7fu, T->d_count, T->d_sorted, 7, &l);
                        if (s < 0) { type = 4; break; }
                        br_take(R, l);
                        if (s < 16) { L[idx++] = (uint8_t)s; continue; }
                        uint32_t rep, val = 0;
                        if (s == 16) { if (idx == 0) { type = 4; break; } val = L[idx - 1]; rep = 3u + br_take(R, 2); }
                        else if (s == 17) rep = 3u + br_take(R, 3);
                        else rep = 11u + br_take(R, 7);
                        if (idx + rep > total) { type = 4; break; }
                        for (uint32_t i = 0; i < rep; ++i) L[idx++] = (uint8_t)val;
                    }
                    if (type == 2) {
                        /* move the distance lengths to their own area (backwards: the areas may overlap), pad both with zeros */
                        for (int i = (int)hdist - 1; i >= 0; --i) L[288 + i] = L[hlit + (uint32_t)i];
                        for (uint32_t i = hdist; i < 32; ++i) L[288 + i] = 0;
                        for (uint32_t i = hlit; i < 288; ++i) L[i] = 0;
                    }
                }
            } else if (type == 3) type = 4;
        }
        type = __shfl_sync(FULL, type, 0);
        bfinal = __shfl_sync(FULL, bfinal, 0);
        if (type == 4) { fail = 1; break; }
        if (type == 0) {
            st_len = __shfl_sync(FULL, st_len, 0);
            st_off = __shfl_sync(FULL, st_off, 0);
            if (produced + st_len > expect || st_off + st_len > in_n) { fail = 2; break; }
            for (uint32_t i = lane; i < st_len; i += 32) o16[produced + i] = (sym_t)in[st_off + i];
            produced += st_len;
            if (lane == 0) br_init(R, in, st_off + st_len, in_n);
            __syncwarp();
            if (INDEP && bfinal) break;
            continue;
        }
        if (type == 1) {   /* fixed code: trees.c static_ltree / RFC 1951 3.2.6 */
            for (uint32_t i = lane; i < 288; i += 32) T->lens[i] = i < 144 ? 8 : (i < 256 ? 9 : (i < 280 ? 7 : 8));
            T->lens[288 + lane] = lane < 30 ? 5 : 0;
        }
        __syncwarp();
        if (lane == 0) {
            inf_sort(T->lens, 288, T->ll_count, T->ll_sorted);
            inf_sort(T->lens + 288, 32, T->d_count, T->d_sorted);
        }
        __syncwarp();
        inf_fill(T->ll_tab, INF_LL_BITS, T->ll_count, T->ll_sorted, lane);
        inf_fill(T->d_tab, INF_D_BITS, T->d_count, T->d_sorted, lane);
        __syncwarp();

        /* ---- symbols: lane 0 decodes — literals straight to the output, as many as follow each other — and
         * hands the next length/distance pair (or the end of the block) to the warp, which copies together */
        for (;;) {
            uint32_t tok = 0, nlit = 0;
            if (lane == 0) {
                sym_t *lp = o16 + produced;
                const uint32_t room = expect - produced;
                for (;;) {
                    br_refill(R);
                    uint32_t e = T->ll_tab[(uint32_t)R.buf & ((1u << INF_LL_BITS) - 1u)], len = e & 15u;
                    int sym = (int)(e >> 4);
                    if (!e) sym = inf_canon((uint32_t)R.buf & 0x7fffu, T->ll_count, T->ll_sorted, 15, &len);
                    if (sym < 0) { tok = INF_TOK_ERR; break; }
                    br_take(R, len);
                    if (sym < 256) {
                        if (nlit >= room) { tok = INF_TOK_ERR; break; }
                        lp[nlit++] = (sym_t)sym;
                        continue;
                    }
                    if (sym == 256) { tok = INF_TOK_EOB; break; }
                    if (sym > 285) { tok = INF_TOK_ERR; break; }
                    const uint32_t ls = (uint32_t)sym - 257u;
                    const uint32_t mlen = c_lbase[ls] + br_take(R, c_lext[ls]);
                    br_refill(R);
                    e = T->d_tab[(uint32_t)R.buf & ((1u << INF_D_BITS) - 1u)]; len = e & 15u;
                    int ds = (int)(e >> 4);
                    if (!e) ds = inf_canon((uint32_t)R.buf & 0x7fffu, T->d_count, T->d_sorted, 15, &len);
                    if (ds < 0 || ds > 29) { tok = INF_TOK_ERR; break; }
                    br_take(R, len);
                    const uint32_t dist = c_dbase[ds] + br_take(R, c_dext[ds]);
                    tok = 0x80000000u | (mlen << 16) | (dist - 1u);
                    break;
                }
            }
            tok = __shfl_sync(FULL, tok, 0);
            nlit = __shfl_sync(FULL, nlit, 0);
            produced += nlit;
            if (tok == INF_TOK_EOB) break;
            if (tok == INF_TOK_ERR) { fail = 3; break; }
            const uint32_t mlen = (tok >> 16) & 0x1ffu, dist = (tok & 0x7fffu) + 1u;
            if (produced + mlen > expect || dist > produced + reach) { fail = 2; break; }
            __syncwarp();    /* lane 0's literals and the previous copy are visible to this copy's loads */
            const int src0 = (int)produced - (int)dist;
            for (uint32_t i = lane; i < mlen; i += 32) {
                const int rel = src0 + (int)(dist >= mlen ? i : i % dist);
                /* before the block: a marker for byte (32768 + rel) of the 32 KiB that precede it */
                const uint32_t v = rel < 0 ? 256u + (uint32_t)((int)B2D_WINDOW + rel) : (uint32_t)__ldcg(o16 + rel);
                o16[produced + i] = (sym_t)v;
            }
            produced += mlen;
            __syncwarp();
        }
        if (INDEP && bfinal) break;                         /* a member is over after its final block, whatever its ISIZE says */
    }
    if (!fail && produced != expect) fail = 2;
    if (fail) {
        /* the scratch is reused across calls: what this block did not decode would otherwise reach the resolving
         * passes as stale symbols (markers included) of an earlier call */
        for (uint32_t i = produced + lane; i < expect; i += 32) o16[i] = 0;
        if (lane == 0) atomicCAS(err, 0u, (fail << 28) | (b & 0x0fffffffu));
    }
    __syncwarp();
    if (INDEP) return;
    if (SUBS) {
        /* the warp that completes the block rewrites the markers of its sub-blocks 1.. in order: "byte k of the 32 KiB
         * before sub-block s" is symbol k of sub-block s - 1, which by then holds bytes or block-level markers only */
        uint32_t done = 0;
        if (lane == 0) { __threadfence(); done = atomicAdd(flags + b, 1u) + 1u; }
        done = __shfl_sync(FULL, done, 0);
        if (done != nsub_b) return;
        __threadfence();
        uint16_t *blk = reinterpret_cast<uint16_t *>(outp) + blk_start;
        for (uint32_t s2 = 1; s2 < nsub_b; ++s2) {
            const uint32_t len2 = blk_len - s2 * B2D_SUB < B2D_SUB ? blk_len - s2 * B2D_SUB : B2D_SUB;
            uint16_t *cur = blk + s2 * B2D_SUB;
            const uint16_t *prev = cur - B2D_SUB;
            /* eight groups of 32 symbols per trip: their loads (and the lookups of the markers among them) are in flight
             * together — one group per trip waited an L2 round trip per 32 symbols, ~0.9 ms per block on the way to its
             * ready flag */
            for (uint32_t i0 = 0; i0 < len2; i0 += 8u * 32u) {
                uint32_t v[8];
#pragma unroll
                for (uint32_t q = 0; q < 8; ++q) { const uint32_t i = i0 + 32u * q + lane; v[q] = i < len2 ? (uint32_t)__ldcg(cur + i) : 0u; }
#pragma unroll
                for (uint32_t q = 0; q < 8; ++q) if (v[q] >= 256u) v[q] = 0x10000u | (uint32_t)__ldcg(prev + (v[q] - 256u));
#pragma unroll
                for (uint32_t q = 0; q < 8; ++q) if (v[q] & 0x10000u) cur[i0 + 32u * q + lane] = (uint16_t)v[q];
            }
            __syncwarp();
            __threadfence();
        }
    }
    /* the block's symbols are in memory: k_inf_tails (running beside this kernel) may pick its tail up */
    if (lane == 0) { __threadfence(); *reinterpret_cast<volatile uint32_t *>(flags + b) = INF_READY; }
}

/* the byte a marker stands for: `k` indexes the 32 KiB that precede the block starting at output byte bs */
__device__ __forceinline__ uint32_t inf_behind(const uint8_t *out8, const uint8_t *dict, uint32_t dict_n, uint64_t bs, uint32_t k) {
    /* the decoder only emits markers with 32768 - reach <= k < 32768; the tests below keep a marker that did not come
     * from it (a failed block, see k_inf_decode) inside the buffers */
    if (k >= B2D_WINDOW) return 0u;
    if (bs == 0) return k + dict_n >= B2D_WINDOW ? dict[dict_n - B2D_WINDOW + k] : 0u;
    return __ldcg(out8 + (bs - B2D_WINDOW + k));
}

/* eight symbols -> eight bytes; markers are looked up in the resolved 32 KiB before the block at bs */
__device__ __forceinline__ uint2 inf_resolve8(uint4 q, const uint8_t *out8, const uint8_t *dict, uint32_t dict_n, uint64_t bs) {
    uint32_t s[8] = {q.x & 0xffffu, q.x >> 16, q.y & 0xffffu, q.y >> 16, q.z & 0xffffu, q.z >> 16, q.w & 0xffffu, q.w >> 16};
    if ((q.x | q.y | q.z | q.w) & 0xff00ff00u) {
#pragma unroll
        for (int i = 0; i < 8; ++i) if (s[i] >= 256u) s[i] = inf_behind(out8, dict, dict_n, bs, s[i] - 256u);
    }
    uint2 o;
    o.x = s[0] | (s[1] << 8) | (s[2] << 16) | (s[3] << 24);
    o.y = s[4] | (s[5] << 8) | (s[6] << 16) | (s[7] << 24);
    return o;
}

/* The chain over the blocks is serial, and one SM needs ~5 us of instruction issue to look 32768 symbols up.  A
 * thread-block cluster spreads every step over INF_CLUSTER SMs: a thread owns eight symbols of each tail, the next
 * block's symbols are loaded while this block's markers are looked up, and a step ends with one cluster barrier
 * (release / acquire at cluster scope: block b's tail is final before block b+1 looks into it). */
#define INF_CLUSTER 8
#define INF_TAIL_THREADS (B2D_WINDOW / 8 / INF_CLUSTER)      /* 512 */
#define INF_READY_GROUP 32u                                  /* decode flags are polled for this many blocks at a time */
__global__ void __launch_bounds__(INF_TAIL_THREADS)
k_inf_tails(const uint16_t *out16, uint8_t *out8, const uint8_t *__restrict__ dict, uint32_t dict_n, uint32_t b0, uint32_t b1,
            uint32_t nblocks, uint64_t raw_n, const uint32_t *flags, uint32_t *err) {
    /* blocks [b0, b1) of the stream's nblocks: the tail of block b0 - 1 was resolved by the launch before this one */
    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t t = blockIdx.x * INF_TAIL_THREADS + threadIdx.x;          /* 0 .. 4095: one group of 8 symbols */
    const uint32_t nfull = (uint32_t)(raw_n / B2D_BLOCK);          /* blocks of 128 KiB: their tails are aligned */
    const uint64_t toff = (B2D_BLOCK - B2D_WINDOW) + 8u * t;
    /* This kernel is launched AFTER k_inf_decode, on a second stream, and follows it block by block: blocks
     * [first, last) must have been decoded before their symbols are read.  The decode never waits for this kernel,
     * so the flags always arrive — at once when the two kernels cannot run side by side (profilers,
     * CUDA_LAUNCH_BLOCKING: the decode has finished by the time this one starts).  The bound below is a hang guard
     * for a decode that died; it sets err[1], which the host answers by running this pass again after the decode
     * (never by reporting bad data). */
    auto wait_ready = [&](uint32_t first, uint32_t last) {
        bool ok = true;
        if (threadIdx.x < last - first && *reinterpret_cast<volatile uint32_t *>(err + 1) == 0u) {
            const volatile uint32_t *f = flags + first + threadIdx.x;
            uint32_t spins = 0;
            while ((*f & INF_READY) == 0u) {
                __nanosleep(1000);
                if (++spins > 4000000u) { ok = false; break; }
            }
        }
        if (!__syncthreads_and(ok) && threadIdx.x == 0) atomicCAS(err + 1, 0u, 1u + first);
        __threadfence();
    };
    auto load = [&](uint32_t b) { return __ldcg(reinterpret_cast<const uint4 *>(out16 + (uint64_t)b * B2D_BLOCK + toff)); };
    uint4 cur = make_uint4(0, 0, 0, 0), nxt = cur;
    const uint32_t e_full = b1 < nfull ? b1 : nfull;              /* end of the full blocks of this launch */
    if (b0 < e_full) {
        const uint32_t g = (b0 / INF_READY_GROUP + 1u) * INF_READY_GROUP;
        wait_ready(b0, g < e_full ? g : e_full);
        cur = load(b0);
    }
    for (uint32_t b = b0; b < e_full; ++b) {
        const uint64_t bs = (uint64_t)b * B2D_BLOCK;
        if (b + 1 < e_full) {
            if ((b + 1) % INF_READY_GROUP == 0) wait_ready(b + 1, b + 1 + INF_READY_GROUP < e_full ? b + 1 + INF_READY_GROUP : e_full);
            nxt = load(b + 1);
        }
        *reinterpret_cast<uint2 *>(out8 + bs + toff) = inf_resolve8(cur, out8, dict, dict_n, bs);
        cluster.sync();
        cur = nxt;
    }
    if (b1 == nblocks && nfull < nblocks) {   /* the last, shorter block: nobody depends on it */
        wait_ready(nfull, nblocks);
        const uint64_t bs = (uint64_t)nfull * B2D_BLOCK;
        const uint64_t ts = raw_n - bs > B2D_WINDOW ? raw_n - B2D_WINDOW : bs;
        for (uint64_t p = ts + t; p < raw_n; p += INF_CLUSTER * INF_TAIL_THREADS) {
            const uint32_t s = __ldcg(out16 + p);
            out8[p] = (uint8_t)(s < 256u ? s : inf_behind(out8, dict, dict_n, bs, s - 256u));
        }
    }
}

__global__ void __launch_bounds__(256)
k_inf_finish(const uint16_t *__restrict__ out16, uint8_t *out8, const uint8_t *__restrict__ dict, uint32_t dict_n,
             uint64_t raw_n, uint32_t b0) {
    const uint64_t bs = (uint64_t)(b0 + blockIdx.x) * B2D_BLOCK;
    const uint64_t be = raw_n - bs < B2D_BLOCK ? raw_n : bs + B2D_BLOCK;
    const uint64_t ts = be - bs > B2D_WINDOW ? be - B2D_WINDOW : bs;      /* [ts, be) was written by k_inf_tails */
    for (uint64_t p = bs + 8ull * threadIdx.x; p < ts; p += 8ull * 256) {
        /* a group of 8 that straddles ts rewrites a few tail bytes with the same values */
        const uint4 q = *reinterpret_cast<const uint4 *>(out16 + p);
        *reinterpret_cast<uint2 *>(out8 + p) = inf_resolve8(q, out8, dict, dict_n, bs);
    }
}

static cudaError_t launch_tails(const uint16_t *d_out16, uint8_t *d_out, const uint8_t *d_dict, uint32_t dict_n, uint32_t b0, uint32_t b1,
                                uint32_t nblocks, uint64_t raw_n, const uint32_t *d_flags, uint32_t *d_err, cudaStream_t st) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(INF_CLUSTER); cfg.blockDim = dim3(INF_TAIL_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = INF_CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, k_inf_tails, d_out16, d_out, d_dict, dict_n, b0, b1, nblocks, raw_n, d_flags, d_err);
}

/* Dynamic shared memory that k_inf_decode asks for without using it: with it 12 of its CTAs (48 warps) fit an SM
 * instead of 13, which leaves the 16 warps one CTA of the tails cluster needs free on every SM — the serial pass
 * finds room beside the decode whenever the device can run two kernels at once. */
#define INF_DECODE_PAD 1024

/* One chunk: blocks [b0, b0 + nb) of a stream of nblocks; chunks are issued in order (the tails of chunk c build on
 * those of chunk c - 1).  The caller has cleared d_flags / d_err for the whole stream and made both streams wait for that. */
cudaError_t b2d_launch_inflate(const uint8_t *d_in, uint64_t in_n, const uint64_t *d_index, const uint32_t *d_sub_bits, uint32_t b0, uint32_t nb,
                               uint32_t nblocks, const uint8_t *d_dict, uint32_t dict_n, uint16_t *d_out16, uint8_t *d_out,
                               uint64_t raw_n, uint32_t *d_flags, uint32_t *d_err, cudaStream_t st_dec, cudaStream_t st_res, int mode) {
    if (nb == 0) return cudaSuccess;
    cudaError_t e;
    if (mode != B2D_INF_RESOLVE_ONLY) {
        /* The decode goes first (in host order) and waits for nobody. */
        if (d_sub_bits)
            k_inf_decode<INF_SUBS><<<(nb * B2D_NSUB + INF_WARPS - 1) / INF_WARPS, INF_WARPS * 32, INF_DECODE_PAD, st_dec>>>(
                d_in, in_n, d_index, nullptr, d_sub_bits, b0, nb, raw_n, dict_n, d_out16, d_flags, d_err);
        else
            k_inf_decode<INF_BLOCKS><<<(nb + INF_WARPS - 1) / INF_WARPS, INF_WARPS * 32, INF_DECODE_PAD, st_dec>>>(
                d_in, in_n, d_index, nullptr, nullptr, b0, nb, raw_n, dict_n, d_out16, d_flags, d_err);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
        if (mode == B2D_INF_DECODE_ONLY) return cudaSuccess;
    }
    /* The serial pass over the tails is launched right behind it — on its own (high-priority) stream when the device
     * runs kernels side by side, where it follows the decode through d_flags — and the rest of each block is resolved
     * on that stream too.  B2D_INF_SERIAL / B2D_INF_RESOLVE_ONLY: the caller passes st_res == st_dec. */
    if ((e = launch_tails(d_out16, d_out, d_dict, dict_n, b0, b0 + nb, nblocks, raw_n, d_flags, d_err, st_res)) != cudaSuccess) return e;
    k_inf_finish<<<nb, 256, 0, st_res>>>(d_out16, d_out, d_dict, dict_n, raw_n, b0);
    return cudaGetLastError();
}

/* independent members (BGZF): every block decodes straight to bytes, nothing to resolve */
cudaError_t b2d_launch_inflate_members(const uint8_t *d_in, uint64_t in_n, const uint64_t *d_index, const uint64_t *d_out_index,
                                       uint32_t nblocks, uint8_t *d_out, uint64_t raw_n, uint32_t *d_err, cudaStream_t st) {
    if (nblocks == 0) return cudaSuccess;
    k_inf_decode<INF_MEMBERS><<<(nblocks + INF_WARPS - 1) / INF_WARPS, INF_WARPS * 32, 0, st>>>(d_in, in_n, d_index, d_out_index, nullptr, 0, nblocks,
                                                                                              raw_n, 0, d_out, nullptr, d_err);
    return cudaGetLastError();
}
