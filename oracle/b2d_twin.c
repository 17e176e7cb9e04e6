/* oracle/b2d_twin.c — TEST INFRASTRUCTURE ONLY.
 *
 * CPU twin of the B200 pipeline (python-zlib-ng_b200/csrc/b2d_kernels.cu):
 * the same six stages, executed sequentially, sharing the integer helpers in
 * b2d_common.h / b2d_huff.h, so that a GPU run can be compared with it
 * byte-for-byte and stage-by-stage (chains, matches, tokens, plans, bytes).
 * It is NOT a restatement of the reference (that is oracle/oracle.c and
 * oracle/_ref) and it is never imported by the product: only tests/ use it.
 *
 *   stage 1  chains    prev-distance per position (hash of 5 bytes, 2^14 heads)
 *   stage 2  matches   best (len, dist) per position, chain depth by level
 *   stage 3  parse     lazy parse per 1 KiB unit -> tokens + per-sub-block histograms
 *   stage 4  plan      Huffman lengths / header / stored-fixed-dynamic choice
 *   stage 5  scan      bit offsets of sub-blocks, byte offsets of blocks
 *   stage 6  emit      bit packing, sync-flush marker per device block
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../python-zlib-ng_b200/csrc/b2d_common.h"
#include "../python-zlib-ng_b200/csrc/b2d_huff.h"

static inline uint32_t ld32(const uint8_t *p) { uint32_t v; memcpy(&v, p, 4); return v; }

/* stage 1: block b of the job buf[0 .. D+n), data starts at D */
static void twin_chains(const uint8_t *buf, size_t D, size_t n, size_t b, uint16_t *prevd, uint16_t *head, int hash_bits) {
    size_t job_end = D + n;
    size_t start = D + b * (size_t)B2D_BLOCK;
    size_t end = start + B2D_BLOCK < job_end ? start + B2D_BLOCK : job_end;
    size_t halo = start < B2D_WINDOW ? start : B2D_WINDOW;
    size_t span_start = start - halo;
    memset(head, 0xff, sizeof(uint16_t) * B2D_HASH_SIZE_MAX);
    for (size_t p = span_start; p < end; p++) {
        uint32_t r = (uint32_t)(p - span_start);
        uint16_t out = 0;
        if (p + 5 <= end) {            /* the hash never looks past the block: blocks are independent of what follows */
            uint32_t h = b2d_hash5(ld32(buf + p), buf[p + 4], hash_bits);
            uint32_t d = (r - head[h]) & 0xffffu;
            if (d != 0 && d <= r && d <= B2D_WINDOW) out = (uint16_t)d;
            head[h] = (uint16_t)(r & 0xffffu);
        }
        prevd[r] = out;      /* per-block links, indexed from span_start: a block depends on nothing but its own span */
    }
}

/* stage 2 */
static void twin_matches(const uint8_t *buf, size_t D, size_t n, size_t b, const uint16_t *prevd,
                         uint32_t *match, b2d_policy pol, uint32_t max_dist) {
    size_t job_end = D + n;
    size_t start = D + b * (size_t)B2D_BLOCK;
    size_t end = start + B2D_BLOCK < job_end ? start + B2D_BLOCK : job_end;
    size_t span_start = start - (start < B2D_WINDOW ? start : B2D_WINDOW);
    prevd -= span_start;                               /* prevd[q] for absolute q >= span_start */
    for (size_t p = start; p < end; p++) {
        size_t uend = ((p - D) / B2D_UNIT + 1) * (size_t)B2D_UNIT + D;   /* no match crosses a parse unit */
        size_t lim = uend < end ? uend : end;
        uint32_t maxl = lim - p < B2D_MAX_MATCH ? (uint32_t)(lim - p) : B2D_MAX_MATCH;
        uint32_t best = 0, bd = 0;
        if (maxl >= B2D_MIN_MATCH) {
            size_t cand = p;
            uint32_t d = pol.search == B2D_SEARCH_RLE ? (p >= 1 ? 1u : 0u) : prevd[p];
            for (int step = 0; step < pol.depth && d != 0; step++) {
                cand -= d;
                size_t dist = p - cand;
                if (dist > max_dist) break;
                uint32_t l = 0;
                while (l < maxl && buf[cand + l] == buf[p + l]) l++;
                if (l > best) { best = l; bd = (uint32_t)dist; if (l >= (uint32_t)pol.nice || l == maxl) break; }
                d = pol.search == B2D_SEARCH_RLE ? 0u : prevd[cand];
            }
        }
        match[p - D] = best >= B2D_MIN_MATCH ? b2d_pack_match(best, bd, buf[p]) : (uint32_t)buf[p];
    }
}

/* stage 3: sub-block k of the job is parsed on its own (every sub-block is a
 * whole number of parse units): tokens go to tok[k*SUB ...], the summary to
 * info[k], histograms to hist[k*316 ...] (286 lit/len then 30 dist) */
static void twin_parse(const uint8_t *buf, size_t D, size_t n, size_t k, const uint32_t *match,
                       uint32_t *tok, b2d_subinfo *info, uint32_t *hist, b2d_policy pol) {
    size_t start = k * (size_t)B2D_SUB;               /* relative to D */
    uint32_t len = (uint32_t)(n - start < B2D_SUB ? n - start : B2D_SUB);
    const uint32_t *m = match + start;
    uint32_t *t = tok + start;
    uint32_t *H = hist + k * (size_t)B2D_HIST;
    uint32_t nt = 0;
    memset(H, 0, sizeof(uint32_t) * B2D_HIST);
    for (uint32_t u = 0; u < len; u += B2D_UNIT) {
        uint32_t uend = u + B2D_UNIT < len ? u + B2D_UNIT : len;
        uint32_t i = u;
        while (i < uend) {
            uint32_t l0 = b2d_match_len(m[i]);
            uint32_t l1 = i + 1 < len ? b2d_match_len(m[i + 1]) : 0;
            uint32_t l2 = i + 2 < len ? b2d_match_len(m[i + 2]) : 0;
            if (l0 < (uint32_t)pol.min_take) l0 = 0;      /* Z_FILTERED: short matches count as none */
            if (l1 < (uint32_t)pol.min_take) l1 = 0;
            if (l2 < (uint32_t)pol.min_take) l2 = 0;
            uint32_t step = b2d_parse_step(l0, l1, l2, pol.lazy);
            if (step > 1) {
                uint32_t dist = b2d_match_dist(m[i]);
                uint32_t ls, le, lv, ds, de, dv;
                b2d_len_sym(step - 3, &ls, &le, &lv);
                b2d_dist_sym(dist - 1, &ds, &de, &dv);
                H[257 + ls]++; H[B2D_LL_CODES + ds]++;
                t[nt++] = b2d_tok_match(step, dist);
            } else {
                uint8_t c = buf[D + start + i];
                H[c]++;
                t[nt++] = c;
            }
            i += step;
        }
    }
    info[k].in_len = len; info[k].tok_count = nt; info[k].bit_off = 0; info[k].pad = 0;
    if (nt) H[256] = 1;
}

/* ---- stage 6 helpers: LSB-first bit writer into a zeroed byte buffer ---- */
typedef struct { uint8_t *out; uint64_t bitpos; } bitw;
static void put(bitw *w, uint64_t value, uint32_t count) {
    for (uint32_t i = 0; i < count; i++) {
        if ((value >> i) & 1u) w->out[w->bitpos >> 3] |= (uint8_t)(1u << (w->bitpos & 7));
        w->bitpos++;
    }
}

typedef struct {
    uint16_t *prevd;      /* [BLOCK + WINDOW] links of the block being processed */
    uint32_t *match;      /* [n]                */
    uint32_t *tok;        /* [nsub * SUB]       */
    b2d_subinfo *info;    /* [nsub]             */
    uint32_t *hist;       /* [nsub * 316]       */
    b2d_plan *plan;       /* [nsub]             */
    uint64_t *block_off;  /* [nblocks + 1] byte offsets */
    size_t nblocks;
} twin_state;

/* Run all stages. buf = [dict (D bytes) | data (n bytes)], readable for B2D_IN_PAD more.
 * Returns compressed size, or 0 if cap is too small. If st_out != NULL the
 * intermediate arrays are handed to the caller (free with b2d_twin_free). */
size_t b2d_twin_deflate_ex(const uint8_t *buf, size_t D, size_t n, int level, uint32_t max_dist, uint8_t *out, size_t cap,
                           twin_state **st_out) {
    b2d_policy pol = b2d_level_policy(level);
    size_t nb = (n + B2D_BLOCK - 1) / B2D_BLOCK;
    size_t nsub = (n + B2D_SUB - 1) / B2D_SUB;
    twin_state *S = (twin_state *)calloc(1, sizeof *S);
    S->nblocks = nb;
    S->prevd = (uint16_t *)calloc(B2D_BLOCK + B2D_WINDOW + 8, sizeof(uint16_t));   /* one block's span at a time */
    S->match = (uint32_t *)calloc(n + 8, sizeof(uint32_t));
    S->tok = (uint32_t *)calloc(nb * (size_t)B2D_BLOCK + 8, sizeof(uint32_t));
    S->info = (b2d_subinfo *)calloc(nb * B2D_NSUB + 1, sizeof(b2d_subinfo));
    S->hist = (uint32_t *)calloc((nb * B2D_NSUB + 1) * (size_t)(B2D_LL_CODES + B2D_D_CODES), sizeof(uint32_t));
    S->plan = (b2d_plan *)calloc(nb * B2D_NSUB + 1, sizeof(b2d_plan));
    S->block_off = (uint64_t *)calloc(nb + 1, sizeof(uint64_t));
    uint16_t *head = (uint16_t *)malloc(sizeof(uint16_t) * B2D_HASH_SIZE_MAX);
    b2d_huff_scratch *scr = (b2d_huff_scratch *)malloc(sizeof *scr);
    if (pol.search != B2D_SEARCH_NONE) {
        for (size_t b = 0; b < nb; b++) {
            if (pol.search == B2D_SEARCH_CHAINS) twin_chains(buf, D, n, b, S->prevd, head, pol.hash_bits);
            twin_matches(buf, D, n, b, S->prevd, S->match, pol, max_dist);
        }
    }
    for (size_t k = 0; k < nsub; k++) twin_parse(buf, D, n, k, S->match, S->tok, S->info, S->hist, pol);
    for (size_t k = 0; k < nsub; k++) {
        if (!S->info[k].tok_count) continue;
        const uint32_t *H = S->hist + k * (size_t)B2D_HIST;
        b2d_plan_subblock(H, H + B2D_LL_CODES, S->info[k].in_len, &S->plan[k], scr, pol.blocks);
    }
    /* stage 5: sizes */
    uint64_t off = 0;
    for (size_t b = 0; b < nb; b++) {
        S->block_off[b] = off;
        uint64_t bits = 0;
        for (uint32_t s = 0; s < B2D_NSUB; s++) {
            size_t k = b * B2D_NSUB + s;
            if (k >= nsub || !S->info[k].tok_count) continue;
            S->info[k].bit_off = (uint32_t)bits;
            bits += S->plan[k].hdr_bits;
            if (S->plan[k].type == 0) bits = ((bits + 7) & ~7ull) + 32 + 8ull * S->info[k].in_len;
            else bits += S->plan[k].body_bits;
        }
        bits += 3; bits = (bits + 7) & ~7ull; bits += 32;
        off += bits / 8;
    }
    S->block_off[nb] = off;
    size_t result = 0;
    if (off <= cap) {
        memset(out, 0, off);
        /* stage 6 */
        for (size_t b = 0; b < nb; b++) {
            bitw w = { out, S->block_off[b] * 8 };
            for (uint32_t s = 0; s < B2D_NSUB; s++) {
                size_t k = b * B2D_NSUB + s;
                if (k >= nsub) continue;
                const b2d_subinfo *I = &S->info[k];
                const b2d_plan *P = &S->plan[k];
                if (!I->tok_count) continue;
                for (uint32_t i = 0; i < P->hdr_bits; i++) put(&w, (P->hdr[i >> 5] >> (i & 31)) & 1u, 1);
                if (P->type == 0) {
                    w.bitpos = (w.bitpos + 7) & ~7ull;
                    put(&w, I->in_len, 16); put(&w, (~I->in_len) & 0xffffu, 16);
                    memcpy(out + (w.bitpos >> 3), buf + D + k * (size_t)B2D_SUB, I->in_len);
                    w.bitpos += 8ull * I->in_len;
                    continue;
                }
                const uint32_t *t = S->tok + k * (size_t)B2D_SUB;
                for (uint32_t i = 0; i < I->tok_count; i++) {
                    uint32_t x = t[i];
                    if (!b2d_tok_is_match(x)) { uint32_t c = P->ll_code[x]; put(&w, c & 0xffffu, c >> 16); continue; }
                    uint32_t ls, le, lv, ds, de, dv;
                    b2d_len_sym(b2d_tok_len3(x), &ls, &le, &lv);
                    b2d_dist_sym(b2d_tok_dist1(x), &ds, &de, &dv);
                    uint32_t c = P->ll_code[257 + ls]; put(&w, c & 0xffffu, c >> 16); put(&w, lv, le);
                    c = P->d_code[ds]; put(&w, c & 0xffffu, c >> 16); put(&w, dv, de);
                }
                { uint32_t c = P->ll_code[256]; put(&w, c & 0xffffu, c >> 16); }
            }
            put(&w, 0, 3); w.bitpos = (w.bitpos + 7) & ~7ull;
            put(&w, 0x0000u, 16); put(&w, 0xffffu, 16);
        }
        result = off;
    }
    free(head); free(scr);
    if (st_out) *st_out = S;
    else {
        free(S->prevd); free(S->match); free(S->tok); free(S->info); free(S->hist); free(S->plan); free(S->block_off); free(S);
    }
    return result;
}

size_t b2d_twin_deflate(const uint8_t *buf, size_t D, size_t n, int level, uint8_t *out, size_t cap) {
    return b2d_twin_deflate_ex(buf, D, n, level, B2D_WINDOW, out, cap, NULL);
}

void b2d_twin_free(twin_state *S) {
    if (!S) return;
    free(S->prevd); free(S->match); free(S->tok); free(S->info); free(S->hist); free(S->plan); free(S->block_off); free(S);
}
/* accessors for ctypes-side stage comparison */
const uint32_t *b2d_twin_match(const twin_state *S) { return S->match; }
const uint32_t *b2d_twin_tok(const twin_state *S) { return S->tok; }
const b2d_subinfo *b2d_twin_info(const twin_state *S) { return S->info; }
const uint32_t *b2d_twin_hist(const twin_state *S) { return S->hist; }
const b2d_plan *b2d_twin_plan(const twin_state *S) { return S->plan; }
const uint64_t *b2d_twin_block_off(const twin_state *S) { return S->block_off; }
