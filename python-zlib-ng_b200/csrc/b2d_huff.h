/* b2d_huff.h — per-sub-block Huffman planning: length-limited code lengths,
 * canonical codes, the dynamic-block header, and the stored / fixed / dynamic
 * choice.  Single-thread integer code shared by the CUDA Huffman stage and the
 * test-only CPU twin, so both produce identical plans.
 *
 * What it must reproduce from the reference is the *format*, not the
 * algorithm: zlib-ng builds the trees with a heap and an overflow fix-up
 * (trees.c:186-406); here lengths come from a sort + two-queue merge and a
 * Kraft-sum repair, which yields the same optimal cost whenever no code
 * exceeds the limit.  Format references: send_all_trees / send_tree
 * (trees.c:460-588), bl_order (trees.h:30-31), static tables
 * (trees_tbl.h:8-74), block choice (trees.c:665-700).
 */
#ifndef B2D_HUFF_H
#define B2D_HUFF_H

#include "b2d_common.h"

typedef struct {
    uint32_t key[288];      /* (freq << 9) | symbol, sorted ascending   */
    uint32_t w[576];        /* node weights: leaves then internal nodes */
    uint16_t parent[576];
    uint16_t depth[576];
    uint8_t  ll_len[B2D_LL_CODES];
    uint8_t  d_len[B2D_D_CODES];
    uint8_t  bl_len[B2D_BL_CODES];
    uint8_t  seq_sym[B2D_LL_CODES + B2D_D_CODES];   /* code-length alphabet symbols */
    uint8_t  seq_ext[B2D_LL_CODES + B2D_D_CODES];   /* their extra-bit values       */
    uint32_t bl_freq[B2D_BL_CODES];
    uint32_t bl_code[B2D_BL_CODES];
    uint32_t dfix[B2D_D_CODES];
    int hlit, hdist, hclen, nseq;   /* carried between the phases of b2d_plan_subblock */
} b2d_huff_scratch;

/* in-place heapsort of n unique u32 keys, ascending */
B2D_HDN void b2d_sort_keys(uint32_t *a, int n) {
    for (int start = n / 2 - 1; start >= 0; start--) {
        int root = start;
        for (;;) {
            int child = 2 * root + 1;
            if (child >= n) break;
            if (child + 1 < n && a[child] < a[child + 1]) child++;
            if (a[root] >= a[child]) break;
            uint32_t t = a[root]; a[root] = a[child]; a[child] = t;
            root = child;
        }
    }
    for (int end = n - 1; end > 0; end--) {
        uint32_t t = a[0]; a[0] = a[end]; a[end] = t;
        int root = 0;
        for (;;) {
            int child = 2 * root + 1;
            if (child >= end) break;
            if (child + 1 < end && a[child] < a[child + 1]) child++;
            if (a[root] >= a[child]) break;
            uint32_t t2 = a[root]; a[root] = a[child]; a[child] = t2;
            root = child;
        }
    }
}

/* Code lengths for n symbols (n <= 286) with every length <= maxbits.
 * At least two symbols always receive a code (zlib-ng trees.c:341-350 does the
 * same) so the code is never incomplete. */
/* phase 1: one key per used symbol, (freq << 9) | symbol; returns their number (>= 2) */
B2D_HDN int b2d_huff_keys(const uint32_t *freq, int n, uint8_t *lens, b2d_huff_scratch *S) {
    int m = 0;
    for (int i = 0; i < n; i++) { lens[i] = 0; if (freq[i]) S->key[m++] = (freq[i] << 9) | (uint32_t)i; }
    if (m == 0) { S->key[0] = (1u << 9) | 0u; S->key[1] = (1u << 9) | 1u; m = 2; }
    else if (m == 1) {
        uint32_t s = S->key[0] & 511u;
        S->key[1] = (1u << 9) | (s == 0u ? 1u : 0u);
        m = 2;
    }
    return m;
}

/* phase 3 (phase 2 = sort S->key[0..m) ascending; keys are unique, so any sort gives the same order) */
B2D_HDN void b2d_huff_from_sorted(int m, int maxbits, uint8_t *lens, b2d_huff_scratch *S) {
    for (int i = 0; i < m; i++) S->w[i] = S->key[i] >> 9;
    int li = 0, ni = m, nn = m;
    while (nn < 2 * m - 1) {
        int a, b;
        if (li < m && (ni >= nn || S->w[li] <= S->w[ni])) a = li++; else a = ni++;
        if (li < m && (ni >= nn || S->w[li] <= S->w[ni])) b = li++; else b = ni++;
        S->w[nn] = S->w[a] + S->w[b];
        S->parent[a] = (uint16_t)nn; S->parent[b] = (uint16_t)nn;
        nn++;
    }
    S->depth[nn - 1] = 0;
    for (int i = nn - 2; i >= 0; i--) S->depth[i] = (uint16_t)(S->depth[S->parent[i]] + 1);
    uint32_t cnt[16];
    for (int i = 0; i < 16; i++) cnt[i] = 0;
    for (int i = 0; i < m; i++) { int d = S->depth[i]; if (d > maxbits) d = maxbits; cnt[d]++; }
    uint32_t total = 0;
    for (int i = maxbits; i > 0; i--) total += cnt[i] << (maxbits - i);
    while (total != (1u << maxbits)) {   /* only entered when some depth was clamped */
        cnt[maxbits]--;
        for (int i = maxbits - 1; i > 0; i--) if (cnt[i]) { cnt[i]--; cnt[i + 1] += 2; break; }
        total--;
    }
    int k = 0;   /* least frequent symbols take the longest codes */
    for (int b = maxbits; b >= 1; b--)
        for (uint32_t c = 0; c < cnt[b]; c++) lens[S->key[k++] & 511u] = (uint8_t)b;
}

B2D_HDN void b2d_huff_lengths(const uint32_t *freq, int n, int maxbits, uint8_t *lens,
                              b2d_huff_scratch *S) {
    int m = b2d_huff_keys(freq, n, lens, S);
    b2d_sort_keys(S->key, m);
    b2d_huff_from_sorted(m, maxbits, lens, S);
}

/* canonical codes, bit-reversed, packed (len << 16) | code (RFC 1951 3.2.2) */
B2D_HDN void b2d_canonical(const uint8_t *lens, int n, int maxbits, uint32_t *out) {
    uint32_t bl_count[16], next_code[16];
    for (int i = 0; i < 16; i++) bl_count[i] = 0;
    for (int i = 0; i < n; i++) bl_count[lens[i]]++;
    bl_count[0] = 0;
    uint32_t code = 0;
    next_code[0] = 0;
    for (int b = 1; b <= maxbits; b++) { code = (code + bl_count[b - 1]) << 1; next_code[b] = code; }
    for (int i = 0; i < n; i++) {
        int l = lens[i];
        out[i] = l ? (((uint32_t)l << 16) | b2d_bitrev(next_code[l]++, l)) : 0u;
    }
}

B2D_HD void b2d_put_bits(uint32_t *words, uint32_t *nbits, uint32_t value, uint32_t count) {
    uint32_t pos = *nbits, w = pos >> 5, s = pos & 31u;
    words[w] |= value << s;
    if (s + count > 32u) words[w + 1] |= value >> (32u - s);
    *nbits = pos + count;
}

B2D_HD uint32_t b2d_fixed_ll_len(uint32_t sym) { return sym < 144u ? 8u : sym < 256u ? 9u : sym < 280u ? 7u : 8u; }

/* Build the plan of one sub-block from its histograms.
 * ll_freq must already count the end-of-block symbol (ll_freq[256] == 1).
 * in_len = input bytes the sub-block covers (for the stored alternative).
 * blocks = B2D_BLOCKS_*: level 0 forces stored (deflate_stored.c in the reference), Z_FIXED forbids dynamic. */
/* after ll_len / d_len: trailing-zero trimming, run-length coding of the length sequence, bl_freq */
B2D_HDN void b2d_plan_mid(b2d_huff_scratch *S) {
    int hlit = B2D_LL_CODES; while (hlit > 257 && S->ll_len[hlit - 1] == 0) hlit--;
    int hdist = B2D_D_CODES; while (hdist > 1 && S->d_len[hdist - 1] == 0) hdist--;
    /* run-length encode the concatenated length sequence */
    int nseq = 0, total = hlit + hdist;
    for (int i = 0; i < B2D_BL_CODES; i++) S->bl_freq[i] = 0;
    for (int i = 0; i < total;) {
        int v = i < hlit ? S->ll_len[i] : S->d_len[i - hlit];
        int r = 1;
        while (i + r < total && (i + r < hlit ? S->ll_len[i + r] : S->d_len[i + r - hlit]) == v) r++;
        i += r;
        if (v == 0) {
            while (r >= 11) { int t = r > 138 ? 138 : r; S->seq_sym[nseq] = 18; S->seq_ext[nseq++] = (uint8_t)(t - 11); S->bl_freq[18]++; r -= t; }
            if (r >= 3) { S->seq_sym[nseq] = 17; S->seq_ext[nseq++] = (uint8_t)(r - 3); S->bl_freq[17]++; r = 0; }
            while (r > 0) { S->seq_sym[nseq] = 0; S->seq_ext[nseq++] = 0; S->bl_freq[0]++; r--; }
        } else {
            S->seq_sym[nseq] = (uint8_t)v; S->seq_ext[nseq++] = 0; S->bl_freq[v]++; r--;
            while (r >= 3) { int t = r > 6 ? 6 : r; S->seq_sym[nseq] = 16; S->seq_ext[nseq++] = (uint8_t)(t - 3); S->bl_freq[16]++; r -= t; }
            while (r > 0) { S->seq_sym[nseq] = (uint8_t)v; S->seq_ext[nseq++] = 0; S->bl_freq[v]++; r--; }
        }
    }
    S->hlit = hlit; S->hdist = hdist; S->nseq = nseq;
}

/* after bl_len: costs, stored / fixed / dynamic choice, codes and header bits */
B2D_HDN void b2d_plan_end(const uint32_t *ll_freq, const uint32_t *d_freq, uint32_t in_len,
                          b2d_plan *P, b2d_huff_scratch *S, int blocks) {
    const int force_stored = blocks == B2D_BLOCKS_STORED, no_dyn = blocks == B2D_BLOCKS_NO_DYN;
    const uint8_t bl_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    const int hlit = S->hlit, hdist = S->hdist, nseq = S->nseq;
    int hclen = B2D_BL_CODES; while (hclen > 4 && S->bl_len[bl_order[hclen - 1]] == 0) hclen--;
    uint32_t dyn_hdr = 3u + 5u + 5u + 4u + 3u * (uint32_t)hclen;
    for (int i = 0; i < nseq; i++) {
        uint32_t s = S->seq_sym[i];
        dyn_hdr += S->bl_len[s] + (s == 16u ? 2u : s == 17u ? 3u : s == 18u ? 7u : 0u);
    }
    uint32_t dyn_body = 0, fix_body = 0;
    for (int i = 0; i < B2D_LL_CODES; i++) {
        uint32_t f = ll_freq[i];
        if (!f) continue;
        uint32_t ex = i >= 257 ? b2d_len_extra_bits((uint32_t)i - 257u) : 0u;
        dyn_body += f * (S->ll_len[i] + ex);
        fix_body += f * (b2d_fixed_ll_len((uint32_t)i) + ex);
    }
    for (int i = 0; i < B2D_D_CODES; i++) {
        uint32_t f = d_freq[i];
        if (!f) continue;
        uint32_t ex = b2d_dist_extra_bits((uint32_t)i);
        dyn_body += f * (S->d_len[i] + ex);
        fix_body += f * (5u + ex);
    }
    uint32_t dyn_cost = dyn_hdr + dyn_body, fix_cost = 3u + fix_body;
    uint32_t sto_cost = 3u + 7u + 32u + 8u * in_len;   /* worst-case alignment pad */

    for (int i = 0; i < B2D_HDR_WORDS; i++) P->hdr[i] = 0;
    P->hdr_bits = 0; P->pad = 0;
    if ((force_stored || ((sto_cost <= dyn_cost || no_dyn) && sto_cost <= fix_cost)) && in_len <= 65535u) {
        P->type = 0; P->body_bits = 0;
        b2d_put_bits(P->hdr, &P->hdr_bits, 0u, 3);          /* BFINAL=0, BTYPE=00 */
        for (int i = 0; i < B2D_LL_CODES; i++) P->ll_code[i] = 0;
        for (int i = 0; i < B2D_D_CODES; i++) P->d_code[i] = 0;
    } else if (fix_cost <= dyn_cost || no_dyn) {
        P->type = 1; P->body_bits = fix_body;
        b2d_put_bits(P->hdr, &P->hdr_bits, 2u, 3);          /* BFINAL=0, BTYPE=01 */
        for (int i = 0; i < B2D_LL_CODES; i++) {
            uint32_t l = b2d_fixed_ll_len((uint32_t)i);
            uint32_t c = i < 144 ? 0x30u + (uint32_t)i : i < 256 ? 0x190u + (uint32_t)(i - 144)
                       : i < 280 ? (uint32_t)(i - 256) : 0xC0u + (uint32_t)(i - 280);
            P->ll_code[i] = (l << 16) | b2d_bitrev(c, (int)l);
        }
        for (int i = 0; i < B2D_D_CODES; i++) P->d_code[i] = (5u << 16) | b2d_bitrev((uint32_t)i, 5);
    } else {
        P->type = 2; P->body_bits = dyn_body;
        b2d_canonical(S->ll_len, B2D_LL_CODES, 15, P->ll_code);
        b2d_canonical(S->d_len, B2D_D_CODES, 15, P->d_code);
        b2d_canonical(S->bl_len, B2D_BL_CODES, 7, S->bl_code);
        b2d_put_bits(P->hdr, &P->hdr_bits, 4u, 3);          /* BFINAL=0, BTYPE=10 */
        b2d_put_bits(P->hdr, &P->hdr_bits, (uint32_t)(hlit - 257), 5);
        b2d_put_bits(P->hdr, &P->hdr_bits, (uint32_t)(hdist - 1), 5);
        b2d_put_bits(P->hdr, &P->hdr_bits, (uint32_t)(hclen - 4), 4);
        for (int i = 0; i < hclen; i++) b2d_put_bits(P->hdr, &P->hdr_bits, S->bl_len[bl_order[i]], 3);
        for (int i = 0; i < nseq; i++) {
            uint32_t s = S->seq_sym[i], c = S->bl_code[s];
            b2d_put_bits(P->hdr, &P->hdr_bits, c & 0xffffu, c >> 16);
            if (s == 16u) b2d_put_bits(P->hdr, &P->hdr_bits, S->seq_ext[i], 2);
            else if (s == 17u) b2d_put_bits(P->hdr, &P->hdr_bits, S->seq_ext[i], 3);
            else if (s == 18u) b2d_put_bits(P->hdr, &P->hdr_bits, S->seq_ext[i], 7);
        }
    }
}

B2D_HDN void b2d_plan_subblock(const uint32_t *ll_freq, const uint32_t *d_freq, uint32_t in_len,
                               b2d_plan *P, b2d_huff_scratch *S, int blocks) {
    b2d_huff_lengths(ll_freq, B2D_LL_CODES, 15, S->ll_len, S);
    b2d_huff_lengths(d_freq, B2D_D_CODES, 15, S->d_len, S);
    b2d_plan_mid(S);
    b2d_huff_lengths(S->bl_freq, B2D_BL_CODES, 7, S->bl_len, S);
    b2d_plan_end(ll_freq, d_freq, in_len, P, S, blocks);
}

#endif /* B2D_HUFF_H */
