/* tools/lzmodel.c — DEV TOOL (not product, not oracle).
 *
 * CPU model of the planned GPU pipeline, used to tune ratio knobs before
 * spending GPU time: exact hash chains (head table 2^HB, u16 prev distances),
 * all-position match search with depth D, local lazy parse, fixed-size
 * deflate sub-blocks with length-limited Huffman, exact bit cost incl. the
 * dynamic-block header, min(stored, fixed, dynamic) per sub-block.
 * Prints the total compressed size in bytes for the given file.
 *
 * usage: lzmodel file [key=value ...]
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static int P_block = 131072, P_dict = 32768, P_hb = 15, P_hbytes = 4, P_depth = 16,
           P_nice = 258, P_lazy = 1, P_sub = 32768, P_min4 = 32768, P_min5 = 32768,
           P_min3 = 0, P_verbose = 0, P_good = 258, P_lazy2 = 0, P_subtok = 0;

static const int extra_lbits[29] = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
static const int extra_dbits[30] = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,12,12,13,13};
static int base_length[29], base_dist[30];
static uint8_t len_code[259];
static uint8_t dist_code_lo[512];

static void init_tables(void) {
    int l = 3;
    for (int c = 0; c < 28; c++) { base_length[c] = l; for (int k = 0; k < (1 << extra_lbits[c]); k++) if (l <= 258) len_code[l++] = c; }
    base_length[28] = 258; len_code[258] = 28;
    int d = 1;
    for (int c = 0; c < 30; c++) { base_dist[c] = d; d += 1 << extra_dbits[c]; }
}
static inline int dcode(int dist) {
    int c = 0;
    for (c = 29; c > 0; c--) if (dist >= base_dist[c]) break;
    return c;
}

/* ---- length-limited Huffman (sort + two-queue build + miniz-style limit) ---- */
typedef struct { uint32_t f; int s; } SF;
static int cmp_sf(const void *a, const void *b) {
    const SF *x = a, *y = b;
    if (x->f != y->f) return x->f < y->f ? -1 : 1;
    return x->s - y->s;
}
static void huff_lengths(const uint32_t *freq, int n, int maxbits, uint8_t *lens) {
    SF sf[288]; int m = 0;
    memset(lens, 0, n);
    for (int i = 0; i < n; i++) if (freq[i]) { sf[m].f = freq[i]; sf[m].s = i; m++; }
    if (m == 0) return;
    if (m == 1) { lens[sf[0].s] = 1; return; }
    qsort(sf, m, sizeof(SF), cmp_sf);
    /* two-queue Huffman: leaves sorted, internal nodes appended in order */
    uint64_t w[576]; int parent[576];
    for (int i = 0; i < m; i++) w[i] = sf[i].f;
    int li = 0, ni = m, nn = m;
    while (nn < 2 * m - 1) {
        int a, b;
        if (li < m && (ni >= nn || w[li] <= w[ni])) a = li++; else a = ni++;
        if (li < m && (ni >= nn || w[li] <= w[ni])) b = li++; else b = ni++;
        w[nn] = w[a] + w[b]; parent[a] = nn; parent[b] = nn; nn++;
    }
    int depth[576]; depth[nn - 1] = 0;
    for (int i = nn - 2; i >= 0; i--) depth[i] = depth[parent[i]] + 1;
    int cnt[64]; memset(cnt, 0, sizeof cnt);
    for (int i = 0; i < m; i++) { int d = depth[i]; if (d > maxbits) d = maxbits; cnt[d]++; }
    uint32_t total = 0;
    for (int i = maxbits; i > 0; i--) total += (uint32_t)cnt[i] << (maxbits - i);
    while (total != (1u << maxbits)) {
        cnt[maxbits]--;
        for (int i = maxbits - 1; i > 0; i--) if (cnt[i]) { cnt[i]--; cnt[i + 1] += 2; break; }
        total--;
    }
    /* assign: least frequent get the longest codes */
    int k = 0;
    for (int b = maxbits; b >= 1; b--) for (int c = 0; c < cnt[b]; c++) lens[sf[k++].s] = b;
}

/* cost in bits of a dynamic block given freqs (ll[286], d[30]) */
static const uint8_t bl_order[19] = {16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15};
static uint64_t dyn_cost(const uint32_t *llf, const uint32_t *df, uint64_t *hdr_bits_out) {
    uint8_t ll[286], dl[30];
    uint32_t dfx[30]; memcpy(dfx, df, sizeof dfx);
    huff_lengths(llf, 286, 15, ll);
    int nd = 0; for (int i = 0; i < 30; i++) if (dfx[i]) nd++;
    if (nd == 0) { dfx[0] = 1; }
    huff_lengths(dfx, 30, 15, dl);
    int hlit = 286; while (hlit > 257 && ll[hlit - 1] == 0) hlit--;
    int hdist = 30; while (hdist > 1 && dl[hdist - 1] == 0) hdist--;
    uint8_t seq[316]; int n = 0;
    for (int i = 0; i < hlit; i++) seq[n++] = ll[i];
    for (int i = 0; i < hdist; i++) seq[n++] = dl[i];
    uint32_t blf[19]; memset(blf, 0, sizeof blf);
    int extra = 0;
    for (int i = 0; i < n;) {
        int v = seq[i], r = 1;
        while (i + r < n && seq[i + r] == v) r++;
        i += r;
        if (v == 0) {
            while (r >= 11) { int t = r > 138 ? 138 : r; blf[18]++; extra += 7; r -= t; }
            if (r >= 3) { blf[17]++; extra += 3; r = 0; }
            blf[0] += r;
        } else {
            blf[v]++; r--;
            while (r >= 3) { int t = r > 6 ? 6 : r; blf[16]++; extra += 2; r -= t; }
            blf[v] += r;
        }
    }
    uint8_t bl[19]; huff_lengths(blf, 19, 7, bl);
    int hclen = 19; while (hclen > 4 && bl[bl_order[hclen - 1]] == 0) hclen--;
    uint64_t hdr = 3 + 5 + 5 + 4 + 3 * hclen + extra;
    for (int i = 0; i < 19; i++) hdr += (uint64_t)blf[i] * bl[i];
    uint64_t body = 0;
    for (int i = 0; i < 286; i++) body += (uint64_t)llf[i] * ll[i];
    for (int i = 257; i < 286; i++) body += (uint64_t)llf[i] * extra_lbits[i - 257];
    for (int i = 0; i < 30; i++) body += (uint64_t)df[i] * (dl[i] + extra_dbits[i]);
    if (hdr_bits_out) *hdr_bits_out = hdr;
    return hdr + body;
}
static uint64_t fixed_cost(const uint32_t *llf, const uint32_t *df) {
    uint64_t c = 3;
    for (int i = 0; i < 286; i++) {
        int l = i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8;
        c += (uint64_t)llf[i] * l;
    }
    for (int i = 257; i < 286; i++) c += (uint64_t)llf[i] * extra_lbits[i - 257];
    for (int i = 0; i < 30; i++) c += (uint64_t)df[i] * (5 + extra_dbits[i]);
    return c;
}

static inline uint32_t ld32(const uint8_t *p) { uint32_t v; memcpy(&v, p, 4); return v; }
static inline uint64_t ld64(const uint8_t *p) { uint64_t v; memcpy(&v, p, 8); return v; }
static inline uint32_t hashf(const uint8_t *p) {
    if (P_hbytes == 4) return (ld32(p) * 2654435761u) >> (32 - P_hb);
    uint64_t v = ld64(p) << (64 - 8 * P_hbytes);
    return (uint32_t)((v * 0x9E3779B185EBCA87ull) >> (64 - P_hb));
}

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage\n"); return 2; }
    for (int i = 2; i < argc; i++) {
        char *eq = strchr(argv[i], '='); if (!eq) continue; *eq = 0; int v = atoi(eq + 1);
#define K(n, var) if (!strcmp(argv[i], n)) var = v;
        K("block", P_block) K("dict", P_dict) K("hb", P_hb) K("hbytes", P_hbytes) K("depth", P_depth)
        K("nice", P_nice) K("lazy", P_lazy) K("sub", P_sub) K("min4", P_min4) K("min5", P_min5)
        K("min3", P_min3) K("v", P_verbose) K("good", P_good) K("lazy2", P_lazy2) K("subtok", P_subtok)
    }
    init_tables();
    FILE *f = fopen(argv[1], "rb"); if (!f) { perror("open"); return 1; }
    fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
    uint8_t *buf = malloc(n + 16); if (fread(buf, 1, n, f) != (size_t)n) return 1; fclose(f);
    memset(buf + n, 0, 16);

    uint32_t *head = malloc(sizeof(uint32_t) << P_hb);
    uint16_t *prevd = malloc(sizeof(uint16_t) * (P_block + P_dict + 16));
    uint32_t *mlen = malloc(sizeof(uint32_t) * (P_block + 16));
    uint32_t *mdist = malloc(sizeof(uint32_t) * (P_block + 16));
    uint64_t total_bits = 0, n_tok = 0, n_match = 0, n_stored = 0, n_fixed = 0, n_dyn = 0, hdr_total = 0;
    uint64_t chain_steps = 0;

    for (long start = 0; start < n; start += P_block) {
        long len = n - start < P_block ? n - start : P_block;
        long dl = start < P_dict ? start : P_dict;
        const uint8_t *base = buf + start - dl; /* span origin */
        long span = dl + len;
        /* K1: chains */
        memset(head, 0xff, sizeof(uint32_t) << P_hb);
        for (long p = 0; p < span; p++) {
            if (p + P_hbytes > span) { prevd[p] = 0; continue; }
            uint32_t h = hashf(base + p);
            uint32_t q = head[h];
            long d = (q == 0xffffffffu) ? 0 : p - q;
            prevd[p] = (d > 0 && d <= 65535) ? (uint16_t)d : 0;
            head[h] = (uint32_t)p;
        }
        /* K2: all-position match search */
        for (long i = 0; i < len; i++) {
            long p = dl + i;
            int maxl = len - i < 258 ? (int)(len - i) : 258;
            int best = 0, bd = 0;
            long cand = p; int depth = P_depth;
            uint32_t d = prevd[p];
            while (d && depth-- > 0) {
                cand -= d;
                long dist = p - cand;
                if (dist > 32768 || cand < 0) break;
                chain_steps++;
                /* quick reject */
                if (best < 4 || base[cand + best] == base[p + best]) {
                    int l = 0;
                    while (l < maxl && base[cand + l] == base[p + l]) l++;
                    if (l > best) { best = l; bd = (int)dist; if (l >= P_nice || l >= maxl) break; if (best >= P_good && depth > P_depth / 4) depth = depth / 4; }
                }
                d = prevd[cand];
            }
            int ok = best >= 4;
            if (best == 3) ok = bd <= P_min3;
            if (best == 4 && bd > P_min4) ok = 0;
            if (best == 5 && bd > P_min5) ok = 0;
            if (best < 3) ok = 0;
            mlen[i] = ok ? best : 0; mdist[i] = bd;
        }
        mlen[len] = 0; mlen[len + 1] = 0;
        /* K3: parse + histograms per sub-block */
        uint32_t llf[286], df[30];
        long sub_start = 0; long i = 0; uint64_t sub_tok = 0;
        memset(llf, 0, sizeof llf); memset(df, 0, sizeof df);
        while (1) {
            int flush = 0;
            if (i >= len) flush = 1;
            else if (P_subtok ? (sub_tok >= (uint64_t)P_subtok) : (i - sub_start >= P_sub)) flush = 1;
            if (flush && i > sub_start) {
                llf[256] = 1;
                uint64_t hb; uint64_t dc = dyn_cost(llf, df, &hb), fc = fixed_cost(llf, df);
                uint64_t sc = 0; { long sl = i - sub_start; long nb = (sl + 65534) / 65535; sc = 8ull * sl + 40ull * nb; /* approx incl. align */ }
                uint64_t c = dc; int kind = 2;
                if (fc <= c) { c = fc; kind = 1; }
                if (sc < c) { c = sc; kind = 0; }
                if (kind == 0) n_stored++; else if (kind == 1) n_fixed++; else { n_dyn++; hdr_total += hb; }
                total_bits += c;
                memset(llf, 0, sizeof llf); memset(df, 0, sizeof df);
                sub_start = i; sub_tok = 0;
            }
            if (i >= len) break;
            int l = mlen[i];
            int take = l > 0;
            if (take && P_lazy) {
                int l1 = mlen[i + 1];
                if (l1 > l) take = 0;
                else if (P_lazy2 && i + 2 < len && (int)mlen[i + 2] > l + 1) take = 0;
            }
            if (take) { llf[257 + len_code[l]]++; df[dcode(mdist[i])]++; i += l; n_match++; }
            else { llf[base[dl + i]]++; i++; }
            n_tok++; sub_tok++;
        }
        /* sync flush marker: align + 00 00 ff ff (empty stored block: 3 bits + pad + 4 bytes) */
        total_bits += 3; total_bits = (total_bits + 7) & ~7ull; total_bits += 32;
    }
    printf("%lu\n", (unsigned long)(total_bits / 8));
    if (P_verbose)
        fprintf(stderr, "in=%ld out=%lu ratio=%.4f tok=%lu match=%lu stored=%lu fixed=%lu dyn=%lu hdr_bytes=%lu chain_steps/byte=%.2f\n",
                n, (unsigned long)(total_bits / 8), (double)(total_bits / 8) / n, (unsigned long)n_tok, (unsigned long)n_match,
                (unsigned long)n_stored, (unsigned long)n_fixed, (unsigned long)n_dyn, (unsigned long)(hdr_total / 8), (double)chain_steps / n);
    return 0;
}
