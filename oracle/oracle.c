/* oracle/oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C restatement of the reference's arithmetic for the hot path's
 * checkable results.  It is the checker in tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline leg; the product never links or calls it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here
 * against (a) the reference's own known-answer vectors
 * (tests/test_zlib_compliance.py:94-137, zlib-ng test/test_crc32.cc:27-181,
 * test/test_adler32.cc:201-344, committed under tests/golden/) and (b) the
 * reference itself compiled from /root/reference into oracle/_ref by
 * oracle/Makefile.
 *
 * What is restated, and from where (paths under /root/reference/src/zlib_ng/):
 *   orc_crc32          zlib-ng/crc32_braid.c:114-267 computes CRC-32/ISO-HDLC
 *                      (reflected polynomial 0xEDB88320, crc32_braid_p.h:46,
 *                      pre- and post-inverted); restated byte-at-a-time.
 *   orc_crc32_combine  zlib-ng/crc32_braid_comb.c:17-19 with multmodp /
 *                      x2nmodp of crc32_braid_comb_p.h:8-40: crc(A||B) =
 *                      crc(A) * x^(8 len B) mod P  xor  crc(B); the reference
 *                      reads x^(2^k) from x2n_table (crc32_braid_tbl.h:9437),
 *                      here the powers are produced by repeated squaring.
 *   orc_adler32        zlib-ng/adler32.c:11-54, adler32_p.h:11-12 (BASE 65521,
 *                      sums reduced at least every NMAX=5552 bytes).
 *   orc_adler32_combine zlib-ng/adler32.c:78-100.
 *   orc_inflate_raw    the decoder the reference round-trips through
 *                      (zlib-ng/inflate.c:477-1150 state machine, inftrees.c)
 *                      restated as a direct RFC 1951 decoder: stored / fixed /
 *                      dynamic blocks, canonical-code decode by length counts,
 *                      32 KiB window incl. a preset dictionary, and the same
 *                      rejection of over-subscribed or incomplete code sets
 *                      (inftrees.c:112-131).
 */
#include "oracle.h"
#include <string.h>

#define POLY 0xEDB88320u

static uint32_t crc_table[256];
static int crc_table_ready;

static void crc_init(void) {
    for (uint32_t i = 0; i < 256; i++) {
        uint32_t c = i;
        for (int k = 0; k < 8; k++) c = (c & 1u) ? (c >> 1) ^ POLY : c >> 1;
        crc_table[i] = c;
    }
    crc_table_ready = 1;
}

uint32_t orc_crc32(uint32_t start, const uint8_t *buf, size_t len) {
    if (!crc_table_ready) crc_init();
    uint32_t c = ~start;
    for (size_t i = 0; i < len; i++) c = crc_table[(c ^ buf[i]) & 0xffu] ^ (c >> 8);
    return ~c;
}

/* a(x) * b(x) mod P(x), reflected bit order: bit 31 is x^0 */
static uint32_t gf_mul(uint32_t a, uint32_t b) {
    uint32_t p = 0;
    for (uint32_t m = 1u << 31; m; m >>= 1) {
        if (a & m) p ^= b;
        b = (b & 1u) ? (b >> 1) ^ POLY : b >> 1;
    }
    return p;
}

uint32_t orc_crc32_combine(uint32_t crc1, uint32_t crc2, int64_t len2) {
    /* x^(8*len2) by square-and-multiply; x^1 is 1<<30 in reflected order */
    uint32_t result = 1u << 31;           /* x^0 */
    uint32_t sq = 1u << 30;               /* x^(2^0) */
    uint64_t e = (uint64_t)len2;
    int carry_bits = 3;                   /* exponent is len2 * 8 */
    for (int i = 0; i < carry_bits; i++) sq = gf_mul(sq, sq);
    while (e) {
        if (e & 1u) result = gf_mul(sq, result);
        sq = gf_mul(sq, sq);
        e >>= 1;
    }
    return gf_mul(result, crc1) ^ crc2;
}

#define BASE 65521u
#define NMAX 5552u

uint32_t orc_adler32(uint32_t start, const uint8_t *buf, size_t len) {
    uint32_t s1 = start & 0xffffu, s2 = (start >> 16) & 0xffffu;
    while (len) {
        size_t n = len < NMAX ? len : NMAX;
        for (size_t i = 0; i < n; i++) { s1 += buf[i]; s2 += s1; }
        s1 %= BASE; s2 %= BASE;
        buf += n; len -= n;
    }
    return s1 | (s2 << 16);
}

uint32_t orc_adler32_combine(uint32_t adler1, uint32_t adler2, int64_t len2) {
    if (len2 < 0) return 0xffffffffu;
    uint64_t rem = (uint64_t)len2 % BASE;
    uint64_t a1 = adler1 & 0xffffu, b1 = (adler1 >> 16) & 0xffffu;
    uint64_t a2 = adler2 & 0xffffu, b2 = (adler2 >> 16) & 0xffffu;
    /* second stream started from s1=1: remove that 1 from both sums */
    uint64_t s1 = (a1 + a2 + BASE - 1) % BASE;
    uint64_t s2 = (rem * a1 + b1 + b2 + BASE - rem) % BASE;
    return (uint32_t)(s1 | (s2 << 16));
}

/* ------------------------------ inflate ---------------------------------- */
typedef struct {
    const uint8_t *in; size_t in_len, pos;
    uint64_t bitbuf; int bitcnt;
} bitr;

static int need(bitr *b, int n) {
    while (b->bitcnt < n) {
        if (b->pos >= b->in_len) return -1;
        b->bitbuf |= (uint64_t)b->in[b->pos++] << b->bitcnt;
        b->bitcnt += 8;
    }
    return 0;
}
static int getbits(bitr *b, int n, uint32_t *v) {
    if (n == 0) { *v = 0; return 0; }
    if (need(b, n)) return -1;
    *v = (uint32_t)(b->bitbuf & ((1ull << n) - 1));
    b->bitbuf >>= n; b->bitcnt -= n;
    return 0;
}

typedef struct { uint16_t count[16]; uint16_t symbol[288]; } hufft;

/* returns 0 complete, >0 incomplete, <0 over-subscribed */
static int build(hufft *h, const uint8_t *lens, int n) {
    uint16_t offs[16];
    memset(h->count, 0, sizeof h->count);
    for (int i = 0; i < n; i++) h->count[lens[i]]++;
    if (h->count[0] == n) return 0;
    int left = 1;
    for (int l = 1; l < 16; l++) { left <<= 1; left -= h->count[l]; if (left < 0) return left; }
    offs[1] = 0;
    for (int l = 1; l < 15; l++) offs[l + 1] = (uint16_t)(offs[l] + h->count[l]);
    for (int i = 0; i < n; i++) if (lens[i]) h->symbol[offs[lens[i]]++] = (uint16_t)i;
    return left;
}
static int decode(bitr *b, const hufft *h) {
    int code = 0, first = 0, index = 0;
    for (int l = 1; l < 16; l++) {
        uint32_t bit;
        if (getbits(b, 1, &bit)) return -1;
        code |= (int)bit;
        int count = h->count[l];
        if (code - count < first) return h->symbol[index + (code - first)];
        index += count; first += count; first <<= 1; code <<= 1;
    }
    return -2;
}

static const uint16_t LBASE[29] = {3,4,5,6,7,8,9,10,11,13,15,17,19,23,27,31,35,43,51,59,67,83,99,115,131,163,195,227,258};
static const uint8_t LEXT[29] = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
static const uint16_t DBASE[30] = {1,2,3,4,5,7,9,13,17,25,33,49,65,97,129,193,257,385,513,769,1025,1537,2049,3073,4097,6145,8193,12289,16385,24577};
static const uint8_t DEXT[30] = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,12,12,13,13};

int orc_inflate_raw(const uint8_t *in, size_t in_len, const uint8_t *dict, size_t dict_len,
                    uint8_t *out, size_t out_cap, size_t *out_len, size_t *consumed,
                    int stop_at_final) {
    bitr b = { in, in_len, 0, 0, 0 };
    size_t op = 0;
    static const uint8_t order[19] = {16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15};
    if (dict_len > 32768) { dict += dict_len - 32768; dict_len = 32768; }
    for (;;) {
        uint32_t last, type;
        if (!stop_at_final && b.bitcnt < 8 && b.pos >= b.in_len) {
            /* clean end on a byte boundary after a flush marker */
            break;
        }
        if (getbits(&b, 1, &last) || getbits(&b, 2, &type)) return -3;
        if (type == 0) {
            b.bitbuf >>= (b.bitcnt & 7); b.bitcnt -= (b.bitcnt & 7);
            uint32_t len, nlen;
            if (getbits(&b, 16, &len) || getbits(&b, 16, &nlen)) return -3;
            if ((len ^ 0xffffu) != nlen) return -4;
            /* bitbuf holds whole bytes now */
            for (uint32_t i = 0; i < len; i++) {
                uint32_t v;
                if (getbits(&b, 8, &v)) return -3;
                if (op >= out_cap) return -5;
                out[op++] = (uint8_t)v;
            }
        } else if (type == 3) {
            return -6;
        } else {
            hufft ll, dd;
            uint8_t lens[320];
            if (type == 1) {
                int i = 0;
                for (; i < 144; i++) lens[i] = 8;
                for (; i < 256; i++) lens[i] = 9;
                for (; i < 280; i++) lens[i] = 7;
                for (; i < 288; i++) lens[i] = 8;
                build(&ll, lens, 288);
                for (i = 0; i < 30; i++) lens[i] = 5;
                build(&dd, lens, 30);
            } else {
                uint32_t hlit, hdist, hclen, v;
                if (getbits(&b, 5, &hlit) || getbits(&b, 5, &hdist) || getbits(&b, 4, &hclen)) return -3;
                hlit += 257; hdist += 1; hclen += 4;
                if (hlit > 286 || hdist > 30) return -7;
                uint8_t cl[19]; memset(cl, 0, sizeof cl);
                for (uint32_t i = 0; i < hclen; i++) { if (getbits(&b, 3, &v)) return -3; cl[order[i]] = (uint8_t)v; }
                hufft cc;
                if (build(&cc, cl, 19) != 0) return -8;      /* code-length code must be complete */
                uint32_t idx = 0;
                while (idx < hlit + hdist) {
                    int sym = decode(&b, &cc);
                    if (sym < 0) return -3;
                    if (sym < 16) lens[idx++] = (uint8_t)sym;
                    else {
                        uint32_t rep; uint8_t val = 0;
                        if (sym == 16) { if (idx == 0) return -9; val = lens[idx - 1]; if (getbits(&b, 2, &rep)) return -3; rep += 3; }
                        else if (sym == 17) { if (getbits(&b, 3, &rep)) return -3; rep += 3; }
                        else { if (getbits(&b, 7, &rep)) return -3; rep += 11; }
                        if (idx + rep > hlit + hdist) return -9;
                        while (rep--) lens[idx++] = val;
                    }
                }
                if (lens[256] == 0) return -10;
                int r = build(&ll, lens, (int)hlit);
                if (r < 0 || (r > 0 && !(hlit - ll.count[0] == 1 && ll.count[1] == 1))) return -11;
                r = build(&dd, lens + hlit, (int)hdist);
                if (r < 0 || (r > 0 && !(hdist - dd.count[0] == 1 && dd.count[1] == 1))) return -12;
            }
            for (;;) {
                int sym = decode(&b, &ll);
                if (sym < 0) return -3;
                if (sym < 256) { if (op >= out_cap) return -5; out[op++] = (uint8_t)sym; }
                else if (sym == 256) break;
                else {
                    sym -= 257;
                    if (sym >= 29) return -13;
                    uint32_t eb, len, dist;
                    if (getbits(&b, LEXT[sym], &eb)) return -3;
                    len = LBASE[sym] + eb;
                    int ds = decode(&b, &dd);
                    if (ds < 0) return -3;
                    if (ds >= 30) return -14;
                    if (getbits(&b, DEXT[ds], &eb)) return -3;
                    dist = DBASE[ds] + eb;
                    if (dist > op + dict_len) return -15;    /* too far back */
                    if (op + len > out_cap) return -5;
                    for (uint32_t i = 0; i < len; i++) {
                        size_t src = op - dist;               /* may wrap below 0 into dict */
                        out[op] = (dist > op) ? dict[dict_len - (dist - op)] : out[src];
                        op++;
                    }
                }
            }
        }
        if (last) break;
    }
    /* give back whole unread bytes */
    while (b.bitcnt >= 8) { b.pos--; b.bitcnt -= 8; }
    *out_len = op;
    if (consumed) *consumed = b.pos;
    return 0;
}
