/* b2d_checksum.cu — CRC-32 and Adler-32 over device-resident bytes (sm_100a).
 *
 * What is computed is fixed by the reference:
 *   crc32    CRC-32/ISO-HDLC, reflected polynomial 0xEDB88320, pre/post inverted,
 *            chained through a start value       (zlib-ng crc32_braid.c:114-267,
 *            crc32_braid_p.h:46; zlib_ngmodule.c:1517-1562)
 *   combine  crc(A||B) = crc(A) * x^(8|B|) mod P  xor  crc(B)
 *                                                 (crc32_braid_comb.c:17-19,
 *            crc32_braid_comb_p.h:8-40)
 *   adler32  s1 = 1 + sum b, s2 = sum s1, both mod 65521 (adler32.c:11-54)
 * How it is computed is ours.  The CPU code folds 64 B per step with carry-less
 * multiplies; a GPU has no such instruction but has 148 x 32 shared-memory
 * banks, so the bulk kernel is a *strided* slice-by-4:
 *
 *   - the buffer is cut into rows of 1024 x 16 B; thread t of a CTA reads the
 *     t-th uint4 of every row of its segment (fully coalesced 16-byte loads)
 *     and keeps one 32-bit remainder per word lane j:
 *         acc_j <- acc_j * x^(8*ROW)  xor  word_j
 *     i.e. Horner's rule with the row stride as the step.  The multiply by the
 *     constant x^(8*ROW) is four byte-indexed lookups;
 *   - the four 256-entry tables are replicated 32x in shared memory
 *     (4 x 256 x 32 x 4 B = 128 KiB) so that lane l only ever touches bank l:
 *     every lookup is conflict-free, 32 lookups per cycle per SM;
 *   - at the end thread t multiplies acc_j by x^(32*(S - 4t - j)) (its distance
 *     from the segment end) and the CTA xor-reduces.  The result is the *raw*
 *     remainder L(seg) = seg(x) * x^32 mod P; the host turns raw remainders of
 *     the segments into the final CRC with log-time multiplies.
 *
 * Algorithmic bytes per launch = n (each input byte read once); bound = HBM.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "b2d_kernels.h"

#define CRC_POLY 0xEDB88320u
#define CK_THREADS 1024u
#define CK_ROW (CK_THREADS * 16u)          /* 16 KiB */
#define CK_S (CK_ROW / 4u)                 /* words per row */
#define CK_MAX_SEG 1024u                   /* max CTAs of the bulk kernels */
#define ADLER_BASE 65521u

/* ---- GF(2)[x] mod P, reflected: bit 31 is x^0 --------------------------- */
__host__ __device__ static inline uint32_t gf_mulmod(uint32_t a, uint32_t b) {
    uint32_t p = 0;
    for (int i = 31; i >= 0; --i) {
        if ((a >> i) & 1u) p ^= b;
        b = (b >> 1) ^ ((b & 1u) ? CRC_POLY : 0u);
    }
    return p;
}
/* x^(8*nbytes) mod P */
__host__ __device__ static inline uint32_t gf_xpow8(uint64_t nbytes) {
    uint32_t r = 0x80000000u;          /* x^0 */
    uint32_t sq = 0x00800000u;         /* x^8 */
    while (nbytes) {
        if (nbytes & 1u) r = gf_mulmod(sq, r);
        sq = gf_mulmod(sq, sq);
        nbytes >>= 1;
    }
    return r;
}

uint32_t b2d_host_crc32_combine(uint32_t crc1, uint32_t crc2, uint64_t len2) {
    return gf_mulmod(gf_xpow8(len2), crc1) ^ crc2;
}

uint32_t b2d_host_adler32_combine(uint32_t a1, uint32_t a2, uint64_t len2) {
    const uint64_t rem = len2 % ADLER_BASE;
    const uint64_t s1a = a1 & 0xffffu, s2a = a1 >> 16, s1b = a2 & 0xffffu, s2b = a2 >> 16;
    const uint64_t s1 = (s1a + s1b + ADLER_BASE - 1u) % ADLER_BASE;
    const uint64_t s2 = (rem * s1a + s2a + s2b + ADLER_BASE - rem) % ADLER_BASE;
    return (uint32_t)(s1 | (s2 << 16));
}

/* ---- bulk CRC kernel ----------------------------------------------------- */
__device__ __forceinline__ uint32_t crc_step(uint32_t v, const uint8_t *tb) {
    /* tb = table 0 replica of this lane; tables are 32 KiB apart */
    const uint32_t a0 = *reinterpret_cast<const uint32_t *>(tb + ((v << 7) & 0x7f80u));
    const uint32_t a1 = *reinterpret_cast<const uint32_t *>(tb + 32768u + ((v >> 1) & 0x7f80u));
    const uint32_t a2 = *reinterpret_cast<const uint32_t *>(tb + 65536u + ((v >> 9) & 0x7f80u));
    const uint32_t a3 = *reinterpret_cast<const uint32_t *>(tb + 98304u + ((v >> 17) & 0x7f80u));
    return (a0 ^ a1) ^ (a2 ^ a3);
}

__global__ void __launch_bounds__(CK_THREADS, 1)
k_crc_bulk(const uint4 *__restrict__ rows, uint64_t nrows, uint32_t rows_per_cta,
           const uint32_t *__restrict__ tabA, const uint32_t *__restrict__ xp32, uint32_t *__restrict__ raw) {
    extern __shared__ __align__(16) uint8_t s_tab[];       /* [4][256][32] u32 */
    __shared__ uint32_t s_red[CK_THREADS / 32];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    {
        uint32_t *t32 = reinterpret_cast<uint32_t *>(s_tab);
#pragma unroll 4
        for (uint32_t i = 0; i < 32; ++i) { const uint32_t idx = i * CK_THREADS + tid; t32[idx] = __ldg(tabA + (idx >> 5)); }
    }
    __syncthreads();
    const uint8_t *tb = s_tab + 4u * lane;
    const uint64_t r0 = (uint64_t)blockIdx.x * rows_per_cta;
    uint64_t r1 = r0 + rows_per_cta; if (r1 > nrows) r1 = nrows;
    const uint4 *p = rows + r0 * CK_THREADS + tid;
    uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
    uint64_t r = r0;
    for (; r + 4 <= r1; r += 4) {
        const uint4 v0 = __ldcs(p), v1 = __ldcs(p + CK_THREADS), v2 = __ldcs(p + 2 * CK_THREADS), v3 = __ldcs(p + 3 * CK_THREADS);
        p += 4 * CK_THREADS;
        a0 = crc_step(a0, tb) ^ v0.x; a1 = crc_step(a1, tb) ^ v0.y; a2 = crc_step(a2, tb) ^ v0.z; a3 = crc_step(a3, tb) ^ v0.w;
        a0 = crc_step(a0, tb) ^ v1.x; a1 = crc_step(a1, tb) ^ v1.y; a2 = crc_step(a2, tb) ^ v1.z; a3 = crc_step(a3, tb) ^ v1.w;
        a0 = crc_step(a0, tb) ^ v2.x; a1 = crc_step(a1, tb) ^ v2.y; a2 = crc_step(a2, tb) ^ v2.z; a3 = crc_step(a3, tb) ^ v2.w;
        a0 = crc_step(a0, tb) ^ v3.x; a1 = crc_step(a1, tb) ^ v3.y; a2 = crc_step(a2, tb) ^ v3.z; a3 = crc_step(a3, tb) ^ v3.w;
    }
    for (; r < r1; ++r) {
        const uint4 v = __ldcs(p); p += CK_THREADS;
        a0 = crc_step(a0, tb) ^ v.x; a1 = crc_step(a1, tb) ^ v.y; a2 = crc_step(a2, tb) ^ v.z; a3 = crc_step(a3, tb) ^ v.w;
    }
    /* weight by the distance to the segment end: word (t, j) of the last row is
     * S - 4t - j words (incl. itself) before the end */
    const uint32_t e = CK_S - 4u * tid;
    uint32_t c = gf_mulmod(a0, __ldg(xp32 + e)) ^ gf_mulmod(a1, __ldg(xp32 + e - 1)) ^
                 gf_mulmod(a2, __ldg(xp32 + e - 2)) ^ gf_mulmod(a3, __ldg(xp32 + e - 3));
#pragma unroll
    for (uint32_t o = 16; o; o >>= 1) c ^= __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) s_red[tid >> 5] = c;
    __syncthreads();
    if (tid < 32) {
        c = s_red[tid];
#pragma unroll
        for (uint32_t o = 16; o; o >>= 1) c ^= __shfl_xor_sync(0xffffffffu, c, o);
        if (tid == 0) raw[blockIdx.x] = c;
    }
}

/* ---- small CRC kernel: CTA c handles byte range rng[c] (any alignment) ----- */
struct ck_range { const uint8_t *p; uint64_t len; };

__global__ void __launch_bounds__(256)
k_crc_small(ck_range r0, ck_range r1, const uint32_t *__restrict__ tab8, uint32_t *__restrict__ raw) {
    __shared__ uint32_t s_t[256];
    __shared__ uint32_t s_red[8];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const ck_range R = blockIdx.x == 0 ? r0 : r1;
    s_t[tid] = tab8[tid];
    __syncthreads();
    uint32_t c = 0;
    for (uint64_t off = 64ull * tid; off < R.len; off += 64ull * 256) {
        const uint64_t m = R.len - off < 64 ? R.len - off : 64;
        uint32_t v = 0;
        for (uint32_t i = 0; i < m; ++i) v = s_t[(v ^ R.p[off + i]) & 0xffu] ^ (v >> 8);
        c ^= gf_mulmod(v, gf_xpow8(R.len - off - m));
    }
#pragma unroll
    for (uint32_t o = 16; o; o >>= 1) c ^= __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) s_red[tid >> 5] = c;
    __syncthreads();
    if (tid == 0) { uint32_t x = 0; for (int i = 0; i < 8; ++i) x ^= s_red[i]; raw[blockIdx.x] = x; }
}

/* ---- Adler-32 ------------------------------------------------------------- */
/* per CTA: S1 = sum of bytes, P = sum over 16-byte pieces of ((n - o) mod M) * sum(piece),
 * Q = sum over pieces of sum(idx * byte); the host finishes
 *   s1 = a + S1,  s2 = b + n*a + P - Q   (mod 65521)                                    */
struct adler_part { unsigned long long s1, p, q; };

__global__ void __launch_bounds__(1024)
k_adler_bulk(const uint4 *__restrict__ pieces, uint64_t npieces, uint64_t per_cta, uint64_t n_after_first,
             adler_part *__restrict__ parts) {
    /* n_after_first = (n - o) for piece 0 */
    __shared__ unsigned long long s_red[3][32];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint64_t i0 = (uint64_t)blockIdx.x * per_cta;
    uint64_t i1 = i0 + per_cta; if (i1 > npieces) i1 = npieces;
    unsigned long long S1 = 0, P = 0, Q = 0;
    uint64_t i = i0 + tid;
    /* (n - o_i) mod M, stepping by 16*1024 bytes per iteration */
    uint32_t nmo = i < i1 ? (uint32_t)((n_after_first - 16ull * i) % ADLER_BASE) : 0u;
    const uint32_t stepm = (16u * 1024u) % ADLER_BASE;
    for (; i < i1; i += 1024) {
        const uint4 v = __ldcs(pieces + i);
        uint32_t a = __dp4a(v.x, 0x01010101u, 0u);
        a = __dp4a(v.y, 0x01010101u, a); a = __dp4a(v.z, 0x01010101u, a); a = __dp4a(v.w, 0x01010101u, a);
        uint32_t w = __dp4a(v.x, 0x03020100u, 0u);
        w = __dp4a(v.y, 0x07060504u, w); w = __dp4a(v.z, 0x0b0a0908u, w); w = __dp4a(v.w, 0x0f0e0d0cu, w);
        S1 += a; P += (unsigned long long)nmo * a; Q += w;
        nmo = nmo >= stepm ? nmo - stepm : nmo + ADLER_BASE - stepm;
    }
#pragma unroll
    for (uint32_t o = 16; o; o >>= 1) {
        S1 += __shfl_xor_sync(0xffffffffu, S1, o); P += __shfl_xor_sync(0xffffffffu, P, o); Q += __shfl_xor_sync(0xffffffffu, Q, o);
    }
    if (lane == 0) { s_red[0][warp] = S1; s_red[1][warp] = P; s_red[2][warp] = Q; }
    __syncthreads();
    if (tid == 0) {
        unsigned long long a = 0, b = 0, c = 0;
        for (int k = 0; k < 32; ++k) { a += s_red[0][k]; b += s_red[1][k]; c += s_red[2][k]; }
        parts[blockIdx.x].s1 = a; parts[blockIdx.x].p = b % ADLER_BASE; parts[blockIdx.x].q = c;
    }
}

/* bytes outside the aligned pieces: range 0 = head, range 1 = tail; one thread each.
 * o0/o1 = offset of the range's first byte from the start of the buffer. */
__global__ void k_adler_edges(ck_range r0, uint64_t o0, ck_range r1, uint64_t o1, uint64_t n, adler_part *__restrict__ parts) {
    const ck_range R = threadIdx.x == 0 ? r0 : r1;
    const uint64_t o = threadIdx.x == 0 ? o0 : o1;
    unsigned long long S1 = 0, P = 0;
    for (uint64_t i = 0; i < R.len; ++i) { const uint32_t b = R.p[i]; S1 += b; P += ((n - (o + i)) % ADLER_BASE) * b; }
    parts[threadIdx.x].s1 = S1; parts[threadIdx.x].p = P % ADLER_BASE; parts[threadIdx.x].q = 0;
}

/* ---- host side -------------------------------------------------------------- */
struct b2d_cksum_ws {
    uint32_t *d_tabA, *d_xp32, *d_tab8, *d_raw;
    uint32_t *h_raw;                  /* pinned */
    adler_part *d_parts, *h_parts;
    uint32_t adler_g;                 /* partial sums the last adler launch wrote */
    /* description of the last launch, for finish() */
    uint64_t seg_len[CK_MAX_SEG + 2];
    uint32_t nseg;
    int sm_count;
};

cudaError_t b2d_cksum_ws_create(b2d_cksum_ws **out) {
    b2d_cksum_ws *w = (b2d_cksum_ws *)calloc(1, sizeof *w);
    if (!w) return cudaErrorMemoryAllocation;
    cudaError_t e;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&w->sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (w->sm_count <= 0) w->sm_count = 148;
    uint32_t *tabA = (uint32_t *)malloc(4 * 256 * 4), *xp = (uint32_t *)malloc((CK_S + 1) * 4), tab8[256];
    const uint32_t xrow = gf_xpow8(CK_ROW);
    for (int i = 0; i < 4; ++i) for (uint32_t b = 0; b < 256; ++b) tabA[i * 256 + b] = gf_mulmod(b << (8 * i), xrow);
    const uint32_t x32 = gf_xpow8(4);
    xp[0] = 0x80000000u;
    for (uint32_t m = 1; m <= CK_S; ++m) xp[m] = gf_mulmod(xp[m - 1], x32);
    const uint32_t x8 = gf_xpow8(1);
    for (uint32_t b = 0; b < 256; ++b) tab8[b] = gf_mulmod(b, x8);
#define CK(x) do { e = (x); if (e != cudaSuccess) { free(tabA); free(xp); b2d_cksum_ws_destroy(w); return e; } } while (0)
    CK(cudaMalloc(&w->d_tabA, 4 * 256 * 4));
    CK(cudaMalloc(&w->d_xp32, (CK_S + 1) * 4));
    CK(cudaMalloc(&w->d_tab8, 256 * 4));
    CK(cudaMalloc(&w->d_raw, (CK_MAX_SEG + 2) * 4));
    CK(cudaMalloc(&w->d_parts, (CK_MAX_SEG + 2) * sizeof(adler_part)));
    CK(cudaMallocHost(&w->h_raw, (CK_MAX_SEG + 2) * 4));
    CK(cudaMallocHost(&w->h_parts, (CK_MAX_SEG + 2) * sizeof(adler_part)));
    CK(cudaMemcpy(w->d_tabA, tabA, 4 * 256 * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(w->d_xp32, xp, (CK_S + 1) * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(w->d_tab8, tab8, 256 * 4, cudaMemcpyHostToDevice));
    CK(cudaFuncSetAttribute(k_crc_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 256 * 32 * 4));
#undef CK
    free(tabA); free(xp);
    *out = w;
    return cudaSuccess;
}

void b2d_cksum_ws_destroy(b2d_cksum_ws *w) {
    if (!w) return;
    cudaFree(w->d_tabA); cudaFree(w->d_xp32); cudaFree(w->d_tab8); cudaFree(w->d_raw); cudaFree(w->d_parts);
    if (w->h_raw) cudaFreeHost(w->h_raw);
    if (w->h_parts) cudaFreeHost(w->h_parts);
    free(w);
}

/* Segment layout: [head < 16 B][bulk rows, G segments][tail]; raw[0] = head, raw[1] = tail, raw[2+g] = bulk g */
cudaError_t b2d_crc32_dev_launch(b2d_cksum_ws *w, const uint8_t *d, uint64_t n, cudaStream_t st) {
    uint64_t head = (16 - ((uintptr_t)d & 15)) & 15;
    if (head > n) head = n;
    const uint64_t nrows = (n - head) / CK_ROW;
    const uint64_t tail = n - head - nrows * CK_ROW;
    uint32_t G = 0, rpc = 0;
    if (nrows) {
        const uint64_t want = (uint64_t)w->sm_count;
        G = (uint32_t)(nrows < want ? nrows : want);
        rpc = (uint32_t)((nrows + G - 1) / G);
        G = (uint32_t)((nrows + rpc - 1) / rpc);
        k_crc_bulk<<<G, CK_THREADS, 4 * 256 * 32 * 4, st>>>(reinterpret_cast<const uint4 *>(d + head), nrows, rpc,
                                                            w->d_tabA, w->d_xp32, w->d_raw + 2);
    }
    ck_range r0 = { d, head }, r1 = { d + head + nrows * CK_ROW, tail };
    k_crc_small<<<2, 256, 0, st>>>(r0, r1, w->d_tab8, w->d_raw);
    w->nseg = 0;
    w->seg_len[w->nseg++] = head;
    for (uint32_t g = 0; g < G; ++g) {
        uint64_t a = (uint64_t)g * rpc, b = a + rpc; if (b > nrows) b = nrows;
        w->seg_len[w->nseg++] = (b - a) * CK_ROW;
    }
    w->seg_len[w->nseg++] = tail;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    return cudaMemcpyAsync(w->h_raw, w->d_raw, (2 + G) * 4, cudaMemcpyDeviceToHost, st);
}

uint32_t b2d_crc32_dev_finish(b2d_cksum_ws *w, uint64_t n, uint32_t start) {
    /* raw L of the whole message by Horner over the segments */
    uint32_t L = 0;
    const uint32_t nseg = w->nseg;
    for (uint32_t s = 0; s < nseg; ++s) {
        const uint64_t len = w->seg_len[s];
        if (!len) continue;
        const uint32_t raw = s == 0 ? w->h_raw[0] : s == nseg - 1 ? w->h_raw[1] : w->h_raw[2 + (s - 1)];
        L = gf_mulmod(L, gf_xpow8(len)) ^ raw;
    }
    /* crc32(0, msg) = L xor crc32(0, zeros(n)); zeros: state 0xffffffff advanced n bytes, inverted */
    const uint32_t z = ~gf_mulmod(0xffffffffu, gf_xpow8(n));
    const uint32_t crc0 = n ? (L ^ z) : 0u;
    return b2d_host_crc32_combine(start, crc0, n);
}

cudaError_t b2d_crc32_dev(b2d_cksum_ws *w, const uint8_t *d, uint64_t n, uint32_t start, uint32_t *out, cudaStream_t st) {
    if (n == 0) { *out = start; return cudaSuccess; }
    cudaError_t e = b2d_crc32_dev_launch(w, d, n, st);
    if (e != cudaSuccess) return e;
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return e;
    *out = b2d_crc32_dev_finish(w, n, start);
    return cudaSuccess;
}

cudaError_t b2d_adler32_dev_launch(b2d_cksum_ws *w, const uint8_t *d, uint64_t n, cudaStream_t st) {
    w->adler_g = 0;
    if (n == 0) return cudaSuccess;
    uint64_t head = (16 - ((uintptr_t)d & 15)) & 15;
    if (head > n) head = n;
    const uint64_t npieces = (n - head) / 16;
    const uint64_t tail = n - head - npieces * 16;
    uint32_t G = 0;
    if (npieces) {
        const uint64_t want = (uint64_t)w->sm_count * 2;
        uint64_t per = (npieces + want - 1) / want;
        per = (per + 1023) / 1024 * 1024;
        G = (uint32_t)((npieces + per - 1) / per);
        k_adler_bulk<<<G, 1024, 0, st>>>(reinterpret_cast<const uint4 *>(d + head), npieces, per, n - head, w->d_parts + 2);
    }
    ck_range r0 = { d, head }, r1 = { d + head + npieces * 16, tail };
    k_adler_edges<<<1, 2, 0, st>>>(r0, 0, r1, head + npieces * 16, n, w->d_parts);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    w->adler_g = 2 + G;
    return cudaMemcpyAsync(w->h_parts, w->d_parts, (2 + G) * sizeof(adler_part), cudaMemcpyDeviceToHost, st);
}

/* after the stream has passed the launch: fold the partial sums into adler32(start, data[0..n)) */
uint32_t b2d_adler32_dev_finish(b2d_cksum_ws *w, uint64_t n, uint32_t start) {
    if (n == 0) return start;
    uint64_t S1 = 0, P = 0, Q = 0;
    for (uint32_t i = 0; i < w->adler_g; ++i) {
        S1 = (S1 + w->h_parts[i].s1 % ADLER_BASE) % ADLER_BASE;
        P = (P + w->h_parts[i].p) % ADLER_BASE;
        Q = (Q + w->h_parts[i].q % ADLER_BASE) % ADLER_BASE;
    }
    const uint64_t a = start & 0xffffu, b = (start >> 16) & 0xffffu;
    const uint64_t s1 = (a + S1) % ADLER_BASE;
    const uint64_t s2 = (b + (n % ADLER_BASE) * a % ADLER_BASE + P + ADLER_BASE - Q) % ADLER_BASE;
    return (uint32_t)(s1 | (s2 << 16));
}

cudaError_t b2d_adler32_dev(b2d_cksum_ws *w, const uint8_t *d, uint64_t n, uint32_t start, uint32_t *out, cudaStream_t st) {
    if (n == 0) { *out = start; return cudaSuccess; }
    cudaError_t e = b2d_adler32_dev_launch(w, d, n, st);
    if (e != cudaSuccess) return e;
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return e;
    *out = b2d_adler32_dev_finish(w, n, start);
    return cudaSuccess;
}

/* ---- Adler-32 without any allocation (b2d_adler32_device_noalloc, include/b2d.h) -----------------------------
 * For callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  No counterpart in the reference.  The per-CTA partial
 * sums that k_adler_bulk / k_adler_edges write — the scratch the other entry points keep in a cudaMalloc'ed
 * workspace — live in a file-scope device array instead: B2D_NOALLOC_PARTS partial sums, 24 KiB, held by every GPU
 * that loads the library until the process exits.  Every call shares the array and it keeps the last call's
 * contents; calls are serialised inside the library, but calls that overlap on the GPU from several processes'
 * worth of streams would race (there is one array per device, not per caller). */
#define B2D_NOALLOC_PARTS (CK_MAX_SEG + 2u)
static __device__ __align__(256) adler_part g_noalloc_parts[B2D_NOALLOC_PARTS];

#include <mutex>
#define B2D_NOALLOC_MAX_DEVICES 64
static std::mutex g_noalloc_mu;
static adler_part *g_noalloc_ptr[B2D_NOALLOC_MAX_DEVICES];      /* device address of the array, per device */
static int g_noalloc_sms[B2D_NOALLOC_MAX_DEVICES];

/* 0 = ok; on failure *why names the step.  `max_bytes` is enforced by the caller (b2d_api.cu). */
cudaError_t b2d_adler32_noalloc(int device, const uint8_t *d, uint64_t n, uint32_t start, uint32_t *out, const char **why) {
    *why = "";
    if (device < 0 || device >= B2D_NOALLOC_MAX_DEVICES) { *why = "device index"; return cudaErrorInvalidDevice; }
    if (n == 0) { *out = start; return cudaSuccess; }
    std::lock_guard<std::mutex> g(g_noalloc_mu);                /* one call at a time: the scratch is shared */
    cudaError_t e;
    if ((e = cudaSetDevice(device)) != cudaSuccess) { *why = "cudaSetDevice"; return e; }
    if (!g_noalloc_ptr[device]) {
        /* once per device, with that device current: in host code the array's name is only a stand-in for the
         * symbol, not a device address */
        adler_part *p = nullptr;
        if ((e = cudaGetSymbolAddress((void **)&p, g_noalloc_parts)) != cudaSuccess) { *why = "cudaGetSymbolAddress"; return e; }
        int sms = 0;
        if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) { *why = "cudaDeviceGetAttribute"; return e; }
        g_noalloc_sms[device] = sms > 0 ? sms : 148;
        g_noalloc_ptr[device] = p;
    }
    adler_part *parts = g_noalloc_ptr[device];
    cudaStream_t st = cudaStreamPerThread;
    uint64_t head = (16 - ((uintptr_t)d & 15)) & 15;
    if (head > n) head = n;
    const uint64_t npieces = (n - head) / 16;
    const uint64_t tail = n - head - npieces * 16;
    uint32_t G = 0;
    uint64_t per = 0;
    if (npieces) {
        uint64_t want = (uint64_t)g_noalloc_sms[device] * 2;
        if (want > CK_MAX_SEG) want = CK_MAX_SEG;
        per = (npieces + want - 1) / want;
        per = (per + 1023) / 1024 * 1024;
        G = (uint32_t)((npieces + per - 1) / per);
    }
    /* the array keeps the last call's sums: clear the entries this call reads back */
    if ((e = cudaMemsetAsync(parts, 0, (2 + G) * sizeof(adler_part), st)) != cudaSuccess) { *why = "cudaMemsetAsync"; return e; }
    if (G) k_adler_bulk<<<G, 1024, 0, st>>>(reinterpret_cast<const uint4 *>(d + head), npieces, per, n - head, parts + 2);
    ck_range r0 = { d, head }, r1 = { d + head + npieces * 16, tail };
    k_adler_edges<<<1, 2, 0, st>>>(r0, 0, r1, head + npieces * 16, n, parts);
    if ((e = cudaGetLastError()) != cudaSuccess) { *why = "kernel launch"; return e; }
    adler_part h[B2D_NOALLOC_PARTS];                            /* on the stack: nothing is allocated */
    if ((e = cudaMemcpyAsync(h, parts, (2 + G) * sizeof(adler_part), cudaMemcpyDeviceToHost, st)) != cudaSuccess) { *why = "cudaMemcpyAsync"; return e; }
    if ((e = cudaStreamSynchronize(st)) != cudaSuccess) { *why = "cudaStreamSynchronize"; return e; }
    uint64_t S1 = 0, P = 0, Q = 0;
    for (uint32_t i = 0; i < 2 + G; ++i) {
        S1 = (S1 + h[i].s1 % ADLER_BASE) % ADLER_BASE;
        P = (P + h[i].p) % ADLER_BASE;
        Q = (Q + h[i].q % ADLER_BASE) % ADLER_BASE;
    }
    const uint64_t a = start & 0xffffu, b = (start >> 16) & 0xffffu;
    const uint64_t s1 = (a + S1) % ADLER_BASE;
    const uint64_t s2 = (b + (n % ADLER_BASE) * a % ADLER_BASE + P + ADLER_BASE - Q) % ADLER_BASE;
    *out = (uint32_t)(s1 | (s2 << 16));
    return cudaSuccess;
}
