"""Deterministic synthetic corpora for tests and bench.

Test/bench infrastructure, not product code.  Two generators, both seeded
``numpy.random.Generator(PCG64(seed))`` and fully vectorised so GiB-sized
inputs are produced in seconds on the host:

* ``fastq_like(n, seed)``   4-line records modelled on the reference fixture
  ``tests/data/test.fastq.gz`` (152-base reads, base frequencies
  T .292 / A .292 / G .208 / C .208, ~35 quality symbols peaked at '?',
  30 % of reads near-duplicates of a read from the previous 2,000 records).
* ``mixed_entropy(n, seed)``  log-uniform 4 KiB-4 MiB segments cycling over
  text / FASTQ-like / random / runs / binary records / base64.

Nothing here reads /root/reference: the English-like text is synthesised
from a seeded Zipf vocabulary.
"""
from __future__ import annotations

import numpy as np

READ_LEN = 152
_BASES = np.frombuffer(b"TAGC", dtype=np.uint8)
_BASE_P = np.array([0.292, 0.292, 0.208, 0.208])


def _rng(seed: int) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64(seed))


def _digits(vals: np.ndarray, width: int) -> np.ndarray:
    """Right-aligned decimal digits, leading zeros replaced by 0 (=absent)."""
    n = vals.shape[0]
    out = np.zeros((n, width), dtype=np.uint8)
    v = vals.astype(np.int64).copy()
    for k in range(width - 1, -1, -1):
        out[:, k] = (v % 10) + 48
        v //= 10
    # blank leading zeros (keep last digit)
    nz = np.cumsum(out != 48, axis=1) > 0
    nz[:, -1] = True
    out[~nz] = 0
    return out


def _fastq_records(nrec: int, rng: np.random.Generator, first_index: int) -> np.ndarray:
    """Return the flat bytes of ``nrec`` records."""
    idx = np.arange(first_index, first_index + nrec, dtype=np.int64)
    chrom = rng.integers(1, 23, nrec)
    pos = rng.integers(1, 240_000_000, nrec)
    ins = rng.integers(150, 400, nrec)
    # header: @chr{c}_{p}_{p+q}_{i%2}_{i}/1\n
    W = 4 + 2 + 1 + 9 + 1 + 9 + 1 + 1 + 1 + 10 + 3
    hdr = np.zeros((nrec, W), dtype=np.uint8)
    c = 0
    hdr[:, 0:4] = np.frombuffer(b"@chr", dtype=np.uint8); c = 4
    hdr[:, c:c + 2] = _digits(chrom, 2); c += 2
    hdr[:, c] = 95; c += 1
    hdr[:, c:c + 9] = _digits(pos, 9); c += 9
    hdr[:, c] = 95; c += 1
    hdr[:, c:c + 9] = _digits(pos + ins, 9); c += 9
    hdr[:, c] = 95; c += 1
    hdr[:, c] = 48 + (idx % 2); c += 1
    hdr[:, c] = 95; c += 1
    hdr[:, c:c + 10] = _digits(idx, 10); c += 10
    hdr[:, c:c + 3] = np.frombuffer(b"/1\n", dtype=np.uint8); c += 3
    # bases: own random read, or near-copy of an earlier record's base read
    base = _BASES[rng.choice(4, size=(nrec, READ_LEN), p=_BASE_P)]
    dup = rng.random(nrec) < 0.30
    back = rng.integers(1, 2001, nrec)
    src = np.maximum(np.arange(nrec) - back, 0)
    seq = np.where(dup[:, None], base[src], base)
    sub = rng.random((nrec, READ_LEN)) < 0.01
    seq = np.where(sub & dup[:, None], _BASES[rng.integers(0, 4, (nrec, READ_LEN))], seq)
    # quality: smoothed discretised normal around '?' (Phred 30), clipped '1'..'I'
    draw = rng.normal(63.0, 4.0, (nrec, READ_LEN))
    q = np.empty_like(draw)
    q[:, 0] = draw[:, 0]
    # first-order smoothing q_t = .5 q_{t-1} + .5 draw_t, done as an IIR via cumulative form
    # (exact recursion, vectorised over records)
    for t in range(1, READ_LEN):
        q[:, t] = 0.5 * q[:, t - 1] + 0.5 * draw[:, t]
    qual = np.clip(np.rint(q), 49, 73).astype(np.uint8)
    rec = np.zeros((nrec, W + READ_LEN + 1 + 2 + READ_LEN + 1), dtype=np.uint8)
    rec[:, :W] = hdr
    c = W
    rec[:, c:c + READ_LEN] = seq; c += READ_LEN
    rec[:, c] = 10; c += 1
    rec[:, c:c + 2] = np.frombuffer(b"+\n", dtype=np.uint8); c += 2
    rec[:, c:c + READ_LEN] = qual; c += READ_LEN
    rec[:, c] = 10
    return rec[rec != 0]


def fastq_like(nbytes: int, seed: int = 1) -> np.ndarray:
    """``nbytes`` of FASTQ-like data as a uint8 array (truncated at exact size)."""
    rng = _rng(seed)
    out = np.empty(nbytes, dtype=np.uint8)
    filled = 0
    index = 0
    batch = 1 << 16
    while filled < nbytes:
        need = nbytes - filled
        nrec = min(batch, need // 340 + 2)
        chunk = _fastq_records(nrec, rng, index)
        index += nrec
        take = min(need, chunk.shape[0])
        out[filled:filled + take] = chunk[:take]
        filled += take
    return out


_SYLL = ["a", "e", "i", "o", "u", "th", "er", "on", "an", "re", "he", "in", "ed", "nd",
         "ha", "at", "en", "es", "of", "or", "nt", "ea", "ti", "to", "it", "st", "io",
         "le", "is", "ou", "ar", "as", "de", "rt", "ve", "ss", "ly", "ch", "wh", "qu"]


def _text(nbytes: int, rng: np.random.Generator) -> np.ndarray:
    """English-like text: Zipf-distributed pseudo-words, punctuation, ~72-col lines."""
    nvocab = 4096
    vr = _rng(12345)  # fixed vocabulary independent of the stream seed
    words = []
    for _ in range(nvocab):
        k = int(vr.integers(1, 5))
        words.append("".join(_SYLL[int(j)] for j in vr.integers(0, len(_SYLL), k)))
    maxw = max(len(w) for w in words) + 2
    table = np.zeros((nvocab, maxw), dtype=np.uint8)
    for i, w in enumerate(words):
        b = np.frombuffer(w.encode(), dtype=np.uint8)
        table[i, :len(b)] = b
    nwords = nbytes // 4 + 16
    ranks = rng.zipf(1.15, nwords)
    ranks = np.where(ranks > nvocab, rng.integers(1, nvocab + 1, nwords), ranks) - 1
    mat = table[ranks].copy()
    wl = (mat != 0).sum(axis=1)
    r = rng.random(nwords)
    sep = np.full(nwords, 32, dtype=np.uint8)
    punct = np.zeros(nwords, dtype=np.uint8)
    punct[r < 0.08] = 44  # ','
    punct[r < 0.03] = 46  # '.'
    sep[np.cumsum(wl + 1) % 72 < (wl + 1)] = 10
    rows = np.arange(nwords)
    mat[rows, wl] = np.where(punct != 0, punct, sep)
    mat[rows, wl + 1] = np.where(punct != 0, sep, 0)
    cap = np.roll(punct == 46, 1)
    mat[cap, 0] -= 32
    flat = mat[mat != 0]
    if flat.shape[0] < nbytes:
        flat = np.resize(flat, nbytes)
    return flat[:nbytes]


_B64 = np.frombuffer(b"ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/",
                     dtype=np.uint8)


def _segment(kind: int, n: int, rng: np.random.Generator) -> np.ndarray:
    if kind == 0:
        return _text(n, rng)
    if kind == 1:
        chunk = _fastq_records(n // 340 + 2, rng, int(rng.integers(0, 1 << 30)))
        return chunk[:n]
    if kind == 2:
        return rng.integers(0, 256, n, dtype=np.uint8)
    if kind == 3:
        out = np.zeros(n, dtype=np.uint8)
        # single-byte runs of random lengths
        nruns = max(1, n // 4096)
        cuts = np.sort(rng.integers(0, n, nruns))
        vals = rng.integers(0, 256, nruns + 1, dtype=np.uint8)
        vals[rng.random(nruns + 1) < 0.5] = 0
        out[:] = vals[np.searchsorted(cuts, np.arange(n), side="right")]
        return out
    if kind == 4:
        nrec = n // 16 + 1
        a = (1_000_000 + np.cumsum(rng.integers(0, 3, nrec))).astype("<u4")
        b = (np.cumsum(rng.integers(-2, 3, nrec)) + 50_000).astype("<u4")
        c = rng.integers(0, 16, nrec).astype("<u4")
        d = (np.arange(nrec) * 16).astype("<u4")
        rec = np.stack([a, b, c, d], axis=1).view(np.uint8).reshape(-1)
        return rec[:n]
    sym = _B64[rng.integers(0, 64, n)]
    sym[76::77] = 10
    return sym


_KIND_P = np.array([0.30, 0.25, 0.15, 0.05, 0.15, 0.10])


def mixed_entropy(nbytes: int, seed: int = 2) -> np.ndarray:
    """``nbytes`` of mixed-entropy data (text / FASTQ / random / runs / records / base64)."""
    rng = _rng(seed)
    out = np.empty(nbytes, dtype=np.uint8)
    filled = 0
    while filled < nbytes:
        seglen = int(np.exp(rng.uniform(np.log(4096), np.log(4 << 20))))
        seglen = min(seglen, nbytes - filled)
        # weights are by BYTES; segment lengths are iid so choosing kinds iid with
        # _KIND_P gives the target byte mix in expectation
        kind = int(rng.choice(6, p=_KIND_P))
        out[filled:filled + seglen] = _segment(kind, seglen, rng)
        filled += seglen
    return out


def make(kind: str, nbytes: int, seed: int) -> np.ndarray:
    if kind == "fastq":
        return fastq_like(nbytes, seed)
    if kind == "mixed":
        return mixed_entropy(nbytes, seed)
    raise ValueError(kind)
