"""Host-side shard plan for multi-GPU jobs.

Blocks are independent given the raw input, so a stream is cut into contiguous
ranges aligned to the 128 KiB device block; each range is compressed (by its own
GPU / process) primed with the 32 KiB of input before it, and the host
concatenates the raw-deflate pieces in order and folds the per-range CRCs with
crc32_combine — the same fold the reference's writer thread performs
(src/zlib_ng/gzip_ng_threaded.py:394).  No collective touches the data path.
"""
from __future__ import annotations

BLOCK = 131072
WINDOW = 32768


def plan_shards(n: int, world: int, block: int = BLOCK):
    """-> [(start, end)] * world, contiguous, block-aligned except the last end; may be empty ranges."""
    if world < 1:
        raise ValueError("world must be at least 1")
    nblocks = -(-n // block)
    out, b0 = [], 0
    for r in range(world):
        nb = nblocks // world + (1 if r < nblocks % world else 0)
        start, end = b0 * block, min((b0 + nb) * block, n)
        out.append((min(start, n), end))
        b0 += nb
    return out


def shard_dictionary(data, start: int, zdict: bytes = b"") -> bytes:
    """The 32 KiB that prime the shard starting at `start` (the stream's own dictionary for shard 0)."""
    if start == 0:
        return bytes(zdict[-WINDOW:])
    lo = max(0, start - WINDOW)
    head = bytes(data[lo:start])
    if len(head) < WINDOW and zdict:
        head = bytes(zdict[-(WINDOW - len(head)):]) + head
    return head


def fold(parts, crc32_combine):
    """parts = [(deflate_bytes, crc, length)] in shard order -> (stream, crc, total_length)."""
    crc, total, chunks = 0, 0, []
    for blob, c, ln in parts:
        if ln == 0:
            continue
        crc = crc32_combine(crc, c, ln)
        total += ln
        chunks.append(blob)
    return b"".join(chunks), crc, total
