"""Block-parallel gzip writer on B200 — drop-in for ``zlib_ng.gzip_ng_threaded``
(reference: src/zlib_ng/gzip_ng_threaded.py).

``open(filename, mode, compresslevel, ..., threads=N, block_size=B)`` keeps the
reference's signature (gzip_ng_threaded.py:22-24).  What changes is what sits
behind it.  The reference hands ``block_size`` pieces to N CPU threads, each
calling ``compress_and_crc`` (gzip_ng_threaded.py:299-322, 361-380).  Here
the writer gathers a batch of the caller's bytes straight into *pinned* host
memory (the only host copy; none at all for buffers from ``pinned_buffer()``),
and hands it to a multi-device context (``b2d_compress_and_crc_multi``): the
batch is dealt over ``min(threads, GPUs)`` devices primed with the 32 KiB before
each piece, the pieces come back in order in one buffer, and the member's CRC is
folded with crc32_combine exactly as the reference's writer thread does
(:382-398).  Header, terminator ``03 00`` and trailer are the reference's
(:269-284, :324-338), including its header quirk of packing the OS byte before
XFL.

What ``close()`` leaves on disk differs from the reference in one way: after the
reference's two members come a few *empty* gzip members whose FEXTRA
fields hold a block index (``B2D_WRITE_INDEX=0`` switches them off).  Every gzip
reader skips them; this module's reader uses them to inflate the file on the
GPU ("rb" mode).  BGZF files (``BC`` extra subfields) are read on the GPU as
well; any other gzip stream goes to CPython's gzip module.
"""
from __future__ import annotations

import builtins
import ctypes
import gzip as _stdgzip
import io
import multiprocessing
import os
import queue
import struct
import sys
import threading
import time

from . import _native, zlib_ng

DEFLATE_WINDOW_SIZE = 2 ** 15
_COMPRESS_LEVEL_TRADEOFF = 6
_MIN_BATCH = 128 * 1024
# What one native call covers, per GPU.  The reference keeps threads * block_size bytes in flight; a GPU wants more
# per launch, so batches grow to this size (times the number of GPUs) unless the caller asked for less than one
# 128 KiB device block in total (then the reference's sizes, and its OverflowError on oversized blocks, are kept).
_STREAM_BATCH = int(os.environ.get("B2D_STREAM_BATCH_MIB", "128")) << 20
# what one GPU context works on at a time (its batch): a call's bytes are dealt to the devices in pieces of this size,
# four of which are in flight per device
_CTX_BATCH = int(os.environ.get("B2D_MAX_BATCH_MIB", "64")) << 20
# native calls in flight (each on its own multi-device context): call j+1's first H2D and call j-1's last D2H
# overlap call j's kernels.  Inside a call the library already pipelines 3 slots x 2 lanes per device.
_STREAM_DEPTH = max(int(os.environ.get("B2D_STREAM_DEPTH", "3")), 1)
_COPY_CHUNK = 4 << 20
# Block index (for the GPU reader): one data member, then the empty member the reference's close() leaves
# (gzip_ng_threaded.py:340-355), then empty members whose FEXTRA field ("B2" subfields, RFC 1952 2.3.1.1) lists
# the compressed size of every 128 KiB device block (and the bit offsets of its 32 KiB sub-blocks, so that a reader can
# decode four warps per block), then a fixed-size locator member.  Other gzip readers skip
# extra fields and empty members, so the file stays an ordinary multi-member gzip file.
_WRITE_INDEX = os.environ.get("B2D_WRITE_INDEX", "1") != "0"
_WRITE_SUBS = os.environ.get("B2D_WRITE_SUBS", "1") != "0"        # the index also holds the bit offset of every 32 KiB sub-block
_GPU_READ = os.environ.get("B2D_GPU_READ", "1") != "0"
_DEVICE_BLOCK = 128 * 1024
_INDEX_CHUNK = 16000                      # block sizes per index member (XLEN is 16 bits)
_LOCATOR = struct.Struct("<4sIQQQI")      # magic, version, data_start, raw size, index offset, blocks
_LOCATOR_MAGIC = b"B2DX"
_LOCATOR_MEMBER = 10 + 2 + 4 + _LOCATOR.size + 2 + 8
_READER_TRACE = os.environ.get("B2D_READER_TRACE", "") not in ("", "0")       # stage times on stderr when a reader closes
_READ_DEPTH = min(max(int(os.environ.get("B2D_READ_DEPTH", "2")), 1), 4)      # windows of a reader whose decode is queued on the GPU at once
_READ_WINDOW_BLOCKS = max(int(os.environ.get("B2D_READ_WINDOW_MIB", "512")), 1) * 8    # device blocks the reader decodes per GPU call
_BGZF_HEAD = 128 << 10        # room in front of a BGZF window for what the window before left over (a member is at most 64 KiB)
_BGZF_WINDOW = max(int(os.environ.get("B2D_BGZF_WINDOW_MIB", "128")), 1) << 20      # compressed bytes of a BGZF file per GPU call


def _extra_member(payload: bytes) -> bytes:
    """An empty gzip member whose header carries one "B2" extra subfield."""
    sub = b"B2" + struct.pack("<H", len(payload)) + payload
    return (struct.pack("<BBBBIBB", 0x1F, 0x8B, 8, 4, 0, 0, 0xFF) + struct.pack("<H", len(sub)) + sub
            + b"\x03\x00" + b"\x00" * 8)


def _parse_extra_member(buf: bytes, off: int):
    """(payload, next offset) of an empty "B2" member at buf[off:], or None."""
    if len(buf) < off + 12 or buf[off:off + 4] != b"\x1f\x8b\x08\x04":
        return None
    xlen = struct.unpack_from("<H", buf, off + 10)[0]
    end = off + 12 + xlen
    if xlen < 4 or len(buf) < end + 10 or buf[off + 12:off + 14] != b"B2":
        return None
    n = struct.unpack_from("<H", buf, off + 14)[0]
    if n + 4 != xlen or buf[end:end + 10] != b"\x03\x00" + b"\x00" * 8:
        return None
    return buf[off + 16:off + 16 + n], end + 10


_COPY_THREADS = max(int(os.environ.get("B2D_COPY_THREADS", "16")), 1)


def _parallel_copy(dst_addr: int, src, nbytes: int) -> None:
    """memcpy into pinned memory, split over native threads by the library (the GIL is dropped for the call)."""
    saddr, _, keep = _native._addr(src)
    if nbytes < 2 * _COPY_CHUNK:
        ctypes.memmove(dst_addr, saddr, nbytes)
    else:
        _native.check(_native.lib().b2d_host_copy(dst_addr, saddr, nbytes, _COPY_THREADS))
    del keep


class _PinnedPool:
    """Page-locked host buffers are expensive to create (the driver pins every page), so writers and readers hand
    them back here instead of freeing them.  At most _POOL_KEEP bytes stay pooled; what comes back beyond that is
    freed (one writer on one GPU uses about 0.5 GiB: 3 input batches and 3 output buffers of 64 MiB and up)."""

    def __init__(self):
        self._lock = threading.Lock()
        self._free = {}
        self._kept = 0

    def get(self, size):
        with self._lock:
            lst = self._free.get(size)
            if lst:
                self._kept -= size
                return lst.pop()
        ptr = _native.pinned_alloc(size)
        view = memoryview((ctypes.c_char * size).from_address(ptr)).cast("B")
        return ptr, view

    def put(self, size, ptr, view):
        global _POOL_KEEP
        if _POOL_KEEP is None:
            env = os.environ.get("B2D_PINNED_POOL_MIB")
            _POOL_KEEP = (int(env) << 20) if env else _pool_keep_default()
        with self._lock:
            if self._kept + size <= _POOL_KEEP:
                self._kept += size
                self._free.setdefault(size, []).append((ptr, view))
                return
        view.release()
        _native.pinned_free(ptr)


def _pool_keep_default():
    """4 GiB per GPU the process can see (a writer over 8 GPUs cycles through ~14 GiB of batches), at most an eighth of
    the machine's memory"""
    try:
        ndev = max(_native.lib().b2d_device_count(), 1)
    except Exception:
        ndev = 1
    try:
        phys = os.sysconf("SC_PAGE_SIZE") * os.sysconf("SC_PHYS_PAGES")
    except (ValueError, OSError):
        phys = 64 << 30
    return min(ndev * (4 << 30), phys // 8)


_POOL_KEEP = None          # bytes the pool may keep; set on first use (B2D_PINNED_POOL_MIB overrides)
_pinned_pool = _PinnedPool()


class PinnedBuffer:
    """Page-locked host memory a caller can produce its data into: ``write()`` of (a slice of) such a buffer goes to
    the GPUs by DMA straight from it, without the writer's copy into its own pinned batch.  Supports the buffer
    protocol through ``.view`` (a writable memoryview of bytes) and ``numpy.frombuffer(buf.view, ...)``.
    The reference has no counterpart: its write() always copies (``data = bytes(b)``, gzip_ng_threaded.py:315)."""

    def __init__(self, nbytes: int):
        self.nbytes = int(nbytes)
        self._ptr = _native.pinned_alloc(max(self.nbytes, 1))
        self.view = memoryview((ctypes.c_char * max(self.nbytes, 1)).from_address(self._ptr)).cast("B")[:self.nbytes]

    @property
    def address(self) -> int:
        return self._ptr

    def close(self):
        if self._ptr:
            try:
                self.view.release()
            except BufferError:
                raise BufferError("PinnedBuffer.close(): views of the buffer are still exported") from None
            _native.pinned_free(self._ptr)
            self._ptr = 0

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            if self._ptr:
                _native.pinned_free(self._ptr)
                self._ptr = 0
        except Exception:
            pass


def pinned_buffer(nbytes: int) -> PinnedBuffer:
    """A page-locked buffer of nbytes (see PinnedBuffer): data written from it is not copied on the host."""
    return PinnedBuffer(nbytes)


def open(filename, mode="rb", compresslevel=_COMPRESS_LEVEL_TRADEOFF, encoding=None, errors=None, newline=None,
         *, threads=1, block_size=1024 * 1024):
    """Same contract as the reference's open(): returns a BufferedReader,
    a (flushable) BufferedWriter or a TextIOWrapper.  threads == 0 means the
    serial writer (gzip_ng.open in the reference); threads < 0 means "all"."""
    if threads == 0:
        from . import gzip_ng
        return gzip_ng.open(filename, mode, compresslevel, encoding, errors, newline)
    if threads < 0:
        try:
            threads = len(os.sched_getaffinity(0))
        except Exception:
            try:
                threads = multiprocessing.cpu_count()
            except Exception:
                threads = 1
    if "r" in mode:
        gzip_file = io.BufferedReader(_ThreadedGzipReader(filename, block_size=block_size))
    else:
        gzip_file = FlushableBufferedWriter(
            _ThreadedGzipWriter(filename, mode.replace("t", "b"), block_size=block_size, level=compresslevel,
                                threads=threads),
            buffer_size=block_size)
    if "t" in mode:
        return io.TextIOWrapper(gzip_file, encoding, errors, newline)
    return gzip_file


def open_as_binary_stream(filename, open_mode):
    if isinstance(filename, (str, bytes)) or hasattr(filename, "__fspath__"):
        return builtins.open(filename, open_mode), True
    if hasattr(filename, "read") or hasattr(filename, "write"):
        return filename, False
    raise TypeError("filename must be a str or bytes object, or a file")


def _reader_windows(nblocks: int, window: int):
    """(first block, number of blocks) of the GPU calls a reader makes for a file of `nblocks` device blocks: whole windows
    of `window` blocks, but the first two smaller (a quarter, a half), so that the first bytes arrive sooner."""
    wins, first, ramp = [], 0, max(window // 4, 1)
    while first < nblocks:
        nb = min(ramp, window, nblocks - first)
        wins.append((first, nb))
        first += nb
        ramp *= 2
    return wins


class _ReaderContexts:
    """Pairs of contexts for readers (a reader keeps a decode queued on one while it resolves on the other, so the pair is
    its own for as long as it reads); a finished reader's pair, device buffers included, serves the next one."""

    def __init__(self):
        self._lock = threading.Lock()
        self._idle = []

    def get(self):
        with self._lock:
            if self._idle:
                return self._idle.pop()
        return [_native.Context(0, 1 << 20) for _ in range(_READ_DEPTH)]      # they only inflate: no deflate batch to size for

    def put(self, pair):
        with self._lock:
            if len(self._idle) < 2:
                self._idle.append(pair)
                return
        for c in pair:
            c.close()


_reader_contexts = _ReaderContexts()


class _ThreadedGzipReader(io.RawIOBase):
    """Streamed gzip reader.  Files written by this module's writer carry a block index (see _WRITE_INDEX): those
    are inflated on the GPU, a window of device blocks per call (b2d_inflate_host: one warp per 128 KiB block),
    with the member's CRC32 and size checked at the end like gzip does.  Everything else goes through CPython's
    GzipFile — the reference reads with one zng_inflate thread as well (gzip_ng_threaded.py:90-167)."""

    def __init__(self, filename, queue_size=2, block_size=1024 * 1024):
        self.raw, self.closefd = open_as_binary_stream(filename, "rb")
        self.pos = 0
        self._closed = False
        self._gpu = None
        if _GPU_READ:
            try:
                self._gpu = self._find_index() or self._find_bgzf()
            except Exception:
                self._gpu = None
        self.fileobj = None if self._gpu else _stdgzip.GzipFile(fileobj=self.raw, mode="rb")

    # ---- GPU path
    def _find_index(self):
        raw = self.raw
        if not (hasattr(raw, "seekable") and raw.seekable()):
            return None
        here = raw.tell()
        try:
            end = raw.seek(0, io.SEEK_END)
            if end - here < _LOCATOR_MEMBER + 20:
                return None
            raw.seek(end - _LOCATOR_MEMBER)
            got = _parse_extra_member(raw.read(_LOCATOR_MEMBER), 0)
            if not got or len(got[0]) != _LOCATOR.size:
                return None
            magic, version, start, raw_n, index_off, nblocks = _LOCATOR.unpack(got[0])
            if (magic != _LOCATOR_MAGIC or version != 1 or start != here or not here < index_off < end
                    or nblocks != (raw_n + _DEVICE_BLOCK - 1) // _DEVICE_BLOCK or nblocks == 0):
                return None
            raw.seek(index_off)
            buf = raw.read(end - _LOCATOR_MEMBER - index_off)
            sizes, subs, off = [], [], 0
            while len(sizes) < nblocks:
                got = _parse_extra_member(buf, off)
                if not got:
                    return None
                payload, off = got
                kind, _, _, cnt = struct.unpack_from("<BBHI", payload, 0)
                if kind == 1 and len(payload) == 8 + 4 * cnt:          # compressed size of every block
                    sizes.extend(struct.unpack_from(f"<{cnt}I", payload, 8))
                elif kind == 2 and len(payload) == 8 + 16 * cnt:       # ... and the bit offsets of its sub-blocks 1..3
                    rows = struct.unpack_from(f"<{4 * cnt}I", payload, 8)
                    sizes.extend(rows[0::4])
                    for i in range(cnt):
                        subs.extend((0, rows[4 * i + 1], rows[4 * i + 2], rows[4 * i + 3]))
                else:
                    return None
            if len(sizes) != nblocks or off != len(buf) or (subs and len(subs) != 4 * nblocks):
                return None
            raw.seek(here)
            head = raw.read(10)
            # the data member's header is the fixed one this writer emits (no flags, mtime 0, OS byte 0xff first)
            if len(head) != 10 or head[:8] != b"\x1f\x8b\x08\x00\x00\x00\x00\x00" or head[8] != 0xFF or head[9] not in (0, 2, 4):
                return None
            # data member: 10-byte header, the blocks, 03 00, CRC32, ISIZE; then exactly the empty member close() leaves
            # (gzip_ng_threaded.py:340-355 in the reference), then the index.  Anything else between the data and the
            # index (further members, say) would be skipped by this path and read by every other gzip reader: refuse.
            data_end = here + 10 + sum(sizes) + 2 + 8
            if index_off != data_end + 20:
                return None
            raw.seek(data_end)
            empty = raw.read(20)
            if (len(empty) != 20 or empty[:8] != head[:8] or empty[8] != 0xFF or empty[10:] != b"\x03\x00" + b"\x00" * 8):
                return None
            _native.default_context()      # no usable GPU: the CPU reader takes over (raises here)
            return {"sizes": sizes, "subs": subs or None, "raw_n": raw_n, "cpos": here + 10}
        finally:
            if self._gpu is None:
                raw.seek(here)

    def _gpu_file_reader(self):
        """Compressed bytes of window after window into pinned buffers (the file is only touched by this thread)."""
        g = self._gpu
        sizes = g["sizes"]
        try:
            cpos = g["cpos"]
            self.raw.seek(cpos)
            for first, nb in g["windows"]:
                if g["stop"]:
                    break
                comp_n = sum(sizes[first:first + nb])
                comp = None
                while comp is None:
                    try:
                        comp = g["comp_free"].get(timeout=0.05)
                    except queue.Empty:
                        if g["stop"]:
                            return
                view = comp[1]
                t0 = time.perf_counter()
                if not self._pread_window(view, cpos, comp_n):
                    got = 0
                    while got < comp_n:                           # raw streams may return short reads
                        more = self.raw.readinto(view[got:comp_n])
                        if not more:
                            raise EOFError("Compressed file ended before the end-of-stream marker was reached")
                        got += more
                cpos += comp_n
                g["t_file"] = g.get("t_file", 0.0) + time.perf_counter() - t0
                g["comp_ready"].put((comp, comp_n, first, nb))
            if not g["stop"]:
                g["comp_ready"].put(self.raw.read(10))            # 03 00, CRC32, ISIZE
        except BaseException as e:
            g["comp_ready"].put(e)

    def _pread_window(self, view, pos, n) -> bool:
        """A window of a real file is read by four threads at once (os.preadv drops the GIL; one thread copies ~7 GB/s
        out of the page cache, which is less than the GPU inflates).  False: not a plain file, read it the ordinary way."""
        base = self.raw.raw if isinstance(self.raw, io.BufferedReader) else self.raw
        if n < (16 << 20) or not isinstance(base, io.FileIO):
            return False
        try:
            fd = base.fileno()
        except (OSError, ValueError):
            return False
        parts = 4
        per = (n + parts - 1) // parts
        errs = []

        def job(lo, hi):
            try:
                while lo < hi:
                    k = os.preadv(fd, [view[lo:hi]], pos + lo)
                    if k <= 0:
                        raise EOFError("Compressed file ended before the end-of-stream marker was reached")
                    lo += k
            except BaseException as e:
                errs.append(e)
        ths = [threading.Thread(target=job, args=(i * per, min((i + 1) * per, n))) for i in range(1, parts)]
        for t in ths:
            t.start()
        job(0, min(per, n))
        for t in ths:
            t.join()
        if errs:
            raise errs[0]
        self.raw.seek(pos + n)
        return True

    def _gpu_worker(self):
        """Decodes window after window into pinned buffers, ahead of the consumer.  Two contexts take turns: the decode of
        window w + 1 is queued (b2d_inflate_host_decode: no dictionary needed, nothing waited for) before window w is
        resolved against the 32 KiB before it and brought back (b2d_inflate_host_resolve), so the device always has the
        next window's warps to run while the last ones of this window finish and its bytes travel."""
        g = self._gpu
        sizes = g["sizes"]
        ctxs = _reader_contexts.get()
        L = _native.lib()
        pending = []                   # (context, compressed buffer, raw bytes) of the windows whose decode is queued, oldest first
        try:
            tail, crc = b"", ctypes.c_uint32(0)

            def finish(p):
                """False: the reader was closed while waiting for an output buffer"""
                nonlocal tail
                ctx, comp, raw_n = p
                out = None
                while out is None:
                    try:
                        out = g["free"].get(timeout=0.05)
                    except queue.Empty:
                        if g["stop"]:
                            return False
                out_ptr, out_view = out
                t0 = time.perf_counter()
                rc = L.b2d_inflate_host_resolve(ctx.handle, tail if tail else None, len(tail), out_ptr, raw_n, ctypes.byref(crc))
                g["t_gpu"] = g.get("t_gpu", 0.0) + time.perf_counter() - t0
                g["comp_free"].put(comp)
                if rc:
                    g["free"].put(out)
                    raise _stdgzip.BadGzipFile(f"Error {rc} while decompressing data: {_native.last_error()}")
                tail = (tail + bytes(out_view[max(raw_n - DEFLATE_WINDOW_SIZE, 0):raw_n]))[-DEFLATE_WINDOW_SIZE:]
                g["ready"].put((out, raw_n))
                return True

            k = 0
            while not g["stop"]:
                try:
                    item = g["comp_ready"].get(timeout=0.05)
                except queue.Empty:
                    continue
                if isinstance(item, BaseException):
                    raise item
                if isinstance(item, bytes):                       # the member's end
                    while pending:
                        if not finish(pending[0]):
                            return
                        pending.pop(0)
                    if len(item) != 10:
                        raise EOFError("Compressed file ended before the end-of-stream marker was reached")
                    if item[:2] != b"\x03\x00":
                        raise _stdgzip.BadGzipFile("Bad stream end")
                    want_crc, isize = struct.unpack("<II", item[2:10])
                    if want_crc != crc.value:
                        raise _stdgzip.BadGzipFile(f"CRC check failed {hex(want_crc)} != {hex(crc.value)}")
                    if isize != (g["raw_n"] & 0xFFFFFFFF):
                        raise _stdgzip.BadGzipFile("Incorrect length of data produced")
                    g["ready"].put(None)
                    return
                comp, comp_n, first, nb = item
                raw_n = min(nb * _DEVICE_BLOCK, g["raw_n"] - first * _DEVICE_BLOCK)
                offs = (ctypes.c_uint64 * nb)()
                acc = 0
                for i in range(nb):
                    offs[i] = acc
                    acc += sizes[first + i]
                subs = None
                if g.get("subs"):                     # one warp per 32 KiB sub-block instead of one per block
                    subs = (ctypes.c_uint32 * (4 * nb))(*g["subs"][4 * first:4 * (first + nb)])
                ctx = ctxs[k % len(ctxs)]
                k += 1
                # every window but the first follows at least one whole block: its dictionary will be 32 KiB long
                dict_n = DEFLATE_WINDOW_SIZE if first else 0
                t0 = time.perf_counter()
                rc = L.b2d_inflate_host_decode(ctx.handle, comp[0], comp_n, offs, subs, nb, dict_n, raw_n)
                g["t_gpu"] = g.get("t_gpu", 0.0) + time.perf_counter() - t0
                if rc:
                    g["comp_free"].put(comp)
                    raise _stdgzip.BadGzipFile(f"Error {rc} while decompressing data: {_native.last_error()}")
                pending.append((ctx, comp, raw_n))
                while len(pending) >= len(ctxs):      # every context has a decode queued: the oldest window is due
                    if not finish(pending[0]):
                        return
                    pending.pop(0)
        except BaseException as e:       # surfaces in the reading thread
            g["ready"].put(e)
        finally:
            for p in pending:            # closed or failed with decodes still queued: give them up
                L.b2d_inflate_host_resolve(p[0].handle, None, 0, None, 0, None)
            _reader_contexts.put(ctxs)

    # ---- BGZF: blocked gzip as written by bgzip / htslib.  Every member carries its compressed size in a "BC" extra
    # subfield and is an independent DEFLATE stream of at most 64 KiB, so the members of a window decode side by side
    # (one warp each) without any index of ours.  The reference reads these files with its one inflate thread
    # (tests/test_gzip_ng.py:466-473, fixture tests/data/test.fastq.bgzip.gz).
    def _find_bgzf(self):
        raw = self.raw
        if not (hasattr(raw, "seekable") and raw.seekable()):
            return None
        here = raw.tell()
        try:
            head = raw.read(18)
        finally:
            raw.seek(here)
        if len(head) < 18 or head[:4] != b"\x1f\x8b\x08\x04":
            return None
        xlen = struct.unpack_from("<H", head, 10)[0]
        # the first header must hold a BC subfield (bgzip writes it first); whether the member is complete is found out
        # while reading
        o, ok = 12, False
        while o + 4 <= 12 + xlen and o + 4 <= len(head):
            si, slen = head[o:o + 2], struct.unpack_from("<H", head, o + 2)[0]
            if si == b"BC" and slen == 2:
                ok = True
                break
            o += 4 + slen
        if not ok:
            return None
        _native.default_context()      # no usable GPU: the CPU reader takes over (raises here)
        return {"kind": "bgzf", "raw_n": None}

    def _bgzf_worker(self):
        """Reads a window of compressed bytes, finds its complete members (b2d_bgzf_scan), inflates them in one call
        (b2d_inflate_members_host: one warp per member, CRCs checked on the device) into a pinned buffer; whatever is
        left over (a partial member) starts the next window."""
        g = self._gpu
        L, ctx = _native.lib(), _native.default_context()
        win, cap, head = g["win"], g["out_cap"], _BGZF_HEAD
        bufs = g["comp"]                              # two (ptr, view) of head + win bytes: data is read in behind `head`
        eof = False

        def fill(view, lo, hi, res):
            """raw.readinto until view[lo:hi] is full or the file ends; res = [bytes read, at the end of the file]"""
            got, end = 0, False
            try:
                while lo + got < hi:
                    k = self.raw.readinto(view[lo + got:hi])
                    if not k:
                        end = True
                        break
                    got += k
            except BaseException as e:
                res.append(e)
                return
            res.extend((got, end))

        try:
            maxm = (win + head) // 28 + 16
            po, pl = (ctypes.c_uint64 * maxm)(), (ctypes.c_uint32 * maxm)()
            rl, cr = (ctypes.c_uint32 * maxm)(), (ctypes.c_uint32 * maxm)()
            cur, start, keep = 0, head, 0             # the window is bufs[cur][start : start + keep]
            while not g["stop"]:
                ptr, view = bufs[cur]
                if keep < win and not eof:            # (not after a prefetch, which filled the window already)
                    res = []
                    fill(view, start + keep, start + win, res)
                    if isinstance(res[0], BaseException):
                        raise res[0]
                    keep += res[0]
                    eof = res[1]
                if keep == 0 and eof:
                    g["ready"].put(None)
                    return
                base = ptr + start
                cnt, used = ctypes.c_uint64(0), ctypes.c_uint64(0)
                _native.check(L.b2d_bgzf_scan(base, keep, maxm, po, pl, rl, cr, ctypes.byref(cnt), ctypes.byref(used)))
                n = cnt.value
                if n == 0:
                    # not (or no longer) BGZF: a partial member at the end of the file, trailing zero padding, or an
                    # ordinary gzip member appended to the file — CPython's reader takes the rest, errors included
                    rest = bytes(view[start:start + keep])
                    g["ready"].put(("cpu", _stdgzip.GzipFile(fileobj=_Prefixed(rest, self.raw), mode="rb")))
                    return
                # as many members as fit the output buffer
                total, take = 0, 0
                while take < n and total + rl[take] <= cap:
                    total += rl[take]
                    take += 1
                consumed = used.value if take == n else po[take - 1] + pl[take - 1] + 8
                rest = keep - consumed
                # What is left over is at most a partial member (< 64 KiB) unless the output buffer cut the window short:
                # then the next window's bytes can be read into the other buffer, behind room for the leftover, while the
                # GPU works on this one (a 128 MiB read from the page cache takes about as long as its inflate call).
                pre, pre_res = None, []
                if take == n and not eof and rest <= head:
                    pre = threading.Thread(target=fill, args=(bufs[cur ^ 1][1], head, head + win - rest, pre_res), daemon=True)
                    pre.start()
                out = None
                while out is None:
                    try:
                        out = g["free"].get(timeout=0.05)
                    except queue.Empty:
                        if g["stop"]:
                            if pre is not None:
                                pre.join()
                            return
                out_n = ctypes.c_uint64(0)
                rc = L.b2d_inflate_members_host(ctx.handle, base, keep, po, pl, rl, cr, take, out[0], cap, ctypes.byref(out_n))
                if pre is not None:
                    pre.join()
                if rc:
                    g["free"].put(out)
                    msg = _native.last_error()
                    if "CRC" in msg:
                        raise _stdgzip.BadGzipFile("CRC check failed")
                    raise _stdgzip.BadGzipFile(f"Error {rc} while decompressing data: {msg}")
                if out_n.value:
                    g["ready"].put((out, out_n.value))
                else:
                    g["free"].put(out)
                if pre is not None:
                    if isinstance(pre_res[0], BaseException):
                        raise pre_res[0]
                    optr = bufs[cur ^ 1][0]
                    if rest:
                        ctypes.memmove(optr + head - rest, base + consumed, rest)
                    cur ^= 1
                    start, keep, eof = head - rest, rest + pre_res[0], pre_res[1]
                else:
                    if rest:
                        ctypes.memmove(ptr + head, base + consumed, rest)
                    start, keep = head, rest
        except BaseException as e:       # surfaces in the reading thread
            g["ready"].put(e)

    def _bgzf_start(self):
        g = self._gpu
        win = _BGZF_WINDOW
        try:                                   # small files get small buffers
            here = self.raw.tell()
            left = self.raw.seek(0, io.SEEK_END) - here
            self.raw.seek(here)
            win = min(win, max(1 << 20, 1 << max(left - 1, 1).bit_length()))
        except (OSError, ValueError):
            pass
        g["win"] = win
        g["out_cap"] = win * 4
        g["pinned"] = []
        g["free"], g["ready"], g["stop"] = queue.Queue(), queue.Queue(), False
        g["comp"] = []
        for _ in range(2):
            ent = (_BGZF_HEAD + g["win"] + 16,) + _pinned_pool.get(_BGZF_HEAD + g["win"] + 16)
            g["pinned"].append(ent)
            g["comp"].append(ent[1:])
        for _ in range(2):
            ent = (g["out_cap"] + 16,) + _pinned_pool.get(g["out_cap"] + 16)
            g["pinned"].append(ent)
            g["free"].put(ent[1:])
        g["cur"], g["cur_n"], g["off"], g["eof"] = None, 0, 0, False
        g["threads"] = [threading.Thread(target=self._bgzf_worker, daemon=True)]
        g["threads"][0].start()

    def _gpu_start(self):
        g = self._gpu
        if g.get("kind") == "bgzf":
            return self._bgzf_start()
        sizes = g["sizes"]
        t_setup = time.perf_counter()

        def bucket(n):       # few distinct sizes, so that the pinned pool can hand buffers to the next reader
            if n <= 32 << 20:
                return max(1 << (n - 1).bit_length(), 1 << 20)
            return (n + (32 << 20) - 1) // (32 << 20) * (32 << 20)
        wins = g["windows"] = _reader_windows(len(sizes), _READ_WINDOW_BLOCKS)
        win_raw = bucket(min(max(nb for _, nb in wins) * _DEVICE_BLOCK, g["raw_n"]))
        win_comp = bucket(max(sum(sizes[f:f + nb]) for f, nb in wins))
        several = len(wins) > 1
        g["pinned"] = []
        g["free"], g["ready"], g["comp_free"], g["comp_ready"], g["stop"] = queue.Queue(), queue.Queue(), queue.Queue(), queue.Queue(), False
        # compressed windows: one per queued decode and one being read; output windows: one being filled, one being
        # handed out, one spare
        for _ in range(_READ_DEPTH + 1 if several else 1):
            ent = (win_comp + 16,) + _pinned_pool.get(win_comp + 16)
            g["pinned"].append(ent)
            g["comp_free"].put(ent[1:])
        for _ in range(3 if several else 1):
            ent = (win_raw + 16,) + _pinned_pool.get(win_raw + 16)
            g["pinned"].append(ent)
            g["free"].put(ent[1:])
        g["cur"], g["cur_n"], g["off"], g["eof"] = None, 0, 0, False
        g["t_setup"] = time.perf_counter() - t_setup
        g["threads"] = [threading.Thread(target=self._gpu_file_reader, daemon=True),
                        threading.Thread(target=self._gpu_worker, daemon=True)]
        for t in g["threads"]:
            t.start()

    def _gpu_stop(self):
        g = self._gpu
        if not g or "threads" not in g:
            return
        g["stop"] = True
        for t in g["threads"]:
            t.join()
        if _READER_TRACE:
            sys.stderr.write("reader: file %.1f ms, gpu calls %.1f ms, copies out %.1f ms, setup %.1f ms\n" % (
                1e3 * g.get("t_file", 0.0), 1e3 * g.get("t_gpu", 0.0), 1e3 * g.get("t_copy", 0.0), 1e3 * g.get("t_setup", 0.0)))
        for size, ptr, view in g["pinned"]:
            _pinned_pool.put(size, ptr, view)
        g["pinned"] = []

    def readinto(self, b):
        if self._closed:
            raise ValueError("I/O operation on closed file")
        if self._gpu is None:
            n = self.fileobj.readinto(b)
            self.pos += n
            return n
        g = self._gpu
        if "threads" not in g:
            self._gpu_start()
        if g.get("cpu") is not None:             # the rest of a BGZF file that stopped being BGZF
            n = g["cpu"].readinto(b)
            self.pos += n
            return n
        if g["cur"] is None or g["off"] >= g["cur_n"]:
            if g["cur"] is not None:
                g["free"].put(g["cur"])
                g["cur"] = None
            if g.get("error") is not None:
                raise g["error"]                 # a damaged file stays damaged: later reads raise again
            if g["eof"]:
                return 0
            item = g["ready"].get()
            if item is None:
                g["eof"] = True
                return 0
            if isinstance(item, BaseException):
                g["error"] = item
                raise item
            if item[0] == "cpu":
                g["cpu"] = item[1]
                return self.readinto(b)
            g["cur"], g["cur_n"] = item
            g["off"] = 0
        mv = memoryview(b).cast("B")
        n = min(len(mv), g["cur_n"] - g["off"])
        if n >= 2 * _COPY_CHUNK:
            # big reads (BufferedReader hands read(n) its own fresh bytes object): copy on native threads, which also
            # spreads the page faults of a destination nobody has touched yet
            daddr, _, keep = _native._addr(mv)
            t0 = time.perf_counter()
            _native.check(_native.lib().b2d_host_copy(daddr, g["cur"][0] + g["off"], n, _COPY_THREADS))
            g["t_copy"] = g.get("t_copy", 0.0) + time.perf_counter() - t0
            del keep
        else:
            mv[:n] = g["cur"][1][g["off"]:g["off"] + n]
        g["off"] += n
        self.pos += n
        return n

    def readall(self):
        """read() without a size: RawIOBase would collect 8 KiB pieces; the size is known here, so the windows are
        copied straight into one (uninitialised) bytes object."""
        if self._closed:
            raise ValueError("I/O operation on closed file")
        if self._gpu is None:
            data = self.fileobj.read()
            self.pos += len(data)
            return data
        if self._gpu["raw_n"] is None:           # BGZF: the size is not known in advance
            return self._readall_unknown_size()
        left = self._gpu["raw_n"] - self.pos
        if left <= 0:
            self.readinto(bytearray(1))      # runs the end-of-member checks
            return b""
        out = zlib_ng._OutBytes(left)
        view = memoryview((ctypes.c_char * left).from_address(out.addr)).cast("B")
        got = 0
        while got < left:
            k = self.readinto(view[got:])
            if not k:
                break
            got += k
        if got == left:
            self.readinto(bytearray(1))
        view.release()
        return out.finish(got)

    def _readall_unknown_size(self):
        """The rest of a BGZF file into one bytes object: a guess of eight times the compressed bytes left is reserved
        (untouched pages cost nothing) and filled window by window; if the data outgrows it, the pieces are joined."""
        guess = 64 << 20
        try:
            here = self.raw.tell()
            guess = max((self.raw.seek(0, io.SEEK_END) - here) * 8 + (4 << 20), 1 << 20)
            self.raw.seek(here)
        except (OSError, ValueError):
            pass
        if "threads" in self._gpu:             # reading has begun: the file position says nothing about what is left
            guess = 1 << 30
        done = []
        while True:
            out = zlib_ng._OutBytes(guess)
            view = memoryview((ctypes.c_char * guess).from_address(out.addr)).cast("B")
            got = 0
            while got < guess:
                k = self.readinto(view[got:])
                if not k:
                    break
                got += k
            view.release()
            done.append(out.finish(got))
            if got < guess:
                return done[0] if len(done) == 1 else b"".join(done)

    def readable(self) -> bool:
        return True

    def tell(self) -> int:
        if self._closed:
            raise ValueError("I/O operation on closed file")
        return self.pos

    def close(self) -> None:
        if self._closed:
            return
        self._gpu_stop()
        if self.fileobj is not None:
            self.fileobj.close()
        if self.closefd:
            self.raw.close()
        self._closed = True

    @property
    def closed(self) -> bool:
        return self._closed


class _Prefixed(io.RawIOBase):
    """bytes already taken from a stream, then the rest of the stream"""

    def __init__(self, head: bytes, raw):
        self._head, self._raw = head, raw

    def readable(self):
        return True

    def readinto(self, b):
        mv = memoryview(b).cast("B")
        if self._head:
            k = min(len(mv), len(self._head))
            mv[:k] = self._head[:k]
            self._head = self._head[k:]
            return k
        return self._raw.readinto(mv) or 0


class FlushableBufferedWriter(io.BufferedWriter):
    def flush(self):
        super().flush()
        self.raw.flush()


class _PinnedBuffer:
    """[32 KiB dictionary area | batch bytes] in page-locked host memory."""

    def __init__(self, batch):
        self.size = DEFLATE_WINDOW_SIZE + batch
        self.ptr, self.view = _pinned_pool.get(self.size)
        self.fill = 0          # data bytes
        self.dict_n = 0        # dictionary bytes stored right before the data

    def free(self):
        if self.ptr:
            _pinned_pool.put(self.size, self.ptr, self.view)
            self.ptr, self.view = 0, None


class _ThreadedGzipWriter(io.RawIOBase):
    """Batched GPU submitter behind the reference's writer interface."""

    def __init__(self, filename, mode: str = "wb", level: int = zlib_ng.Z_DEFAULT_COMPRESSION, threads: int = 1,
                 queue_size: int = 1, block_size: int = 1024 * 1024):
        self._closed = True
        if "t" in mode or "r" in mode:
            raise ValueError("Only binary writing is supported")
        if "b" not in mode:
            mode += "b"
        if threads < 1:
            raise ValueError(f"threads should be at least 1, got {threads}")
        self.lock = threading.Lock()
        self._calling_thread = threading.current_thread()
        self.exception = None
        self.level = level
        self._zlevel = zlib_ng._level(level)          # raises zlib_ng.error("Bad compression level")
        self.block_size = block_size
        self.threads = threads
        # The reference's N compression threads (gzip_ng_threaded.py:233-246) become min(N, GPUs) devices behind
        # one multi-device context; `depth` such contexts take batches round-robin so that consecutive native
        # calls overlap (what is left of N after one worker per GPU, at most _STREAM_DEPTH).
        devices = _devices(threads)
        depth = max(1, min(threads // len(devices), _STREAM_DEPTH))
        self._contexts = [_multi_context_for(devices, slot) for slot in range(depth)]
        nw = len(self._contexts)
        # one batch = what the reference would have in flight across its threads, grown to a GPU-sized piece per device
        nominal = max(block_size * threads // nw, 1)
        # (128 MiB per device and call with three calls in flight: profiles/writer_sweep_r2.log)
        per_dev = _STREAM_BATCH
        self.batch = nominal if block_size * threads < _MIN_BATCH else max(nominal, per_dev * len(devices))
        self._out_cap = self.batch + max(self.batch // 10, 500)
        # the first batches of a file go out early (a quarter, then half of the full size), so that the GPUs start
        # while the caller is still writing and short files do not wait for a batch that never fills
        self._target = self.batch
        if self.batch >= 4 * _CTX_BATCH * len(devices):
            self._target = self.batch // 4 // _DEVICE_BLOCK * _DEVICE_BLOCK
        self.input_queues = [queue.Queue(max(queue_size, 1)) for _ in range(nw)]
        self.output_queues = [queue.Queue(max(queue_size, 1)) for _ in range(nw)] if nw > 1 else []
        self._free = queue.Queue()
        self._buffers = []
        self._max_buffers = nw * (max(queue_size, 1) + 1) + 1
        self._cur = None
        self._last_tail = b""
        self.index = 0
        self._crc = 0
        self._size = 0
        # block index: only for a file this writer starts itself and fills with whole device blocks
        self._indexing = _WRITE_INDEX and "a" not in mode and "+" not in mode and self.batch % _DEVICE_BLOCK == 0
        self._blk_sizes = []          # compressed size of every device block of the current member
        self._members = []            # (start offset, raw size, block sizes) of finished members that hold data
        self.running = False
        if nw > 1:
            self.compression_workers = [threading.Thread(target=self._compress, args=(i,)) for i in range(nw)]
            self.output_worker = threading.Thread(target=self._write)
        else:
            self.compression_workers = []
            self.output_worker = threading.Thread(target=self._compress_and_write)
        self.raw, self.closefd = open_as_binary_stream(filename, mode)
        self._closed = False
        self._member_start = None
        if self._indexing:
            try:
                self._member_start = self.raw.tell()
            except Exception:
                self._indexing = False
        self._write_gzip_header()
        self.start()

    # ---- framing (gzip_ng_threaded.py:269-284, 324-338 in the reference)
    def _write_gzip_header(self):
        if self.level == zlib_ng.Z_BEST_COMPRESSION:
            xfl = 2
        elif self.level == zlib_ng.Z_BEST_SPEED:
            xfl = 4
        else:
            xfl = 0
        # the reference packs (…, mtime, os=0xff, xfl): byte 8 = 0xff, byte 9 = xfl
        self.raw.write(struct.pack("<BBBBIBB", 0x1F, 0x8B, 8, 0, 0, 0xFF, xfl))

    def _note_blocks(self, sizes):
        if sizes is None:
            self._indexing = False
        elif self._indexing:
            self._blk_sizes.extend(sizes)

    def _write_index(self):
        """After the last member: the block index of a file that holds exactly one data member (see _WRITE_INDEX)."""
        if not self._indexing or len(self._members) != 1:
            return
        start, raw_n, sizes = self._members[0]
        if len(sizes) != (raw_n + _DEVICE_BLOCK - 1) // _DEVICE_BLOCK:
            return
        with_subs = all(isinstance(x, tuple) for x in sizes)
        if not with_subs:
            sizes = [x[0] if isinstance(x, tuple) else x for x in sizes]
        if max(x[0] if with_subs else x for x in sizes) >= 1 << 32:
            return
        index_off = self.raw.tell()
        chunk = _INDEX_CHUNK // 4 if with_subs else _INDEX_CHUNK
        for i in range(0, len(sizes), chunk):
            part = sizes[i:i + chunk]
            if with_subs:       # kind 2: (compressed size, bit offsets of sub-blocks 1..3) per block
                flat = [v for row in part for v in row]
                self.raw.write(_extra_member(struct.pack("<BBHI", 2, 0, 0, len(part)) + struct.pack(f"<{len(flat)}I", *flat)))
            else:               # kind 1: compressed sizes only
                self.raw.write(_extra_member(struct.pack("<BBHI", 1, 0, 0, len(part)) + struct.pack(f"<{len(part)}I", *part)))
        self.raw.write(_extra_member(_LOCATOR.pack(_LOCATOR_MAGIC, 1, start, raw_n, index_off, len(sizes))))
        self.raw.flush()

    def _check_closed(self, msg=None):
        if self._closed:
            raise ValueError("I/O operation on closed file")

    def start(self):
        self.running = True
        self.output_worker.start()
        for w in self.compression_workers:
            w.start()

    def stop(self):
        """Stop, but do not care for remaining work"""
        self.running = False
        # the workers poll their queues like the reference's (50 ms time-outs); a sentinel ends them at once
        for q in list(self.input_queues) + list(self.output_queues):
            try:
                q.put_nowait(None)
            except queue.Full:
                pass
        for w in self.compression_workers:
            w.join()
        self.output_worker.join()
        for b in self._buffers:
            b.free()
        self._buffers = []

    # ---- producer side
    def _take_buffer(self):
        while True:
            try:
                buf = self._free.get_nowait() if len(self._buffers) < self._max_buffers else self._free.get(timeout=0.05)
                break
            except queue.Empty:
                if len(self._buffers) < self._max_buffers:       # pinned input buffers are created on demand
                    buf = _PinnedBuffer(self.batch)
                    self._buffers.append(buf)
                    break
                with self.lock:
                    if self.exception:
                        raise self.exception
                if not self.running:
                    raise ValueError("writer is not running")
        tail = self._last_tail
        buf.dict_n = len(tail)
        buf.fill = 0
        if tail:
            buf.view[DEFLATE_WINDOW_SIZE - len(tail):DEFLATE_WINDOW_SIZE] = tail
        return buf

    def _submit_current(self):
        buf = self._cur
        self._cur = None
        if buf is None:
            return
        if buf.fill == 0 or self.exception:
            self._free.put(buf)
            return
        lo = DEFLATE_WINDOW_SIZE
        data = buf.view[lo:lo + buf.fill]
        zdict = buf.view[lo - buf.dict_n:lo]
        # the next batch is primed with the last 32 KiB of the stream so far
        both = buf.view[lo - buf.dict_n:lo + buf.fill]
        self._last_tail = bytes(both[-DEFLATE_WINDOW_SIZE:])
        worker = self.index % len(self.input_queues)
        self.index += 1
        self._put_job(worker, (data, zdict, buf))

    def write(self, b) -> int:
        self._check_closed()
        with self.lock:
            if self.exception:
                raise self.exception
        mv = memoryview(b)
        if mv.ndim != 1 or mv.format != "B":
            mv = mv.cast("B")
        length = mv.nbytes
        if length >= self.batch and self.batch >= _MIN_BATCH and self._write_direct(mv):
            return length
        pos = 0
        while pos < length:
            if self._cur is None:
                self._cur = self._take_buffer()
            buf = self._cur
            k = min(length - pos, self._target - buf.fill)
            o = DEFLATE_WINDOW_SIZE + buf.fill
            if k >= 2 * _COPY_CHUNK:
                _parallel_copy(buf.ptr + o, mv[pos:pos + k], k)
            else:
                buf.view[o:o + k] = mv[pos:pos + k]
            buf.fill += k
            pos += k
            if buf.fill >= self._target:
                self._submit_current()
                self._target = min(self.batch, self._target * 2)
        return length

    def _write_direct(self, mv) -> bool:
        """A write of a whole batch or more is not gathered into the writer's own pinned batches: its pieces go to
        the workers as they lie and the call returns once they are compressed (so the caller may reuse the buffer as
        soon as write() returns, as with any file object).  Page-locked memory (pinned_buffer(), or anything
        cudaHostAlloc'ed / registered) then moves by DMA without any host copy; pageable memory is staged by the
        library's per-device threads, overlapped with the kernels."""
        n = mv.nbytes
        if self._cur is not None and self._cur.fill % _DEVICE_BLOCK:
            return False          # the stream would no longer consist of whole device blocks: gather as usual
        self._submit_current()
        pos = 0
        while pos < n:
            # whole device blocks, except at the very end of the write
            k = min(self.batch, n - pos)
            data = mv[pos:pos + k]
            if pos >= DEFLATE_WINDOW_SIZE:
                zdict = mv[pos - DEFLATE_WINDOW_SIZE:pos]
            else:
                zdict = (self._last_tail + bytes(mv[:pos]))[-DEFLATE_WINDOW_SIZE:]
            worker = self.index % len(self.input_queues)
            self.index += 1
            self._put_job(worker, (data, zdict, None))
            pos += k
        self._last_tail = (self._last_tail + bytes(mv[-DEFLATE_WINDOW_SIZE:]))[-DEFLATE_WINDOW_SIZE:]
        for q in list(self.input_queues) + list(self.output_queues):
            self._join(q)
        with self.lock:
            if self.exception:
                raise self.exception
        return True

    def _put_job(self, worker, job):
        q = self.input_queues[worker]
        while True:
            try:
                q.put(job, timeout=0.05)
                return
            except queue.Full:
                with self.lock:
                    if self.exception:
                        raise self.exception
                if not self.running:
                    raise ValueError("writer is not running")

    def _end_gzip_stream(self):
        self._check_closed()
        self._submit_current()
        for q in list(self.input_queues) + list(self.output_queues):
            self._join(q)
        with self.lock:
            if self.exception:
                return
        self.raw.write(b"\x03\x00")          # = zlib_ng.compress(b"", wbits=-15): BFINAL fixed block with only EOB
        self.raw.write(struct.pack("<II", self._crc, self._size & 0xFFFFFFFF))
        if self._indexing and self._size:
            self._members.append((self._member_start, self._size, self._blk_sizes))
        self._blk_sizes = []
        self._crc = 0
        self._size = 0
        self._last_tail = b""
        self.raw.flush()

    def _join(self, q):
        """queue.join() that gives up once a worker has failed (nobody would drain the queue)"""
        import time
        while q.unfinished_tasks:
            if self.exception or not self.output_worker.is_alive():
                return
            time.sleep(0.0005)

    def flush(self):
        self._end_gzip_stream()
        with self.lock:
            if self.exception:
                raise self.exception
        if self._indexing:
            self._member_start = self.raw.tell()
        self._write_gzip_header()

    def close(self) -> None:
        if self._closed:
            return
        try:
            self._end_gzip_stream()
            if not self.exception:
                self._write_index()
        finally:
            self.stop()
        if self.exception:
            self.raw.close()
            self._closed = True
            raise self.exception
        if self.closefd:
            self.raw.close()
        self._closed = True

    @property
    def closed(self) -> bool:
        return self._closed

    def writable(self) -> bool:
        return True

    # ---- consumer side
    def _alive(self):
        return self.running and self._calling_thread.is_alive()

    def _run_job(self, index, job, out_buf):
        data, zdict = job[0], job[1]
        daddr, n, k1 = _native._addr(data)
        zaddr, zn, k2 = _native._addr(zdict)
        try:
            # one stream over the context's devices; the block sizes come back with the call (contexts are shared
            # by every writer of the process, so they cannot be asked for afterwards)
            out_n, crc, sizes, subs = self._contexts[index].compress_and_crc(self._zlevel, daddr, n, zaddr, zn, out_buf.ptr,
                                                                             self._out_cap, want_sizes=self._indexing,
                                                                             want_subs=True)
        except _native.NativeError as e:
            if e.code == _native.B2D_BUF_ERROR:
                raise OverflowError(f"Compressed output exceeds buffer size of {self._out_cap}") from None
            if e.code == _native.B2D_MEM_ERROR:
                raise MemoryError(e.msg) from None
            raise zlib_ng.error(f"Error {e.code} while compressing data: {e.msg[:200]}") from None
        finally:
            del k1, k2
        if sizes is not None and sum(sizes) != out_n:
            sizes = None
        if sizes is not None and subs is not None and _WRITE_SUBS:
            sizes = [(sizes[i], subs[4 * i + 1], subs[4 * i + 2], subs[4 * i + 3]) for i in range(len(sizes))]
        return out_buf.view[:out_n], crc, n, sizes

    def _release(self, job):
        if len(job) > 2 and job[2] is not None:
            self._free.put(job[2])

    def _compress_and_write(self):
        in_queue = self.input_queues[0]
        out_buf = _OutBuffer(self._out_cap)
        try:
            while True:
                try:
                    job = in_queue.get(timeout=0.05)
                except queue.Empty:
                    if not self._alive():
                        return
                    continue
                if job is None:
                    in_queue.task_done()
                    return
                try:
                    compressed, crc, n, sizes = self._run_job(0, job, out_buf)
                    self._crc = zlib_ng.crc32_combine(self._crc, crc, n)
                    self._size += n
                    self._note_blocks(sizes)
                    self.raw.write(compressed)
                except Exception as e:
                    self._release(job)
                    in_queue.task_done()
                    self._set_error_and_empty_queue(e, in_queue)
                    return
                self._release(job)
                in_queue.task_done()
        finally:
            out_buf.free()

    def _compress(self, index: int):
        in_queue = self.input_queues[index]
        out_queue = self.output_queues[index]
        # two pinned output buffers per worker: the writer thread owns one while the next batch fills the other
        spare = queue.Queue()
        bufs = [_OutBuffer(self._out_cap) for _ in range(2)]
        for b in bufs:
            spare.put(b)
        try:
            while True:
                try:
                    job = in_queue.get(timeout=0.05)
                except queue.Empty:
                    if not self._alive():
                        return
                    continue
                if job is None:
                    in_queue.task_done()
                    return
                out_buf = None
                while out_buf is None:
                    try:
                        out_buf = spare.get(timeout=0.05)
                    except queue.Empty:
                        if not self._alive():
                            self._release(job)
                            in_queue.task_done()
                            return
                try:
                    compressed, crc, n, sizes = self._run_job(index, job, out_buf)
                except Exception as e:
                    self._release(job)
                    in_queue.task_done()
                    self._set_error_and_empty_queue(e, in_queue)
                    return
                self._release(job)
                out_queue.put((compressed, crc, n, spare, out_buf, sizes))
                in_queue.task_done()
        finally:
            # the writer thread may still be writing from one of these: let it drain before they go back to the pool
            if threading.current_thread() is not self.output_worker:
                self.output_worker.join()
            for b in bufs:
                b.free()

    def _write(self):
        index = 0
        nq = len(self.output_queues)
        while True:
            q = self.output_queues[index % nq]
            try:
                item = q.get(timeout=0.05)
            except queue.Empty:
                if not self._alive():
                    return
                continue
            if item is None:
                q.task_done()
                return
            compressed, crc, n, spare, out_buf, sizes = item
            try:
                self._crc = zlib_ng.crc32_combine(self._crc, crc, n)
                self._size += n
                self._note_blocks(sizes)
                self.raw.write(compressed)
            except Exception as e:
                q.task_done()
                self._set_error_and_empty_queue(e, q)
                return
            finally:
                compressed.release()
                spare.put(out_buf)
            q.task_done()
            index += 1

    def _set_error_and_empty_queue(self, error, q):
        with self.lock:
            self.exception = error
            self.running = False
        for qq in list(self.input_queues) + list(self.output_queues):
            while True:
                try:
                    job = qq.get(timeout=0.05)
                    if qq in self.input_queues and job is not None:
                        self._release(job)
                    qq.task_done()
                except queue.Empty:
                    break


class _OutBuffer:
    def __init__(self, cap):
        self.cap = max(cap, 16)
        self.ptr, self.view = _pinned_pool.get(self.cap)

    def free(self):
        if self.ptr:
            _pinned_pool.put(self.cap, self.ptr, self.view)
            self.ptr, self.view = 0, None


_mctx_lock = threading.Lock()
_mctx = {}


def _devices(threads: int):
    """GPUs one writer spreads its batches over.  The reference's ``threads`` asks for that many compression
    workers; a worker here is a GPU, so min(threads, GPUs present) of them are used.  B2D_DEVICES="0,1,..." names the
    candidates (default: all devices the process can see)."""
    spec = os.environ.get("B2D_DEVICES", "all").strip()
    if spec == "all":
        cand = list(range(max(_native.lib().b2d_device_count(), 1)))
    else:
        cand = [int(x) for x in spec.split(",") if x.strip() != ""] or [0]
    return cand[:max(threads, 1)]


def _multi_context_for(devices, slot: int = 0):
    """Process-wide multi-device contexts, created on first use and shared by every writer of the process."""
    key = (tuple(devices), slot)
    m = _mctx.get(key)
    if m is None:
        with _mctx_lock:
            m = _mctx.get(key)
            if m is None:
                try:
                    m = _native.MultiContext(devices, _CTX_BATCH)
                except _native.NativeError as e:
                    raise zlib_ng.error(f"Error {e.code} while opening the B200 device: {e.msg}") from None
                _mctx[key] = m
    return m
