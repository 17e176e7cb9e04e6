"""Drop-in for the compress / checksum half of ``zlib_ng.zlib_ng``
(reference: src/zlib_ng/zlib_ngmodule.c, signatures in src/zlib_ng/zlib_ng.pyi).

Same names, argument meaning and error behaviour as the reference; the work
is done by the sm_100a kernels behind ``libb2d.so``:

=====================  =====================================================
this module            reference entry point it replaces
=====================  =====================================================
crc32 / adler32        zlib_crc32 / zlib_adler32 (zlib_ngmodule.c:1517,1455)
crc32_combine          zlib_crc32_combine (:1585)
compress               zlib_compress_impl (:199-273)
compressobj/_Compress  zlib_compressobj_impl, Compress.compress/flush/copy
                       (:376-430, :530-575, :718-784, :795-844)
_ParallelCompress      ParallelCompress_compress_and_crc (:1696-1782)
=====================  =====================================================

Inflate of arbitrary streams (decompress, decompressobj) is out of scope of the
accelerated path and is served by CPython's own zlib, which is
also the decoder the reference's tests use as the cross-check; files written by
gzip_ng_threaded (block index) and BGZF files are inflated on the GPU by
gzip_ng_threaded's reader.  There is no CPU *compress* fallback: without
libb2d.so or a CUDA device every compress/checksum call raises.

Flush modes: every block this compressor emits ends byte-aligned with a sync
marker, so Z_PARTIAL_FLUSH and Z_BLOCK produce what Z_SYNC_FLUSH produces (all
pending input is emitted and the output ends on a byte boundary).  The
reference distinguishes them (deflate.c:951-968: Z_PARTIAL_FLUSH sends an empty
fixed block instead of the stored marker, Z_BLOCK stops at a block boundary
without padding); both are valid DEFLATE continuations and decode identically,
they differ by at most a few bytes per flush.
"""
from __future__ import annotations

import ctypes
import struct
import threading
import zlib as _stdzlib

from . import _native

MAX_WBITS = 15
DEFLATED = 8
DEF_MEM_LEVEL = 8
DEF_BUF_SIZE = 16 * 1024

Z_NO_COMPRESSION = 0
Z_BEST_SPEED = 1
Z_BEST_COMPRESSION = 9
Z_DEFAULT_COMPRESSION = -1

Z_FILTERED = 1
Z_HUFFMAN_ONLY = 2
Z_RLE = 3
Z_FIXED = 4
Z_DEFAULT_STRATEGY = 0

Z_NO_FLUSH = 0
Z_PARTIAL_FLUSH = 1
Z_SYNC_FLUSH = 2
Z_FULL_FLUSH = 3
Z_FINISH = 4
Z_BLOCK = 5
Z_TREES = 6

ZLIBNG_VERSION = "b200-0.1.0"
ZLIBNG_RUNTIME_VERSION = ZLIBNG_VERSION
ZLIB_VERSION = _stdzlib.ZLIB_VERSION
ZLIB_RUNTIME_VERSION = _stdzlib.ZLIB_RUNTIME_VERSION

error = _stdzlib.error            # one exception type for both halves of the module

_WINDOW = 32768
_BLOCK = 131072                   # device block; streaming cuts happen on multiples of it
_STREAM_THRESHOLD = 32 << 20      # compressobj submits to the GPU once this much input is pending
_UINT32_MAX = 0xFFFFFFFF


def _raise(rc: int, where: str):
    msg = _native.last_error()
    if rc == _native.B2D_MEM_ERROR:
        raise MemoryError(msg)
    if rc == _native.B2D_STREAM_ERROR and "Bad compression level" in msg:
        raise error("Bad compression level")
    raise error(f"Error {rc} {where}: {msg[:200]}")


def _level(level: int) -> int:
    if level == Z_DEFAULT_COMPRESSION:
        return 6
    if not 0 <= level <= 9:
        raise error("Bad compression level")
    return level


def _ctx():
    try:
        return _native.default_context()
    except _native.NativeError as e:
        raise error(f"Error {e.code} while opening the B200 device: {e.msg}") from None


# ------------------------------------------------------------------ checksums
def crc32(data, value=0, /):
    """crc32(data, value=0) -> unsigned CRC-32 (zlib_ngmodule.c:1517-1562)."""
    addr, n, keep = _native._addr(data)
    out = ctypes.c_uint32(0)
    rc = _native.lib().b2d_crc32(_ctx().handle, int(value) & _UINT32_MAX, addr, n, ctypes.byref(out))
    if rc:
        _raise(rc, "while computing crc32")
    del keep
    return out.value


def adler32(data, value=1, /):
    """adler32(data, value=1) -> unsigned Adler-32 (zlib_ngmodule.c:1455-1500)."""
    addr, n, keep = _native._addr(data)
    out = ctypes.c_uint32(0)
    rc = _native.lib().b2d_adler32(_ctx().handle, int(value) & _UINT32_MAX, addr, n, ctypes.byref(out))
    if rc:
        _raise(rc, "while computing adler32")
    del keep
    return out.value


def crc32_combine(crc1, crc2, crc2_length, /):
    """crc32_combine(crc1, crc2, crc2_length) (zlib_ngmodule.c:1585-1596, format "IIn")."""
    for v in (crc1, crc2):
        if not isinstance(v, int):
            raise TypeError("an integer is required")
        if not 0 <= v <= _UINT32_MAX:
            raise OverflowError("Python int too large to convert to C unsigned int" if v > 0
                                else "can't convert negative value to unsigned int")
    if not isinstance(crc2_length, int):
        raise TypeError("an integer is required")
    return int(_native.lib().b2d_crc32_combine(crc1, crc2, crc2_length))


# ------------------------------------------------------------------ raw deflate call
_api = ctypes.pythonapi
_api.PyBytes_FromStringAndSize.restype = ctypes.c_void_p
_api.PyBytes_FromStringAndSize.argtypes = [ctypes.c_void_p, ctypes.c_ssize_t]
_api.PyBytes_AsString.restype = ctypes.c_void_p
_api.PyBytes_AsString.argtypes = [ctypes.c_void_p]
_api._PyBytes_Resize.restype = ctypes.c_int
_api._PyBytes_Resize.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_ssize_t]
_api.Py_DecRef.restype = None
_api.Py_DecRef.argtypes = [ctypes.c_void_p]


try:
    _libc = ctypes.CDLL(None, use_errno=True)
    _libc.madvise.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    _libc.madvise.restype = ctypes.c_int
except Exception:          # pragma: no cover
    _libc = None


class _OutBytes:
    """An uninitialised bytes object the library writes into, trimmed in place afterwards: what
    arrange_output_buffer does in the reference (zlib_ngmodule.c:149-197) — no zero fill of the
    worst-case size and no copy of the result."""

    def __init__(self, cap: int):
        self._p = _api.PyBytes_FromStringAndSize(None, cap)
        if not self._p:
            raise MemoryError
        self.addr = _api.PyBytes_AsString(self._p)
        self.cap = cap
        if cap >= (8 << 20) and _libc is not None:
            # a fresh multi-MiB bytes object is untouched memory: ask for huge pages, so that filling it costs a page
            # fault per 2 MiB instead of per 4 KiB (the faults, not the copy, bound compress() of a GiB)
            lo = (self.addr + 0x1FFFFF) & ~0x1FFFFF
            hi = (self.addr + cap) & ~0x1FFFFF
            if hi > lo:
                try:
                    _libc.madvise(ctypes.c_void_p(lo), ctypes.c_size_t(hi - lo), 14)      # MADV_HUGEPAGE
                except Exception:
                    pass

    def finish(self, size: int) -> bytes:
        pp = ctypes.c_void_p(self._p)
        self._p = None
        if _api._PyBytes_Resize(ctypes.byref(pp), size) != 0:      # frees the object on failure
            raise MemoryError
        obj = ctypes.cast(pp, ctypes.py_object).value              # takes a second reference
        _api.Py_DecRef(pp)                                         # drop ours
        return obj

    def __del__(self):
        if getattr(self, "_p", None):
            _api.Py_DecRef(ctypes.c_void_p(self._p))
            self._p = None


def _deflate(data_addr, n, zdict: bytes, level: int, wbits_abs: int, final: bool, crc=None, adler=None,
             head: bytes = b"", tail_fn=None):
    """-> (head + raw deflate bytes + tail_fn(crc, adler), crc or None, adler or None) in one bytes object"""
    L = _native.lib()
    cap = int(L.b2d_deflate_bound(n)) + 8
    out = _OutBytes(len(head) + cap + 16)
    out_n = ctypes.c_uint64(0)
    c_crc = ctypes.c_uint32(crc) if crc is not None else None
    c_ad = ctypes.c_uint32(adler) if adler is not None else None
    rc = L.b2d_deflate_host(_ctx().handle, level, wbits_abs, data_addr, n, zdict if zdict else None, len(zdict),
                            1 if final else 0, ctypes.c_void_p(out.addr + len(head)), cap, ctypes.byref(out_n),
                            ctypes.byref(c_crc) if c_crc is not None else None,
                            ctypes.byref(c_ad) if c_ad is not None else None)
    if rc:
        _raise(rc, "while compressing data")
    r_crc = c_crc.value if c_crc is not None else None
    r_ad = c_ad.value if c_ad is not None else None
    size = len(head) + out_n.value
    if head:
        ctypes.memmove(out.addr, head, len(head))
    if tail_fn is not None:
        tail = tail_fn(r_crc, r_ad)
        ctypes.memmove(out.addr + size, tail, len(tail))
        size += len(tail)
    return out.finish(size), r_crc, r_ad


def _wrapper(wbits: int):
    """wbits -> (kind, window bits); zlib-ng deflate.c:211-221."""
    if 9 <= wbits <= 15:
        return "zlib", wbits
    if wbits == 8:
        return "zlib", 9
    if -15 <= wbits <= -9:
        return "raw", -wbits
    if wbits == -8:
        return "raw", 9
    if 25 <= wbits <= 31:
        return "gzip", wbits - 16
    if wbits == 24:
        return "gzip", 9
    raise ValueError("Invalid initialization option")


def _zlib_header(level: int, wb: int, zdict: bytes) -> bytes:
    # zlib-ng deflate.c:758-782: CMF = method 8 + window, FLEVEL from the level, FCHECK mod 31
    flevel = 0 if level < 2 else 1 if level < 6 else 2 if level == 6 else 3
    hdr = ((8 + ((wb - 8) << 4)) << 8) | (flevel << 6)
    if zdict:
        hdr |= 0x20
    hdr += 31 - (hdr % 31)
    out = struct.pack(">H", hdr)
    if zdict:
        out += struct.pack(">I", adler32(zdict))            # DICTID = Adler-32 of the preset dictionary
    return out


def _gzip_header(level: int) -> bytes:
    # zlib-ng deflate.c:792-811 with no gz_header: mtime 0, XFL by level, OS_CODE 3 (unix)
    xfl = 2 if level == 9 else 4 if level == 1 else 0
    return struct.pack("<BBBBIBB", 0x1F, 0x8B, 8, 0, 0, xfl, 3)


def compress(data, /, level=Z_DEFAULT_COMPRESSION, wbits=MAX_WBITS):
    """compress(data, /, level=-1, wbits=15) -> bytes (zlib_ngmodule.c:1816-1833, 199-273)."""
    try:
        kind, wb = _wrapper(wbits)
    except ValueError:
        raise error("Bad compression level") from None   # deflateInit2 Z_STREAM_ERROR (zlib_ngmodule.c:224-225)
    lvl = _level(level)
    addr, n, keep = _native._addr(data)
    if kind == "zlib":
        return _deflate(addr, n, b"", lvl, wb, True, adler=1, head=_zlib_header(lvl, wb, b""),
                        tail_fn=lambda c, a: struct.pack(">I", a))[0]
    if kind == "gzip":
        return _deflate(addr, n, b"", lvl, wb, True, crc=0, head=_gzip_header(lvl),
                        tail_fn=lambda c, a: struct.pack("<II", c, n & _UINT32_MAX))[0]
    body, _, _ = _deflate(addr, n, b"", lvl, wb, True)
    del keep
    return body


_PIN_BYTES = 256 << 20            # gather buffer of a long compressobj stream = one GPU call of four pipelined batches (a multiple of _BLOCK)
_pin_free, _pin_lock = [], threading.Lock()


def _pin_take():
    while True:
        with _pin_lock:
            pin = _pin_free.pop() if _pin_free else None
        if pin is None:
            break
        if len(pin[1]) == _PIN_BYTES:
            return pin
        pin[1].release()                   # another size (tests change _PIN_BYTES)
        _native.pinned_free(pin[0])
    ptr = _native.pinned_alloc(_PIN_BYTES)
    return ptr, memoryview((ctypes.c_char * _PIN_BYTES).from_address(ptr)).cast("B")


def _pin_give(pin):
    with _pin_lock:
        if len(_pin_free) < 4:
            _pin_free.append(pin)
            return
    pin[1].release()
    _native.pinned_free(pin[0])


class _Compress:
    """Streaming compressor (compobject, zlib_ngmodule.c:56-66).  Input is gathered
    until a flush or until _STREAM_THRESHOLD bytes are pending, then whole device
    blocks go to the GPU primed with the previous 32 KiB, so the bytes produced do
    not depend on how the caller sliced its writes (test_pair in the reference's
    tests/test_zlib_compliance.py:295-306 relies on compress() == compressobj())."""

    def __init__(self, level, wbits, zdict, strategy=Z_DEFAULT_STRATEGY):
        self._level = level
        self._strategy = strategy          # carried in bits 8..11 of the C ABI's level argument (B2D_LEVEL)
        self._kind, self._wb = _wrapper(wbits)
        self._zdict = bytes(zdict) if zdict is not None else b""
        self._tail = self._zdict[-(1 << self._wb):][-_WINDOW:]
        self._pending = bytearray()
        self._pin = None                   # (ptr, view) of a page-locked gather buffer: streams that reach the threshold get one
        self._pin_fill = 0
        self._started = False
        self._finished = False
        self._crc = 0
        self._adler = 1
        self._total = 0
        self._lock = threading.Lock()

    def _header(self) -> bytes:
        if self._started:
            return b""
        self._started = True
        if self._kind == "zlib":
            return _zlib_header(self._level, self._wb, self._zdict)
        if self._kind == "gzip":
            return _gzip_header(self._level)
        return b""

    def _run(self, addr: int, n: int, final: bool) -> bytes:
        """n bytes at addr -> raw deflate bytes primed with the stream's last 32 KiB; updates checksums and tail"""
        body, crc, ad = _deflate(addr, n, self._tail, self._level | (self._strategy << 8), self._wb, final,
                                 crc=self._crc if self._kind == "gzip" else None,
                                 adler=self._adler if self._kind == "zlib" else None)
        if crc is not None:
            self._crc = crc
        if ad is not None:
            self._adler = ad
        self._total += n
        if n >= _WINDOW:
            self._tail = ctypes.string_at(addr + n - _WINDOW, _WINDOW)
        elif n:
            self._tail = (self._tail + ctypes.string_at(addr, n))[-_WINDOW:]
        return body

    def _submit(self, nbytes: int, final: bool) -> bytes:
        """the first nbytes of the pending buffer, in place (no copy of the input on the Python side)"""
        if self._pin is not None:
            nbytes, self._pin_fill = self._pin_fill, 0          # everything gathered so far (flush / finish)
            if nbytes:
                return self._run(self._pin[0], nbytes, final)
        if nbytes == 0:
            if not final and not self._flush_empty:
                return b""
            return self._run(0, 0, final)
        hold = (ctypes.c_char * len(self._pending)).from_buffer(self._pending)
        try:
            body = self._run(ctypes.addressof(hold), nbytes, final)
        finally:
            del hold                      # releases the export so that the bytearray can shrink again
        del self._pending[:nbytes]
        return body

    _flush_empty = False

    def compress(self, data, /):
        with self._lock:
            if self._finished:
                raise error("Error -2 while compressing data: inconsistent stream state")
            mv = memoryview(data)
            mv = mv.cast("B") if mv.ndim == 1 and mv.c_contiguous else memoryview(bytes(mv))
            out = b""
            if not self._pending and self._pin_fill == 0 and mv.nbytes >= _STREAM_THRESHOLD:
                # a large write on an empty buffer: its whole blocks go straight from the caller's memory
                k = mv.nbytes // _BLOCK * _BLOCK
                addr, _, keep = _native._addr(mv)
                out = self._header() + self._run(addr, k, False)
                del keep
                mv = mv[k:]
            if self._pin is not None:
                return out + self._gather_pinned(mv)
            self._pending += mv
            if len(self._pending) >= _STREAM_THRESHOLD:
                out += self._header() + self._submit(len(self._pending) // _BLOCK * _BLOCK, False)
                # a long stream: from now on the input is gathered in page-locked memory (one copy, like the bytearray's,
                # but the GPU then reads it by DMA without the library staging it again) in calls of _PIN_BYTES
                self._pin = _pin_take()
                rest = len(self._pending)
                self._pin[1][:rest] = self._pending
                self._pin_fill = rest
                self._pending = bytearray()
            return out

    def _gather_pinned(self, mv) -> bytes:
        out, pos, n = b"", 0, mv.nbytes
        ptr, view = self._pin
        cap = len(view)
        while pos < n:
            k = min(n - pos, cap - self._pin_fill)
            if k >= (4 << 20):
                saddr, _, keep = _native._addr(mv[pos:pos + k])
                _native.check(_native.lib().b2d_host_copy(ptr + self._pin_fill, saddr, k, 8))
                del keep
            else:
                view[self._pin_fill:self._pin_fill + k] = mv[pos:pos + k]
            self._pin_fill += k
            pos += k
            if self._pin_fill == cap:
                out += self._header() + self._run(ptr, cap, False)
                self._pin_fill = 0
        return out

    def flush(self, mode=Z_FINISH, /):
        with self._lock:
            if mode == Z_NO_FLUSH:
                return b""
            if self._finished:
                raise error("Error -2 while flushing: inconsistent stream state")
            if mode not in (Z_PARTIAL_FLUSH, Z_SYNC_FLUSH, Z_FULL_FLUSH, Z_FINISH, Z_BLOCK):
                raise error("Error -2 while flushing: inconsistent stream state")
            hdr = self._header()
            if mode == Z_FINISH:
                body = self._submit(len(self._pending), True)
                self._finished = True
                if self._kind == "zlib":
                    body += struct.pack(">I", self._adler)
                elif self._kind == "gzip":
                    body += struct.pack("<II", self._crc, self._total & _UINT32_MAX)
                return hdr + body
            self._flush_empty = True
            try:
                body = self._submit(len(self._pending), False)
            finally:
                self._flush_empty = False
            if mode == Z_FULL_FLUSH:
                self._tail = b""          # later blocks may not refer back across a full flush
            return hdr + body

    def copy(self):
        with self._lock:
            if self._finished:
                raise ValueError("Inconsistent stream state")
            c = _Compress.__new__(_Compress)
            c.__dict__.update({k: v for k, v in self.__dict__.items() if k != "_lock"})
            c._pending = bytearray(self._pending) if self._pin is None else bytearray(self._pin[1][:self._pin_fill])
            c._pin, c._pin_fill = None, 0
            c._lock = threading.Lock()
            return c

    def __del__(self):
        pin, self._pin = getattr(self, "_pin", None), None
        if pin is not None:
            _pin_give(pin)

    __copy__ = copy

    def __deepcopy__(self, memo):
        return self.copy()

    def __reduce__(self):
        raise TypeError(f"cannot pickle '{type(self).__name__}' object")


def compressobj(level=Z_DEFAULT_COMPRESSION, method=DEFLATED, wbits=MAX_WBITS, memLevel=DEF_MEM_LEVEL,
                strategy=Z_DEFAULT_STRATEGY, zdict=None):
    """compressobj(level=-1, method=8, wbits=15, memLevel=8, strategy=0, zdict=None)
    (zlib_ngmodule.c:1908-1931, 376-430)."""
    if zdict is not None:
        try:
            memoryview(zdict)
        except TypeError:
            raise TypeError("zdict argument must support the buffer protocol") from None
    if method != DEFLATED or not 1 <= memLevel <= 9 or not 0 <= strategy <= 4:
        raise ValueError("Invalid initialization option")
    if level != Z_DEFAULT_COMPRESSION and not 0 <= level <= 9:
        raise ValueError("Invalid initialization option")
    _wrapper(wbits)
    _ctx()
    return _Compress(_level(level), wbits, zdict, strategy)


_PC_CONTEXTS = 4
_pc_counter = 0
_pc_lock = threading.Lock()


class _ParallelCompress:
    """_ParallelCompress(buffersize, level=-1).compress_and_crc(data, zdict) -> (bytes, crc32)
    (zlib_ngmodule.c:1598-1797).  One owner thread, no lock, like the reference."""

    def __init__(self, buffersize, level=Z_DEFAULT_COMPRESSION):
        if not isinstance(buffersize, int):
            raise TypeError("an integer is required")
        if buffersize > _UINT32_MAX:
            raise ValueError(f"buffersize must be at most {_UINT32_MAX}, got {buffersize}")
        self._level = _level(level)
        self._cap = int(buffersize)
        self._buf = ctypes.create_string_buffer(max(self._cap, 1))
        # The reference's pattern is N objects on N threads (gzip_ng_threaded.py:231-246), each with its own deflate
        # state.  Calls on one GPU context are serialised, so the objects are dealt over a few contexts of device 0
        # (their kernels then overlap on the device like the reference's threads do on the cores).
        global _pc_counter
        with _pc_lock:
            slot = _pc_counter % _PC_CONTEXTS
            _pc_counter += 1
        try:
            self._ctx = _native.default_context(0, slot)
        except _native.NativeError as e:
            raise error(f"Error {e.code} while opening the B200 device: {e.msg}") from None

    def compress_and_crc(self, data, zdict, /):
        daddr, n, k1 = _native._addr(data)
        zaddr, zn, k2 = _native._addr(zdict)
        if n + zn > _UINT32_MAX:
            raise OverflowError(f"Can only compress {_UINT32_MAX} bytes of data.")
        out_n = ctypes.c_uint64(0)
        crc = ctypes.c_uint32(0)
        rc = _native.lib().b2d_compress_and_crc(self._ctx.handle, self._level, daddr, n, zaddr if zn else None, zn,
                                                ctypes.cast(self._buf, ctypes.c_void_p), self._cap,
                                                ctypes.byref(out_n), ctypes.byref(crc))
        if rc == _native.B2D_BUF_ERROR:
            raise OverflowError(f"Compressed output exceeds buffer size of {self._cap}")
        if rc:
            _raise(rc, "while compressing data")
        del k1, k2
        return ctypes.string_at(self._buf, out_n.value), crc.value


# ------------------------------------------------------------------ inflate half (not accelerated)
decompress = _stdzlib.decompress
decompressobj = _stdzlib.decompressobj
