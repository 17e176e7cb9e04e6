"""ctypes binding of libb2d.so (C ABI in include/b2d.h).

This is the stub a maintainer of the reference would add next to
``src/zlib_ng/zlib_ngmodule.c``: every function here maps 1:1 onto an
``extern "C"`` entry point.  There is no CPU fallback: if the library is
missing, or no CUDA device can be opened, importing works but the first call
raises ``RuntimeError`` / ``zlib_ng.error`` loudly.
"""
from __future__ import annotations

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2D_LIB") or os.path.join(_HERE, "libb2d.so")   # B2D_LIB: another build of the same ABI (A/B runs)

B2D_OK = 0
B2D_STREAM_ERROR = -2
B2D_MEM_ERROR = -4
B2D_BUF_ERROR = -5
B2D_CUDA_ERROR = -100

_u8p = ctypes.c_void_p
_u64 = ctypes.c_uint64
_u32 = ctypes.c_uint32
_P = ctypes.POINTER

# every symbol include/b2d.h declares: (restype, argtypes)
SIGNATURES = {
    "b2d_ctx_create": (ctypes.c_int, [ctypes.c_int, _u64, _P(ctypes.c_void_p)]),
    "b2d_ctx_destroy": (None, [ctypes.c_void_p]),
    "b2d_last_error": (ctypes.c_char_p, []),
    "b2d_device_count": (ctypes.c_int, []),
    "b2d_deflate_bound": (_u64, [_u64]),
    "b2d_compress_and_crc": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _u8p, _u64, _u8p, _u64, _u8p, _u64,
                                            _P(_u64), _P(_u32)]),
    "b2d_deflate_host": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, _u8p, _u64, _u8p, _u64, ctypes.c_int, _u8p,
                                        _u64, _P(_u64), _P(_u32), _P(_u32)]),
    "b2d_crc32": (ctypes.c_int, [ctypes.c_void_p, _u32, _u8p, _u64, _P(_u32)]),
    "b2d_adler32": (ctypes.c_int, [ctypes.c_void_p, _u32, _u8p, _u64, _P(_u32)]),
    "b2d_crc32_combine": (_u32, [_u32, _u32, ctypes.c_int64]),
    "b2d_adler32_combine": (_u32, [_u32, _u32, ctypes.c_int64]),
    "b2d_deflate_device": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _u8p, _u64, _u64, _u8p, _u64, _P(_u64),
                                          _P(_u32), _P(ctypes.c_float)]),
    "b2d_crc32_device": (ctypes.c_int, [ctypes.c_void_p, _u32, _u8p, _u64, _P(_u32), _P(ctypes.c_float)]),
    "b2d_adler32_device": (ctypes.c_int, [ctypes.c_void_p, _u32, _u8p, _u64, _P(_u32), _P(ctypes.c_float)]),
    "b2d_noalloc_max_bytes": (_u64, []),
    "b2d_adler32_device_noalloc": (ctypes.c_int, [ctypes.c_int, _u32, _u8p, _u64, _P(_u32)]),
    "b2d_device_alloc": (ctypes.c_int, [ctypes.c_void_p, _u64, _P(ctypes.c_void_p)]),
    "b2d_device_free": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "b2d_host_alloc": (ctypes.c_int, [ctypes.c_void_p, _u64, _P(ctypes.c_void_p)]),
    "b2d_host_free": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p]),
    "b2d_copy_h2d": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, _u64]),
    "b2d_copy_d2h": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, _u64]),
    "b2d_device_sync": (ctypes.c_int, [ctypes.c_void_p]),
    "b2d_pinned_alloc": (ctypes.c_int, [_u64, _P(ctypes.c_void_p)]),
    "b2d_pinned_free": (ctypes.c_int, [ctypes.c_void_p]),
    "b2d_host_is_pinned": (ctypes.c_int, [ctypes.c_void_p]),
    "b2d_host_copy": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, _u64, ctypes.c_int]),
    "b2d_launch_count": (_u64, [ctypes.c_void_p]),
    "b2d_block_index": (ctypes.c_int, [ctypes.c_void_p, _P(_u64), _u64, _P(_u64)]),
    "b2d_compress_and_crc_indexed": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _u8p, _u64, _u8p, _u64, _u8p, _u64,
                                                    _P(_u64), _P(_u32), _P(_u64), _u64, _P(_u64)]),
    "b2d_inflate_device": (ctypes.c_int, [ctypes.c_void_p, _u8p, _u64, _P(_u64), _u64, _u8p, _u64, _u8p, _u64,
                                          _P(ctypes.c_float)]),
    "b2d_inflate_host": (ctypes.c_int, [ctypes.c_void_p, _u8p, _u64, _P(_u64), _u64, _u8p, _u64, _u8p, _u64, _P(_u32)]),
    "b2d_mctx_create": (ctypes.c_int, [_P(ctypes.c_int), ctypes.c_int, _u64, _P(ctypes.c_void_p)]),
    "b2d_mctx_destroy": (None, [ctypes.c_void_p]),
    "b2d_mctx_devices": (ctypes.c_int, [ctypes.c_void_p]),
    "b2d_mctx_launch_count": (_u64, [ctypes.c_void_p]),
    "b2d_deflate_host_multi": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, _u8p, _u64, _u8p, _u64, ctypes.c_int,
                                              _u8p, _u64, _P(_u64), _P(_u32), _P(_u32), _P(_u64), _u64, _P(_u64)]),
    "b2d_compress_and_crc_multi": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _u8p, _u64, _u8p, _u64, _u8p, _u64,
                                                  _P(_u64), _P(_u32), _P(_u64), _u64, _P(_u64)]),
    "b2d_bgzf_scan": (ctypes.c_int, [_u8p, _u64, _u64, _P(_u64), _P(_u32), _P(_u32), _P(_u32), _P(_u64), _P(_u64)]),
    "b2d_inflate_members_host": (ctypes.c_int, [ctypes.c_void_p, _u8p, _u64, _P(_u64), _P(_u32), _P(_u32), _P(_u32), _u64, _u8p, _u64,
                                                _P(_u64)]),
    "b2d_subblock_bits": (ctypes.c_int, [ctypes.c_void_p, _P(_u32), _u64, _P(_u64)]),
    "b2d_inflate_device_sub": (ctypes.c_int, [ctypes.c_void_p, _u8p, _u64, _P(_u64), _P(_u32), _u64, _u8p, _u64, _u8p, _u64,
                                              _P(ctypes.c_float)]),
    "b2d_inflate_host_sub": (ctypes.c_int, [ctypes.c_void_p, _u8p, _u64, _P(_u64), _P(_u32), _u64, _u8p, _u64, _u8p, _u64, _P(_u32)]),
    "b2d_inflate_host_decode": (ctypes.c_int, [ctypes.c_void_p, _u8p, _u64, _P(_u64), _P(_u32), _u64, _u64, _u64]),
    "b2d_inflate_host_resolve": (ctypes.c_int, [ctypes.c_void_p, _u8p, _u64, _u8p, _u64, _P(_u32)]),
    "b2d_compress_and_crc_multi_sub": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _u8p, _u64, _u8p, _u64, _u8p, _u64,
                                                      _P(_u64), _P(_u32), _P(_u64), _P(_u32), _u64, _P(_u64)]),
    "b2d_profile": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "b2d_profile_read": (ctypes.c_int, [ctypes.c_void_p, _P(ctypes.c_double), _P(_u64), ctypes.c_int]),
}

_lib = None
_lib_lock = threading.Lock()


def lib() -> ctypes.CDLL:
    """Load libb2d.so (once) and attach the prototypes."""
    global _lib
    if _lib is None:
        with _lib_lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        f"{LIB_PATH} is missing: build it with `make -C python-zlib-ng_b200/csrc` "
                        "(there is no CPU fallback)")
                L = ctypes.CDLL(LIB_PATH)
                for name, (res, args) in SIGNATURES.items():
                    fn = getattr(L, name)
                    fn.restype = res
                    fn.argtypes = args
                _lib = L
    return _lib


def last_error() -> str:
    return (lib().b2d_last_error() or b"").decode("utf-8", "replace")


class NativeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libb2d error {code}: {msg}")
        self.code = code
        self.msg = msg


def check(rc: int) -> None:
    if rc != B2D_OK:
        raise NativeError(rc, last_error())


class Context:
    """One GPU context (b2d_ctx)."""

    def __init__(self, device: int = 0, max_batch: int = 0):
        self._h = ctypes.c_void_p()
        check(lib().b2d_ctx_create(device, max_batch, ctypes.byref(self._h)))
        self.device = device

    @property
    def handle(self):
        if not self._h:
            raise ValueError("context is closed")
        return self._h

    def close(self):
        if self._h:
            lib().b2d_ctx_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- helpers used by bench / tests for device-resident runs
    def device_alloc(self, n: int) -> int:
        p = ctypes.c_void_p()
        check(lib().b2d_device_alloc(self.handle, n, ctypes.byref(p)))
        return p.value

    def device_free(self, p: int):
        check(lib().b2d_device_free(self.handle, ctypes.c_void_p(p)))

    def host_alloc(self, n: int) -> int:
        p = ctypes.c_void_p()
        check(lib().b2d_host_alloc(self.handle, n, ctypes.byref(p)))
        return p.value

    def host_free(self, p: int):
        check(lib().b2d_host_free(self.handle, ctypes.c_void_p(p)))

    def h2d(self, d: int, h, n: int):
        check(lib().b2d_copy_h2d(self.handle, ctypes.c_void_p(d), h, n))

    def d2h(self, h, d: int, n: int):
        check(lib().b2d_copy_d2h(self.handle, h, ctypes.c_void_p(d), n))

    def sync(self):
        check(lib().b2d_device_sync(self.handle))

    def launch_count(self) -> int:
        return int(lib().b2d_launch_count(self.handle))


def pinned_alloc(n: int) -> int:
    p = ctypes.c_void_p()
    check(lib().b2d_pinned_alloc(n, ctypes.byref(p)))
    return p.value


def pinned_free(p: int) -> None:
    check(lib().b2d_pinned_free(ctypes.c_void_p(p)))


class MultiContext:
    """Several GPU contexts working on one stream (b2d_mctx): batch j of a call runs on devices[j % len(devices)]."""

    def __init__(self, devices, max_batch: int = 0):
        self.devices = list(devices)
        arr = (ctypes.c_int * len(self.devices))(*self.devices)
        self._h = ctypes.c_void_p()
        check(lib().b2d_mctx_create(arr, len(self.devices), max_batch, ctypes.byref(self._h)))
        self.device = self.devices[0]         # where page-locked buffers of this context's users are allocated from

    @property
    def handle(self):
        if not self._h:
            raise ValueError("context is closed")
        return self._h

    def close(self):
        if self._h:
            lib().b2d_mctx_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def launch_count(self) -> int:
        return int(lib().b2d_mctx_launch_count(self.handle))

    def compress_and_crc(self, level, daddr, n, zaddr, zn, out_addr, out_cap, want_sizes=False, want_subs=False):
        """-> (out_n, crc, [compressed size of every 128 KiB block] or None[, bit offsets of the sub-blocks]); raises NativeError"""
        out_n, crc, cnt = _u64(0), _u32(0), _u64(0)
        nb = (n + 131071) // 131072
        sizes = (_u64 * max(nb, 1))() if want_sizes else None
        subs = (_u32 * max(4 * nb, 1))() if (want_sizes and want_subs) else None
        check(lib().b2d_compress_and_crc_multi_sub(self.handle, level, daddr, n, zaddr if zn else None, zn, out_addr, out_cap,
                                                   ctypes.byref(out_n), ctypes.byref(crc), sizes, subs, nb if want_sizes else 0,
                                                   ctypes.byref(cnt)))
        if want_subs:
            return out_n.value, crc.value, (list(sizes[:nb]) if want_sizes else None), (list(subs[:4 * nb]) if subs is not None else None)
        return out_n.value, crc.value, (list(sizes[:nb]) if want_sizes else None)


_default = {}
_default_lock = threading.Lock()


def default_context(device: int = 0, slot: int = 0) -> Context:
    """Process-wide contexts, created on first use.  Slot 0 of a device serves the one-shot calls; the
    streaming writer also opens slots 1.. so that several batches are in flight per GPU."""
    key = (device, slot)
    ctx = _default.get(key)
    if ctx is None:
        with _default_lock:
            ctx = _default.get(key)
            if ctx is None:
                mb = int(os.environ.get("B2D_MAX_BATCH_MIB", "64")) << 20
                ctx = Context(device, mb)
                _default[key] = ctx
    return ctx


def _addr(buf):
    """(address, length, keepalive) of a bytes-like object, zero-copy."""
    if isinstance(buf, bytes):
        return ctypes.cast(ctypes.c_char_p(buf), ctypes.c_void_p).value or 0, len(buf), buf
    mv = memoryview(buf)
    if not mv.c_contiguous:
        raise BufferError("memoryview: underlying buffer is not C-contiguous")
    n = mv.nbytes
    if n == 0:
        return 0, 0, mv
    if mv.readonly:
        # ctypes' from_buffer wants a writable export: ask the buffer protocol directly (PyObject_GetBuffer, what the
        # reference does with "y*", zlib_ngmodule.c:1819) and release the view when the caller drops `keep`
        view = _PyBuffer()
        if ctypes.pythonapi.PyObject_GetBuffer(ctypes.py_object(mv), ctypes.byref(view), 0) != 0:   # PyBUF_SIMPLE
            raise BufferError("object does not export a simple buffer")
        return view.buf or 0, n, (_BufferHold(view), mv)
    c = (ctypes.c_char * n).from_buffer(mv)
    return ctypes.addressof(c), n, (c, mv)


class _PyBuffer(ctypes.Structure):
    """Py_buffer (cpython/object.h)"""
    _fields_ = [("buf", ctypes.c_void_p), ("obj", ctypes.c_void_p), ("len", ctypes.c_ssize_t), ("itemsize", ctypes.c_ssize_t),
                ("readonly", ctypes.c_int), ("ndim", ctypes.c_int), ("format", ctypes.c_char_p),
                ("shape", ctypes.POINTER(ctypes.c_ssize_t)), ("strides", ctypes.POINTER(ctypes.c_ssize_t)),
                ("suboffsets", ctypes.POINTER(ctypes.c_ssize_t)), ("internal", ctypes.c_void_p)]


ctypes.pythonapi.PyObject_GetBuffer.argtypes = [ctypes.py_object, ctypes.POINTER(_PyBuffer), ctypes.c_int]
ctypes.pythonapi.PyObject_GetBuffer.restype = ctypes.c_int
ctypes.pythonapi.PyBuffer_Release.argtypes = [ctypes.POINTER(_PyBuffer)]
ctypes.pythonapi.PyBuffer_Release.restype = None


class _BufferHold:
    def __init__(self, view):
        self._view = view

    def __del__(self):
        if self._view is not None:
            ctypes.pythonapi.PyBuffer_Release(ctypes.byref(self._view))
            self._view = None
