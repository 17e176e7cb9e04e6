"""Test-side loaders for the checkers under oracle/ (tests only — the product
package never imports this module or anything under oracle/)."""
import base64
import ctypes
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
GOLDEN = os.path.join(ROOT, "tests", "golden")
_u8p = ctypes.c_char_p


def _ensure(lib):
    path = os.path.join(ORACLE_DIR, lib)
    if not os.path.exists(path):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, lib])
    return path


_oracle = None


def oracle():
    """ctypes handle on oracle/liboracle.so (plain-C restatement of the reference)."""
    global _oracle
    if _oracle is None:
        L = ctypes.CDLL(_ensure("liboracle.so"))
        L.orc_crc32.restype = ctypes.c_uint32
        L.orc_crc32.argtypes = [ctypes.c_uint32, _u8p, ctypes.c_size_t]
        L.orc_adler32.restype = ctypes.c_uint32
        L.orc_adler32.argtypes = [ctypes.c_uint32, _u8p, ctypes.c_size_t]
        L.orc_crc32_combine.restype = ctypes.c_uint32
        L.orc_crc32_combine.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_int64]
        L.orc_adler32_combine.restype = ctypes.c_uint32
        L.orc_adler32_combine.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_int64]
        L.orc_inflate_raw.restype = ctypes.c_int
        L.orc_inflate_raw.argtypes = [_u8p, ctypes.c_size_t, _u8p, ctypes.c_size_t, _u8p, ctypes.c_size_t,
                                      ctypes.POINTER(ctypes.c_size_t), ctypes.POINTER(ctypes.c_size_t),
                                      ctypes.c_int]
        _oracle = L
    return _oracle


def orc_inflate(stream: bytes, out_cap: int, zdict: bytes = b"", need_final: bool = True) -> bytes:
    L = oracle()
    out = ctypes.create_string_buffer(max(out_cap, 1))
    n = ctypes.c_size_t(0)
    used = ctypes.c_size_t(0)
    rc = L.orc_inflate_raw(stream, len(stream), zdict, len(zdict), out, out_cap, ctypes.byref(n),
                           ctypes.byref(used), 1 if need_final else 0)
    if rc != 0:
        raise ValueError(f"oracle inflate failed rc={rc}")
    return out.raw[:n.value]


_twin = None


def twin():
    """ctypes handle on oracle/libb2dtwin.so (CPU twin of the GPU pipeline)."""
    global _twin
    if _twin is None:
        L = ctypes.CDLL(_ensure("libb2dtwin.so"))
        L.b2d_twin_deflate.restype = ctypes.c_size_t
        L.b2d_twin_deflate.argtypes = [_u8p, ctypes.c_size_t, ctypes.c_size_t, ctypes.c_int, _u8p,
                                       ctypes.c_size_t]
        _twin = L
    return _twin


def twin_deflate(data: bytes, level: int, zdict: bytes = b"") -> bytes:
    if not data:
        return b"\x00\x00\x00\xff\xff"       # the empty job is only the sync marker
    zdict = zdict[-32768:]
    buf = zdict + data + b"\0" * 64
    cap = len(data) + (len(data) >> 10) + 1024
    out = ctypes.create_string_buffer(cap)
    n = twin().b2d_twin_deflate(buf, len(zdict), len(data), level, out, cap)
    return out.raw[:n]


def reference():
    """The real reference (python-zlib-ng built into oracle/_ref), or None if absent."""
    pkg = os.path.join(ORACLE_DIR, "_ref", "pkg")
    if not os.path.isdir(os.path.join(pkg, "zlib_ng")):
        return None
    if pkg not in sys.path:
        sys.path.insert(0, pkg)
    try:
        from zlib_ng import zlib_ng  # type: ignore
        return zlib_ng
    except Exception:
        return None


def reference_pkg():
    """(zlib_ng, gzip_ng, gzip_ng_threaded) of the real reference (oracle/_ref), or None if absent."""
    if reference() is None:
        return None
    from zlib_ng import gzip_ng, gzip_ng_threaded, zlib_ng  # type: ignore
    return zlib_ng, gzip_ng, gzip_ng_threaded


def golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def b64(s):
    return base64.b64decode(s)
