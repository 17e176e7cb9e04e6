"""The ``gzip`` facade of the reference (src/zlib_ng/gzip_ng.py:51-205) on the B200 compressor.

That module is serial glue around ``zlib_ng.compressobj`` and outside the accelerated path, so nearly
all of it is left to the code it was derived from, CPython's ``gzip``:

* ``GzipNGFile`` is ``gzip.GzipFile`` with the deflate object swapped for this package's (everything else, the
  write path's checks and CRC included, inherited); ``open()`` builds one, wrapped for text modes;
* ``compress()`` is ``zlib_ng.compress(wbits=31)`` with the header's mtime / OS bytes set (gzip_ng.py:182-195);
* ``decompress()`` inflates indexed / BGZF data on the GPU and leaves anything else to CPython.

The reference's command line (gzip_ng.py:208-317) is not rebuilt.
"""
from __future__ import annotations

import gzip
import io
import os
import struct
import time

from . import zlib_ng

__all__ = ["GzipFile", "GzipNGFile", "open", "compress", "decompress", "BadGzipFile", "READ_BUFFER_SIZE"]

_COMPRESS_LEVEL_FAST = zlib_ng.Z_BEST_SPEED
_COMPRESS_LEVEL_TRADEOFF = zlib_ng.Z_DEFAULT_COMPRESSION
_COMPRESS_LEVEL_BEST = zlib_ng.Z_BEST_COMPRESSION
READ_BUFFER_SIZE = 512 * 1024
BadGzipFile = gzip.BadGzipFile


def open(filename, mode="rb", compresslevel=_COMPRESS_LEVEL_TRADEOFF, encoding=None, errors=None, newline=None):
    """gzip.open()'s contract, returning GzipNGFile objects (what gzip_ng_threaded.open(threads=0) hands out too,
    reference tests/test_gzip_ng_threaded.py:53-57)."""
    text = "t" in mode
    if text and "b" in mode:
        raise ValueError(f"Invalid mode: {mode!r}")
    if not text:
        for name, value in (("encoding", encoding), ("errors", errors), ("newline", newline)):
            if value is not None:
                raise ValueError(f"Argument '{name}' not supported in binary mode")
    if isinstance(filename, (str, bytes, os.PathLike)):
        gz = GzipNGFile(filename, mode.replace("t", ""), compresslevel)
    elif hasattr(filename, "read") or hasattr(filename, "write"):
        gz = GzipNGFile(None, mode.replace("t", ""), compresslevel, filename)
    else:
        raise TypeError("filename must be a str or bytes object, or a file")
    return io.TextIOWrapper(gz, encoding, errors, newline) if text else gz


class GzipNGFile(gzip.GzipFile):
    """gzip.GzipFile whose deflate runs on the GPU: only the compressor object differs from the parent class."""

    def __init__(self, filename=None, mode=None, compresslevel=_COMPRESS_LEVEL_BEST, fileobj=None, mtime=None):
        super().__init__(filename, mode, compresslevel, fileobj, mtime)
        if self.mode == gzip.WRITE:
            self.compress = zlib_ng.compressobj(compresslevel, zlib_ng.DEFLATED, -zlib_ng.MAX_WBITS, zlib_ng.DEF_MEM_LEVEL, 0)

    def __repr__(self):
        return f"<gzip_ng {repr(self.fileobj)[1:-1]} {hex(id(self))}>"


GzipFile = GzipNGFile


def compress(data, compresslevel=_COMPRESS_LEVEL_BEST, *, mtime=None):
    """One gzip member: the GPU's wbits=31 stream with mtime and the OS byte (255, "unknown", like gzip.compress) set."""
    blob = zlib_ng.compress(data, level=compresslevel, wbits=31)
    stamp = int(time.time() if mtime is None else mtime)
    return blob[:4] + struct.pack("<L", stamp) + blob[8:9] + b"\xff" + blob[10:]


def decompress(data):
    """gzip.decompress(); data with a block index of gzip_ng_threaded's writer, or BGZF data, is inflated on the GPU."""
    from . import gzip_ng_threaded
    if len(data) >= 28:
        reader = gzip_ng_threaded._ThreadedGzipReader(io.BytesIO(data))
        try:
            if reader._gpu is not None:
                return reader.readall()
        finally:
            reader.close()
    return gzip.decompress(data)
