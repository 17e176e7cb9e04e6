"""gzip module facade on the B200 compressor — mirrors ``zlib_ng.gzip_ng``
(reference: src/zlib_ng/gzip_ng.py:51-205): ``open``, ``GzipNGFile``,
``compress``, ``decompress``.  Writing goes through
``zlib_ng_b200.zlib_ng.compressobj`` (GPU); reading is CPython's gzip (inflate
is outside the accelerated path).  The CLI of the reference (gzip_ng.py:208-317)
is out of scope."""
from __future__ import annotations

import gzip
import io
import struct
import time

from . import zlib_ng

__all__ = ["GzipFile", "open", "compress", "decompress", "BadGzipFile", "READ_BUFFER_SIZE"]

_COMPRESS_LEVEL_FAST = zlib_ng.Z_BEST_SPEED
_COMPRESS_LEVEL_TRADEOFF = zlib_ng.Z_DEFAULT_COMPRESSION
_COMPRESS_LEVEL_BEST = zlib_ng.Z_BEST_COMPRESSION
READ_BUFFER_SIZE = 512 * 1024
READ, WRITE = gzip.READ, gzip.WRITE
BadGzipFile = gzip.BadGzipFile


def open(filename, mode="rb", compresslevel=_COMPRESS_LEVEL_TRADEOFF, encoding=None, errors=None, newline=None):
    """gzip.open with the GPU compressor behind it (reference gzip_ng.py:51-95)."""
    if "t" in mode:
        if "b" in mode:
            raise ValueError("Invalid mode: %r" % (mode,))
    else:
        if encoding is not None:
            raise ValueError("Argument 'encoding' not supported in binary mode")
        if errors is not None:
            raise ValueError("Argument 'errors' not supported in binary mode")
        if newline is not None:
            raise ValueError("Argument 'newline' not supported in binary mode")
    gz_mode = mode.replace("t", "")
    if isinstance(filename, (str, bytes)) or hasattr(filename, "__fspath__"):
        binary_file = GzipNGFile(filename, gz_mode, compresslevel)
    elif hasattr(filename, "read") or hasattr(filename, "write"):
        binary_file = GzipNGFile(None, gz_mode, compresslevel, filename)
    else:
        raise TypeError("filename must be a str or bytes object, or a file")
    if "t" in mode:
        return io.TextIOWrapper(binary_file, encoding, errors, newline)
    return binary_file


class GzipNGFile(gzip.GzipFile):
    """gzip.GzipFile whose deflate and CRC run on the GPU.  The compressor object
    keeps the member CRC itself (computed on the device next to the deflate
    kernels), so write() makes one native call, not two."""

    def __init__(self, filename=None, mode=None, compresslevel=_COMPRESS_LEVEL_BEST, fileobj=None, mtime=None):
        super().__init__(filename, mode, compresslevel, fileobj, mtime)
        if self.mode == WRITE:
            self.compress = zlib_ng.compressobj(compresslevel, zlib_ng.DEFLATED, -zlib_ng.MAX_WBITS,
                                                zlib_ng.DEF_MEM_LEVEL, 0)
            self.compress._kind_crc = True

    def __repr__(self):
        s = repr(self.fileobj)
        return "<gzip_ng " + s[1:-1] + " " + hex(id(self)) + ">"

    # GzipFile.close() writes self.crc after flushing the compressor; by then the
    # compressor has seen every byte, so its running CRC is the member CRC.
    @property
    def crc(self):
        c = getattr(self, "compress", None)
        if isinstance(c, zlib_ng._Compress):
            return c._crc
        return self.__dict__.get("_crc_value", 0)

    @crc.setter
    def crc(self, value):
        self.__dict__["_crc_value"] = value

    def write(self, data):
        self._check_not_closed()
        if self.mode != WRITE:
            import errno
            raise OSError(errno.EBADF, "write() on read-only GzipNGFile object")
        if self.fileobj is None:
            raise ValueError("write() on closed GzipNGFile object")
        if isinstance(data, bytes):
            length = len(data)
        else:
            data = memoryview(data)
            length = data.nbytes
        if length > 0:
            self.fileobj.write(self.compress.compress(data))
            self.size += length
            self.offset += length
        return length


GzipFile = GzipNGFile


def compress(data, compresslevel=_COMPRESS_LEVEL_BEST, *, mtime=None):
    """One-shot gzip (reference gzip_ng.py:182-195): zlib_ng.compress(wbits=31) with
    the mtime and OS bytes of the header patched."""
    gzip_data = zlib_ng.compress(data, level=compresslevel, wbits=31)
    if mtime is None:
        mtime = time.time()
    header = struct.pack("<4sLBB", gzip_data[:4], int(mtime), gzip_data[8], 255)
    return header + gzip_data[10:]


def decompress(data):
    """gzip_ng.decompress (reference gzip_ng.py:181-205).  Data written by gzip_ng_threaded's writer carries a block
    index and is inflated on the GPU (gzip_ng_threaded._ThreadedGzipReader); anything else goes to CPython's gzip."""
    from . import gzip_ng_threaded
    if len(data) > gzip_ng_threaded._LOCATOR_MEMBER + 20:
        reader = gzip_ng_threaded._ThreadedGzipReader(io.BytesIO(data))
        try:
            if reader._gpu is not None:
                return reader.readall()
        finally:
            reader.close()
    return gzip.decompress(data)
