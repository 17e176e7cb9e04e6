"""B200-native drop-in for the compress/checksum half of python-zlib-ng.

Modules mirror the reference package ``zlib_ng`` (src/zlib_ng/__init__.py):
``zlib_ng`` (functions of the C extension), ``gzip_ng_threaded`` (the
block-parallel writer) and ``gzip_ng``.  Everything on the hot path runs in
hand-written sm_100a kernels behind the C ABI of ``libb2d.so``.
"""
__version__ = "0.1.0"
