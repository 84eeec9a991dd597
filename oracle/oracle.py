"""ctypes bindings for the CPU checker libraries (TEST INFRASTRUCTURE ONLY).

Two libraries share one call shape:
  * ``port``      -- oracle/libb2oracle.so, the plain-C restatement (b2o_*)
  * ``reference`` -- oracle/_ref/libb2ref.so, the reference's own sources (b2r_*)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.  The product never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "libb2oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libb2ref.so")

AZ09, RAND62, WORDS = 0, 1, 2
CORPORA = {"az09": AZ09, "rand62": RAND62, "words": WORDS}

_u8p = C.POINTER(C.c_uint8)
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)
_i32p = C.POINTER(C.c_int32)


def build(ref: bool = True) -> None:
    """Compile the checker libraries (building the checker is not using it)."""
    targets = ["liboracle"] + (["ref"] if ref else [])
    subprocess.run(["make", "-s", "-C", HERE] + targets, check=True)


def _ptr(a: np.ndarray, typ):
    return a.ctypes.data_as(typ)


def _as_u8(b) -> np.ndarray:
    if isinstance(b, np.ndarray):
        assert b.dtype == np.uint8
        return np.ascontiguousarray(b)
    return np.frombuffer(bytes(b), dtype=np.uint8)


class Codec:
    """One CPU codec library (port or reference) behind a uniform Python API."""

    def __init__(self, path: str, prefix: str, kind: str):
        self.kind = kind
        self.path = path
        self.lib = C.CDLL(path)
        p = prefix
        L = self.lib

        def fn(name, res, args):
            f = getattr(L, p + name)
            f.restype = res
            f.argtypes = args
            return f

        self._max = fn("snappy_max_compressed_length", C.c_size_t, [C.c_size_t])
        self._comp = fn("snappy_compress", C.c_size_t, [_u8p, C.c_size_t, _u8p])
        self._ulen_l = fn("snappy_read_ulen_lenient", C.c_int, [_u8p, C.c_size_t, _u32p])
        self._ulen_s = fn("snappy_read_ulen_strict", C.c_int, [_u8p, C.c_size_t, _u32p])
        self._uncomp = fn("snappy_uncompress", C.c_int, [_u8p, C.c_size_t, _u8p, C.c_size_t, _u32p])
        self._valid = fn("snappy_is_valid", C.c_int, [_u8p, C.c_size_t])
        self._extend = fn("crc32c_extend", C.c_uint32, [C.c_uint32, _u8p, C.c_size_t])
        self._mask = fn("crc32c_mask", C.c_uint32, [C.c_uint32])
        self._unmask = fn("crc32c_unmask", C.c_uint32, [C.c_uint32])
        self._b_crc = fn("batch_crc32c", C.c_double, [_u8p, _u64p, _u32p, C.c_uint32, _u32p, C.c_int])
        self._b_comp = fn("batch_compress", C.c_double,
                          [_u8p, _u64p, _u32p, C.c_uint32, _u8p, _u64p, _u32p, C.c_int])
        self._b_uncomp = fn("batch_uncompress", C.c_double,
                            [_u8p, _u64p, _u32p, C.c_uint32, _u8p, _u64p, _u32p, _u32p, _i32p, C.c_int])

    # ---- single message ------------------------------------------------
    def max_compressed_length(self, n: int) -> int:
        return self._max(n)

    def compress(self, data) -> bytes:
        a = _as_u8(data)
        out = np.empty(self._max(a.size), dtype=np.uint8)
        n = self._comp(_ptr(a, _u8p), a.size, _ptr(out, _u8p))
        return out[:n].tobytes()

    def uncompress(self, data, cap: int | None = None):
        """Returns (ok, bytes).  Lenient preamble (Source-based path)."""
        a = _as_u8(data)
        ulen = C.c_uint32(0)
        k = self._ulen_l(_ptr(a, _u8p), a.size, C.byref(ulen))
        if not k:
            return False, b""
        if cap is None:
            cap = ulen.value
        if cap > (1 << 31):
            return False, b""
        out = np.empty(max(cap, 1), dtype=np.uint8)
        ok = self._uncomp(_ptr(a, _u8p), a.size, _ptr(out, _u8p), cap, C.byref(ulen))
        return bool(ok), (out[: ulen.value].tobytes() if ok else b"")

    def is_valid(self, data) -> bool:
        a = _as_u8(data)
        return bool(self._valid(_ptr(a, _u8p), a.size))

    def ulen_lenient(self, data):
        a = _as_u8(data)
        v = C.c_uint32(0)
        k = self._ulen_l(_ptr(a, _u8p), a.size, C.byref(v))
        return (v.value if k else None)

    def ulen_strict(self, data):
        a = _as_u8(data)
        v = C.c_uint32(0)
        k = self._ulen_s(_ptr(a, _u8p), a.size, C.byref(v))
        return (v.value if k else None)

    def crc32c_extend(self, init: int, data) -> int:
        a = _as_u8(data)
        return self._extend(init & 0xFFFFFFFF, _ptr(a, _u8p), a.size)

    def crc32c(self, data) -> int:
        return self.crc32c_extend(0, data)

    def mask(self, c: int) -> int:
        return self._mask(c)

    def unmask(self, c: int) -> int:
        return self._unmask(c)

    # ---- batches ---------------------------------------------------------
    def batch_crc32c(self, data: np.ndarray, off: np.ndarray, length: np.ndarray, threads: int = 1):
        n = len(length)
        crc = np.empty(n, dtype=np.uint32)
        dt = self._b_crc(_ptr(data, _u8p), _ptr(off, _u64p), _ptr(length, _u32p), n, _ptr(crc, _u32p), threads)
        return crc, dt

    def batch_compress(self, data, off, length, threads: int = 1):
        """Returns (out_arena, out_off, out_len, seconds); message i at out_off[i]."""
        n = len(length)
        cap = (32 + length.astype(np.uint64) + length.astype(np.uint64) // 6)
        out_off = np.zeros(n, dtype=np.uint64)
        if n:
            np.cumsum(cap[:-1], out=out_off[1:])
        total = int(cap.sum()) if n else 0
        out = np.empty(max(total, 1), dtype=np.uint8)
        out_len = np.empty(n, dtype=np.uint32)
        dt = self._b_comp(_ptr(data, _u8p), _ptr(off, _u64p), _ptr(length, _u32p), n,
                          _ptr(out, _u8p), _ptr(out_off, _u64p), _ptr(out_len, _u32p), threads)
        return out, out_off, out_len, dt

    def batch_uncompress(self, data, off, length, out_off, out_cap, threads: int = 1):
        n = len(length)
        total = int(out_off[-1] + out_cap[-1]) if n else 0
        out = np.empty(max(total, 1), dtype=np.uint8)
        ulen = np.empty(n, dtype=np.uint32)
        ok = np.empty(n, dtype=np.int32)
        dt = self._b_uncomp(_ptr(data, _u8p), _ptr(off, _u64p), _ptr(length, _u32p), n, _ptr(out, _u8p),
                            _ptr(out_off, _u64p), _ptr(out_cap, _u32p), _ptr(ulen, _u32p), _ptr(ok, _i32p), threads)
        return out, ulen, ok, dt


_port = None
_ref = None


def port() -> Codec:
    global _port
    if _port is None:
        if not os.path.exists(PORT_SO):
            build(ref=False)
        _port = Codec(PORT_SO, "b2o_", "port")
        L = _port.lib
        L.b2o_corpus_fill.restype = None
        L.b2o_corpus_fill.argtypes = [C.c_int, C.c_uint64, _u8p, C.c_size_t]
        L.b2o_corpus_fill_batch.restype = None
        L.b2o_corpus_fill_batch.argtypes = [C.c_int, C.c_uint64, C.c_uint32, _u64p, _u32p, _u8p, C.c_int]
        L.b2o_snappy_compress_fragment.restype = C.c_size_t
        L.b2o_snappy_compress_fragment.argtypes = [_u8p, C.c_size_t, _u8p]
        L.b2o_batch_equal.restype = C.c_uint32
        L.b2o_batch_equal.argtypes = [_u8p, _u64p, _u8p, _u64p, _u32p, C.c_uint32, _i32p, _u32p, C.c_int]
        L.b2o_crc32c_extend_tables.restype = C.c_uint32
        L.b2o_crc32c_extend_tables.argtypes = [C.c_uint32, _u8p, C.c_size_t]
    return _port


def reference() -> Codec | None:
    """The reference's own code, or None when oracle/_ref was never built."""
    global _ref
    if _ref is None:
        if not os.path.exists(REF_SO):
            if os.path.isdir("/root/reference/src/butil"):
                build(ref=True)
            else:
                return None
        _ref = Codec(REF_SO, "b2r_", "reference")
        L = _ref.lib
        L.b2r_snappy_compress_chunked.restype = C.c_size_t
        L.b2r_snappy_compress_chunked.argtypes = [_u8p, C.c_size_t, C.c_size_t, _u8p]
        L.b2r_snappy_uncompress_sink.restype = C.c_int
        L.b2r_snappy_uncompress_sink.argtypes = [_u8p, C.c_size_t, C.c_size_t, _u8p, C.c_size_t,
                                                 C.POINTER(C.c_size_t)]
        L.b2r_snappy_uncompress_string.restype = C.c_int
        L.b2r_snappy_uncompress_string.argtypes = [_u8p, C.c_size_t]
        L.b2r_snappy_uncompress_as_much.restype = C.c_size_t
        L.b2r_snappy_uncompress_as_much.argtypes = [_u8p, C.c_size_t, _u8p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.b2r_crc32c_is_fast.restype = C.c_int
        L.b2r_crc32c_is_fast.argtypes = []
    return _ref


def ref_uncompress_as_much(data):
    """The reference's snappy::UncompressAsMuchAsPossible on one stream: (returned value, bytes its sink received).
    Reference library only (the port has no such entry)."""
    ref = reference()
    a = _as_u8(data) if len(data) else np.zeros(1, dtype=np.uint8)
    cap = max(64, len(data) * 22 + 70000)
    out = np.empty(cap, dtype=np.uint8)
    sunk = C.c_size_t(0)
    r = ref.lib.b2r_snappy_uncompress_as_much(_ptr(a, _u8p), len(data), _ptr(out, _u8p), cap, C.byref(sunk))
    return int(r), out[: sunk.value].tobytes()


def best() -> Codec:
    """Reference library when present, else the port."""
    r = reference()
    return r if r is not None else port()


def batch_equal(a: np.ndarray, a_off: np.ndarray, b: np.ndarray, b_off: np.ndarray, length: np.ndarray,
                threads: int = 0):
    """Compare message i of arena `a` (at a_off[i]) with message i of arena `b` (at b_off[i]), length[i] bytes
    each, on `threads` host threads.  Returns (number of differing messages, index of the first or None)."""
    n = len(length)
    if n == 0:
        return 0, None
    a_off = np.ascontiguousarray(a_off, dtype=np.uint64)
    b_off = np.ascontiguousarray(b_off, dtype=np.uint64)
    length = np.ascontiguousarray(length, dtype=np.uint32)
    ok = np.empty(n, dtype=np.int32)
    first = C.c_uint32(0)
    if threads <= 0:
        threads = min(os.cpu_count() or 1, 64)
    bad = port().lib.b2o_batch_equal(_ptr(a, _u8p), _ptr(a_off, _u64p), _ptr(b, _u8p), _ptr(b_off, _u64p),
                                     _ptr(length, _u32p), n, _ptr(ok, _i32p), C.byref(first), threads)
    return int(bad), (int(first.value) if bad else None)


# ---- corpora -------------------------------------------------------------
def corpus(kind, n: int, index: int = 0) -> np.ndarray:
    k = CORPORA[kind] if isinstance(kind, str) else kind
    out = np.empty(max(n, 1), dtype=np.uint8)
    port().lib.b2o_corpus_fill(k, index, _ptr(out, _u8p), n)
    return out[:n]


def corpus_batch(kind, lengths: np.ndarray, first_index: int = 0, align: int = 16, threads: int = 0):
    """Build a packed arena of messages.  Returns (arena u8, off u64[n], len u32[n])."""
    k = CORPORA[kind] if isinstance(kind, str) else kind
    lengths = np.ascontiguousarray(lengths, dtype=np.uint32)
    n = len(lengths)
    padded = (lengths.astype(np.uint64) + (align - 1)) // align * align
    off = np.zeros(n, dtype=np.uint64)
    if n:
        np.cumsum(padded[:-1], out=off[1:])
    total = int(padded.sum()) if n else 0
    arena = np.zeros(total + 64, dtype=np.uint8)
    if threads <= 0:
        threads = min(os.cpu_count() or 1, 64)
    port().lib.b2o_corpus_fill_batch(k, first_index, n, _ptr(off, _u64p), _ptr(lengths, _u32p),
                                     _ptr(arena, _u8p), threads)
    return arena, off, lengths
