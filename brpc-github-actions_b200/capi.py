"""ctypes binding of include/b2c_codec.h (libb2c.so).

This is harness plumbing for tests and bench.py: it declares the C ABI, moves
numpy arrays in and out, and nothing else.  There is no CPU fallback: if the
library is missing or no GPU is usable, construction fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import nvcc_build as _build

OK, CORRUPT, TOO_SMALL, CRC_MISMATCH = 0, 1, 2, 3
E_CUDA, E_ARG, E_NOMEM, E_CAPACITY, E_NODEVICE = -1, -2, -3, -4, -5

_u8p = C.POINTER(C.c_uint8)
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)
_i32p = C.POINTER(C.c_int32)
_vp = C.c_void_p


class IoVec(C.Structure):
    _fields_ = [("iov_base", C.c_void_p), ("iov_len", C.c_size_t)]


_lib = None

# name -> (restype, argtypes); every symbol include/b2c_codec.h declares
SYMBOLS = {
    "b2c_ctx_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "b2c_ctx_destroy": (None, [_vp]),
    "b2c_last_error": (C.c_char_p, [_vp]),
    "b2c_ctx_device": (C.c_int, [_vp]),
    "b2c_ctx_stream": (_vp, [_vp]),
    "b2c_ctx_sync": (C.c_int, [_vp]),
    "b2c_ctx_set_option": (C.c_int, [_vp, C.c_int, C.c_int]),
    "b2c_ctx_set_tunable": (C.c_int, [_vp, C.c_char_p, C.c_long]),
    "b2c_ctx_launch_count": (C.c_uint64, [_vp]),
    "b2c_dev_alloc": (C.c_int, [_vp, C.c_size_t, C.POINTER(_vp)]),
    "b2c_dev_free": (C.c_int, [_vp, _vp]),
    "b2c_host_alloc_pinned": (C.c_int, [_vp, C.c_size_t, C.POINTER(_vp)]),
    "b2c_host_free_pinned": (C.c_int, [_vp, _vp]),
    "b2c_memcpy_h2d": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "b2c_memcpy_d2h": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "b2c_memset_dev": (C.c_int, [_vp, _vp, C.c_int, C.c_size_t]),
    "b2c_event_create": (C.c_int, [_vp, C.POINTER(_vp)]),
    "b2c_event_destroy": (None, [_vp]),
    "b2c_event_record": (C.c_int, [_vp, _vp]),
    "b2c_event_elapsed_ms": (C.c_int, [_vp, _vp, C.POINTER(C.c_float)]),
    "b2c_prof_enable": (C.c_int, [_vp, C.c_int]),
    "b2c_prof_total_ms": (C.c_int, [_vp, C.c_int, C.POINTER(C.c_float), _u32p]),
    "b2c_snappy_max_compressed_length": (C.c_size_t, [C.c_size_t]),
    "b2c_snappy_get_uncompressed_length": (C.c_int, [_vp, C.c_size_t, _u32p]),
    "b2c_snappy_read_preamble": (C.c_int, [_vp, C.c_size_t, _u32p]),
    "b2c_crc32c_mask": (C.c_uint32, [C.c_uint32]),
    "b2c_crc32c_unmask": (C.c_uint32, [C.c_uint32]),
    "b2c_layout_create": (C.c_int, [_vp, C.c_uint32, _u64p, _u32p, C.POINTER(_vp)]),
    "b2c_layout_destroy": (None, [_vp]),
    "b2c_layout_count": (C.c_uint32, [_vp]),
    "b2c_layout_payload_bytes": (C.c_uint64, [_vp]),
    "b2c_layout_arena_bytes": (C.c_uint64, [_vp]),
    "b2c_layout_max_compressed_bytes": (C.c_uint64, [_vp]),
    "b2c_crc32c_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "b2c_crc32c_noalloc_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "b2c_snappy_compress_dev": (C.c_int, [_vp, _vp, _vp, _vp, C.c_uint64, _vp]),
    "b2c_snappy_uncompressed_length_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "b2c_snappy_decompress_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b2c_snappy_validate_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp]),
    "b2c_snappy_decodable_prefix_dev": (C.c_int, [_vp, _vp, _vp, _vp]),
    "b2c_snappy_decompress_verify_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b2c_multi_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(_vp)]),
    "b2c_multi_destroy": (None, [_vp]),
    "b2c_multi_device_count": (C.c_int, [_vp]),
    "b2c_multi_last_error": (C.c_char_p, [_vp]),
    "b2c_multi_crc32c_batch": (C.c_int, [_vp, C.POINTER(IoVec), _u32p, C.c_uint32, _u32p, _u32p]),
    "b2c_multi_snappy_compress_crc_batch": (C.c_int, [_vp, C.POINTER(IoVec), _u32p, C.c_uint32, _vp, C.c_uint64, _u64p, _u32p,
                                                      _u32p, C.c_int]),
    "b2c_multi_snappy_decompress_verify_batch": (C.c_int, [_vp, C.POINTER(IoVec), _u32p, C.c_uint32, _u32p, _vp, C.c_uint64,
                                                           _u64p, _i32p]),
    "b2c_snappy_compress_batch": (C.c_int, [_vp, C.POINTER(IoVec), _u32p, C.c_uint32, _vp, C.c_uint64, _u64p]),
    "b2c_snappy_decompress_batch": (C.c_int, [_vp, C.POINTER(IoVec), _u32p, C.c_uint32, _vp, C.c_uint64, _u64p, _i32p]),
    "b2c_snappy_compress_crc_batch": (C.c_int, [_vp, C.POINTER(IoVec), _u32p, C.c_uint32, _vp, C.c_uint64, _u64p, _u32p,
                                                C.c_int]),
    "b2c_snappy_decompress_verify_batch": (C.c_int, [_vp, C.POINTER(IoVec), _u32p, C.c_uint32, _u32p, _vp, C.c_uint64,
                                                     _u64p, _i32p]),
    "b2c_snappy_decompress_iov_batch": (C.c_int, [_vp, C.POINTER(IoVec), _u32p, C.c_uint32, C.POINTER(IoVec), _u32p,
                                                  _i32p]),
    "b2c_crc32c_batch": (C.c_int, [_vp, C.POINTER(IoVec), _u32p, C.c_uint32, _u32p, _u32p]),
    "b2c_last_transfer_bytes": (None, [_vp, _u64p, _u64p]),
}


def lib_path() -> str:
    return _build.LIB


def load():
    """dlopen libb2c.so and declare every entry point.  No GPU needed for this."""
    global _lib
    if _lib is None:
        path = lib_path()
        if not os.path.exists(path):
            raise RuntimeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(the codec has no CPU fallback)")
        L = C.CDLL(path)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


class B2CError(RuntimeError):
    pass


def _ptr(a: np.ndarray, typ):
    return a.ctypes.data_as(typ)


class DevMem:
    def __init__(self, codec: "Codec", nbytes: int):
        self.codec = codec
        self.nbytes = int(nbytes)
        p = _vp()
        codec._check(codec.lib.b2c_dev_alloc(codec.ctx, self.nbytes, C.byref(p)))
        self.ptr = p.value

    def free(self):
        if self.ptr and self.codec.ctx:
            self.codec.lib.b2c_dev_free(self.codec.ctx, self.ptr)
        self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Layout:
    def __init__(self, codec: "Codec", off: np.ndarray, length: np.ndarray):
        self.codec = codec
        self.off = np.ascontiguousarray(off, dtype=np.uint64)
        self.len = np.ascontiguousarray(length, dtype=np.uint32)
        assert self.off.shape == self.len.shape
        h = _vp()
        codec._check(codec.lib.b2c_layout_create(codec.ctx, len(self.len), _ptr(self.off, _u64p),
                                                 _ptr(self.len, _u32p), C.byref(h)))
        self.h = h.value

    @property
    def n(self):
        return len(self.len)

    @property
    def payload_bytes(self):
        return self.codec.lib.b2c_layout_payload_bytes(self.h)

    @property
    def arena_bytes(self):
        return self.codec.lib.b2c_layout_arena_bytes(self.h)

    @property
    def max_compressed_bytes(self):
        return self.codec.lib.b2c_layout_max_compressed_bytes(self.h)

    def destroy(self):
        if self.h and self.codec.ctx:
            self.codec.lib.b2c_layout_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


class Event:
    def __init__(self, codec: "Codec"):
        self.codec = codec
        h = _vp()
        codec._check(codec.lib.b2c_event_create(codec.ctx, C.byref(h)))
        self.h = h.value

    def record(self):
        self.codec._check(self.codec.lib.b2c_event_record(self.codec.ctx, self.h))

    def elapsed_ms_since(self, start: "Event") -> float:
        ms = C.c_float(0)
        rc = self.codec.lib.b2c_event_elapsed_ms(start.h, self.h, C.byref(ms))
        if rc:
            raise B2CError("event_elapsed failed")
        return float(ms.value)

    def __del__(self):
        try:
            if self.h:
                self.codec.lib.b2c_event_destroy(self.h)
        except Exception:
            pass


def make_iovecs(arena: np.ndarray, off: np.ndarray, length: np.ndarray, seg_bytes: int | None = None,
                cuts: np.ndarray | None = None):
    """Describe messages inside `arena` as block lists.

    seg_bytes: split every message into blocks of that many bytes (like IOBuf's
    8160-byte payload blocks); cuts[i]: instead split message i once at cuts[i].
    Returns (segs array of IoVec, seg_begin u32[n+1]).  `arena` must stay alive."""
    n = len(length)
    base = arena.ctypes.data
    off = off.astype(np.uint64)
    ln = length.astype(np.uint64)
    if cuts is not None:
        c = np.minimum(cuts.astype(np.uint64), ln)
        two = (c > 0) & (c < ln)
        cnt = np.where(two, 2, 1).astype(np.uint64)
    elif seg_bytes:
        cnt = np.maximum(1, (ln + np.uint64(seg_bytes - 1)) // np.uint64(seg_bytes))
    else:
        cnt = np.ones(n, dtype=np.uint64)
    seg_begin = np.zeros(n + 1, dtype=np.uint32)
    seg_begin[1:] = np.cumsum(cnt).astype(np.uint32)
    total = int(seg_begin[-1])
    raw = np.zeros((total, 2), dtype=np.uint64)  # (iov_base, iov_len) pairs == struct b2c_iovec on LP64
    first = seg_begin[:-1].astype(np.int64)
    if cuts is not None:
        raw[first, 0] = base + off
        raw[first, 1] = np.where(two, c, ln)
        idx = first[two] + 1
        raw[idx, 0] = base + off[two] + c[two]
        raw[idx, 1] = ln[two] - c[two]
    elif seg_bytes:
        msg = np.repeat(np.arange(n), cnt.astype(np.int64))
        k = (np.arange(total) - first[msg]).astype(np.uint64)
        start = k * np.uint64(seg_bytes)
        raw[:, 0] = base + off[msg] + start
        raw[:, 1] = np.minimum(np.uint64(seg_bytes), ln[msg] - np.minimum(start, ln[msg]))
    else:
        raw[:, 0] = base + off
        raw[:, 1] = ln
    return raw, seg_begin


class Codec:
    """One b2c_ctx bound to one GPU."""

    def __init__(self, device: int = 0):
        self.lib = load()
        h = _vp()
        rc = self.lib.b2c_ctx_create(device, C.byref(h))
        if rc != OK:
            raise B2CError(f"b2c_ctx_create(device={device}) failed with {rc}"
                           + (" (no usable CUDA device; this codec has no CPU fallback)" if rc == E_NODEVICE else ""))
        self.ctx = h.value
        self.device = device

    def close(self):
        if self.ctx:
            self.lib.b2c_ctx_destroy(self.ctx)
        self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int):
        if rc != OK:
            msg = self.lib.b2c_last_error(self.ctx).decode(errors="replace") if self.ctx else ""
            raise B2CError(f"libb2c error {rc}: {msg}")

    # ---- plumbing ----------------------------------------------------------
    OPT_INPUT_PINNED = 1
    OPT_HOST_THREADS = 2

    def set_option(self, option: int, value: int):
        self._check(self.lib.b2c_ctx_set_option(self.ctx, option, value))

    def set_tunable(self, name: str, value: int):
        self._check(self.lib.b2c_ctx_set_tunable(self.ctx, name.encode(), value))

    def sync(self):
        self._check(self.lib.b2c_ctx_sync(self.ctx))

    @property
    def launch_count(self) -> int:
        return self.lib.b2c_ctx_launch_count(self.ctx)

    def alloc(self, nbytes: int) -> DevMem:
        return DevMem(self, nbytes)

    def pinned(self, nbytes: int) -> np.ndarray:
        p = _vp()
        self._check(self.lib.b2c_host_alloc_pinned(self.ctx, max(nbytes, 1), C.byref(p)))
        buf = (C.c_uint8 * max(nbytes, 1)).from_address(p.value)
        a = np.frombuffer(buf, dtype=np.uint8)
        return a

    def h2d(self, dev: DevMem, host: np.ndarray, dev_offset: int = 0):
        host = np.ascontiguousarray(host)
        assert dev_offset + host.nbytes <= dev.nbytes
        self._check(self.lib.b2c_memcpy_h2d(self.ctx, dev.ptr + dev_offset, host.ctypes.data, host.nbytes))
        self.sync()  # pageable source: keep it simple and safe

    def d2h(self, dev: DevMem, nbytes: int, dtype=np.uint8, dev_offset: int = 0) -> np.ndarray:
        out = np.empty(nbytes // np.dtype(dtype).itemsize, dtype=dtype)
        if nbytes:
            self._check(self.lib.b2c_memcpy_d2h(self.ctx, out.ctypes.data, dev.ptr + dev_offset, nbytes))
            self.sync()
        return out

    def to_device(self, host: np.ndarray, pad: int = 64) -> DevMem:
        host = np.ascontiguousarray(host)
        d = self.alloc(host.nbytes + pad)
        if host.nbytes:
            self.h2d(d, host)
        return d

    def layout(self, off, length) -> Layout:
        return Layout(self, off, length)

    def event(self) -> Event:
        return Event(self)

    # ---- device-resident ops ---------------------------------------------
    def crc32c_dev(self, layout: Layout, d_data: DevMem, d_crc: DevMem, d_init: DevMem | None = None):
        self._check(self.lib.b2c_crc32c_dev(self.ctx, layout.h, d_data.ptr, d_init.ptr if d_init else None, d_crc.ptr))

    def crc32c_noalloc_dev(self, layout: Layout, d_data: DevMem, d_crc: DevMem, d_init: DevMem | None = None):
        self._check(self.lib.b2c_crc32c_noalloc_dev(self.ctx, layout.h, d_data.ptr, d_init.ptr if d_init else None,
                                                    d_crc.ptr))

    def compress_dev(self, layout: Layout, d_in: DevMem, d_out: DevMem, d_out_off: DevMem):
        self._check(self.lib.b2c_snappy_compress_dev(self.ctx, layout.h, d_in.ptr, d_out.ptr, d_out.nbytes,
                                                     d_out_off.ptr))

    def ulen_dev(self, layout: Layout, d_in: DevMem, d_ulen: DevMem, d_status: DevMem | None = None):
        self._check(self.lib.b2c_snappy_uncompressed_length_dev(self.ctx, layout.h, d_in.ptr, d_ulen.ptr,
                                                                d_status.ptr if d_status else None))

    def decompress_dev(self, in_layout: Layout, d_in: DevMem, out_layout: Layout, d_out: DevMem, d_status: DevMem,
                       d_ulen: DevMem | None = None):
        self._check(self.lib.b2c_snappy_decompress_dev(self.ctx, in_layout.h, d_in.ptr, out_layout.h, d_out.ptr,
                                                       d_ulen.ptr if d_ulen else None, d_status.ptr))

    def validate_dev(self, layout: Layout, d_in: DevMem, d_status: DevMem, d_ulen: DevMem | None = None):
        self._check(self.lib.b2c_snappy_validate_dev(self.ctx, layout.h, d_in.ptr, d_ulen.ptr if d_ulen else None,
                                                     d_status.ptr))

    def decompress_verify_dev(self, in_layout: Layout, d_in: DevMem, out_layout: Layout, d_out: DevMem,
                              d_masked_crc: DevMem, d_status: DevMem, d_ulen: DevMem | None = None):
        self._check(self.lib.b2c_snappy_decompress_verify_dev(self.ctx, in_layout.h, d_in.ptr, out_layout.h, d_out.ptr,
                                                              d_masked_crc.ptr, d_ulen.ptr if d_ulen else None,
                                                              d_status.ptr))

    def decodable_prefix(self, arena: np.ndarray, off, length) -> np.ndarray:
        """[n][5] u32: returned size, sunk bytes, consumed, partial literal bytes, partial literal source"""
        lay = self.layout(off, length)
        d = self.to_device(arena)
        d_out = self.alloc(20 * max(lay.n, 1))
        self._check(self.lib.b2c_snappy_decodable_prefix_dev(self.ctx, lay.h, d.ptr, d_out.ptr))
        return self.d2h(d_out, 20 * lay.n, np.uint32).reshape(-1, 5)

    def validate(self, arena: np.ndarray, off, length):
        lin = self.layout(off, length)
        d = self.to_device(arena)
        ds = self.alloc(4 * max(lin.n, 1))
        du = self.alloc(4 * max(lin.n, 1))
        self.validate_dev(lin, d, ds, du)
        return self.d2h(ds, 4 * lin.n, np.int32), self.d2h(du, 4 * lin.n, np.uint32)

    def decompress_verify(self, arena: np.ndarray, off, length, out_off, out_cap, masked_crc):
        lin = self.layout(off, length)
        lout = self.layout(out_off, out_cap)
        d = self.to_device(arena)
        do = self.alloc(int(lout.arena_bytes) + 64)
        ds = self.alloc(4 * max(lin.n, 1))
        dm = self.to_device(np.ascontiguousarray(masked_crc, dtype=np.uint32))
        self.decompress_verify_dev(lin, d, lout, do, dm, ds)
        return self.d2h(do, int(lout.arena_bytes)), self.d2h(ds, 4 * lin.n, np.int32)

    # ---- convenience: numpy in, numpy out, through the device-resident ABI ----
    def crc32c(self, arena: np.ndarray, off, length, init=None) -> np.ndarray:
        lay = self.layout(off, length)
        d = self.to_device(arena)
        dc = self.alloc(4 * max(lay.n, 1))
        di = self.to_device(np.ascontiguousarray(init, dtype=np.uint32)) if init is not None else None
        self.crc32c_dev(lay, d, dc, di)
        return self.d2h(dc, 4 * lay.n, np.uint32)

    def compress(self, arena: np.ndarray, off, length):
        """Returns (packed bytes u8, out_off u64[n+1])."""
        lay = self.layout(off, length)
        d = self.to_device(arena)
        do = self.alloc(lay.max_compressed_bytes + 64)
        doff = self.alloc(8 * (lay.n + 1))
        self.compress_dev(lay, d, do, doff)
        out_off = self.d2h(doff, 8 * (lay.n + 1), np.uint64)
        total = int(out_off[-1]) if lay.n else 0
        return self.d2h(do, total), out_off

    def decompress(self, arena: np.ndarray, off, length, out_off, out_cap):
        """Returns (out arena u8, ulen u32[n], status i32[n])."""
        lin = self.layout(off, length)
        lout = self.layout(out_off, out_cap)
        d = self.to_device(arena)
        total = int(lout.arena_bytes)
        do = self.alloc(total + 64)
        ds = self.alloc(4 * max(lin.n, 1))
        du = self.alloc(4 * max(lin.n, 1))
        self.decompress_dev(lin, d, lout, do, ds, du)
        return self.d2h(do, total), self.d2h(du, 4 * lin.n, np.uint32), self.d2h(ds, 4 * lin.n, np.int32)

    def uncompressed_lengths(self, arena: np.ndarray, off, length):
        lin = self.layout(off, length)
        d = self.to_device(arena)
        du = self.alloc(4 * max(lin.n, 1))
        ds = self.alloc(4 * max(lin.n, 1))
        self.ulen_dev(lin, d, du, ds)
        return self.d2h(du, 4 * lin.n, np.uint32), self.d2h(ds, 4 * lin.n, np.int32)

    # ---- host (end-to-end) ops: iovec block lists in, flat buffers out ----------
    def _segs(self, raw: np.ndarray):
        return raw.ctypes.data_as(C.POINTER(IoVec))

    def crc32c_batch(self, raw_segs: np.ndarray, seg_begin: np.ndarray, init=None) -> np.ndarray:
        n = len(seg_begin) - 1
        crc = np.empty(n, dtype=np.uint32)
        ini = np.ascontiguousarray(init, dtype=np.uint32) if init is not None else None
        self._check(self.lib.b2c_crc32c_batch(self.ctx, self._segs(raw_segs), _ptr(seg_begin, _u32p), n,
                                              _ptr(ini, _u32p) if ini is not None else None, _ptr(crc, _u32p)))
        return crc

    def compress_batch(self, raw_segs: np.ndarray, seg_begin: np.ndarray, out: np.ndarray):
        n = len(seg_begin) - 1
        out_off = np.empty(n + 1, dtype=np.uint64)
        self._check(self.lib.b2c_snappy_compress_batch(self.ctx, self._segs(raw_segs), _ptr(seg_begin, _u32p), n,
                                                       out.ctypes.data, out.nbytes, _ptr(out_off, _u64p)))
        return out_off

    def decompress_batch(self, raw_segs: np.ndarray, seg_begin: np.ndarray, out: np.ndarray):
        n = len(seg_begin) - 1
        out_off = np.empty(n + 1, dtype=np.uint64)
        status = np.empty(n, dtype=np.int32)
        self._check(self.lib.b2c_snappy_decompress_batch(self.ctx, self._segs(raw_segs), _ptr(seg_begin, _u32p), n,
                                                         out.ctypes.data, out.nbytes, _ptr(out_off, _u64p),
                                                         _ptr(status, _i32p)))
        return out_off, status

    def compress_crc_batch(self, raw_segs: np.ndarray, seg_begin: np.ndarray, out: np.ndarray, masked: bool = True):
        """Sender side in one pass: (out_off, crc) -- the compressed streams and each body's (masked) CRC32C."""
        n = len(seg_begin) - 1
        out_off = np.empty(n + 1, dtype=np.uint64)
        crc = np.empty(n, dtype=np.uint32)
        self._check(self.lib.b2c_snappy_compress_crc_batch(self.ctx, self._segs(raw_segs), _ptr(seg_begin, _u32p), n,
                                                           out.ctypes.data, out.nbytes, _ptr(out_off, _u64p),
                                                           _ptr(crc, _u32p), 1 if masked else 0))
        return out_off, crc

    def decompress_verify_batch(self, raw_segs: np.ndarray, seg_begin: np.ndarray, masked_crc: np.ndarray,
                                out: np.ndarray):
        """Receiver side in one pass: decode and check every body against its masked CRC32C."""
        n = len(seg_begin) - 1
        out_off = np.empty(n + 1, dtype=np.uint64)
        status = np.empty(n, dtype=np.int32)
        mc = np.ascontiguousarray(masked_crc, dtype=np.uint32)
        self._check(self.lib.b2c_snappy_decompress_verify_batch(self.ctx, self._segs(raw_segs), _ptr(seg_begin, _u32p), n,
                                                                _ptr(mc, _u32p), out.ctypes.data, out.nbytes,
                                                                _ptr(out_off, _u64p), _ptr(status, _i32p)))
        return out_off, status

    def decompress_iov_batch(self, raw_segs: np.ndarray, seg_begin: np.ndarray, out_segs: np.ndarray,
                             out_seg_begin: np.ndarray) -> np.ndarray:
        """RawUncompressToIOVec for a batch: message i is scattered over out_segs[out_seg_begin[i]:out_seg_begin[i+1]]."""
        n = len(seg_begin) - 1
        status = np.empty(n, dtype=np.int32)
        self._check(self.lib.b2c_snappy_decompress_iov_batch(self.ctx, self._segs(raw_segs), _ptr(seg_begin, _u32p), n,
                                                             self._segs(out_segs), _ptr(out_seg_begin, _u32p),
                                                             _ptr(status, _i32p)))
        return status

    def prof_enable(self, on: bool = True):
        self._check(self.lib.b2c_prof_enable(self.ctx, 1 if on else 0))

    def prof_total_ms(self, which: int):
        ms, cnt = C.c_float(0), C.c_uint32(0)
        self._check(self.lib.b2c_prof_total_ms(self.ctx, which, C.byref(ms), C.byref(cnt)))
        return float(ms.value), int(cnt.value)

    def last_transfer_bytes(self):
        a, b = C.c_uint64(0), C.c_uint64(0)
        self.lib.b2c_last_transfer_bytes(self.ctx, C.byref(a), C.byref(b))
        return a.value, b.value


class MultiCodec:
    """One b2c_multi: a batch split over several GPUs by bytes (b2c_multi_*), results in message order."""

    def __init__(self, devices):
        self.lib = load()
        devs = (C.c_int * len(devices))(*devices)
        h = _vp()
        rc = self.lib.b2c_multi_create(devs, len(devices), C.byref(h))
        if rc != OK:
            raise B2CError(f"b2c_multi_create({list(devices)}) failed with {rc}")
        self.h = h.value
        self.devices = list(devices)

    def close(self):
        if self.h:
            self.lib.b2c_multi_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != OK:
            raise B2CError(f"libb2c multi error {rc}: {self.lib.b2c_multi_last_error(self.h).decode(errors='replace')}")

    def _segs(self, raw):
        return raw.ctypes.data_as(C.POINTER(IoVec))

    def crc32c_batch(self, raw_segs, seg_begin, init=None):
        n = len(seg_begin) - 1
        crc = np.empty(n, dtype=np.uint32)
        ini = np.ascontiguousarray(init, dtype=np.uint32) if init is not None else None
        self._check(self.lib.b2c_multi_crc32c_batch(self.h, self._segs(raw_segs), _ptr(seg_begin, _u32p), n,
                                                    _ptr(ini, _u32p) if ini is not None else None, _ptr(crc, _u32p)))
        return crc

    def compress_crc_batch(self, raw_segs, seg_begin, out, masked=True, want_crc=True):
        """(out_off[n], out_len[n], crc[n] or None): stream i = out[out_off[i] : out_off[i] + out_len[i]]"""
        n = len(seg_begin) - 1
        out_off = np.empty(n, dtype=np.uint64)
        out_len = np.empty(n, dtype=np.uint32)
        crc = np.empty(n, dtype=np.uint32) if want_crc else None
        self._check(self.lib.b2c_multi_snappy_compress_crc_batch(
            self.h, self._segs(raw_segs), _ptr(seg_begin, _u32p), n, out.ctypes.data, out.nbytes, _ptr(out_off, _u64p),
            _ptr(out_len, _u32p), _ptr(crc, _u32p) if want_crc else None, 1 if masked else 0))
        return out_off, out_len, crc

    def decompress_verify_batch(self, raw_segs, seg_begin, masked_crc, out):
        """(out_off[n+1], status[n]); masked_crc None = plain decompress"""
        n = len(seg_begin) - 1
        out_off = np.empty(n + 1, dtype=np.uint64)
        status = np.empty(n, dtype=np.int32)
        mc = np.ascontiguousarray(masked_crc, dtype=np.uint32) if masked_crc is not None else None
        self._check(self.lib.b2c_multi_snappy_decompress_verify_batch(
            self.h, self._segs(raw_segs), _ptr(seg_begin, _u32p), n, _ptr(mc, _u32p) if mc is not None else None,
            out.ctypes.data, out.nbytes, _ptr(out_off, _u64p), _ptr(status, _i32p)))
        return out_off, status
