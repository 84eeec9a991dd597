"""CPU tests of the boundary: libb2c.so loads, exports every symbol the header
declares, refuses to run without a GPU, and its pure-host helpers match the oracle."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import brpc_b200
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cuda():
    import torch
    return torch.cuda.is_available()


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b2c_codec.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(b2c_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    lib = brpc_b200.load()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/b2c_codec.h but not exported"
    assert declared == set(brpc_b200.SYMBOLS), declared ^ set(brpc_b200.SYMBOLS)


def test_no_cpu_fallback_without_device():
    if _cuda():
        pytest.skip("a GPU is present")
    with pytest.raises(brpc_b200.B2CError):
        brpc_b200.Codec(0)


def test_host_helpers_match_oracle(golden):
    lib = brpc_b200.load()
    p = O.port()
    for n in (0, 1, 5, 6, 1024, 4096, 65536, 1 << 20):
        assert lib.b2c_snappy_max_compressed_length(n) == p.max_compressed_length(n)
    for e in golden["malformed"]:
        b = bytes.fromhex(e["stream"])
        if not b:
            continue
        buf = (C.c_uint8 * len(b)).from_buffer_copy(b)
        v = C.c_uint32(0)
        ok = lib.b2c_snappy_get_uncompressed_length(buf, len(b), C.byref(v))
        assert (v.value if ok else None) == e["ulen_strict"], e
        k = lib.b2c_snappy_read_preamble(buf, len(b), C.byref(v))
        assert (v.value if k else None) == e["ulen_lenient"], e
    c = p.crc32c(b"foo")
    assert lib.b2c_crc32c_mask(c) == p.mask(c) == golden["crc"]["mask_foo"]
    assert lib.b2c_crc32c_unmask(lib.b2c_crc32c_mask(c)) == c


def test_make_iovecs_describes_the_same_bytes():
    rng = np.random.default_rng(0)
    ln = np.array([0, 1, 100, 8160, 8161, 20000], dtype=np.uint32)
    arena, off, ln = O.corpus_batch("words", ln)
    for kw in ({}, {"seg_bytes": 8160}, {"seg_bytes": 7}, {"cuts": rng.integers(0, 9000, len(ln))}):
        raw, sb = brpc_b200.make_iovecs(arena, off, ln, **kw)
        assert sb[0] == 0 and sb[-1] == len(raw)
        for i in range(len(ln)):
            parts = []
            for s in range(sb[i], sb[i + 1]):
                a = int(raw[s, 0]) - arena.ctypes.data
                parts.append(arena[a:a + int(raw[s, 1])].tobytes())
            assert b"".join(parts) == arena[int(off[i]):int(off[i]) + int(ln[i])].tobytes()
