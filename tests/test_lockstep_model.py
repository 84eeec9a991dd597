"""The lock-step lane code of the small-body kernels (csrc/snappy_lockstep.cuh) compiled for the HOST and run
trip by trip as a warp would, against the oracle: byte-exact CompressFragment for bodies <= 1 KiB.
Runs without a GPU (the same source is what the CUDA kernels execute)."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

from oracle import oracle as O
from tests.util import pack_tight

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "lockstep_model.cpp")
LIB = os.path.join(ROOT, "tests", "cpp", "liblockstep_model.so")
CSRC = os.path.join(ROOT, "brpc-github-actions_b200", "csrc")


@pytest.fixture(scope="module")
def model():
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++")
    subprocess.run([cxx, "-O2", "-std=c++17", "-shared", "-fPIC", "-I", CSRC, "-o", LIB, SRC], check=True)
    lib = C.CDLL(LIB)
    u8p, u32p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)
    lib.ls_model_decompress.restype = C.c_int
    lib.ls_model_decompress.argtypes = [u8p, u64p, u32p, C.c_uint32, u32p, u8p, C.c_uint32, u32p, C.POINTER(C.c_int32), u64p]
    lib.ls_model_compress.restype = C.c_int
    lib.ls_model_compress.argtypes = [u8p, u64p, u32p, C.c_uint32, u8p, C.c_uint32, u32p, u64p, u64p]

    def compress(bodies):
        """list of <= 32 bodies -> (list of fragment byte strings, trips, lane-trips probing / extending)"""
        arena, off, ln = pack_tight(bodies)
        stride = 1280
        out = np.zeros(32 * stride, dtype=np.uint8)
        clen = np.zeros(32, dtype=np.uint32)
        trips = C.c_uint64(0)
        modes = np.zeros(2, dtype=np.uint64)
        rc = lib.ls_model_compress(arena.ctypes.data_as(u8p), off.ctypes.data_as(u64p), ln.ctypes.data_as(u32p),
                                   len(bodies), out.ctypes.data_as(u8p), stride, clen.ctypes.data_as(u32p),
                                   C.byref(trips), modes.ctypes.data_as(u64p))
        assert rc == 0
        return [out[i * stride:i * stride + int(clen[i])].tobytes() for i in range(len(bodies))], trips.value, modes

    def decompress(streams, caps):
        """list of <= 32 streams -> (list of (status, bytes), trips)"""
        arena, off, ln = pack_tight(streams)
        caps = np.ascontiguousarray(caps, dtype=np.uint32)
        stride = 1040
        out = np.zeros(32 * stride, dtype=np.uint8)
        ulen = np.zeros(32, dtype=np.uint32)
        status = np.zeros(32, dtype=np.int32)
        trips = C.c_uint64(0)
        rc = lib.ls_model_decompress(arena.ctypes.data_as(u8p), off.ctypes.data_as(u64p), ln.ctypes.data_as(u32p),
                                     len(streams), caps.ctypes.data_as(u32p), out.ctypes.data_as(u8p), stride,
                                     ulen.ctypes.data_as(u32p), status.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(trips))
        assert rc == 0
        return [(int(status[i]), out[i * stride:i * stride + int(ulen[i])].tobytes() if status[i] == 0 else b"")
                for i in range(len(streams))], trips.value

    compress.decompress = decompress
    return compress


def oracle_fragment(body: bytes) -> bytes:
    lib = O.port().lib
    a = np.frombuffer(body, dtype=np.uint8) if body else np.zeros(1, dtype=np.uint8)
    out = np.zeros(32 + len(body) + len(body) // 6 + 16, dtype=np.uint8)
    u8p = C.POINTER(C.c_uint8)
    n = lib.b2o_snappy_compress_fragment(a.ctypes.data_as(u8p), len(body), out.ctypes.data_as(u8p))
    return out[:n].tobytes()


def check(model, bodies):
    got, trips, modes = model(bodies)
    for i, b in enumerate(bodies):
        want = oracle_fragment(b)
        assert got[i] == want, f"body {i} (n={len(b)}): got {got[i][:24].hex()}.. want {want[:24].hex()}.."
    return trips, modes


def test_corpora_all_sizes(model):
    for kind in ("words", "az09", "rand62"):
        sizes = list(range(0, 40)) + [63, 64, 65, 127, 128, 255, 256, 257, 511, 512, 513, 777, 1000, 1023, 1024]
        for base in range(0, len(sizes), 32):
            chunk = sizes[base:base + 32]
            check(model, [O.corpus(kind, n, index=100 + base + j).tobytes() for j, n in enumerate(chunk)])


def test_c2_shape_trip_count(model):
    bodies = [O.corpus("words", 1024, index=i).tobytes() for i in range(32)]
    trips, modes = check(model, bodies)
    print("trips", trips, "lane-trips per body (probe, ext)", modes / 32)
    # the warp's main loop is as long as its slowest lane
    assert trips < 450, trips


def test_runs_long_matches_and_literals(model):
    rng = np.random.default_rng(5)
    bodies = [bytes(1024), b"a" * 1000, b"ab" * 500, b"abc" * 300, bytes(range(256)) * 4,
              bytes(rng.integers(0, 256, 1024, dtype=np.uint8)),
              bytes(rng.integers(0, 256, 500, dtype=np.uint8)) + b"x" * 200 + bytes(rng.integers(0, 256, 300, dtype=np.uint8)),
              b"0123456789" * 7 + bytes(rng.integers(0, 256, 70, dtype=np.uint8)) + b"0123456789" * 7,
              ]
    for ln in (64, 65, 66, 67, 68, 69, 70, 127, 128, 129, 131, 132, 133):  # EmitCopy's 64/60-byte pieces
        bodies.append(b"q" * 5 + b"z" * ln + bytes(rng.integers(0, 256, 40, dtype=np.uint8)))
    for base in range(0, len(bodies), 32):
        check(model, bodies[base:base + 32])


def test_random_mixtures(model):
    rng = np.random.default_rng(11)
    for trial in range(40):
        bodies = []
        for _ in range(32):
            n = int(rng.integers(0, 1025))
            alpha = int(rng.choice([2, 4, 16, 64, 256]))
            b = rng.integers(0, alpha, n, dtype=np.uint8)
            if n > 64 and rng.random() < 0.5:  # paste earlier pieces to create matches of many lengths
                for _ in range(int(rng.integers(1, 12))):
                    a = int(rng.integers(0, n - 8))
                    l = int(rng.integers(4, min(200, n - a)))
                    d = int(rng.integers(0, n - l))
                    b[d:d + l] = b[a:a + l].copy()
            bodies.append(b.tobytes())
        check(model, bodies)


# ---------------------------------------------------------------- the lock-step decoder
def check_decode(model, streams, caps=None):
    cpu = O.port()
    if caps is None:
        caps = [min(1024, cpu.ulen_lenient(s) or 0) for s in streams]
    got, trips = model.decompress(streams, caps)
    for i, s in enumerate(streams):
        ulen = cpu.ulen_lenient(s)
        if ulen is None:
            want = (1, b"")
        elif ulen > caps[i]:
            want = (2, b"")
        else:
            ok, body = cpu.uncompress(s, cap=ulen)
            want = (0, body) if ok else (1, b"")
        assert got[i] == want, f"stream {i} ({len(s)} B): got status {got[i][0]} want {want[0]}"
    return trips


def test_decoder_round_trips_all_corpora(model):
    cpu = O.port()
    for kind in ("words", "az09", "rand62"):
        sizes = list(range(0, 40)) + [63, 64, 65, 127, 128, 255, 256, 257, 511, 512, 513, 777, 1000, 1023, 1024]
        for base in range(0, len(sizes), 32):
            bodies = [O.corpus(kind, n, index=300 + base + j).tobytes() for j, n in enumerate(sizes[base:base + 32])]
            check_decode(model, [cpu.compress(b) for b in bodies])
    trips = check_decode(model, [cpu.compress(O.corpus("words", 1024, index=i).tobytes()) for i in range(32)])
    print("decode trips for 32 x 1 KiB words:", trips)
    assert trips < 300


def test_decoder_runs_and_foreign_streams(model):
    cpu = O.port()
    rng = np.random.default_rng(8)
    bodies = [bytes(1024), b"a" * 1000, b"ab" * 500, b"abc" * 300, b"abcde" * 200, b"abcdef" * 170, b"abcdefg" * 140,
              bytes(range(256)) * 4, bytes(rng.integers(0, 256, 1024, dtype=np.uint8))]
    streams = [cpu.compress(b) for b in bodies]

    def lit(b):
        n1 = len(b) - 1
        if n1 < 60:
            return bytes([n1 << 2]) + b
        return bytes([62 << 2]) + n1.to_bytes(3, "little") + b  # a 3-byte length, as a foreign encoder may write

    def copy4(offset, length):
        return bytes([3 | ((length - 1) << 2)]) + offset.to_bytes(4, "little")

    def copy2(offset, length):
        return bytes([2 | ((length - 1) << 2)]) + offset.to_bytes(2, "little")

    for o in range(1, 12):  # overlapping copies of every small offset, lengths straddling the 8-byte chunk
        for ln in (4, 7, 8, 9, 15, 16, 17, 33, 64):
            body_len = o + ln
            streams.append(bytes([body_len]) + lit(bytes(range(65, 65 + o))) + copy2(o, ln))
    streams.append(bytes([100]) + lit(b"x" * 50) + copy4(50, 50))
    streams.append(bytes([0x80, 0x02]) + lit(bytes(rng.integers(0, 256, 200, dtype=np.uint8))) + copy4(200, 56))
    for base in range(0, len(streams), 32):
        check_decode(model, streams[base:base + 32])


def test_decoder_rejects_what_the_reference_rejects(model):
    cpu = O.port()
    rng = np.random.default_rng(21)
    streams = []
    for t in range(640):
        body = O.corpus(("words", "az09", "rand62")[t % 3], int(rng.integers(1, 1025)), index=t).tobytes()
        z = bytearray(cpu.compress(body))
        how = t % 5
        if how == 0 and len(z) > 3:
            z = z[: int(rng.integers(1, len(z)))]  # truncated
        elif how == 1:
            z[int(rng.integers(0, len(z)))] ^= 1 << int(rng.integers(0, 8))  # a flipped bit
        elif how == 2:
            z += bytes(rng.integers(0, 256, int(rng.integers(1, 6)), dtype=np.uint8))  # trailing bytes
        elif how == 3:
            z[0] = (z[0] + int(rng.integers(1, 5))) & 0x7f  # the preamble lies
        streams.append(bytes(z))
    for base in range(0, len(streams), 32):
        chunk = streams[base:base + 32]
        caps = [1024] * len(chunk)
        check_decode(model, chunk, caps)
