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
