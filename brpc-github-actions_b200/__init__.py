"""B200 batch payload codec for bRPC: snappy + CRC32C over many message bodies.

The product is the C-ABI library libb2c.so (csrc/, include/b2c_codec.h) and the
C++ host shim in host/.  This Python package is only the test/bench harness
around that ABI.
"""
from .nvcc_build import build  # noqa: F401
from .capi import (B2CError, Codec, MultiCodec, CORRUPT, CRC_MISMATCH, OK, TOO_SMALL, SYMBOLS, lib_path, load,  # noqa: F401
                   make_iovecs)
from . import sharding  # noqa: F401,E402
