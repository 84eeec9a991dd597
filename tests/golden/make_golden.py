"""Generate tests/golden/codec_golden.json from the REFERENCE's own code.

Run in the build container only (needs /root/reference to build oracle/_ref):
    python tests/golden/make_golden.py
The reference has no compressed-byte goldens of its own (its snappy tests are
round-trip only, test/brpc_snappy_compress_unittest.cpp:80-255), so these
vectors are outputs of the reference library itself on deterministic inputs.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402


def main():
    ref = O.reference()
    assert ref is not None, "reference library not built"
    vec = []
    sizes = [0, 1, 3, 4, 14, 15, 16, 17, 31, 32, 33, 59, 60, 61, 62, 100, 128, 255, 256, 257, 782, 1000,
             1023, 1024, 1025, 2047, 2048, 4095, 4096, 4097, 8160, 8161, 12435, 16384, 32768, 65535,
             65536, 65537, 70000, 131072, 200000, 1048576]
    for kind in ("az09", "rand62", "words"):
        for n in sizes:
            for index in ((0, 7) if n <= 4097 else (0,)):
                d = O.corpus(kind, n, index)
                z = ref.compress(d)
                e = {"corpus": kind, "n": n, "index": index, "crc": ref.crc32c(d), "clen": len(z),
                     "ccrc": ref.crc32c(z), "head": z[:16].hex()}
                if n <= 256:
                    e["z"] = z.hex()
                vec.append(e)
    # special inputs
    special = {
        "this is a test": b"this is a test",
        "zeros65536": bytes(65536),
        "ff4096": b"\xff" * 4096,
        "abab": b"ab" * 3000,
        "a26x782": (b"abcdefghijklmnopqrstuvwxyz" * 31)[:782],
        "run_then_random": bytes(300) + O.corpus("rand62", 700, 3).tobytes() + bytes(300),
    }
    spec = []
    for name, b in special.items():
        z = ref.compress(b)
        spec.append({"name": name, "in": b.hex() if len(b) <= 64 else None, "n": len(b), "crc": ref.crc32c(b),
                     "clen": len(z), "ccrc": ref.crc32c(z), "z": z.hex() if len(z) <= 512 else None})
    # malformed / foreign streams, answers from the reference
    streams = ["", "00", "0510" + b"hello".hex(), "0510" + b"hel".hex(), "0510" + b"hello".hex() + "58",
               "0610" + b"hello".hex(), "080c616263640100", "080c616263640105", "080c616263640104",
               "0800610d01", "080c616263640f04000000", "060c616263640104", "ffffffffff01", "8080808010",
               "05f004" + b"hello".hex(), "05f4040068656c6c6f", "05f804000068656c6c6f", "05fc0400000068656c6c6f",
               "05fcffffffff", "0100", "01", "0200", "0a0c61626364" + "0e0400", "0a0c616263640e04",
               "0a0c61626364fe0400", "400c61626364" + "ee0400", "03fcffffffff616263",
               "8080808000", "808080800f", "80808080ff"]
    mal = []
    for h in streams:
        b = bytes.fromhex(h)
        a = np.frombuffer(b, dtype=np.uint8)
        ok, out = ref.uncompress(a) if len(b) else (False, b"")
        mal.append({"stream": h, "uncompress": bool(ok), "out": out.hex() if ok else None,
                    "is_valid": ref.is_valid(a) if len(b) else False,
                    "string_api": bool(ref.lib.b2r_snappy_uncompress_string(
                        a.ctypes.data_as(O._u8p), a.size)) if len(b) else False,
                    "ulen_lenient": ref.ulen_lenient(a) if len(b) else None,
                    "ulen_strict": ref.ulen_strict(a) if len(b) else None})
    # crc vectors: the reference's own known answers (test/crc32c_unittest.cc:18-62) + extend chains
    crc = {
        "rfc3720": [
            {"in": (bytes(32)).hex(), "crc": 0x8A9136AA},
            {"in": (b"\xff" * 32).hex(), "crc": 0x62A8AB43},
            {"in": bytes(range(32)).hex(), "crc": 0x46DD794E},
            {"in": bytes(range(31, -1, -1)).hex(), "crc": 0x113FDB5C},
            {"in": ("01c00000" + "00" * 12 + "14000000" + "00000400" + "00000014" + "00000018" + "28000000"
                    + "00" * 4 + "02000000" + "00" * 4), "crc": 0xD9963A56},
        ],
        "value_123456789": ref.crc32c(b"123456789"),
        "hello_world": ref.crc32c(b"hello world"),
        "hello_": ref.crc32c(b"hello "),
        "extend": [],
        "mask_foo": ref.mask(ref.crc32c(b"foo")),
    }
    for e in crc["rfc3720"]:
        assert ref.crc32c(bytes.fromhex(e["in"])) == e["crc"], e
    rng = np.random.default_rng(20261106)
    for _ in range(24):
        n = int(rng.integers(0, 5000))
        cut = int(rng.integers(0, n + 1))
        init = int(rng.integers(0, 2**32))
        d = rng.integers(0, 256, n, dtype=np.uint8)
        crc["extend"].append({"seed_n": n, "cut": cut, "init": init, "data_crc": ref.crc32c(d),
                              "extend_init": ref.crc32c_extend(init, d),
                              "data": d.tobytes().hex() if n <= 64 else None, "rng_index": _})
    out = {"vectors": vec, "special": spec, "malformed": mal, "crc": crc,
           "generator": "tests/golden/make_golden.py over oracle/_ref/libb2ref.so (bRPC 1.4.0 sources)"}
    path = os.path.join(ROOT, "tests", "golden", "codec_golden.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=0, separators=(",", ":"))
    print("wrote", path, os.path.getsize(path), "bytes;", len(vec), "vectors")


if __name__ == "__main__":
    main()
