"""CPU tests: pin the oracle (oracle/*.c) against the reference's golden vectors,
the committed fixtures (tests/golden, generated from the reference's own code) and,
where oracle/_ref exists, the reference library itself."""
import numpy as np
import pytest

from oracle import oracle as O

CODECS = [("port", O.port)]
if O.reference() is not None:
    CODECS.append(("reference", O.reference))


@pytest.fixture(params=CODECS, ids=[c[0] for c in CODECS])
def codec(request):
    return request.param[1]()


def test_crc_known_answers(codec, golden):
    # reference: test/crc32c_unittest.cc:18-53 (RFC 3720 B.4)
    for e in golden["crc"]["rfc3720"]:
        assert codec.crc32c(bytes.fromhex(e["in"])) == e["crc"]
    assert codec.crc32c(b"123456789") == 0xE3069283 == golden["crc"]["value_123456789"]
    # Values / Extend / Mask: crc32c_unittest.cc:55-70
    assert codec.crc32c(b"a") != codec.crc32c(b"foo")
    assert codec.crc32c(b"hello world") == codec.crc32c_extend(codec.crc32c(b"hello "), b"world")
    c = codec.crc32c(b"foo")
    assert c != codec.mask(c) and c != codec.mask(codec.mask(c))
    assert c == codec.unmask(codec.mask(c))
    assert c == codec.unmask(codec.unmask(codec.mask(codec.mask(c))))
    assert codec.mask(c) == golden["crc"]["mask_foo"]


def test_crc_extend_chain(codec, golden):
    rng = np.random.default_rng(20261106)
    for e in golden["crc"]["extend"]:
        n = int(rng.integers(0, 5000))
        cut = int(rng.integers(0, n + 1))
        init = int(rng.integers(0, 2**32))
        d = rng.integers(0, 256, n, dtype=np.uint8)
        assert (n, cut, init) == (e["seed_n"], e["cut"], e["init"])
        assert codec.crc32c(d) == e["data_crc"]
        assert codec.crc32c_extend(init, d) == e["extend_init"]
        assert codec.crc32c_extend(codec.crc32c(d[:cut]), d[cut:]) == e["data_crc"]


def test_crc_table_path_equals_fast_path():
    p = O.port()
    rng = np.random.default_rng(1)
    for n in (0, 1, 7, 8, 15, 16, 17, 63, 1000, 4099):
        d = rng.integers(0, 256, n + 3, dtype=np.uint8)
        for s in range(3):  # exercise the alignment prologue
            a = d[s:s + n]
            a = np.ascontiguousarray(a)
            assert p.lib.b2o_crc32c_extend_tables(5, a.ctypes.data_as(O._u8p), n) == p.crc32c_extend(5, a)


def test_snappy_golden_vectors(codec, golden):
    for e in golden["vectors"]:
        d = O.corpus(e["corpus"], e["n"], e["index"])
        assert codec.crc32c(d) == e["crc"], e
        z = codec.compress(d)
        assert len(z) == e["clen"], e
        assert codec.crc32c(z) == e["ccrc"], e
        assert z[:16].hex() == e["head"]
        if "z" in e:
            assert z.hex() == e["z"]
        ok, back = codec.uncompress(z)
        assert ok and back == d.tobytes()
        assert codec.is_valid(z)
        assert codec.ulen_strict(z) == e["n"] == codec.ulen_lenient(z)


def test_snappy_special_inputs(codec, golden):
    special = {
        "this is a test": b"this is a test",
        "zeros65536": bytes(65536),
        "ff4096": b"\xff" * 4096,
        "abab": b"ab" * 3000,
        "a26x782": (b"abcdefghijklmnopqrstuvwxyz" * 31)[:782],
        "run_then_random": bytes(300) + O.corpus("rand62", 700, 3).tobytes() + bytes(300),
    }
    for e in golden["special"]:
        b = special[e["name"]]
        z = codec.compress(b)
        assert (len(b), codec.crc32c(b), len(z), codec.crc32c(z)) == (e["n"], e["crc"], e["clen"], e["ccrc"])
        if e["z"]:
            assert z.hex() == e["z"]
        ok, back = codec.uncompress(z)
        assert ok and back == b


def test_snappy_malformed_known_answers(codec, golden):
    # SURVEY.md Appendix B; answers recorded from the reference library
    for e in golden["malformed"]:
        b = bytes.fromhex(e["stream"])
        if not b:
            continue
        ok, out = codec.uncompress(b)
        assert ok == e["uncompress"], e
        if ok:
            assert out.hex() == e["out"], e
        assert codec.is_valid(b) == e["is_valid"], e
        assert codec.ulen_lenient(b) == e["ulen_lenient"], e
        assert codec.ulen_strict(b) == e["ulen_strict"], e


def test_max_compressed_length(codec):
    # snappy.cc:55-77
    for n in (0, 1, 5, 6, 1024, 4096, 65536, 1 << 20):
        assert codec.max_compressed_length(n) == 32 + n + n // 6
    assert codec.max_compressed_length(65536) == 76490
    assert codec.max_compressed_length(4096) == 4810


@pytest.mark.skipif(O.reference() is None, reason="oracle/_ref not built")
def test_port_equals_reference_random_sweep():
    p, r = O.port(), O.reference()
    rng = np.random.default_rng(7)
    for t in range(300):
        kind = ("az09", "rand62", "words")[t % 3]
        n = int(rng.integers(0, 70000 if t % 10 == 0 else 3000))
        d = O.corpus(kind, n, t)
        if t % 4 == 0 and n > 8:  # splice a long run and a repeated block in
            d = d.copy()
            a = int(rng.integers(0, n // 2))
            d[a:a + n // 4] = d[0]
        zp, zr = p.compress(d), r.compress(d)
        assert zp == zr, (kind, n, t)
        assert p.crc32c(d) == r.crc32c(d)
        okp, bp = p.uncompress(zr)
        assert okp and bp == d.tobytes()
    # random binary with low entropy (many short matches, overlapping copies)
    for t in range(100):
        n = int(rng.integers(1, 20000))
        d = rng.integers(0, 4, n, dtype=np.uint8)
        assert p.compress(d) == r.compress(d)


@pytest.mark.skipif(O.reference() is None, reason="oracle/_ref not built")
def test_corrupted_streams_agree_with_reference():
    p, r = O.port(), O.reference()
    rng = np.random.default_rng(11)
    base = [r.compress(O.corpus(k, n, i)) for k in ("words", "az09", "rand62") for i, n in enumerate((40, 300, 5000))]
    for t in range(1500):
        z = bytearray(base[t % len(base)])
        mode = t % 4
        if mode == 0:
            z[int(rng.integers(0, len(z)))] = int(rng.integers(0, 256))
        elif mode == 1:
            del z[int(rng.integers(1, len(z))):]
        elif mode == 2:
            z += bytes(rng.integers(0, 256, int(rng.integers(1, 6)), dtype=np.uint8))
        else:
            for _ in range(3):
                z[int(rng.integers(0, len(z)))] ^= 1 << int(rng.integers(0, 8))
        z = bytes(z)
        okr, outr = r.uncompress(z, cap=1 << 20)
        okp, outp = p.uncompress(z, cap=1 << 20)
        assert okr == okp, (t, z[:32].hex())
        if okr:
            assert outr == outp
        assert r.is_valid(z) == p.is_valid(z)


@pytest.mark.skipif(O.reference() is None, reason="oracle/_ref not built")
def test_block_layout_independence():
    # Compress(Source*,Sink*) output ignores how the Source chunks its bytes
    # (snappy.cc:887-921), so the oracle may take the concatenated payload.
    r = O.reference()
    for kind, n in (("words", 200000), ("az09", 70000), ("rand62", 65537), ("words", 8161)):
        d = O.corpus(kind, n)
        flat = r.compress(d)
        out = np.empty(r.max_compressed_length(n), dtype=np.uint8)
        for chunk in (1, 7, 100, 8160, 65536, 70000):
            k = r.lib.b2r_snappy_compress_chunked(d.ctypes.data_as(O._u8p), n, chunk, out.ctypes.data_as(O._u8p))
            assert out[:k].tobytes() == flat
        # and the sink-based decoder (scattered writer + RefillTag stitching) round-trips
        z = np.frombuffer(flat, dtype=np.uint8)
        back = np.empty(n, dtype=np.uint8)
        import ctypes as C
        for chunk in (1, 3, 100, 8160):
            prod = C.c_size_t(0)
            ok = r.lib.b2r_snappy_uncompress_sink(z.ctypes.data_as(O._u8p), z.size, chunk,
                                                  back.ctypes.data_as(O._u8p), n, C.byref(prod))
            assert ok == 1 and prod.value == n and back.tobytes() == d.tobytes()


def test_batch_drivers_roundtrip(codec):
    lengths = np.array([0, 1, 14, 15, 16, 100, 1024, 4096, 70000, 5], dtype=np.uint32)
    arena, off, ln = O.corpus_batch("words", lengths, first_index=3)
    crc, _ = codec.batch_crc32c(arena, off, ln, threads=3)
    for i in range(len(ln)):
        assert crc[i] == codec.crc32c(arena[int(off[i]):int(off[i]) + int(ln[i])])
    out, out_off, out_len, _ = codec.batch_compress(arena, off, ln, threads=2)
    for i in range(len(ln)):
        z = out[int(out_off[i]):int(out_off[i]) + int(out_len[i])].tobytes()
        assert z == codec.compress(arena[int(off[i]):int(off[i]) + int(ln[i])])
    back, ulen, ok, _ = codec.batch_uncompress(out, out_off, out_len, off, ln, threads=2)
    assert ok.all() and (ulen == ln).all()
    for i in range(len(ln)):
        a, b = int(off[i]), int(off[i]) + int(ln[i])
        assert back[a:b].tobytes() == arena[a:b].tobytes()
