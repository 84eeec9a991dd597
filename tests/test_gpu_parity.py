"""GPU parity tests (run on the B200 box with -m gpu): the CUDA path, called
through the C ABI (libb2c.so), against the CPU oracle on the same inputs."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from tests.util import assert_bytes_equal, pack, pack_tight, slices

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def codec():
    import brpc_b200
    c = brpc_b200.Codec(0)
    yield c
    c.close()


@pytest.fixture(scope="module")
def cpu():
    return O.best()


SIZES = [0, 1, 3, 4, 5, 14, 15, 16, 17, 31, 32, 33, 59, 60, 61, 62, 63, 64, 65, 100, 127, 128, 129, 255, 256, 257, 782,
         1000, 1023, 1024, 1025, 2047, 2048, 2049, 4095, 4096, 4097, 8160, 8161, 12435, 16384, 32768, 65535, 65536,
         65537, 70000, 131072, 200000]


# ---------------------------------------------------------------- CRC32C
def test_crc_known_answers(codec, cpu, golden):
    msgs = [bytes.fromhex(e["in"]) for e in golden["crc"]["rfc3720"]] + [b"123456789", b"hello world", b"a", b"foo", b""]
    want = [e["crc"] for e in golden["crc"]["rfc3720"]] + [0xE3069283, golden["crc"]["hello_world"],
                                                            cpu.crc32c(b"a"), cpu.crc32c(b"foo"), 0]
    arena, off, ln = pack(msgs)
    got = codec.crc32c(arena, off, ln)
    assert [int(x) for x in got] == want


def test_crc_parity_sweep_with_init_and_misalignment(codec, cpu):
    rng = np.random.default_rng(3)
    lens = list(range(0, 200)) + SIZES + [1 << 20, (4 << 20) + 3]
    msgs = [rng.integers(0, 256, n, dtype=np.uint8).tobytes() for n in lens]
    for align in (16, 1):
        arena, off, ln = pack(msgs, align=align)
        if align == 1:  # also shift the whole arena by 3 bytes
            arena = np.concatenate([np.zeros(3, np.uint8), arena])
            off = off + np.uint64(3)
        init = rng.integers(0, 2**32, len(msgs), dtype=np.uint32)
        got0 = codec.crc32c(arena, off, ln)
        got1 = codec.crc32c(arena, off, ln, init=init)
        for i, m in enumerate(msgs):
            assert int(got0[i]) == cpu.crc32c(m), (align, i, len(m))
            assert int(got1[i]) == cpu.crc32c_extend(int(init[i]), m), (align, i, len(m))


def test_crc_extend_composition(codec, cpu):
    # crc32c_unittest.cc:59-62: Extend(Value(A), B) == Value(A||B), on the GPU
    rng = np.random.default_rng(4)
    A = [rng.integers(0, 256, n, dtype=np.uint8).tobytes() for n in (6, 100, 5000, 70000)]
    B = [rng.integers(0, 256, n, dtype=np.uint8).tobytes() for n in (5, 4097, 63, 9000)]
    a_arena, a_off, a_len = pack(A)
    b_arena, b_off, b_len = pack(B)
    ab_arena, ab_off, ab_len = pack([x + y for x, y in zip(A, B)])
    ca = codec.crc32c(a_arena, a_off, a_len)
    cab = codec.crc32c(b_arena, b_off, b_len, init=ca)
    whole = codec.crc32c(ab_arena, ab_off, ab_len)
    assert (cab == whole).all()
    assert int(codec.crc32c(*pack([b"hello world"]))[0]) == cpu.crc32c_extend(cpu.crc32c(b"hello "), b"world")


def test_crc_c1_shape(codec, cpu):
    # BASELINE configs[0]: 4096 x 4 KB echo payloads
    arena, off, ln = O.corpus_batch("words", np.full(4096, 4096, dtype=np.uint32))
    got = codec.crc32c(arena, off, ln)
    want, _ = cpu.batch_crc32c(arena, off, ln, threads=8)
    assert (got == want).all()


# ---------------------------------------------------------------- compress
def _check_compress(codec, cpu, arena, off, ln, label):
    packed, out_off = codec.compress(arena, off, ln)
    w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, ln, threads=8)
    got_len = np.diff(out_off).astype(np.uint32)
    bad = np.nonzero(got_len != w_len)[0]
    for i in range(len(ln)):
        g = packed[int(out_off[i]):int(out_off[i + 1])].tobytes()
        w = w_out[int(w_off[i]):int(w_off[i]) + int(w_len[i])].tobytes()
        assert_bytes_equal(g, w, f"{label} msg {i} len {int(ln[i])} (len mismatches: {len(bad)})")
    return packed, out_off


@pytest.mark.parametrize("kind", ["az09", "rand62", "words"])
def test_compress_parity_sizes(codec, cpu, kind):
    arena, off, ln = O.corpus_batch(kind, np.array(SIZES + [1 << 20, (1 << 20) + 77], dtype=np.uint32), first_index=5)
    _check_compress(codec, cpu, arena, off, ln, kind)


def test_compress_golden_fixtures(codec, cpu, golden):
    # committed fixtures generated from the reference's own code (tests/golden/make_golden.py)
    vec = [e for e in golden["vectors"] if e["n"] <= 70000]
    msgs = [O.corpus(e["corpus"], e["n"], e["index"]).tobytes() for e in vec]
    arena, off, ln = pack(msgs)
    packed, out_off = codec.compress(arena, off, ln)
    crc = codec.crc32c(arena, off, ln)
    for i, e in enumerate(vec):
        z = packed[int(out_off[i]):int(out_off[i + 1])].tobytes()
        assert len(z) == e["clen"], e
        assert z[:16].hex() == e["head"], e
        assert cpu.crc32c(z) == e["ccrc"], e
        assert int(crc[i]) == e["crc"], e
        if "z" in e:
            assert z.hex() == e["z"]


def test_compress_low_entropy_and_runs(codec, cpu):
    rng = np.random.default_rng(9)
    msgs = [bytes(65536), b"\xff" * 4096, b"ab" * 3000, bytes(300) + O.corpus("rand62", 700, 3).tobytes() + bytes(300)]
    for t in range(120):
        n = int(rng.integers(1, 9000 if t % 8 else 70000))
        msgs.append(rng.integers(0, int(rng.integers(2, 6)), n, dtype=np.uint8).tobytes())
    for t in range(40):
        n = int(rng.integers(1, 5000))
        d = rng.integers(0, 256, n, dtype=np.uint8)
        if n > 64:
            a = int(rng.integers(0, n // 2))
            d[a:a + n // 3] = d[0]  # long run inside random bytes
        msgs.append(d.tobytes())
    arena, off, ln = pack(msgs, align=1)  # arbitrary alignment of every message
    _check_compress(codec, cpu, arena, off, ln, "low-entropy")


def test_compress_c1_shape(codec, cpu):
    for kind in ("words", "az09", "rand62"):
        arena, off, ln = O.corpus_batch(kind, np.full(4096, 4096, dtype=np.uint32))
        _check_compress(codec, cpu, arena, off, ln, f"C1 {kind}")


def test_compress_c2_shape_sample(codec, cpu):
    arena, off, ln = O.corpus_batch("words", np.full(65536, 1024, dtype=np.uint32), first_index=1000)
    _check_compress(codec, cpu, arena, off, ln, "C2 words")


def _full_parity(codec, cpu, arena, off, ln, label):
    """crc, compressed bytes and decoded bytes of EVERY message against the CPU checker (threaded memcmp)."""
    threads = os.cpu_count() or 1
    crc = codec.crc32c(arena, off, ln)
    want_crc, _ = cpu.batch_crc32c(arena, off, ln, threads=threads)
    assert (crc == want_crc).all(), f"{label}: crc"
    packed, out_off = codec.compress(arena, off, ln)
    c_len = np.diff(out_off).astype(np.uint32)
    w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, ln, threads=threads)
    assert (c_len == w_len).all(), f"{label}: compressed lengths"
    bad, first = O.batch_equal(packed, out_off[:-1], w_out, w_off, w_len)
    assert bad == 0, f"{label}: compressed bytes differ in {bad} messages, first {first}"
    del w_out
    back, ulen, st = codec.decompress(packed, out_off[:-1], c_len, off, ln)
    assert (st == 0).all() and (ulen == ln).all(), f"{label}: decompress status"
    ref_back, ref_ulen, ref_ok, _ = cpu.batch_uncompress(packed, out_off[:-1], c_len, off, ln, threads=threads)
    assert ref_ok.all() and (ref_ulen == ln).all()
    bad, first = O.batch_equal(back, off, ref_back, off, ln)
    assert bad == 0, f"{label}: decoded bytes differ from the checker's in {bad} messages, first {first}"
    bad, first = O.batch_equal(back, off, arena, off, ln)
    assert bad == 0, f"{label}: round trip differs in {bad} messages, first {first}"
    return packed, out_off, crc


def test_compress_c3_shape(codec, cpu):
    # BASELINE configs[2] at full size: 256 x 1 MiB, every message compared
    arena, off, ln = O.corpus_batch("words", np.full(256, 1 << 20, dtype=np.uint32), first_index=9)
    packed, out_off, _ = _full_parity(codec, cpu, arena, off, ln, "C3 words")
    assert packed[:3].tobytes() == bytes([0x80, 0x80, 0x40])  # varint(1 MiB)


def test_compress_empty_batch_and_empty_messages(codec, cpu):
    packed, out_off = codec.compress(np.zeros(64, np.uint8), np.zeros(0, np.uint64), np.zeros(0, np.uint32))
    assert len(packed) == 0 and list(out_off) == [0]
    arena, off, ln = pack([b"", b"", b"x", b""])
    packed, out_off = codec.compress(arena, off, ln)
    assert packed.tobytes() == b"\x00" + b"\x00" + b"\x01\x00x" + b"\x00"


# ---------------------------------------------------------------- decompress
def _decompress(codec, streams, caps=None):
    arena, off, ln = pack_tight(streams)
    if caps is None:
        caps = []
        for s in streams:
            v = O.port().ulen_lenient(s) if len(s) else None
            caps.append(v if v is not None and v < (1 << 26) else 0)
    caps = np.array(caps, dtype=np.uint32)
    o_off = np.zeros(len(caps), dtype=np.uint64)
    if len(caps):
        o_off[1:] = np.cumsum(caps[:-1].astype(np.uint64))
    out, ulen, status = codec.decompress(arena, off, ln, o_off, caps)
    return out, o_off, ulen, status


@pytest.mark.parametrize("kind", ["az09", "rand62", "words"])
def test_decompress_roundtrip_sizes(codec, cpu, kind):
    msgs = [O.corpus(kind, n, 11 + i).tobytes() for i, n in enumerate(SIZES + [1 << 20])]
    streams = [cpu.compress(m) for m in msgs]
    out, o_off, ulen, status = _decompress(codec, streams)
    assert (status == 0).all(), np.nonzero(status)[0]
    for i, m in enumerate(msgs):
        assert int(ulen[i]) == len(m)
        assert_bytes_equal(out[int(o_off[i]):int(o_off[i]) + len(m)].tobytes(), m, f"{kind} msg {i} n={len(m)}")


def test_decompress_malformed_known_answers(codec, golden):
    # SURVEY Appendix B + extras; expected answers recorded from the reference library
    es = [e for e in golden["malformed"] if e["stream"]]
    streams = [bytes.fromhex(e["stream"]) for e in es]
    out, o_off, ulen, status = _decompress(codec, streams)
    for i, e in enumerate(es):
        if e["uncompress"]:
            assert status[i] == 0, e
            n = len(bytes.fromhex(e["out"]))
            assert out[int(o_off[i]):int(o_off[i]) + n].tobytes().hex() == e["out"], e
        else:
            assert status[i] != 0, e


def test_decompress_corrupted_streams_agree_with_oracle(codec, cpu):
    rng = np.random.default_rng(11)
    base = [cpu.compress(O.corpus(k, n, i)) for k in ("words", "az09", "rand62") for i, n in enumerate((40, 300, 5000, 70000))]
    streams = []
    for t in range(2000):
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
        streams.append(bytes(z))
    out, o_off, ulen, status = _decompress(codec, streams)
    n_ok = 0
    for i, s in enumerate(streams):
        cap = O.port().ulen_lenient(s)
        if cap is None or cap >= (1 << 26):
            assert status[i] != 0
            continue
        ok, want = cpu.uncompress(s, cap=cap)
        assert (status[i] == 0) == ok, (i, s[:24].hex(), int(status[i]), ok)
        if ok:
            n_ok += 1
            assert_bytes_equal(out[int(o_off[i]):int(o_off[i]) + len(want)].tobytes(), want, f"corrupted {i}")
    assert 0 < n_ok < len(streams)


def test_decompress_runs_of_every_period(codec, cpu):
    """Runs: consecutive copies with one offset are decoded as one period extension; every period from 1 byte to more
    than a batch of copies produces, in slots of every decoder class, must still give the reference's bytes."""
    rng = np.random.default_rng(5)
    msgs = []
    for period in (1, 2, 3, 4, 5, 7, 16, 36, 63, 64, 65, 100, 255, 1000, 2047, 2048, 5000):
        pat = rng.integers(0, 256, period, dtype=np.uint8).tobytes()
        for total in (900, 7000, 66000, 300000):
            body = (pat * (total // period + 1))[:total]
            if total > 7000:  # a run interrupted by other data now and then
                body = body[:total // 2] + O.corpus("words", 300, period).tobytes() + body[total // 2:]
            msgs.append(body)
    streams = [cpu.compress(np.frombuffer(m, dtype=np.uint8)) for m in msgs]
    out, o_off, ulen, status = _decompress(codec, streams)
    assert (status == 0).all(), np.nonzero(status)[0]
    for i, m in enumerate(msgs):
        assert int(ulen[i]) == len(m)
        assert_bytes_equal(out[int(o_off[i]):int(o_off[i]) + len(m)].tobytes(), m, f"run {i}")


def test_decompress_too_small_and_ulen(codec, cpu):
    streams = [cpu.compress(O.corpus("words", n, n)) for n in (100, 1000, 10000)]
    out, o_off, ulen, status = _decompress(codec, streams, caps=[100, 999, 20000])
    assert list(status) == [0, 2, 0] and list(ulen) == [100, 1000, 10000]
    arena, off, ln = pack_tight(streams + [b"\xff\xff\xff\xff\xff\x01", b""])
    u, st = codec.uncompressed_lengths(arena, off, ln)
    assert list(u[:3]) == [100, 1000, 10000] and list(st) == [0, 0, 0, 1, 1]


def test_decompress_foreign_encoder_features(codec):
    # copy-4 tags, offsets beyond 64 KiB and overlapping (RLE) copies — never emitted by the
    # reference compressor but accepted by its decoder (snappy.h:199-201)
    lit = bytes(range(256)) * 300  # 76800 bytes
    def literal(b):
        n = len(b) - 1
        return bytes([62 << 2]) + n.to_bytes(3, "little") + b
    body = literal(lit) + bytes([3 | (63 << 2)]) + (70000).to_bytes(4, "little") + bytes([2 | (63 << 2)]) + (1).to_bytes(2, "little")
    total = len(lit) + 128
    pre = bytearray()
    v = total
    while v >= 128:
        pre.append((v & 0x7f) | 0x80)
        v >>= 7
    pre.append(v)
    s = bytes(pre) + body
    ok, want = O.port().uncompress(s)
    assert ok
    out, o_off, ulen, status = _decompress(codec, [s])
    assert status[0] == 0
    assert_bytes_equal(out[:total].tobytes(), want, "foreign")


# ---------------------------------------------------------------- end-to-end host API
def test_e2e_iovec_batches(codec, cpu):
    import brpc_b200
    rng = np.random.default_rng(21)
    # C2 shape: 1 KiB bodies cut once at a random place; C3 shape: 1 MiB in 8160-byte blocks
    arena2, off2, ln2 = O.corpus_batch("words", np.full(20000, 1024, dtype=np.uint32), first_index=77)
    segs2, sb2 = brpc_b200.make_iovecs(arena2, off2, ln2, cuts=rng.integers(0, 1025, len(ln2)))
    arena3, off3, ln3 = O.corpus_batch("words", np.array([1 << 20, 12435, 0, 1, (1 << 20) + 5], dtype=np.uint32))
    segs3, sb3 = brpc_b200.make_iovecs(arena3, off3, ln3, seg_bytes=8160)
    for arena, off, ln, segs, sb in ((arena2, off2, ln2, segs2, sb2), (arena3, off3, ln3, segs3, sb3)):
        crc = codec.crc32c_batch(segs, sb)
        want_crc, _ = cpu.batch_crc32c(arena, off, ln, threads=8)
        assert (crc == want_crc).all()
        cap = int(sum(32 + int(l) + int(l) // 6 + 5 for l in ln))
        out = np.empty(cap, dtype=np.uint8)
        out_off = codec.compress_batch(segs, sb, out)
        # same call with a page-locked result buffer (the library then DMAs straight into it)
        out_p = codec.pinned(cap)
        out_off_p = codec.compress_batch(segs, sb, out_p)
        assert (out_off_p == out_off).all()
        assert out_p[:int(out_off[-1])].tobytes() == out[:int(out_off[-1])].tobytes()
        w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, ln, threads=8)
        assert (np.diff(out_off).astype(np.uint32) == w_len).all()
        for i in range(0, len(ln), max(1, len(ln) // 500)):
            assert_bytes_equal(out[int(out_off[i]):int(out_off[i + 1])].tobytes(),
                               w_out[int(w_off[i]):int(w_off[i]) + int(w_len[i])].tobytes(), f"e2e msg {i}")
        # feed the compressed streams back, in 100-byte blocks
        csegs, csb = brpc_b200.make_iovecs(out, out_off[:-1], np.diff(out_off).astype(np.uint32), seg_bytes=100)
        back = np.empty(int(ln.astype(np.uint64).sum()) + 64, dtype=np.uint8)
        b_off, status = codec.decompress_batch(csegs, csb, back)
        assert (status == 0).all()
        back_p = codec.pinned(len(back))
        b_off_p, status_p = codec.decompress_batch(csegs, csb, back_p)
        assert (b_off_p == b_off).all() and (status_p == 0).all()
        assert back_p[:int(b_off[-1])].tobytes() == back[:int(b_off[-1])].tobytes()
        assert (np.diff(b_off).astype(np.uint32) == ln).all()
        for i in range(0, len(ln), max(1, len(ln) // 500)):
            a, b = int(off[i]), int(off[i]) + int(ln[i])
            assert_bytes_equal(back[int(b_off[i]):int(b_off[i + 1])].tobytes(), arena[a:b].tobytes(), f"e2e back {i}")
        h2d, d2h = codec.last_transfer_bytes()
        assert h2d > 0 and d2h > 0


def test_e2e_fused_sender_and_receiver_calls(codec, cpu):
    """compress+crc in one pass, decompress+verify in one pass: same bytes and checksums as the separate calls
    (and as the CPU checker), and a flipped checksum / a damaged stream are told apart."""
    import brpc_b200
    rng = np.random.default_rng(5)
    lens = np.concatenate([np.full(6000, 1024), rng.integers(0, 3000, 500), [0, 1, 70000, 1 << 20]]).astype(np.uint32)
    arena, off, ln = O.corpus_batch("words", lens, first_index=3)
    segs, sb = brpc_b200.make_iovecs(arena, off, ln, seg_bytes=8160)
    cap = int(sum(32 + int(l) + int(l) // 6 + 5 for l in ln))
    out, out2 = np.empty(cap, dtype=np.uint8), codec.pinned(cap)
    ref_off = codec.compress_batch(segs, sb, out)
    want_crc, _ = cpu.batch_crc32c(arena, off, ln, threads=8)
    for masked in (False, True):
        o_off, crc = codec.compress_crc_batch(segs, sb, out2, masked=masked)
        assert (o_off == ref_off).all()
        assert out2[:int(o_off[-1])].tobytes() == out[:int(ref_off[-1])].tobytes()
        want = np.array([cpu.mask(int(c)) for c in want_crc], dtype=np.uint32) if masked else want_crc
        assert (crc == want).all()
    # receiver: streams back in 100-byte blocks, one checksum flipped, one stream damaged
    clen = np.diff(ref_off).astype(np.uint32)
    damaged = out.copy()
    victim = 17
    damaged[int(ref_off[victim]) + 3] ^= 0xFF  # 2 preamble bytes, the literal's tag, then its first byte: decodes, to other bytes
    csegs, csb = brpc_b200.make_iovecs(damaged, ref_off[:-1], clen, seg_bytes=100)
    bad_crc = crc.copy()
    bad_crc[40] ^= 1
    back = codec.pinned(int(ln.astype(np.uint64).sum()) + 64)
    b_off, status = codec.decompress_verify_batch(csegs, csb, bad_crc, back)
    expect = np.zeros(len(ln), dtype=np.int32)
    expect[40] = 3  # B2C_CRC_MISMATCH
    ok = [i for i in range(len(ln)) if i != victim]
    assert (status[ok] == expect[ok]).all()
    assert status[victim] == 3
    assert (np.diff(b_off).astype(np.uint32) == ln).all()
    for i in list(range(0, len(ln), 97)) + [len(ln) - 1, len(ln) - 2]:
        if i == victim:
            continue
        a, b = int(off[i]), int(off[i]) + int(ln[i])
        assert_bytes_equal(back[int(b_off[i]):int(b_off[i + 1])].tobytes(), arena[a:b].tobytes(), f"fused back {i}")


def test_e2e_decompress_into_iovecs(codec, cpu):
    """RawUncompressToIOVec for a batch: output scattered over 8160-byte blocks (and odd little ones), too little
    room reported as the reference's writer would fail, corrupt streams rejected, spare room left untouched."""
    import brpc_b200
    rng = np.random.default_rng(77)
    lens = np.concatenate([np.full(3000, 1024), rng.integers(0, 40000, 300), [0, 1, 65536, 200000]]).astype(np.uint32)
    arena, off, ln = O.corpus_batch("words", lens, first_index=11)
    cap = int(sum(32 + int(l) + int(l) // 6 + 5 for l in ln))
    comp = np.empty(cap, dtype=np.uint8)
    isegs, isb = brpc_b200.make_iovecs(arena, off, ln, seg_bytes=8160)
    c_off = codec.compress_batch(isegs, isb, comp)
    clen = np.diff(c_off).astype(np.uint32)
    comp[int(c_off[5]) + 3] ^= 0x55  # literal byte of message 5: still decodes
    comp[int(c_off[9])] = 0xFF       # preamble of message 9 now claims far more than the stream can hold
    csegs, csb = brpc_b200.make_iovecs(comp, c_off[:-1], clen, seg_bytes=1000)
    # output: every message gets its own region of a big buffer, cut into blocks; message 7 gets one byte too little,
    # message 8 gets 100 spare bytes
    room = ln.astype(np.int64).copy()
    room[7] -= 1
    room[8] += 100
    room[9] += 200  # enough for what message 9's damaged preamble claims (1151): it must fail in the decoder, not on room
    o_off = np.concatenate([[0], np.cumsum(room + 16)]).astype(np.uint64)
    sink = np.full(int(o_off[-1]), 0xEE, dtype=np.uint8)
    osegs, osb = brpc_b200.make_iovecs(sink, o_off[:-1], room.astype(np.uint32), seg_bytes=777)
    status = codec.decompress_iov_batch(csegs, csb, osegs, osb)
    for i in range(len(ln)):
        a = int(o_off[i])
        got = sink[a:a + int(room[i])]
        if i == 7:
            assert status[i] == 2 and (got == 0xEE).all()  # B2C_TOO_SMALL, nothing written
        elif i == 9:
            assert status[i] == 1
        else:
            assert status[i] == 0, i
            want = arena[int(off[i]):int(off[i]) + int(ln[i])]
            if i == 5:
                assert got[:len(want)].tobytes() != want.tobytes()
            else:
                assert_bytes_equal(got[:len(want)].tobytes(), want.tobytes(), f"iov msg {i}")
            assert (got[len(want):] == 0xEE).all()
        assert (sink[a + int(room[i]):int(o_off[i + 1])] == 0xEE).all()  # the gap between regions


def test_e2e_decompress_rejects_like_reference(codec, golden):
    import brpc_b200
    es = [e for e in golden["malformed"] if e["stream"]]
    arena, off, ln = pack_tight([bytes.fromhex(e["stream"]) for e in es])
    segs, sb = brpc_b200.make_iovecs(arena, off, ln)
    out = np.empty(1 << 16, dtype=np.uint8)
    o_off, status = codec.decompress_batch(segs, sb, out)
    for i, e in enumerate(es):
        assert (status[i] == 0) == e["uncompress"], e
        if e["uncompress"]:
            n = len(bytes.fromhex(e["out"]))
            assert out[int(o_off[i]):int(o_off[i]) + n].tobytes().hex() == e["out"]


# ---------------------------------------------------------------- full-size properties
def test_c2_full_size_roundtrip_properties(codec, cpu):
    """BASELINE configs[1] at full size (2^20 x 1 KiB): compress -> decompress round trip,
    CRC of every decoded body equals CRC of the original, and an oracle check on a sample."""
    n = 1 << 20
    arena, off, ln = O.corpus_batch("words", np.full(n, 1024, dtype=np.uint32), first_index=0)
    lay = codec.layout(off, ln)
    d_in = codec.to_device(arena)
    d_crc = codec.alloc(4 * n)
    codec.crc32c_dev(lay, d_in, d_crc)
    d_out = codec.alloc(lay.max_compressed_bytes + 64)
    d_off = codec.alloc(8 * (n + 1))
    codec.compress_dev(lay, d_in, d_out, d_off)
    out_off = codec.d2h(d_off, 8 * (n + 1), np.uint64)
    total = int(out_off[-1])
    clay = codec.layout(out_off[:-1], np.diff(out_off).astype(np.uint32))
    d_back = codec.alloc(int(lay.arena_bytes) + 64)
    d_st = codec.alloc(4 * n)
    codec.decompress_dev(clay, d_out, lay, d_back, d_st)
    d_crc2 = codec.alloc(4 * n)
    codec.crc32c_dev(lay, d_back, d_crc2)
    st = codec.d2h(d_st, 4 * n, np.int32)
    c1 = codec.d2h(d_crc, 4 * n, np.uint32)
    c2 = codec.d2h(d_crc2, 4 * n, np.uint32)
    assert (st == 0).all()
    assert (c1 == c2).all()
    # sample against the oracle: CRCs and compressed bytes of 4096 evenly spaced messages
    idx = np.arange(0, n, n // 4096)
    packed = codec.d2h(d_out, total)
    want_crc, _ = cpu.batch_crc32c(arena, off[idx], ln[idx], threads=8)
    assert (c1[idx] == want_crc).all()
    w_out, w_off, w_len, _ = cpu.batch_compress(arena, off[idx], ln[idx], threads=8)
    for j, i in enumerate(idx):
        assert_bytes_equal(packed[int(out_off[i]):int(out_off[i + 1])].tobytes(),
                           w_out[int(w_off[j]):int(w_off[j]) + int(w_len[j])].tobytes(), f"C2 full msg {i}")


# ---------------------------------------------------------------- validate / decompress+verify / large shapes
def test_validate_agrees_with_oracle(codec, cpu, golden):
    rng = np.random.default_rng(5)
    base = [cpu.compress(O.corpus(k, n, i)) for k in ("words", "az09", "rand62") for i, n in enumerate((40, 300, 5000, 200000))]
    streams = [bytes.fromhex(e["stream"]) for e in golden["malformed"] if e["stream"]] + base
    for t in range(600):
        z = bytearray(base[t % len(base)])
        if t % 3 == 0:
            z[int(rng.integers(0, len(z)))] ^= 1 << int(rng.integers(0, 8))
        elif t % 3 == 1:
            del z[int(rng.integers(1, len(z))):]
        else:
            z += bytes(rng.integers(0, 256, 2, dtype=np.uint8))
        streams.append(bytes(z))
    arena, off, ln = pack_tight(streams)
    st, ulen = codec.validate(arena, off, ln)
    for i, s in enumerate(streams):
        assert (st[i] == 0) == cpu.is_valid(s), (i, s[:24].hex())
        v = O.port().ulen_lenient(s)
        assert int(ulen[i]) == (v or 0)


def test_decompress_verify_masked_crc(codec, cpu):
    # snappy framing's chunk check: masked CRC-32C of the uncompressed data (framing_format.txt:39-59)
    msgs = [O.corpus("words", n, i).tobytes() for i, n in enumerate((0, 1, 100, 1024, 4096, 65536, 70000, 300000))]
    streams = [cpu.compress(m) for m in msgs]
    masked = np.array([cpu.mask(cpu.crc32c(m)) for m in msgs], dtype=np.uint32)
    arena, off, ln = pack_tight(streams)
    caps = np.array([len(m) for m in msgs], dtype=np.uint32)
    o_off = np.zeros(len(caps), dtype=np.uint64)
    o_off[1:] = np.cumsum(caps[:-1].astype(np.uint64))
    out, st = codec.decompress_verify(arena, off, ln, o_off, caps, masked)
    assert (st == 0).all()
    for i, m in enumerate(msgs):
        assert out[int(o_off[i]):int(o_off[i]) + len(m)].tobytes() == m
    bad = masked.copy()
    bad[3] ^= 1
    bad[6] += 7
    _, st = codec.decompress_verify(arena, off, ln, o_off, caps, bad)
    assert [int(x) for x in st] == [0, 0, 0, 3, 0, 0, 3, 0]
    # a corrupt stream stays "corrupt", not "crc mismatch"
    streams[4] = streams[4][:-3]
    arena, off, ln = pack_tight(streams)
    _, st = codec.decompress_verify(arena, off, ln, o_off, caps, masked)
    assert st[4] == 1 and st[5] == 0


def test_c4_shape_mixed_sizes_parity(codec, cpu):
    # BASELINE configs[3]: lognormal 64 B - 4 MiB mix, 1 GB of it here (bench.py gates the full 8 GB the same way)
    # — every size class and both decode modes in one batch, every message compared
    rng = np.random.default_rng(2024)
    ln = np.clip(np.round(np.exp(rng.normal(np.log(16384), 2.0, 40000))), 64, 4 << 20).astype(np.uint32)
    ln = ln[: int(np.searchsorted(np.cumsum(ln.astype(np.uint64)), 1_000_000_000)) + 1]
    assert int(ln.astype(np.uint64).sum()) >= 1_000_000_000
    arena, off, ln = O.corpus_batch("words", ln, first_index=31)
    _full_parity(codec, cpu, arena, off, ln, "C4 mix")


@pytest.mark.parametrize("kind", ["words", "az09", "rand62"])
def test_c5_shape_fragments_decompress_verify(codec, cpu, kind):
    # BASELINE configs[4]: independent 64 KiB fragments (1 GiB = 16,384 of them), decompress + CRC32C verify;
    # all of them compared with the checker
    n = 16384
    arena, off, ln = O.corpus_batch(kind, np.full(n, 65536, dtype=np.uint32), first_index=7)
    packed, out_off, crc = _full_parity(codec, cpu, arena, off, ln, f"C5 {kind}")
    masked = ((crc >> np.uint32(15)) | (crc << np.uint32(17))) + np.uint32(0xa282ead8)  # crc32c.h:44-48
    assert int(masked[5]) == cpu.mask(int(crc[5]))
    out, st = codec.decompress_verify(packed, out_off[:-1], np.diff(out_off).astype(np.uint32), off, ln, masked)
    assert (st == 0).all()
    bad, first = O.batch_equal(out, off, arena, off, ln)
    assert bad == 0, (bad, first)
    wrong = masked.copy()
    wrong[[3, n - 1]] ^= np.uint32(1)
    _, st = codec.decompress_verify(packed, out_off[:-1], np.diff(out_off).astype(np.uint32), off, ln, wrong)
    assert set(np.nonzero(st)[0].tolist()) == {3, n - 1} and int(st[3]) == 3


def test_large_streams_with_literals_of_every_length_and_damage(codec, cpu):
    """Streams of 0.2-2.5 MiB whose elements come in every size the block-boundary pass tells apart (short elements,
    literals of 61-255 / 256-2047 / 2048+ bytes that jump over its near window, its segments and the range of a table
    entry), whole and damaged: status and bytes must equal the reference's."""
    rng = np.random.default_rng(77)
    bodies = []
    for t, total in enumerate((200_000, 700_000, 2_500_000)):
        parts, have = [], 0
        while have < total:
            if rng.random() < 0.5:
                n = int(rng.integers(50, 4000))
                parts.append(O.corpus("words", n, int(rng.integers(0, 1 << 20))).tobytes())
            else:
                n = int(rng.choice([61, 70, 200, 256, 300, 2040, 2100, 9000, 17000, 70000]))
                parts.append(O.corpus("rand62", n, int(rng.integers(0, 1 << 20))).tobytes())
            have += n
        bodies.append(b"".join(parts))
    base = [cpu.compress(np.frombuffer(b, dtype=np.uint8)) for b in bodies]
    streams = list(base)
    for t in range(60):
        z = bytearray(base[t % len(base)])
        mode = t % 3
        if mode == 0:
            z[int(rng.integers(8, len(z)))] ^= 1 << int(rng.integers(0, 8))
        elif mode == 1:
            del z[int(rng.integers(len(z) // 2, len(z))):]
        else:
            at = int(rng.integers(8, len(z) - 4))
            z[at:at + 2] = bytes(rng.integers(0, 256, 2, dtype=np.uint8))
        streams.append(bytes(z))
    caps = [len(bodies[i % len(bodies)]) for i in range(len(streams))]
    out, o_off, ulen, status = _decompress(codec, streams, caps=caps)
    n_ok = 0
    for i, z in enumerate(streams):
        ok, want = cpu.uncompress(z, cap=caps[i])
        assert (status[i] == 0) == ok, (i, int(status[i]), ok)
        if ok:
            n_ok += 1
            assert_bytes_equal(out[int(o_off[i]):int(o_off[i]) + len(want)].tobytes(), want, f"large stream {i}")
    assert n_ok >= 3


def test_large_foreign_stream_falls_back_to_whole_message_decode(codec):
    # a > 64 KiB stream whose copies reach across 64 KiB output boundaries (legal, but never produced by the
    # reference compressor): the boundary scan must refuse to split it and the result must still be exact
    block = O.corpus("rand62", 40000, 1).tobytes()
    def literal(b):
        return bytes([62 << 2]) + (len(b) - 1).to_bytes(3, "little") + b
    def copy4(offset, length):
        return bytes([3 | ((length - 1) << 2)]) + offset.to_bytes(4, "little")
    body = literal(block)
    total = len(block)
    for _ in range(4000):          # 4000 x 64-byte copies from 40000 bytes back: crosses 64 KiB boundaries
        body += copy4(40000, 64)
        total += 64
    pre = bytearray()
    v = total
    while v >= 128:
        pre.append((v & 0x7f) | 0x80)
        v >>= 7
    pre.append(v)
    s = bytes(pre) + body
    ok, want = O.port().uncompress(s)
    assert ok and len(want) == total
    out, o_off, ulen, status = _decompress(codec, [s, s])
    assert list(status) == [0, 0]
    assert_bytes_equal(out[:total].tobytes(), want, "foreign large")
    st, _ = codec.validate(*pack_tight([s]))
    assert st[0] == 0


def test_e2e_zero_copy_gather_from_pinned_blocks(codec, cpu):
    """B2C_OPT_INPUT_PINNED: the input blocks live in page-locked memory (a pinned IOBuf block pool) and
    the device gathers them itself; results must equal the staged path and the oracle."""
    import brpc_b200
    rng = np.random.default_rng(33)
    ln = np.concatenate([np.full(6000, 1024), rng.integers(0, 70000, 300), np.array([0, 1, 1 << 20])]).astype(np.uint32)
    rng.shuffle(ln)
    arena, off, ln = O.corpus_batch("words", ln, first_index=55)
    pinned = codec.pinned(len(arena))
    pinned[:] = arena
    for kw in ({"cuts": rng.integers(0, 5000, len(ln))}, {"seg_bytes": 8160}, {"seg_bytes": 333}):
        segs, sb = brpc_b200.make_iovecs(pinned, off, ln, **kw)
        codec.set_option(codec.OPT_INPUT_PINNED, 1)
        try:
            crc = codec.crc32c_batch(segs, sb)
            out = codec.pinned(int(sum(32 + int(l) + int(l) // 6 + 5 for l in ln)))
            out_off = codec.compress_batch(segs, sb, out)
            csegs, csb = brpc_b200.make_iovecs(out, out_off[:-1], np.diff(out_off).astype(np.uint32), seg_bytes=4096)
            back = codec.pinned(int(ln.astype(np.uint64).sum()) + 64)
            b_off, status = codec.decompress_batch(csegs, csb, back)
        finally:
            codec.set_option(codec.OPT_INPUT_PINNED, 0)
        want_crc, _ = cpu.batch_crc32c(arena, off, ln, threads=8)
        assert (crc == want_crc).all()
        w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, ln, threads=8)
        assert (np.diff(out_off).astype(np.uint32) == w_len).all()
        bad, first = O.batch_equal(out, out_off[:-1], w_out, w_off, w_len)
        assert bad == 0, f"zc: compressed bytes differ in {bad} messages, first {first}"
        assert (status == 0).all() and (np.diff(b_off).astype(np.uint32) == ln).all()
        bad, first = O.batch_equal(back, b_off[:-1], arena, off, ln)
        assert bad == 0, f"zc: decoded bytes differ in {bad} messages, first {first}"


def test_e2e_zero_copy_refuses_pageable_blocks(codec):
    """With B2C_OPT_INPUT_PINNED set, a block that is NOT page-locked must be refused on the host (the gather
    kernel would fault on it), and the ctx must stay usable."""
    import brpc_b200
    arena, off, ln = O.corpus_batch("words", np.full(64, 3000, dtype=np.uint32))
    segs, sb = brpc_b200.make_iovecs(arena, off, ln, seg_bytes=1000)  # numpy memory: pageable
    codec.set_option(codec.OPT_INPUT_PINNED, 1)
    try:
        with pytest.raises(brpc_b200.B2CError, match="page-locked"):
            codec.crc32c_batch(segs, sb)
    finally:
        codec.set_option(codec.OPT_INPUT_PINNED, 0)
    want, _ = O.best().batch_crc32c(arena, off, ln)
    assert (codec.crc32c_batch(segs, sb) == want).all()


# ---------------------------------------------------------------- alternate kernel paths stay exact
@pytest.mark.parametrize("env", [
    {"B2C_WINDOW": "7"},                                   # orbit compressor on every fragment class
    {"B2C_WINDOW": "7", "B2C_WINDOW_STAGE": "0", "B2C_GTAB_WPS": "0"},
    {"B2C_WINDOW_SMEM64": "1"},                            # 64 KiB fragments: table + staged fragment in shared memory
    {"B2C_LOCKSTEP": "1", "B2C_LOCKSTEP_GLOBAL": "4"},     # lock-step compressor with its state in global memory
    {"B2C_SMALL_LPW_C": "2"},                              # two bodies per warp
    {"B2C_SMALL_KERNELS": "0"},                            # warp kernels for small bodies too
    {"B2C_LOCKSTEP": "1"},                                 # lock-step state machines for bodies <= 1 KiB (8 per warp)
    {"B2C_LOCKSTEP": "1", "B2C_LOCKSTEP_LPW": "32", "B2C_LOCKSTEP_D": "1"},  # 32 per warp; lock-step decoder
    {"B2C_GTAB_WPS": "0", "B2C_BATCH_DECODE": "0"},        # shared-memory tables; one tag per step
    {"B2C_BATCH_DECODE_SMALL": "1", "B2C_RING_DECODE": "1", "B2C_BLOCK_DECODE": "0"},
    {"B2C_LIFT_SCAN": "0"},                                # tag-by-tag block-boundary scan
    {"B2C_ASYNC_GATHER": "0", "B2C_NT_GATHER": "0", "B2C_SLAB_MB": "1", "B2C_HOST_THREADS": "3"},
])
def test_alternate_kernel_paths_agree(cpu, env):
    """Every tunable path (the defaults are the fastest measured, the others stay selectable) is byte-exact too."""
    import os
    import brpc_b200
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        c = brpc_b200.Codec(0)  # the ctx reads its tunables when it is created
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    try:
        rng = np.random.default_rng(9)
        lens = np.concatenate([rng.integers(0, 1500, 600), rng.integers(1500, 9000, 60),
                               [65536, 65537, 200000, 0, 14, 15]]).astype(np.uint32)
        for kind in ("words", "az09", "rand62"):
            arena, off, ln = O.corpus_batch(kind, lens, first_index=5)
            packed, c_off = c.compress(arena, off, ln)
            w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, ln, threads=4)
            assert (np.diff(c_off).astype(np.uint32) == w_len).all(), (env, kind)
            for i in range(len(ln)):
                assert_bytes_equal(packed[int(c_off[i]):int(c_off[i + 1])].tobytes(),
                                   w_out[int(w_off[i]):int(w_off[i]) + int(w_len[i])].tobytes(), f"{env} {kind} msg {i}")
            back, ulen, status = c.decompress(packed, c_off[:-1], np.diff(c_off).astype(np.uint32), off, ln)
            assert (status == 0).all() and (ulen == ln).all()
            for i in range(len(ln)):
                a, b = int(off[i]), int(off[i]) + int(ln[i])
                assert_bytes_equal(back[a:b].tobytes(), arena[a:b].tobytes(), f"{env} {kind} back {i}")
        # host-buffer calls on the same ctx
        arena, off, ln = O.corpus_batch("words", lens, first_index=5)
        segs, sb = brpc_b200.make_iovecs(arena, off, ln, seg_bytes=8160)
        out = np.empty(int(sum(32 + int(l) + int(l) // 6 + 5 for l in ln)), dtype=np.uint8)
        o_off, crc = c.compress_crc_batch(segs, sb, out, masked=False)
        want_crc, _ = cpu.batch_crc32c(arena, off, ln, threads=4)
        assert (crc == want_crc).all()
        w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, ln, threads=4)
        assert (np.diff(o_off).astype(np.uint32) == w_len).all()
    finally:
        c.close()


def test_tunables_by_name(cpu):
    """b2c_ctx_set_tunable: the supported way to pick a kernel variant (the environment is only a shortcut to it)."""
    import brpc_b200
    c = brpc_b200.Codec(0)
    try:
        with pytest.raises(brpc_b200.B2CError):
            c.set_tunable("no_such_tunable", 1)
        with pytest.raises(brpc_b200.B2CError):
            c.set_tunable("lockstep_lpw", 5)
        arena, off, ln = O.corpus_batch("words", np.array([1024] * 300 + [5000, 70000], dtype=np.uint32), first_index=3)
        want = [cpu.compress(arena[int(o):int(o) + int(l)]) for o, l in zip(off, ln)]
        for name, value in (("lockstep", 1), ("lockstep", 0), ("window", 7), ("window", 6), ("small_lpw_c", 8)):
            c.set_tunable(name, value)
            packed, c_off = c.compress(arena, off, ln)
            assert [packed[int(c_off[i]):int(c_off[i + 1])].tobytes() for i in range(len(ln))] == want, (name, value)
    finally:
        c.close()


def test_dev_entry_points_on_an_exactly_sized_arena(codec, cpu):
    """The *_dev entry points may read the aligned unit that holds the arena's last byte (documented in
    include/b2c_codec.h: 16 readable bytes past the arena, which b2c_dev_alloc provides): an arena whose last
    message ends exactly at its end, uploaded without any padding of the harness's own."""
    rng = np.random.default_rng(4)
    for last in (1, 3, 1023, 1024, 4097, 65537):
        ln = np.array([700, 1024, 13, last], dtype=np.uint32)
        arena, off, ln = O.corpus_batch("words", ln, align=1)
        arena = np.ascontiguousarray(arena[:int(off[-1]) + last])  # no slack at all
        lay = codec.layout(off, ln)
        d_in = codec.to_device(arena, pad=0)
        d_crc = codec.alloc(4 * len(ln))
        codec.crc32c_dev(lay, d_in, d_crc)
        want, _ = cpu.batch_crc32c(arena, off, ln)
        assert (codec.d2h(d_crc, 4 * len(ln), np.uint32) == want).all()
        d_comp = codec.alloc(lay.max_compressed_bytes)
        d_coff = codec.alloc(8 * (len(ln) + 1))
        codec.compress_dev(lay, d_in, d_comp, d_coff)
        c_off = codec.d2h(d_coff, 8 * (len(ln) + 1), np.uint64)
        packed = codec.d2h(d_comp, int(c_off[-1]))
        for i in range(len(ln)):
            assert packed[int(c_off[i]):int(c_off[i + 1])].tobytes() == cpu.compress(arena[int(off[i]):int(off[i]) + int(ln[i])])
        # decode from an exactly sized arena of streams into an exactly sized output arena
        d_z = codec.to_device(packed, pad=0)
        clay = codec.layout(c_off[:-1], np.diff(c_off).astype(np.uint32))
        d_back = codec.alloc(len(arena))
        d_st = codec.alloc(4 * len(ln))
        codec.decompress_dev(clay, d_z, lay, d_back, d_st)
        assert (codec.d2h(d_st, 4 * len(ln), np.int32) == 0).all()
        assert codec.d2h(d_back, len(arena)).tobytes() == arena.tobytes()


def test_decodable_prefix_matches_uncompress_as_much_as_possible(codec):
    """b2c_snappy_decodable_prefix_dev against the reference's snappy::UncompressAsMuchAsPossible (snappy.cc:1519-1524):
    the value it returns, the number of bytes its sink receives, and — decoding the prefix the kernel describes
    through the normal decoder — those bytes themselves.  Truncated, damaged and lying streams, and the case where the
    reference's writer counts a block twice."""
    if O.reference() is None:
        pytest.skip("needs oracle/_ref (the reference's own code)")
    cpu = O.best()
    rng = np.random.default_rng(17)

    def lit(b):
        n1 = len(b) - 1
        if n1 < 60:
            return bytes([n1 << 2]) + b
        if n1 < 256:
            return bytes([60 << 2, n1]) + b
        if n1 < 65536:
            return bytes([61 << 2]) + n1.to_bytes(2, "little") + b
        return bytes([62 << 2]) + n1.to_bytes(3, "little") + b

    streams = [b"", b"\x00", bytes([100]) + lit(b"a" * 30) + lit(b"b" * 150),               # literal refused after filling a block
               bytes([100]) + lit(b"a" * 30) + bytes([2 | (39 << 2), 10, 0]) + bytes([2 | (59 << 2), 5, 0]),  # copy overruns
               bytes([100]) + lit(b"a" * 30) + bytes([2 | (39 << 2), 0, 0]),                       # offset 0
               bytes([10]) + lit(b"x" * 20), bytes([0x80, 0x80, 0x10]) + lit(b"q" * 70000)[:50000]]  # too long; cut literal
    for t in range(300):
        body = O.corpus(("words", "az09", "rand62")[t % 3], int(rng.integers(1, 200000)), index=t).tobytes()
        z = bytearray(cpu.compress(body))
        how = t % 4
        if how == 0:
            z = z[: int(rng.integers(1, len(z) + 1))]
        elif how == 1:
            z[int(rng.integers(0, len(z)))] ^= 1 << int(rng.integers(0, 8))
        elif how == 2 and len(z) > 8:
            a = int(rng.integers(4, len(z)))
            z[a:a + 3] = bytes(rng.integers(0, 256, len(z[a:a + 3]), dtype=np.uint8))
        streams.append(bytes(z))
    arena, off, ln = pack_tight(streams)
    got = codec.decodable_prefix(arena, off, ln)
    synth, want_bytes = [], []
    for i, s in enumerate(streams):
        want_ret, sunk = O.ref_uncompress_as_much(s)
        ret, nsunk, consumed, partial, psrc = (int(x) for x in got[i])
        assert (ret, nsunk) == (want_ret, len(sunk)), (i, len(s), ret, nsunk, want_ret, len(sunk))
        # the prefix as a stream of its own: varint(sunk) || whole tags || the partially applied literal
        k = 0
        while k < len(s) and s[k] >= 128:
            k += 1
        k += 1
        v, pre = nsunk, bytearray()
        while v >= 128:
            pre.append((v & 0x7f) | 0x80)
            v >>= 7
        pre.append(v)
        z = bytes(pre) + s[k:consumed]
        if partial:
            z += bytes([63 << 2]) + (partial - 1).to_bytes(4, "little") + s[psrc:psrc + partial]
        synth.append(z)
        want_bytes.append(sunk)
    arena2, off2, ln2 = pack_tight(synth)
    caps = np.array([len(b) for b in want_bytes], dtype=np.uint32)
    o_off = np.zeros(len(caps), dtype=np.uint64)
    o_off[1:] = np.cumsum(caps[:-1].astype(np.uint64))
    out, ulen, st = codec.decompress(arena2, off2, ln2, o_off, caps)
    assert (st == 0).all(), np.nonzero(st)[0][:10]
    for i, b in enumerate(want_bytes):
        assert out[int(o_off[i]):int(o_off[i]) + len(b)].tobytes() == b, i
