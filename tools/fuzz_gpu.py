#!/usr/bin/env python
"""Randomised parity runs on the GPU: batches of random shape (sizes from 0 to 3 MiB, mixed content, runs, damaged
streams), every result compared with the CPU checker (status and bytes).  The fixed cases live in tests/; this is for
soaking new kernel paths (`python tools/fuzz_gpu.py --seconds 120 --seed 7`).  Prints one line per round and a summary;
exits non-zero on the first mismatch."""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import brpc_b200  # noqa: E402
from oracle import oracle as O  # noqa: E402
from tests.util import pack, pack_tight  # noqa: E402


def body(rng, n):
    """n bytes: one corpus, a mix of corpora in pieces, or a run with a period."""
    kind = rng.integers(0, 6)
    if kind <= 2:
        return O.corpus(("words", "az09", "rand62")[kind], n, int(rng.integers(0, 1 << 20))).tobytes()[:n]
    if kind == 3:  # pieces of different corpora: literals and copies of every length
        parts, have = [], 0
        while have < n:
            k = int(min(n - have, rng.choice([3, 17, 61, 70, 200, 300, 1500, 2100, 9000, 70000])))
            parts.append(O.corpus(("words", "rand62", "az09")[int(rng.integers(0, 3))], k, int(rng.integers(0, 1 << 20))).tobytes()[:k])
            have += k
        return b"".join(parts)
    if kind == 4:  # a run
        period = int(rng.choice([1, 2, 3, 4, 7, 36, 64, 100, 1000, 5000]))
        pat = rng.integers(0, 256, period, dtype=np.uint8).tobytes()
        return (pat * (n // period + 1))[:n]
    small = rng.integers(0, int(rng.integers(2, 7)), n, dtype=np.uint8)  # tiny alphabet: many short matches
    return small.tobytes()


def sizes(rng):
    shape = rng.integers(0, 5)
    if shape == 0:
        return rng.integers(0, 1500, int(rng.integers(1, 3000)))
    if shape == 1:
        return rng.integers(0, 70000, int(rng.integers(1, 200)))
    if shape == 2:
        return np.clip(np.round(np.exp(rng.normal(np.log(16384), 2.0, int(rng.integers(1, 400))))), 0, 3 << 20).astype(np.int64)
    if shape == 3:
        return rng.choice([0, 1, 14, 15, 16, 1023, 1024, 1025, 4096, 4097, 8192, 8193, 65535, 65536, 65537, 131072, 200000],
                          int(rng.integers(1, 60)))
    return rng.integers(60000, 3 << 20, int(rng.integers(1, 12)))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=60)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    rng = np.random.default_rng(a.seed)
    cpu = O.best()
    codec = brpc_b200.Codec(0)
    threads = os.cpu_count() or 1
    t_end = time.time() + a.seconds
    rounds = msgs = damaged = total = 0
    while time.time() < t_end:
        ln = sizes(rng)
        bodies = [body(rng, int(n)) for n in ln]
        arena, off, lens = pack(bodies)
        # ---- compress + crc
        packed, c_off = codec.compress(arena, off, lens)
        w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, lens, threads=threads)
        assert (np.diff(c_off).astype(np.uint32) == w_len).all(), f"round {rounds}: compressed lengths differ"
        bad, first = O.batch_equal(packed, c_off[:-1].astype(np.uint64), w_out, w_off.astype(np.uint64), w_len, threads=threads)
        assert bad == 0, f"round {rounds}: compressed bytes differ in {bad} messages, first {first} (len {lens[first]})"
        lay = codec.layout(off, lens)
        d_in = codec.to_device(arena)
        d_crc = codec.alloc(4 * max(len(lens), 1))
        codec.crc32c_dev(lay, d_in, d_crc)
        crc = codec.d2h(d_crc, 4 * len(lens), np.uint32)
        want_crc, _ = cpu.batch_crc32c(arena, off, lens, threads=threads)
        assert (crc == want_crc).all(), f"round {rounds}: crc"
        # ---- decompress: the streams, and damaged copies of some of them
        streams = [packed[int(c_off[i]):int(c_off[i + 1])].tobytes() for i in range(len(lens))]
        caps = [len(b) for b in bodies]
        for i in rng.integers(0, len(streams), min(len(streams), 12)):
            z = bytearray(streams[int(i)])
            if len(z) < 4:
                continue
            mode = int(rng.integers(0, 3))
            if mode == 0:
                z[int(rng.integers(0, len(z)))] ^= 1 << int(rng.integers(0, 8))
            elif mode == 1:
                del z[int(rng.integers(1, len(z))):]
            else:
                at = int(rng.integers(0, len(z) - 2))
                z[at:at + 2] = bytes(rng.integers(0, 256, 2, dtype=np.uint8))
            streams.append(bytes(z))
            caps.append(len(bodies[int(i)]))
            damaged += 1
        s_arena, s_off, s_len = pack_tight(streams)
        caps = np.array(caps, dtype=np.uint32)
        o_off = np.zeros(len(caps), dtype=np.uint64)
        o_off[1:] = np.cumsum(caps[:-1].astype(np.uint64))
        out, ulen, status = codec.decompress(s_arena, s_off, s_len, o_off, caps)
        for i, z in enumerate(streams):
            if i < len(bodies):
                assert status[i] == 0 and int(ulen[i]) == len(bodies[i]), f"round {rounds}: stream {i} status {status[i]}"
                got = out[int(o_off[i]):int(o_off[i]) + len(bodies[i])].tobytes()
                assert got == bodies[i], f"round {rounds}: decoded bytes of message {i} (len {len(bodies[i])}) differ"
            else:
                ok, want = cpu.uncompress(z, cap=int(caps[i]))
                assert (status[i] == 0) == ok, f"round {rounds}: damaged stream {i}: status {status[i]}, reference ok={ok}"
                if ok:
                    got = out[int(o_off[i]):int(o_off[i]) + len(want)].tobytes()
                    assert got == want, f"round {rounds}: damaged stream {i} decodes differently"
        for b in (d_in, d_crc):
            b.free()
        del lay
        # ---- the host-buffer calls (block lists of random block size through the slab pipeline), every third round
        if rounds % 3 == 0 and len(lens):
            seg_bytes = int(rng.choice([64, 700, 4096, 8160, 65536, 1 << 20]))
            segs, sb = brpc_b200.make_iovecs(arena, off, lens, seg_bytes=seg_bytes)
            hout = np.empty(int(sum(32 + int(l) + int(l) // 6 + 5 for l in lens)), dtype=np.uint8)
            h_off, h_crc = codec.compress_crc_batch(segs, sb, hout, masked=True)
            assert np.array_equal(h_off, c_off), f"round {rounds}: host-call offsets differ"
            assert np.array_equal(hout[:int(c_off[-1])], packed[:int(c_off[-1])]), f"round {rounds}: host-call bytes differ"
            masked = ((want_crc >> np.uint32(15)) | (want_crc << np.uint32(17))) + np.uint32(0xa282ead8)
            assert np.array_equal(h_crc, masked), f"round {rounds}: host-call checksums differ"
            zsegs, zsb = brpc_b200.make_iovecs(hout, c_off[:-1].astype(np.uint64), np.diff(c_off).astype(np.uint32),
                                               seg_bytes=int(rng.choice([100, 8160, 1 << 20])))
            wrong = masked.copy()
            flip = int(rng.integers(0, len(lens)))
            wrong[flip] ^= np.uint32(1)
            back = np.empty(int(lens.astype(np.uint64).sum()) + 64, dtype=np.uint8)
            b_off, b_st = codec.decompress_verify_batch(zsegs, zsb, wrong, back)
            for i in range(len(lens)):
                if i == flip:
                    assert b_st[i] != 0, f"round {rounds}: a wrong checksum went unnoticed"
                    continue
                assert b_st[i] == 0, f"round {rounds}: host-call receiver status {b_st[i]} for message {i}"
                assert back[int(b_off[i]):int(b_off[i + 1])].tobytes() == bodies[i], f"round {rounds}: host-call body {i}"
        rounds += 1
        msgs += len(lens)
        total += int(lens.sum())
        print(f"round {rounds}: {len(lens)} bodies, {int(lens.sum())} bytes ok", flush=True)
    print(f"fuzz: {rounds} rounds, {msgs} bodies, {total / 1e6:.1f} MB, {damaged} damaged streams: all equal to the reference "
          f"({cpu.kind})")
    codec.close()


if __name__ == "__main__":
    main()
