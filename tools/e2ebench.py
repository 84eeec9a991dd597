#!/usr/bin/env python
"""End-to-end (host buffers in, host buffers out) timing of the sender / receiver calls alone, back to back and
overlapped on two threads, for one workload shape -- a tuning aid for the slab pipeline (bench.py is the contract).

  [B2C_TRACE=1] [B2C_SLAB_MB=..] python tools/e2ebench.py [--workload c2] [--threads N] [--steps K] [--tag NAME]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import brpc_b200  # noqa: E402
from oracle import oracle as O  # noqa: E402


def multi(a):
    """ONE process, ONE batch of N x the per-GPU payload, split over N GPUs by b2c_multi_*."""
    a.messages *= a.multi
    a.bytes *= a.multi
    arena, off, ln = O.corpus_batch(a.corpus, bench.workload_lengths(a))
    n = len(ln)
    payload = int(ln.astype(np.uint64).sum())
    cpu = O.best()
    threads = os.cpu_count() or 1
    segs, sb = brpc_b200.make_iovecs(arena, off, ln, seg_bytes=8160 if a.workload != "c2" else None,
                                     cuts=np.random.default_rng(1).integers(0, 1025, n) if a.workload == "c2" else None)
    mc = brpc_b200.MultiCodec(list(range(a.multi)))
    one = brpc_b200.Codec(0)
    cap = int((32 + ln.astype(np.uint64) + ln.astype(np.uint64) // 6 + 5).sum()) + 256 * a.multi
    out = one.pinned(cap)
    back = one.pinned(payload + 64)
    o_off, o_len, crc = mc.compress_crc_batch(segs, sb, out)
    w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, ln, threads=threads)
    assert (o_len == w_len).all()
    bad, first = O.batch_equal(out, o_off, w_out, w_off, w_len)
    assert bad == 0, (bad, first)
    del w_out
    comp_host = np.empty(int(o_len.astype(np.uint64).sum()), dtype=np.uint8)  # the receiver's pageable input
    c_off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(o_len, out=c_off[1:])
    for r in range(0, n, 1 << 16):  # pack the shards' streams
        for i in range(r, min(n, r + (1 << 16))):
            pass
    idx = np.repeat(o_off - c_off[:-1], o_len.astype(np.int64))
    comp_host[:] = out[(np.arange(len(comp_host), dtype=np.uint64) + idx).astype(np.int64)]
    csegs, csb = brpc_b200.make_iovecs(comp_host, c_off[:-1], o_len, seg_bytes=8160)
    b_off, status = mc.decompress_verify_batch(csegs, csb, crc, back)
    assert not status.any()
    bad, first = O.batch_equal(back, b_off[:-1], arena, off, ln)
    assert bad == 0, (bad, first)

    def timed(fn):
        t0 = time.perf_counter()
        for _ in range(a.steps):
            fn()
        return (time.perf_counter() - t0) / a.steps * 1e3

    send_ms = timed(lambda: mc.compress_crc_batch(segs, sb, out))
    recv_ms = timed(lambda: mc.decompress_verify_batch(csegs, csb, crc, back))
    print(json.dumps({"tag": a.tag, "mode": "b2c_multi, one process", "gpus": a.multi, "workload": a.workload, "corpus": a.corpus,
                      "payload": payload, "send_ms": round(send_ms, 2), "recv_ms": round(recv_ms, 2),
                      "serial_gbs": round(payload / 1e9 / (send_ms + recv_ms) * 1e3, 2)}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--corpus", default="words")
    ap.add_argument("--bytes", type=int, default=1 << 30)
    ap.add_argument("--messages", type=int, default=1 << 20)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--threads", type=int, default=0, help="host threads per ctx (0 = library default)")
    ap.add_argument("--over-threads", type=int, default=0, help="host threads per ctx in the overlapped run (0 = cores / 2)")
    ap.add_argument("--pinned", action="store_true", help="input blocks in page-locked memory + B2C_OPT_INPUT_PINNED")
    ap.add_argument("--multi", type=int, default=0, help="one process drives this many GPUs through b2c_multi_* (one batch per GPU's worth)")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    if a.multi:
        return multi(a)
    arena, off, ln = O.corpus_batch(a.corpus, bench.workload_lengths(a))
    s, r = brpc_b200.Codec(0), brpc_b200.Codec(0)
    B = bench.DeviceBatch(s, arena, off, ln, f"{a.workload}/{a.corpus}")
    H = bench.HostBatch(s, r, B, a.workload, 0, pinned_inputs=a.pinned)
    for c in (s, r):
        c.set_option(c.OPT_HOST_THREADS, a.threads)
    H.send(), H.recv()
    H.check()

    def timed(fn):
        t0 = time.perf_counter()
        for _ in range(a.steps):
            fn()
        return (time.perf_counter() - t0) / a.steps * 1e3

    send_ms, recv_ms = timed(H.send), timed(H.recv)
    over = H.time_overlapped(a.steps, None, a.over_threads or max(2, (os.cpu_count() or 2) // 2)) / a.steps * 1e3
    gb = B.payload / 1e9
    print(json.dumps({"tag": a.tag, "workload": a.workload, "corpus": a.corpus, "threads": a.threads, "pinned_inputs": a.pinned,
                      "send_ms": round(send_ms, 2), "recv_ms": round(recv_ms, 2), "serial_gbs": round(gb / (send_ms + recv_ms) * 1e3, 2),
                      "overlapped_ms": round(over, 2), "overlapped_gbs": round(gb / over * 1e3, 2)}), flush=True)


if __name__ == "__main__":
    main()
