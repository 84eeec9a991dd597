#!/usr/bin/env python
"""Device-resident op timings for one workload shape (kernel tuning aid; bench.py is the contract).

  python tools/devbench.py --workload c2 --corpus words [--bytes N] [--check] [--tag NAME]

Prints one line: per-op ms (crc / compress / pack / decompress / decompress+verify) from the library's own CUDA
events.  --check compares every compressed stream and every decoded body with the CPU checker first."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import brpc_b200  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--corpus", default="words")
    ap.add_argument("--bytes", type=int, default=1 << 30)
    ap.add_argument("--messages", type=int, default=1 << 20)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    lengths = bench.workload_lengths(a)
    arena, off, ln = O.corpus_batch(a.corpus, lengths)
    n = len(ln)
    payload = int(ln.astype(np.uint64).sum())
    codec = brpc_b200.Codec(0)
    lay = codec.layout(off, ln)
    d_in = codec.to_device(arena)
    d_crc = codec.alloc(4 * n)
    d_comp = codec.alloc(lay.max_compressed_bytes + 64)
    d_coff = codec.alloc(8 * (n + 1))
    d_back = codec.alloc(int(lay.arena_bytes) + 64)
    d_st = codec.alloc(4 * n)
    codec.compress_dev(lay, d_in, d_comp, d_coff)
    c_off = codec.d2h(d_coff, 8 * (n + 1), np.uint64)
    comp = int(c_off[-1])
    clay = codec.layout(c_off[:-1], np.diff(c_off).astype(np.uint32))
    codec.crc32c_dev(lay, d_in, d_crc)
    crc = codec.d2h(d_crc, 4 * n, np.uint32)
    masked = ((crc >> np.uint32(15)) | (crc << np.uint32(17))) + np.uint32(0xa282ead8)
    d_masked = codec.to_device(masked.astype(np.uint32))

    def step():
        codec.crc32c_dev(lay, d_in, d_crc)
        codec.compress_dev(lay, d_in, d_comp, d_coff)
        codec.decompress_dev(clay, d_comp, lay, d_back, d_st)

    step()
    codec.sync()
    if a.check:
        cpu = O.best()
        threads = os.cpu_count() or 1
        w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, ln, threads=threads)
        assert (np.diff(c_off).astype(np.uint32) == w_len).all(), "compressed lengths differ"
        packed = codec.d2h(d_comp, comp)
        bad = [i for i in range(n) if packed[int(c_off[i]):int(c_off[i + 1])].tobytes() !=
               w_out[int(w_off[i]):int(w_off[i]) + int(w_len[i])].tobytes()]
        assert not bad, f"compressed bytes differ at {bad[:5]}"
        back = codec.d2h(d_back, int(lay.arena_bytes))
        for i in range(n):
            s, e = int(off[i]), int(off[i]) + int(ln[i])
            assert back[s:e].tobytes() == arena[s:e].tobytes(), f"round trip {i}"
        want_crc, _ = cpu.batch_crc32c(arena, off, ln, threads=threads)
        assert (crc == want_crc).all()
    for _ in range(2):
        step()
    codec.sync()
    codec.prof_enable(True)
    for _ in range(a.steps):
        step()
    codec.sync()
    prof = {nm: codec.prof_total_ms(k) for k, nm in enumerate(("crc", "compress", "decompress", "pack"))}
    codec.prof_enable(False)
    e0, e1 = codec.event(), codec.event()
    codec.decompress_verify_dev(clay, d_comp, lay, d_back, d_masked, d_st)
    codec.sync()
    e0.record()
    for _ in range(a.steps):
        codec.decompress_verify_dev(clay, d_comp, lay, d_back, d_masked, d_st)
    e1.record()
    codec.sync()
    assert (codec.d2h(d_st, 4 * n, np.int32) == 0).all()
    res = {k: round(v[0] / max(v[1], 1), 4) for k, v in prof.items()}
    res["verify"] = round(e1.elapsed_ms_since(e0) / a.steps, 4)
    res["step"] = round(res["crc"] + res["compress"] + res["pack"] + res["decompress"], 4)
    print(json.dumps({"tag": a.tag, "workload": a.workload, "corpus": a.corpus, "bodies": n, "payload": payload,
                      "ratio": round(comp / max(payload, 1), 4), "checked": a.check, "ms": res}), flush=True)
    codec.close()


if __name__ == "__main__":
    main()
