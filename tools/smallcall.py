import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import brpc_b200
from oracle import oracle as O
codec = brpc_b200.Codec(0)
for n, size in ((1, 1024), (32, 1024), (512, 1024), (32, 16384)):
    arena, off, ln = O.corpus_batch('words', np.full(n, size, dtype=np.uint32))
    segs, sb = brpc_b200.make_iovecs(arena, off, ln, seg_bytes=8160)
    out = np.empty(int(sum(32 + int(l) + int(l) // 6 + 5 for l in ln)), dtype=np.uint8)
    for _ in range(20): codec.compress_crc_batch(segs, sb, out, masked=True)
    t0 = time.perf_counter(); K = 300
    for _ in range(K): o_off, crc = codec.compress_crc_batch(segs, sb, out, masked=True)
    t1 = time.perf_counter()
    for _ in range(20): codec.compress_batch(segs, sb, out) if hasattr(codec, 'compress_batch') else None
    print(f"compress_crc_batch n={n} x {size}: {(t1 - t0) / K * 1e6:.0f} us per call")
    if hasattr(codec, 'crc32c_batch'):
        t0 = time.perf_counter()
        for _ in range(K): codec.crc32c_batch(segs, sb)
        print(f"   crc32c_batch: {(time.perf_counter() - t0) / K * 1e6:.0f} us per call")
