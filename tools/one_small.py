import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
import brpc_b200
from oracle import oracle as O
codec = brpc_b200.Codec(0)
arena, off, ln = O.corpus_batch('words', np.full(32, 1024, dtype=np.uint32))
segs, sb = brpc_b200.make_iovecs(arena, off, ln, seg_bytes=8160)
out = np.empty(int(sum(32 + int(l) + int(l) // 6 + 5 for l in ln)), dtype=np.uint8)
for _ in range(3): o_off, crc = codec.compress_crc_batch(segs, sb, out, masked=True)
back = np.empty(40000, dtype=np.uint8)
