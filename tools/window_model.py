# Lane-level model of the window-parallel compressor (one warp per fragment, 32 speculative table
# steps per round, serial chain walk over the precomputed results).  Validated against the oracle
# before the CUDA transcription in csrc/snappy_window.cu.
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import oracle as O
from compress_model import hashb, emit_literal, emit_copy, oracle_frag

def prefix8(inp, a, b):
    m = 0
    while m < 8 and inp[a + m] == inp[b + m]:
        m += 1
    return m

def model(data, stats=None):
    n = len(data); inp = bytes(data) + bytes(16)
    def ld32(p): return int.from_bytes(inp[p:p+4], 'little')
    tsize = 256
    while tsize < 16384 and tsize < n: tsize <<= 1
    shift = 32 - (tsize.bit_length() - 1)
    T = [0] * tsize
    out = []
    next_emit = 0
    rounds = 0
    if n >= 15:
        ip_limit = n - 15
        ip = 1; probing = True; skip = 32
        done = False
        while not done:
            rounds += 1
            # ---- round setup: 32 speculative positions
            if probing:
                steps = [(skip + l) >> 5 for l in range(32)]
                pos = [ip + sum(steps[:l]) for l in range(32)]
            else:
                pos = [ip + l for l in range(32)]
            ok = [p <= ip_limit for p in pos]          # loads stay inside the fragment
            h = [hashb(ld32(p), shift) if ok[l] else -1 for l, p in enumerate(pos)]
            c0 = [T[h[l]] if ok[l] else 0 for l in range(32)]
            m0 = [prefix8(inp, c0[l], pos[l]) if ok[l] else 0 for l in range(32)]
            L = []            # inserts of this round, chronological: (hash, position)
            elems = []        # (lit_start, p, cand, matched)
            p = ip
            while True:
                nxt = p + (skip >> 5)
                if probing and nxt > ip_limit:  # snappy.cc:386-388, needs nothing precomputed
                    done = True; break
                if p not in pos or not ok[pos.index(p)] or len(L) >= 62:
                    break                       # not precomputed: next round starts here
                j = pos.index(p)
                if probing:
                    skip += 1
                hh = h[j]; cand = c0[j]; m = m0[j]
                late = [q for (hq, q) in L if hq == hh]
                if late:
                    cand = late[-1]; m = prefix8(inp, cand, p)
                T[hh] = p & 0xffff; L.append((hh, p))
                if m >= 4:
                    if m == 8:
                        while p + m < n and inp[cand + m] == inp[p + m]: m += 1
                    elems.append((next_emit, p, cand, m))
                    p2 = p + m; next_emit = p2
                    if p2 >= ip_limit:
                        p = p2; done = True; break
                    hq = hashb(ld32(p2 - 1), shift); T[hq] = (p2 - 1) & 0xffff; L.append((hq, p2 - 1))
                    probing = False; p = p2
                else:
                    if probing: p = nxt
                    else: p += 1; probing = True; skip = 32
            ip = p
            for (ls, q, cand, m) in elems:
                if q > ls: emit_literal(out, inp[ls:q], q - ls)
                emit_copy(out, q - cand, m)
    if next_emit < n: emit_literal(out, inp[next_emit:n], n - next_emit)
    if stats is not None: stats.append(rounds)
    return bytes(out)

if __name__ == "__main__":
    rng = np.random.default_rng(5); cases = []
    for kind in ('az09', 'rand62', 'words'):
        for n in (1, 14, 15, 16, 17, 40, 100, 257, 1000, 1024, 1025, 4096, 20000, 65536):
            cases.append(O.corpus(kind, n, n).tobytes())
    cases += [bytes(65536), b'ab' * 3000, bytes(300) + O.corpus('rand62', 700, 3).tobytes() + bytes(300)]
    for t in range(150):
        n = int(rng.integers(1, 3000)); cases.append(rng.integers(0, int(rng.integers(2, 6)), n, dtype=np.uint8).tobytes())
    for t in range(300):
        cases.append(O.corpus('words', int(rng.integers(1, 1025)), t).tobytes())
    bad = sum(model(d) != oracle_frag(d) for d in cases)
    print(len(cases), 'cases bad', bad)
    st = []
    for t in range(50): model(O.corpus('words', 1024, t).tobytes(), st)
    print('rounds per 1 KiB words body: mean', sum(st) / len(st))
