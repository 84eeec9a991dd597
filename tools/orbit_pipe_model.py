# Lane-level model of the ORBIT compressor: one warp per fragment; every round resolves up to 32 table steps
# with no serial walk.  The 32 lanes run the step for 32 positions against a table snapshot; the window is cut
# at the first lane whose hash repeats an earlier lane's (inside the cut the snapshot cannot go stale); the
# chain through the window is then the orbit of lane 0 under next(j) (= j+1 on a miss, the landing lane of the
# copy on a hit), found by pointer doubling.  Checked against the oracle before the CUDA transcription.
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import oracle as O
from compress_model import hashb, emit_literal, emit_copy, oracle_frag

CAP = 40       # speculative per-lane match length limit (>= window size: a capped match always leaves the window)
END = 99

def model_pipe(data, stats=None):
    n = len(data); inp = bytes(data) + bytes(64)
    def ld32(p): return int.from_bytes(inp[p:p+4], 'little')
    tsize = 256
    while tsize < 16384 and tsize < n: tsize <<= 1
    shift = 32 - (tsize.bit_length() - 1)
    T = [0] * tsize
    out = []; next_emit = 0; rounds = 0; visits = 0
    cache = {}; blocking = 0; blocked_rounds = 0
    if n >= 15:
        ip_limit = n - 15
        ip = 1; s = 32; done = False
        while not done:
            rounds += 1
            for q in [q for q in cache if q < ip or q >= ip + 64]: del cache[q]
            K1 = 32 if s <= 32 else max(0, 64 - s)
            consecutive = K1 >= 16
            if consecutive:
                W = K1; step = [1] * 32; pos = [ip + j for j in range(32)]
            else:
                W = 32; step = [max(1, (s + j) >> 5) for j in range(32)]
                pos = [ip + sum(step[:j]) for j in range(32)]
            if consecutive and pos[0] in cache:   # no blocking probe: the window ends at the first position not probed yet
                for j in range(W):
                    if pos[j] + 1 <= ip_limit and pos[j] not in cache:
                        W = j; break
            vis_ok = [j < W and pos[j] + step[j] <= ip_limit for j in range(32)]
            if not vis_ok[0]:
                done = True; break
            h = [0] * 32; c0 = [0] * 32; m = [0] * 32
            if not consecutive: cache.clear()
            blk = 0
            for j in range(32):
                if not vis_ok[j]: continue
                p = pos[j]
                if consecutive and p in cache:
                    h[j], c0[j] = cache[p]
                    assert T[h[j]] == c0[j], 'stale entry used'
                    assert h[j] == hashb(ld32(p), shift)
                else:
                    h[j] = hashb(ld32(p), shift); c0[j] = T[h[j]]; blk += 1
                    if consecutive: cache[p] = (h[j], c0[j])
            blocking += blk; blocked_rounds += 1 if blk else 0
            if consecutive:   # speculative probes of the next 32 positions, issued now (before this round's inserts)
                for q in range(ip, ip + 64):
                    if q not in cache and q + 1 <= ip_limit:
                        hq_ = hashb(ld32(q), shift); cache[q] = (hq_, T[hq_])
            for j in range(32):
                if not vis_ok[j]: continue
                p = pos[j]
                k = 0
                while k < CAP and p + k < n and inp[c0[j] + k] == inp[p + k]: k += 1
                m[j] = k
            # collision cut
            Wc = W
            for j in range(32):
                if vis_ok[j] and any(vis_ok[i] and h[i] == h[j] for i in range(j)):
                    Wc = min(Wc, j); break
            # next lane
            nl = [END] * 32; kind = [None] * 32   # exit kinds: 'copy', 'miss', 'stop'
            for j in range(Wc):
                if not vis_ok[j]: continue
                if m[j] >= 4:
                    t = pos[j] + m[j]
                    if consecutive and m[j] < CAP and t - ip < Wc and t < ip_limit:
                        nl[j] = t - ip           # lands inside the window (t < ip_limit => lane visitable)
                    else:
                        kind[j] = 'copy'
                else:
                    if j + 1 < Wc:
                        if vis_ok[j + 1]: nl[j] = j + 1
                        else: kind[j] = 'stop'   # the loop-top check fails at the next position
                    else:
                        kind[j] = 'miss'
            # orbit of lane 0 (pointer doubling on the GPU; plain walk here)
            V = []; j = 0
            while j != END:
                V.append(j); j = nl[j]
            last = V[-1]
            visits += len(V)
            # full length of a capped / exiting copy
            for j in V:
                if m[j] >= 4 and m[j] == CAP:
                    k = m[j]; p = pos[j]
                    while p + k < n and inp[c0[j] + k] == inp[p + k]: k += 1
                    m[j] = k
            # inserts: visited lanes, then tails (largest position wins per hash)
            ins = set(h[j] for j in V)
            for j in V: T[h[j]] = pos[j] & 0xffff
            tails = {}
            copies = [j for j in V if m[j] >= 4]
            for j in copies:
                e = pos[j] + m[j]
                if e < ip_limit:
                    q = e - 1; hq = hashb(ld32(q), shift)
                    tails[hq] = max(tails.get(hq, -1), q)
            for hq, q in tails.items(): T[hq] = q & 0xffff
            ins |= set(tails.keys())
            for q in [q for q, (hh, cc) in cache.items() if hh in ins]: cache[q] = (cache[q][0], T[cache[q][0]])  # forwarded: what the inserts left in the slot
            # elements
            for j in copies:
                p = pos[j]
                if p > next_emit: emit_literal(out, inp[next_emit:p], p - next_emit)
                emit_copy(out, p - c0[j], m[j])
                next_emit = p + m[j]
            # state
            if m[last] >= 4:
                ip = pos[last] + m[last]; s = 31
                if ip >= ip_limit: done = True
            else:
                after = len([j for j in V if not copies or j > copies[-1]])
                s = (31 + after) if copies else (s + len(V))
                if kind[last] == 'stop': done = True
                ip = pos[last] + step[last]
    if next_emit < n: emit_literal(out, inp[next_emit:n], n - next_emit)
    if stats is not None: stats.append((rounds, visits, blocking, blocked_rounds))
    return bytes(out)


if __name__ == "__main__":
    rng = np.random.default_rng(5); cases = []
    for kind in ('az09', 'rand62', 'words'):
        for n in (1, 14, 15, 16, 17, 40, 100, 257, 1000, 1024, 1025, 4096, 20000, 65536):
            cases.append(O.corpus(kind, n, n).tobytes())
    cases += [bytes(65536), b'ab' * 3000, bytes(300) + O.corpus('rand62', 700, 3).tobytes() + bytes(300)]
    for t in range(150):
        n = int(rng.integers(1, 3000)); cases.append(rng.integers(0, int(rng.integers(2, 6)), n, dtype=np.uint8).tobytes())
    for t in range(200):
        cases.append(O.corpus('words', int(rng.integers(1, 5000)), t).tobytes())
    bad = 0
    for d in cases:
        if model_pipe(d) != oracle_frag(d):
            bad += 1
            if bad < 4: print('BAD', len(d), d[:40])
    print(len(cases), 'cases bad', bad)
    st = []
    for t in range(10): model_pipe(O.corpus('words', 65536, t).tobytes(), st)
    print('words 64K: rounds, visits, blocking probes, rounds with blocking', [sum(x[i] for x in st)/len(st) for i in range(4)])
