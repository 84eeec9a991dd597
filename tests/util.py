"""Shared helpers for the parity tests."""
import numpy as np


def first_diff(a: bytes, b: bytes) -> int:
    n = min(len(a), len(b))
    x = np.frombuffer(a[:n], dtype=np.uint8)
    y = np.frombuffer(b[:n], dtype=np.uint8)
    d = np.nonzero(x != y)[0]
    return int(d[0]) if len(d) else n


def assert_bytes_equal(got: bytes, want: bytes, label: str):
    if got == want:
        return
    i = first_diff(got, want)
    lo = max(0, i - 8)
    raise AssertionError(f"{label}: len got={len(got)} want={len(want)} first diff at {i}: "
                         f"got ..{got[lo:i + 16].hex()} want ..{want[lo:i + 16].hex()}")


def pack(messages, align=16):
    """list of bytes -> (arena u8, off u64[n], len u32[n])"""
    n = len(messages)
    ln = np.array([len(m) for m in messages], dtype=np.uint32)
    padded = (ln.astype(np.uint64) + (align - 1)) // align * align
    off = np.zeros(n, dtype=np.uint64)
    if n:
        off[1:] = np.cumsum(padded[:-1])
    arena = np.zeros(int(padded.sum()) + 64, dtype=np.uint8)
    for i, m in enumerate(messages):
        arena[int(off[i]):int(off[i]) + len(m)] = np.frombuffer(m, dtype=np.uint8)
    return arena, off, ln


def pack_tight(messages):
    return pack(messages, align=1)


def slices(arena, off, ln):
    return [arena[int(o):int(o) + int(l)].tobytes() for o, l in zip(off, ln)]
