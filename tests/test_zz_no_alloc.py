"""GPU test of b2c_crc32c_noalloc_dev, the CRC entry point for callers that must not
allocate device memory on their call path (request-path RPC workers, or services
whose HBM is already fully handed out).  It must give exactly the CRCs of
b2c_crc32c_dev — which are themselves checked against the oracle elsewhere —
on two calls in a row, the second with different data so that whatever the first
call left in the library's fixed scratch array cannot leak into the result.

Kept in a file that sorts last so a run that stops at the first failure has
already run every other test; it also passes when run alone in a fresh process.
"""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _batch(seed):
    rng = np.random.default_rng(seed)
    # multi-chunk messages (these use the scratch slots) mixed with small and empty ones
    lengths = np.concatenate([rng.integers(4097, 300000, 40), rng.integers(0, 4097, 200),
                              np.array([0, 1, 4096, 4097, 8192, 1 << 20, (3 << 20) + 5])]).astype(np.uint32)
    rng.shuffle(lengths)
    assert (lengths > 4096).sum() >= 40  # messages of more than one 4 KiB chunk are the ones that use the scratch array
    return O.corpus_batch(("words", "rand62")[seed % 2], lengths, first_index=seed * 1000)


def test_noalloc_crc_matches_allocating_entry_point_twice():
    import brpc_b200
    codec = brpc_b200.Codec(0)
    try:
        for call, seed in enumerate((1, 2)):
            arena, off, ln = _batch(seed)
            lay = codec.layout(off, ln)          # set-up time: allowed to allocate
            d_data = codec.to_device(arena)
            init = np.random.default_rng(seed).integers(0, 2**32, len(ln), dtype=np.uint32)
            d_init = codec.to_device(init)
            d_ref = codec.alloc(4 * len(ln))
            d_new = codec.alloc(4 * len(ln))
            codec.crc32c_dev(lay, d_data, d_ref, d_init)            # existing entry point
            codec.crc32c_noalloc_dev(lay, d_data, d_new, d_init)    # the one under test
            ref = codec.d2h(d_ref, 4 * len(ln), np.uint32)
            new = codec.d2h(d_new, 4 * len(ln), np.uint32)
            assert (ref == new).all(), (call, np.nonzero(ref != new)[0][:10])
            # and EVERY message, the multi-chunk ones included, agrees with the CPU checker
            cpu = O.best()
            want = np.array([cpu.crc32c_extend(int(init[i]), arena[int(off[i]):int(off[i]) + int(ln[i])])
                             for i in range(len(ln))], dtype=np.uint32)
            assert (new == want).all(), (call, np.nonzero(new != want)[0][:10])
            multi = ln > 4096
            assert multi.any() and (new[multi] == want[multi]).all()
    finally:
        codec.close()


def test_noalloc_crc_refuses_layouts_beyond_its_fixed_scratch():
    import brpc_b200
    codec = brpc_b200.Codec(0)
    try:
        # 300 x 16 MiB would need 1.2M chunk slots (> 2^20); only the layout is built, no data
        n = 300
        ln = np.full(n, 16 << 20, dtype=np.uint32)
        off = np.arange(n, dtype=np.uint64) * np.uint64(16 << 20)
        lay = codec.layout(off, ln)
        d_crc = codec.alloc(4 * n)
        d_fake = codec.alloc(64)
        with pytest.raises(brpc_b200.B2CError, match="MAX_CHUNKS"):
            codec.crc32c_noalloc_dev(lay, d_fake, d_crc)
    finally:
        codec.close()
