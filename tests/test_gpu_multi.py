"""b2c_multi_*: one batch cut by bytes over several devices, results in message order.  Runs on ONE GPU too: the
device list may name the same device several times (every entry is a shard with its own ctx and worker thread),
which exercises the split, the re-based block lists and the placement of results exactly like distinct GPUs do;
when the box has more GPUs they are all used."""
import os

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _device_lists():
    import torch
    n = torch.cuda.device_count()
    lists = [[0], [0, 0, 0]]
    if n >= 2:
        lists.append(list(range(n)))
    return lists


def _batch(seed):
    rng = np.random.default_rng(seed)
    ln = np.concatenate([np.full(5000, 1024), rng.integers(0, 70000, 400), np.array([0, 1, 1 << 20, 0, 3 << 20]),
                         np.full(300, 65536)]).astype(np.uint32)
    rng.shuffle(ln)
    return O.corpus_batch("words", ln, first_index=seed * 100)


def test_multi_matches_the_reference_on_every_device_list():
    import brpc_b200
    cpu = O.best()
    threads = os.cpu_count() or 1
    arena, off, ln = _batch(3)
    n = len(ln)
    segs, sb = brpc_b200.make_iovecs(arena, off, ln, seg_bytes=8160)
    want_crc, _ = cpu.batch_crc32c(arena, off, ln, threads=threads)
    w_out, w_off, w_len, _ = cpu.batch_compress(arena, off, ln, threads=threads)
    masked = ((want_crc >> np.uint32(15)) | (want_crc << np.uint32(17))) + np.uint32(0xa282ead8)
    for devices in _device_lists():
        mc = brpc_b200.MultiCodec(devices)
        try:
            assert (mc.crc32c_batch(segs, sb) == want_crc).all()
            init = np.arange(n, dtype=np.uint32) * np.uint32(2654435761)
            got = mc.crc32c_batch(segs, sb, init)
            assert all(int(got[i]) == cpu.crc32c_extend(int(init[i]), arena[int(off[i]):int(off[i]) + int(ln[i])])
                       for i in range(0, n, 97))
            cap = int((32 + ln.astype(np.uint64) + ln.astype(np.uint64) // 6 + 5).sum()) + 256 * len(devices)
            out = np.zeros(cap, dtype=np.uint8)
            o_off, o_len, crc = mc.compress_crc_batch(segs, sb, out, masked=True)
            assert (o_len == w_len).all() and (crc == masked).all()
            bad, first = O.batch_equal(out, o_off, w_out, w_off, w_len)
            assert bad == 0, (devices, bad, first)
            # the receiver's side: the compressed streams as block lists, decoded + verified across the devices
            csegs, csb = brpc_b200.make_iovecs(out, o_off, o_len, seg_bytes=4096)
            back = np.zeros(int(ln.astype(np.uint64).sum()) + 64, dtype=np.uint8)
            b_off, status = mc.decompress_verify_batch(csegs, csb, masked, back)
            assert not status.any() and (np.diff(b_off).astype(np.uint32) == ln).all()
            bad, first = O.batch_equal(back, b_off[:-1], arena, off, ln)
            assert bad == 0, (devices, bad, first)
            wrong = masked.copy()
            wrong[[0, n // 2, n - 1]] ^= np.uint32(4)
            _, status = mc.decompress_verify_batch(csegs, csb, wrong, back)
            assert set(np.nonzero(status)[0].tolist()) == {0, n // 2, n - 1}
            _, status = mc.decompress_verify_batch(csegs, csb, None, back)  # plain decompress
            assert not status.any()
        finally:
            mc.close()


def test_multi_small_and_empty_batches():
    import brpc_b200
    mc = brpc_b200.MultiCodec([0, 0, 0, 0])
    try:
        arena, off, ln = O.corpus_batch("az09", np.array([700, 0], dtype=np.uint32))  # fewer messages than shards
        segs, sb = brpc_b200.make_iovecs(arena, off, ln)
        want, _ = O.best().batch_crc32c(arena, off, ln)
        assert (mc.crc32c_batch(segs, sb) == want).all()
        out = np.zeros(4096, dtype=np.uint8)
        o_off, o_len, _ = mc.compress_crc_batch(segs, sb, out)
        assert out[int(o_off[0]):int(o_off[0]) + int(o_len[0])].tobytes() == O.best().compress(arena[:700])
        assert out[int(o_off[1]):int(o_off[1]) + int(o_len[1])].tobytes() == b"\x00"
        with pytest.raises(brpc_b200.B2CError):
            mc.compress_crc_batch(segs, sb, np.zeros(64, dtype=np.uint8))  # out_cap too small
    finally:
        mc.close()
