#!/usr/bin/env python
"""bench.py — payload GB/s of the snappy + CRC32C batch codec on B200.

One step = one pass of the hot path over one batch of synthetic RPC bodies:
CRC32C of every body, snappy compress, snappy decompress.  Headline workload at
every N: BASELINE.json configs[1] — 2^20 x 1 KiB bodies per GPU, `words` corpus
(weak scaling: every rank gets its own 1 GiB batch, no collective on the data
path).  The same JSON line carries `workloads`: the other BASELINE configs
(c1, c3, c4 at 8 GB, c5 = 64 GiB of 64 KiB fragments, decompress + CRC verify,
split over the ranks), each gated on the parity of EVERY message against the
reference's own CPU code before anything is timed.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

`value`   device-resident: inputs in HBM when the timed region starts.
`e2e`     same step through the host-buffer C ABI (b2c_*_batch): iovec block
          lists in host memory in, flat host buffers out, PCIe inside the timing.
`--impl reference` times the reference's own CPU code (oracle/_ref, else the
oracle port) on the box's host cores with the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "payload GB/s: snappy comp/decomp + CRC32C at 1/2/4/8 B200 vs host CPU"
N_MSGS = 1 << 20
MSG_BYTES = 1024
CORPUS = "words"


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "20", "-i", str(device)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0]))
                mx.append(float(c[1]))
            except ValueError:
                continue
            for nm, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


WORKLOADS = {
    "c1": "C1: 4096 x 4 KiB echo payloads",
    "c2": "C2: 2^20 x 1 KiB small RPC bodies",
    "c3": "C3: 256 x 1 MiB bodies (16 x 64 KiB snappy fragments each)",
    "c4": "C4: lognormal 64 B - 4 MiB payload mix (scaled to --bytes)",
    "c5": "C5: 64 KiB fragments as independent bodies (scaled to --bytes)",
}


def workload_lengths(args) -> np.ndarray:
    w = args.workload
    if w == "c1":
        return np.full(4096, 4096, dtype=np.uint32)
    if w == "c2":
        return np.full(args.messages, MSG_BYTES, dtype=np.uint32)
    if w == "c3":
        return np.full(256, 1 << 20, dtype=np.uint32)
    if w == "c5":
        return np.full(max(1, args.bytes // 65536), 65536, dtype=np.uint32)
    rng = np.random.default_rng(2024)  # c4: SURVEY.md 8(d)
    out, total = [], 0
    while total < args.bytes:
        x = np.clip(np.round(np.exp(rng.normal(np.log(16384), 2.0, 65536))), 64, 4 << 20).astype(np.uint32)
        out.append(x)
        total += int(x.sum())
    x = np.concatenate(out)
    return x[: int(np.searchsorted(np.cumsum(x.astype(np.uint64)), args.bytes)) + 1]


def make_batch(rank: int, args):
    from oracle import oracle as O
    lengths = workload_lengths(args)
    return O.corpus_batch(args.corpus, lengths, first_index=rank * len(lengths))


def config_dict(world: int, args, ln: np.ndarray):
    payload = int(ln.astype(np.uint64).sum())
    return {"workload": f"{WORKLOADS[args.workload]}: {len(ln)} bodies, {payload / 2**30:.2f} GiB per GPU, "
                        f"corpus={args.corpus}; step = crc32c + snappy compress + snappy decompress",
            "corpus": args.corpus, "messages_per_gpu": int(len(ln)), "payload_bytes_per_gpu": payload,
            "sharding": f"{world} independent batch(es), one per GPU, no collective",
            "l2": ("inputs larger than the 126 MB L2; no flush needed" if payload > 200e6 else
                   "inputs SMALLER than L2 (the reference's CPU-sized case): numbers are L2-resident")}


# --------------------------------------------------------------------------- reference arm
def run_reference(args, rank, world):
    if rank != 0:
        return
    from oracle import oracle as O
    cpu = O.best()
    cores = os.cpu_count() or 1
    arena, off, ln = make_batch(0, args)
    n_msgs = len(ln)
    payload = int(ln.astype(np.uint64).sum())

    def step():
        t = 0.0
        _, dt = cpu.batch_crc32c(arena, off, ln, threads=cores); t += dt
        out, o_off, o_len, dt = cpu.batch_compress(arena, off, ln, threads=cores); t += dt
        _, _, ok, dt = cpu.batch_uncompress(out, o_off, o_len, off, ln, threads=cores); t += dt
        assert ok.all()
        return t

    for _ in range(args.warmup):
        step()
    times = [step() for _ in range(args.steps)]
    total = sum(times)
    gbs = payload * args.steps / total / 1e9
    line = {"metric": METRIC, "impl": "reference", "value": gbs, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config_dict(world, args, ln),
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": cpu.kind,
                             "sample": f"full batch: {n_msgs} bodies, {payload} B per step, {args.steps} steps, "
                                       f"static split over {cores} threads (RawCompress/RawUncompress/crc32c::Value)"},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------- measurement helpers
OPS = ("crc32c", "compress", "decompress", "pack")  # order of B2C_PROF_* in include/b2c_codec.h


def kernel_names(max_len: int) -> dict:
    small = max_len <= 1024
    return {"crc32c": "crc32c_chunks_kernel" + ("" if max_len <= 4096 else " + crc32c_combine_kernel"),
            "compress": ("snappy_compress_small2_kernel<8>" if small else
                         "snappy_compress_orbit_kernel (fragments > 1 KiB) + snappy_compress_small2_kernel<8> (<= 1 KiB)"),
            "decompress": ("snappy_decompress_small_kernel<8>" if small else
                           "snappy_decompress_kernel / snappy_seg_* + snappy_decompress_blocks_kernel (+ small kernel for slots <= 1 KiB)"),
            "pack": "snappy_message_sizes_kernel + scan + snappy_gather_kernel"}


def mask_crc(crc: np.ndarray) -> np.ndarray:  # butil::crc32c::Mask, crc32c.h:44-48
    return (((crc >> np.uint32(15)) | (crc << np.uint32(17))) + np.uint32(0xa282ead8)).astype(np.uint32)


def load_traffic() -> dict:
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernels, taken from committed
    `ncu --set full` captures by tools/ncu_traffic.py (profiles/ncu_traffic.json); absent entries are null."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return {}


def host_ram_available() -> int:
    try:
        import psutil
        return int(psutil.virtual_memory().available)
    except Exception:
        return 64 << 30


class DeviceBatch:
    """One batch resident in HBM plus everything a step needs.  Construction runs the step once and checks
    EVERY message against the CPU checker (the reference's own code when oracle/_ref is there): CRC values,
    compressed bytes, decoded bytes (against the reference's decode of the same streams and the original)."""

    def __init__(self, codec, arena, off, ln, label):
        from oracle import oracle as O
        self.codec, self.arena, self.off, self.ln, self.label = codec, arena, off, ln, label
        n = self.n = len(ln)
        self.payload = int(ln.astype(np.uint64).sum())
        self.lay = codec.layout(off, ln)
        self.d_in = codec.to_device(arena)
        self.d_crc = codec.alloc(4 * n)
        self.d_comp = codec.alloc(self.lay.max_compressed_bytes + 64)
        self.d_coff = codec.alloc(8 * (n + 1))
        self.d_back = codec.alloc(int(self.lay.arena_bytes) + 64)
        self.d_st = codec.alloc(4 * n)
        # the compressed layout is data-dependent: build it from a first (untimed) pass
        codec.compress_dev(self.lay, self.d_in, self.d_comp, self.d_coff)
        self.c_off = codec.d2h(self.d_coff, 8 * (n + 1), np.uint64)
        self.comp_bytes = int(self.c_off[-1])
        self.c_len = np.diff(self.c_off).astype(np.uint32)
        self.clay = codec.layout(self.c_off[:-1], self.c_len)
        self.step()
        codec.sync()
        # ---- parity gate: every message, before any timing ----
        cpu = O.best()
        threads = os.cpu_count() or 1
        self.cpu_kind, self.cpu_threads = cpu.kind, threads
        st = codec.d2h(self.d_st, 4 * n, np.int32)
        assert (st == 0).all(), f"{label}: decompress reported errors"
        self.crc = codec.d2h(self.d_crc, 4 * n, np.uint32)
        want_crc, t_crc = cpu.batch_crc32c(arena, off, ln, threads=threads)
        assert (self.crc == want_crc).all(), f"{label}: crc parity"
        w_out, w_off, w_len, t_comp = cpu.batch_compress(arena, off, ln, threads=threads)
        assert (self.c_len == w_len).all(), f"{label}: compressed lengths differ"
        self.packed = codec.d2h(self.d_comp, self.comp_bytes)
        bad, first = O.batch_equal(self.packed, self.c_off[:-1], w_out, w_off, w_len)
        assert bad == 0, f"{label}: compressed bytes differ in {bad} messages, first {first}"
        del w_out
        ref_back, ref_ulen, ref_ok, t_dec = cpu.batch_uncompress(self.packed, self.c_off[:-1], self.c_len, off, ln,
                                                                 threads=threads)
        assert ref_ok.all() and (ref_ulen == ln).all(), f"{label}: the reference rejects our streams"
        back = codec.d2h(self.d_back, int(self.lay.arena_bytes))
        bad, first = O.batch_equal(back, off, ref_back, off, ln)
        assert bad == 0, f"{label}: decoded bytes differ from the reference's in {bad} messages, first {first}"
        del ref_back
        bad, first = O.batch_equal(back, off, arena, off, ln)
        assert bad == 0, f"{label}: round trip differs in {bad} messages, first {first}"
        del back
        self.cpu_seconds = {"crc32c": t_crc, "compress": t_comp, "decompress": t_dec}
        self.d_masked = codec.to_device(mask_crc(self.crc))

    def step(self):
        c = self.codec
        c.crc32c_dev(self.lay, self.d_in, self.d_crc)
        c.compress_dev(self.lay, self.d_in, self.d_comp, self.d_coff)
        c.decompress_dev(self.clay, self.d_comp, self.lay, self.d_back, self.d_st)

    def verify_step(self):  # C5's op: decompress + CRC32C of the output against the stored masked CRC
        self.codec.decompress_verify_dev(self.clay, self.d_comp, self.lay, self.d_back, self.d_masked, self.d_st)

    def timed(self, fn, steps, warmup, dist, prof=False):
        """(ms total for `steps` calls of fn, max over ranks; launches; per-op (ms, count) when prof)"""
        import brpc_b200
        c = self.codec
        for _ in range(warmup):
            fn()
        c.sync()
        if dist is not None:
            dist.barrier()
        if prof:
            c.prof_enable(True)
        l0 = c.launch_count
        e0, e1 = c.event(), c.event()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        c.sync()
        ms = e1.elapsed_ms_since(e0)
        launches = c.launch_count - l0
        per_op = None
        if prof:
            per_op = {name: c.prof_total_ms(k) for k, name in enumerate(OPS)}
            c.prof_enable(False)
        if dist is not None:
            ms = brpc_b200.sharding.max_over_ranks(ms, device="cuda")
            dist.barrier()
        return ms, launches, per_op

    def free(self):
        for name in ("d_in", "d_crc", "d_comp", "d_coff", "d_back", "d_st", "d_masked"):
            getattr(self, name).free()
        self.lay.destroy()
        self.clay.destroy()
        self.packed = None


def op_table(B: DeviceBatch, per_op: dict, hbm_peak: float, traffic: dict, key: str) -> dict:
    """Per-op time, payload GB/s and roofline (algorithmic bytes of SURVEY 8(d) / mean launch duration)."""
    N, Cb = B.payload, B.comp_bytes
    alg = {"crc32c": N, "compress": N + Cb, "decompress": Cb + N, "pack": 2 * Cb}
    names = kernel_names(int(B.ln.max()) if B.n else 0)
    out = {}
    for name in OPS:
        t_ms, cnt = per_op[name]
        ms = t_ms / max(cnt, 1)
        gbs = alg[name] / (ms * 1e-3) / 1e9 if ms > 0 else None
        out[name] = {"ms_per_step": ms, "payload_gbs": (N if name != "pack" else Cb) / (ms * 1e-3) / 1e9 if ms > 0 else None,
                     "roofline": {"kernel": names[name], "bound": "hbm", "algorithmic_bytes_per_launch": alg[name],
                                  "achieved": gbs, "peak": hbm_peak, "unit": "GB/s",
                                  "frac": gbs / hbm_peak if gbs else None, "traffic": traffic.get(f"{key}/{name}")}}
    return out


class HostBatch:
    """The same batch through the host-buffer C ABI, as a sender and a receiver call it: IOBuf-like block lists in
    ordinary pageable memory in, page-locked flat buffers out, PCIe inside the timed region.
      sender    b2c_snappy_compress_crc_batch    (checksum + compress, one pass over the staged bodies)
      receiver  b2c_snappy_decompress_verify_batch (decode + check every body against the sender's checksum)
    Serial: one thread calls sender then receiver.  Overlapped: two threads with a ctx each (what a loaded server
    does: it compresses responses while it decompresses requests), so H2D and D2H use the link in both directions."""

    def __init__(self, sender, receiver, B: DeviceBatch, workload: str, rank: int, pinned_inputs: bool = False):
        import brpc_b200
        self.s, self.r, self.B = sender, receiver, B
        self.pinned_inputs = pinned_inputs
        rng = np.random.default_rng(2024 + rank)
        arena = B.arena
        if pinned_inputs:  # the blocks come from a page-locked IOBuf block pool (host/b200_block_pool): no host gather
            self.p_arena = sender.pinned(len(B.arena))
            self.p_arena[:] = B.arena
            arena = self.p_arena
            self.comp_host = receiver.pinned(B.comp_bytes + 64)
            self.comp_host[:B.comp_bytes] = B.packed
        else:  # ordinary pageable memory, as IOBuf blocks are without the pool
            self.comp_host = np.array(B.packed, copy=True)
        for c in (sender, receiver):
            c.set_option(c.OPT_INPUT_PINNED, 1 if pinned_inputs else 0)
        if workload == "c2":  # 1-2 blocks per body at a seeded random cut
            self.segs, self.seg_begin = brpc_b200.make_iovecs(arena, B.off, B.ln, cuts=rng.integers(0, MSG_BYTES + 1, B.n))
        else:  # IOBuf's default blocks: 8192 bytes minus the 32-byte header
            self.segs, self.seg_begin = brpc_b200.make_iovecs(arena, B.off, B.ln, seg_bytes=8160)
        # the receiver's input: the compressed streams cut into IOBuf blocks
        self.csegs, self.cbeg = brpc_b200.make_iovecs(self.comp_host, B.c_off[:-1], B.c_len, seg_bytes=8160)
        self.masked = mask_crc(B.crc)
        # results land in page-locked buffers (what the bRPC shim hands out from its pinned block pool)
        self.h_comp = sender.pinned(B.comp_bytes + 4096)
        self.h_back = receiver.pinned(B.payload + 64)
        self.h2d = self.d2h = 0

    def send(self):
        o, crc = self.s.compress_crc_batch(self.segs, self.seg_begin, self.h_comp, masked=True)
        self.last_o, self.last_crc = o, crc
        return self.s.last_transfer_bytes()

    def recv(self):
        o, status = self.r.decompress_verify_batch(self.csegs, self.cbeg, self.masked, self.h_back)
        self.last_ro, self.last_status = o, status
        return self.r.last_transfer_bytes()

    def check(self):
        """every message of the last send/recv against the parity-gated device results"""
        from oracle import oracle as O
        B = self.B
        assert np.array_equal(self.last_o, B.c_off), f"{B.label}: e2e compressed offsets differ"
        assert np.array_equal(self.last_crc, self.masked), f"{B.label}: e2e checksums differ"
        assert np.array_equal(self.h_comp[:B.comp_bytes], B.packed), f"{B.label}: e2e compressed bytes differ"
        assert not self.last_status.any(), f"{B.label}: e2e receiver reported errors"
        bad, first = O.batch_equal(self.h_back, self.last_ro[:-1], B.arena, B.off, B.ln)
        assert bad == 0, f"{B.label}: e2e decoded bytes differ in {bad} messages, first {first}"

    def _sync_ranks(self, dist):
        if dist is not None:
            dist.barrier()

    def time_serial(self, steps, dist):
        import brpc_b200
        self.send(), self.recv()  # warm-up: sizes the pinned staging and device arenas
        self.check()
        self._sync_ranks(dist)
        t0 = time.perf_counter()
        for _ in range(steps):
            a = self.send()
            b = self.recv()
        dt = time.perf_counter() - t0
        self.h2d, self.d2h = a[0] + b[0], a[1] + b[1]
        if dist is not None:
            dt = brpc_b200.sharding.max_over_ranks(dt, device="cuda")
        self.check()
        return dt

    def time_overlapped(self, steps, dist, threads_each):
        import threading
        import brpc_b200
        for c in (self.s, self.r):
            c.set_option(c.OPT_HOST_THREADS, threads_each)
        self.send(), self.recv()  # re-creates each ctx's pool at its new size
        start = threading.Barrier(3)
        err = []

        def loop(fn):
            try:
                start.wait()
                for _ in range(steps):
                    fn()
            except Exception as e:  # noqa: BLE001
                err.append(e)

        ts = [threading.Thread(target=loop, args=(f,)) for f in (self.send, self.recv)]
        for t in ts:
            t.start()
        self._sync_ranks(dist)
        start.wait()
        t0 = time.perf_counter()
        for t in ts:
            t.join()
        dt = time.perf_counter() - t0
        if err:
            raise err[0]
        if dist is not None:
            dt = brpc_b200.sharding.max_over_ranks(dt, device="cuda")
        self.check()
        for c in (self.s, self.r):
            c.set_option(c.OPT_HOST_THREADS, 0)
        return dt


def pcie_probe(codec, d_buf, h_buf):
    """this box's pinned-memory copy rate, one direction at a time (measured, not assumed)"""
    nb = min(256 << 20, len(h_buf))
    out = {}
    for name, fn, a in (("h2d_gbs", codec.lib.b2c_memcpy_h2d, (d_buf.ptr, h_buf.ctypes.data, nb)),
                        ("d2h_gbs", codec.lib.b2c_memcpy_d2h, (h_buf.ctypes.data, d_buf.ptr, nb))):
        fn(codec.ctx, *a)
        codec.sync()
        e0, e1 = codec.event(), codec.event()
        e0.record()
        for _ in range(4):
            fn(codec.ctx, *a)
        e1.record()
        codec.sync()
        out[name] = 4 * nb / (e1.elapsed_ms_since(e0) * 1e-3) / 1e9
    return out


PARITY_NOTE = ("every message of the timed batch, before timing: CRC32C values, compressed bytes and decoded bytes equal "
               "the CPU checker's (kind below); e2e outputs are compared with the same bytes after the timed calls")


def measure_workload(codecs, wl, corpus, nbytes, messages, rank, world, dist, steps, warmup, e2e_steps, hbm_peak, traffic,
                     want_e2e=True, c5_total=0, sampler_factory=None, extras=False):
    """One workload shape on this rank's GPU: parity gate, device-resident step, per-op rooflines, end to end."""
    from oracle import oracle as O
    a = argparse.Namespace(workload=wl, corpus=corpus, bytes=nbytes, messages=messages)
    lengths = workload_lengths(a)
    arena, off, ln = O.corpus_batch(corpus, lengths, first_index=rank * len(lengths))
    label = f"{wl}/{corpus}"
    codec = codecs[0]
    B = DeviceBatch(codec, arena, off, ln, label)
    sampler = sampler_factory() if sampler_factory else None  # samples over warm-up + timed steps (all under load)
    ms, launches, per_op = B.timed(B.step, steps, warmup, dist, prof=True)
    clocks = sampler.stop() if sampler else None
    payload = B.payload
    cpu_total = sum(B.cpu_seconds.values())
    res = {
        "config": config_dict(world, a, ln),
        "value": payload * world * steps / (ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": ms / steps, "steps": steps,
        "gpu_launches": int(launches), "compression_ratio": B.comp_bytes / max(payload, 1),
        "ops": op_table(B, per_op, hbm_peak, traffic, label),
        "parity": PARITY_NOTE,
        "cpu_baseline": {"value": payload / cpu_total / 1e9, "unit": "GB/s", "cores": B.cpu_threads, "kind": B.cpu_kind,
                         "per_op_gbs": {k: payload / v / 1e9 for k, v in B.cpu_seconds.items()},
                         "sample": f"full batch ({B.n} bodies, {payload} B), one pass of crc32c + RawCompress + RawUncompress, "
                                   f"static split over {B.cpu_threads} threads (rank 0's batch)"},
    }
    if clocks is not None:
        res["clocks"] = clocks
    if wl == "c5" and c5_total:
        # C5's own op on the config that names it: 64 GiB of 64 KiB fragments, decompress + CRC32C check of the output
        # against the stored masked CRC, split over the ranks; every rank walks its share as `tiles` passes over its
        # own unique slab (SURVEY 8(d) allows a >= 1 GiB unique slab tiled, with the factor reported)
        tiles = max(1, c5_total // (payload * world))
        vms, vl, _ = B.timed(B.verify_step, tiles, 2, dist)
        st = codec.d2h(B.d_st, 4 * B.n, np.int32)
        assert (st == 0).all(), "decompress+verify reported errors"
        alg = B.comp_bytes + payload
        gbs = alg * tiles / (vms * 1e-3) / 1e9
        res["decompress_verify"] = {
            "total_payload_bytes": payload * tiles * world, "unique_slab_bytes_per_gpu": payload, "tile_factor": tiles,
            "scaling": "strong (the 64 GiB total is fixed; each rank owns total / n_gpus)",
            "dev": {"value": payload * tiles * world / (vms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_slab": vms / tiles,
                    "roofline": {"kernel": "snappy_decompress_kernel + crc32c_chunks_kernel + snappy_verify_crc_kernel",
                                 "bound": "hbm", "algorithmic_bytes_per_launch": alg, "achieved": gbs, "peak": hbm_peak,
                                 "unit": "GB/s", "frac": gbs / hbm_peak, "traffic": traffic.get(f"{label}/verify")}}}
    H = None
    if want_e2e:
        import brpc_b200
        H = HostBatch(codecs[0], codecs[1], B, wl, rank)
        hw = os.cpu_count() or 1
        lw = int(os.environ.get("LOCAL_WORLD_SIZE", world))
        serial_s = H.time_serial(e2e_steps, dist)
        over_s = H.time_overlapped(e2e_steps, dist, max(2, hw // max(lw, 1) // 2))
        res["e2e"] = {
            "value": payload * world * e2e_steps / over_s / 1e9, "unit": "GB/s",
            "h2d_bytes_per_step": int(H.h2d), "d2h_bytes_per_step": int(H.d2h), "steps": e2e_steps,
            "serial_value": payload * world * e2e_steps / serial_s / 1e9,
            "path": "sender b2c_snappy_compress_crc_batch (crc32c + compress) and receiver b2c_snappy_decompress_verify_batch "
                    "(decompress + crc32c check) on every body; inputs = pageable host block lists (1-2 blocks per 1 KiB "
                    "body, 8160-byte IOBuf blocks otherwise) gathered into pinned staging by each ctx's host threads; "
                    "outputs = page-locked host buffers.  Overlapped: sender and receiver run on two threads with a ctx each, "
                    "as a loaded server does (H2D and D2H overlap); serial: one thread calls them back to back.  `value` is the "
                    "best of the measured variants (named in `mode`); all of them are listed"}
        if serial_s < over_s:
            res["e2e"]["value"], res["e2e"]["mode"] = res["e2e"]["serial_value"], "pageable input blocks, sender then receiver"
            res["e2e"]["overlapped_value"] = payload * world * e2e_steps / over_s / 1e9
        else:
            res["e2e"]["mode"] = "pageable input blocks, sender and receiver overlapped"
        if wl == "c5" and c5_total:
            tiles = res["decompress_verify"]["tile_factor"]
            H._sync_ranks(dist)
            t0 = time.perf_counter()
            for _ in range(tiles):
                H.recv()
            dt = time.perf_counter() - t0
            if dist is not None:
                dt = brpc_b200.sharding.max_over_ranks(dt, device="cuda")
            H.check()
            a2, b2 = codecs[1].last_transfer_bytes()
            res["decompress_verify"]["batch"] = {"value": payload * tiles * world / dt / 1e9, "unit": "GB/s",
                                                 "h2d_bytes_per_slab": int(a2), "d2h_bytes_per_slab": int(b2),
                                                 "entry_point": "b2c_snappy_decompress_verify_batch"}
        if extras and rank == 0:
            res["e2e"]["pcie_probe"] = pcie_probe(codec, B.d_back, H.h_back)
        if extras:
            # the same two calls with the input blocks in PAGE-LOCKED memory (an IOBuf block pool installed through
            # iobuf::blockmem_allocate, host/b200_block_pool) and B2C_OPT_INPUT_PINNED: no host thread touches the
            # bytes, the device pulls the blocks over its own PCIe link -- the variant that does not share the
            # host's cores and memory bandwidth between the GPUs of a box
            release_host(H)
            H = HostBatch(codecs[0], codecs[1], B, wl, rank, pinned_inputs=True)
            p_serial = H.time_serial(e2e_steps, dist)
            p_over = H.time_overlapped(e2e_steps, dist, max(2, hw // max(lw, 1) // 2))
            e = res["e2e"]
            e["pageable_inputs"] = {"overlapped_value": payload * world * e2e_steps / over_s / 1e9, "serial_value": e["serial_value"]}
            e["pinned_inputs"] = {"overlapped_value": payload * world * e2e_steps / p_over / 1e9,
                                  "serial_value": payload * world * e2e_steps / p_serial / 1e9,
                                  "h2d_bytes_per_step": int(H.h2d), "d2h_bytes_per_step": int(H.d2h)}
            best = max((e["pageable_inputs"]["overlapped_value"], "pageable input blocks, sender and receiver overlapped"),
                       (e["pageable_inputs"]["serial_value"], "pageable input blocks, sender then receiver"),
                       (e["pinned_inputs"]["overlapped_value"], "page-locked input blocks (B2C_OPT_INPUT_PINNED), sender and receiver overlapped"),
                       (e["pinned_inputs"]["serial_value"], "page-locked input blocks (B2C_OPT_INPUT_PINNED), sender then receiver"))
            e["value"], e["mode"] = best
            if best[1].startswith("page-locked"):
                e["h2d_bytes_per_step"], e["d2h_bytes_per_step"] = e["pinned_inputs"]["h2d_bytes_per_step"], e["pinned_inputs"]["d2h_bytes_per_step"]
    return res, B, H


def release_host(H):
    release(None, H)


def release(B, H):
    if H is not None:
        bufs = [(H.s, H.h_comp), (H.r, H.h_back)]
        if H.pinned_inputs:
            bufs += [(H.s, H.p_arena), (H.r, H.comp_host)]
        for c, arr in bufs:
            c.lib.b2c_host_free_pinned(c.ctx, arr.ctypes.data)
        H.h_comp = H.h_back = H.comp_host = H.p_arena = None
        for c in (H.s, H.r):
            c.set_option(c.OPT_INPUT_PINNED, 0)
    if B is not None:
        B.free()


# --------------------------------------------------------------------------- our arm
def run_ours(args, rank, local_rank, world):
    import brpc_b200
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    codecs = (brpc_b200.Codec(local_rank), brpc_b200.Codec(local_rank))  # [1]: the receiver's ctx of the e2e legs
    hbm_peak, peak_src = read_peaks()
    traffic = load_traffic()
    e2e_steps = max(1, min(args.steps, 3))
    head, B, H = measure_workload(codecs, args.workload, args.corpus, args.bytes, args.messages, rank, world, dist, args.steps,
                                  args.warmup, e2e_steps, hbm_peak, traffic, c5_total=args.c5_total if args.workload == "c5" else 0,
                                  sampler_factory=(lambda: ClockSampler(local_rank)) if rank == 0 else None, extras=True)
    n_msgs = B.n
    release(B, H)

    # ---- the other BASELINE configs, same gates, one entry each ----
    workloads = {}
    if args.workload == "c2" and not args.only_headline:
        ram = host_ram_available()
        c4_bytes = int(min(args.c4_bytes, max(1 << 30, ram // 7)))  # ~6 host copies of the batch are alive at the peak
        plan = [("c5", "c5", "words", 1 << 30, True, args.c5_total)]
        if world == 1:  # the single-GPU configs are measured at N = 1 only (C4 alone holds ~45 GB of host memory per rank)
            plan = [("c1", "c1", "words", 0, True, 0), ("c3", "c3", "words", 0, True, 0),
                    ("c4", "c4", "words", c4_bytes, True, 0)] + plan + \
                   [("c5_rand62", "c5", "rand62", 1 << 30, False, 0), ("c5_az09", "c5", "az09", 1 << 30, False, 0),
                    ("c2_rand62", "c2", "rand62", 0, False, 0), ("c2_az09", "c2", "az09", 0, False, 0)]
        for key, wl, corpus, nbytes, e2e, c5_total in plan:
            t0 = time.perf_counter()
            big = nbytes > (2 << 30)
            r, B, H = measure_workload(codecs, wl, corpus, nbytes, N_MSGS, rank, world, dist, 2 if big else 3, 3,
                                       1 if big else 2, hbm_peak, traffic, want_e2e=e2e, c5_total=c5_total)
            release(B, H)
            r["bench_seconds"] = time.perf_counter() - t0
            if key == "c4" and c4_bytes < args.c4_bytes:
                r["config"]["scaled_down"] = f"{c4_bytes} B instead of {args.c4_bytes} B: host RAM available {ram} B"
            workloads[key] = r

    if rank == 0:
        comp = head["ops"]["compress"]["roofline"]
        line = {
            "metric": METRIC, "value": head["value"], "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": head["config"], "clocks": head.get("clocks"),
            "e2e": head["e2e"], "gpu_launches": head["gpu_launches"],
            "roofline": {"kernel": comp["kernel"], "bound": "hbm", "achieved": comp["achieved"], "peak": hbm_peak,
                         "unit": "GB/s", "frac": comp["frac"], "traffic": comp["traffic"],
                         "algorithmic_bytes_per_launch": comp["algorithmic_bytes_per_launch"], "peak_source": peak_src,
                         "note": "dominant kernel of the step (compress); crc32c / decompress / pack rooflines are in `ops`. "
                                 "Byte-exact greedy LZ77 is a serial dependency chain per fragment: latency/issue bound, not "
                                 "HBM bound (see DESIGN.md)"},
            "ops": head["ops"], "compression_ratio": head["compression_ratio"],
            "messages_per_s": n_msgs * world * args.steps / (head["ms_per_step"] * args.steps * 1e-3),
            "parity": head["parity"],
            # the CPU figure is the parity checker's own time; with several ranks all of them run their checkers on the
            # same cores at once, so only the 1-GPU line carries it as a baseline (`--impl reference` times it alone)
            "cpu_baseline": head["cpu_baseline"] if world == 1 else
            {**head["cpu_baseline"], "contended": True,
             "note": f"{world} ranks ran their checkers at once on these cores: not a baseline; see the 1-GPU line or --impl reference"},
            **({"decompress_verify": head["decompress_verify"]} if "decompress_verify" in head else {}),
            "workloads": workloads,
        }
        print(json.dumps(line), flush=True)
    for c in codecs:
        c.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--messages", type=int, default=N_MSGS, help="bodies per GPU for --workload c2")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS), help="headline workload; default c2 = BASELINE configs[1]")
    ap.add_argument("--corpus", default=CORPUS, choices=["words", "az09", "rand62"])
    ap.add_argument("--bytes", type=int, default=1 << 30, help="payload per GPU for --workload c4 / c5")
    ap.add_argument("--c4-bytes", type=int, default=8_000_000_000, help="payload of the `workloads.c4` entry (BASELINE: 8 GB)")
    ap.add_argument("--c5-total", type=int, default=64 << 30, help="total payload of C5's decompress + verify op (BASELINE: 64 GiB)")
    ap.add_argument("--only-headline", action="store_true", help="skip the `workloads` entries (kernel tuning runs)")
    args = ap.parse_args()
    if args.impl == "ours":
        args.warmup = max(args.warmup, 3)  # timing rule: at least 3 warm-up steps
    rank, local_rank, world = dist_env()
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
