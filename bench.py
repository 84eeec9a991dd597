#!/usr/bin/env python
"""bench.py — payload GB/s of the snappy + CRC32C batch codec on B200.

One step = one pass of the hot path over one batch of synthetic RPC bodies:
CRC32C of every body, snappy compress, snappy decompress.  Workload at every N:
BASELINE.json configs[1] — 2^20 x 1 KiB bodies per GPU, `words` corpus (weak
scaling: every rank gets its own 1 GiB batch, no collective on the data path).

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


# --------------------------------------------------------------------------- our arm
def run_ours(args, rank, local_rank, world):
    import brpc_b200
    from oracle import oracle as O
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    codec = brpc_b200.Codec(local_rank)
    arena, off, ln = make_batch(rank, args)
    n_msgs = len(ln)
    payload = int(ln.astype(np.uint64).sum())
    lay = codec.layout(off, ln)
    d_in = codec.to_device(arena)
    d_crc = codec.alloc(4 * n_msgs)
    d_comp = codec.alloc(lay.max_compressed_bytes + 64)
    d_coff = codec.alloc(8 * (n_msgs + 1))
    d_back = codec.alloc(int(lay.arena_bytes) + 64)
    d_st = codec.alloc(4 * n_msgs)

    # the compressed layout is data-dependent: build it once from a first (untimed) pass
    codec.compress_dev(lay, d_in, d_comp, d_coff)
    c_off = codec.d2h(d_coff, 8 * (n_msgs + 1), np.uint64)
    comp_bytes = int(c_off[-1])
    clay = codec.layout(c_off[:-1], np.diff(c_off).astype(np.uint32))

    def step():
        codec.crc32c_dev(lay, d_in, d_crc)
        codec.compress_dev(lay, d_in, d_comp, d_coff)
        codec.decompress_dev(clay, d_comp, lay, d_back, d_st)

    # ---- parity gate before any timing (sampled against the CPU checker) ----
    step()
    codec.sync()
    st = codec.d2h(d_st, 4 * n_msgs, np.int32)
    assert (st == 0).all(), "decompress reported errors"
    cpu = O.best()
    idx = np.arange(0, n_msgs, max(1, n_msgs // 2048))
    crc = codec.d2h(d_crc, 4 * n_msgs, np.uint32)
    want_crc, _ = cpu.batch_crc32c(arena, off[idx], ln[idx], threads=4)
    assert (crc[idx] == want_crc).all(), "crc parity"
    packed = codec.d2h(d_comp, comp_bytes)
    w_out, w_off, w_len, _ = cpu.batch_compress(arena, off[idx], ln[idx], threads=4)
    for j, i in enumerate(idx):
        assert packed[int(c_off[i]):int(c_off[i + 1])].tobytes() == \
            w_out[int(w_off[j]):int(w_off[j]) + int(w_len[j])].tobytes(), f"compress parity at message {i}"
    back = codec.d2h(d_back, int(lay.arena_bytes))
    assert all(back[int(off[i]):int(off[i]) + int(ln[i])].tobytes() ==
               arena[int(off[i]):int(off[i]) + int(ln[i])].tobytes() for i in idx), "round-trip parity"
    del packed, back

    sampler = ClockSampler(local_rank) if rank == 0 else None  # samples over warm-up + timed steps (all under load)
    for _ in range(args.warmup):
        step()
    codec.sync()
    if dist is not None:
        dist.barrier()
    codec.prof_enable(True)
    launches0 = codec.launch_count
    e0, e1 = codec.event(), codec.event()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    codec.sync()
    ms = e1.elapsed_ms_since(e0)
    launches = codec.launch_count - launches0
    prof = {name: codec.prof_total_ms(k) for k, name in enumerate(("crc32c", "compress", "decompress", "pack"))}
    codec.prof_enable(False)
    clocks = sampler.stop() if sampler else None
    if dist is not None:
        ms = brpc_b200.sharding.max_over_ranks(ms, device="cuda")  # device time, max over ranks
        dist.barrier()

    # ---- the fused receiver op on device-resident data (decompress + CRC32C of the output compared
    #      with the stored masked CRC, snappy-framing semantics), timed on its own ----
    verify_ms = None
    if args.workload == "c5":
        crc_host = codec.d2h(d_crc, 4 * n_msgs, np.uint32)
        masked = ((crc_host >> np.uint32(15)) | (crc_host << np.uint32(17))) + np.uint32(0xa282ead8)  # crc32c.h:44-48
        d_masked = codec.to_device(masked.astype(np.uint32))
        codec.decompress_verify_dev(clay, d_comp, lay, d_back, d_masked, d_st)
        codec.sync()
        assert (codec.d2h(d_st, 4 * n_msgs, np.int32) == 0).all(), "decompress+verify reported errors"
        v0, v1 = codec.event(), codec.event()
        v0.record()
        for _ in range(args.steps):
            codec.decompress_verify_dev(clay, d_comp, lay, d_back, d_masked, d_st)
        v1.record()
        codec.sync()
        verify_ms = v1.elapsed_ms_since(v0) / args.steps
        if dist is not None:
            verify_ms = brpc_b200.sharding.max_over_ranks(verify_ms, device="cuda")

    # ---- end to end through the host-buffer ABI (iovec block lists, PCIe inside) ----
    rng = np.random.default_rng(2024 + rank)
    if args.workload == "c2":  # 1-2 blocks per body at a seeded random cut
        segs, seg_begin = brpc_b200.make_iovecs(arena, off, ln, cuts=rng.integers(0, MSG_BYTES + 1, n_msgs))
    else:  # IOBuf's default blocks: 8192 bytes minus the 32-byte header
        segs, seg_begin = brpc_b200.make_iovecs(arena, off, ln, seg_bytes=8160)
    # result buffers are page-locked (b2c_host_alloc_pinned), as the bRPC shim's output blocks would be, so the
    # library DMAs results straight into them; the INPUT blocks stay ordinary pageable memory like IOBuf blocks
    h_comp = codec.pinned(int(lay.max_compressed_bytes))
    h_back = codec.pinned(payload + 64)
    e2e_steps = max(1, min(args.steps, 3))
    bytes_h2d = bytes_d2h = 0

    cache = {}

    def e2e_step(fused=True):
        nonlocal bytes_h2d, bytes_d2h
        h2d = d2h = 0
        if fused:  # sender: checksum + compress in one pass over the staged bodies
            o, crc = codec.compress_crc_batch(segs, seg_begin, h_comp, masked=True)
            a, b = codec.last_transfer_bytes(); h2d += a; d2h += b
        else:
            codec.crc32c_batch(segs, seg_begin)
            a, b = codec.last_transfer_bytes(); h2d += a; d2h += b
            o = codec.compress_batch(segs, seg_begin, h_comp)
            a, b = codec.last_transfer_bytes(); h2d += a; d2h += b
        if "o" not in cache or o[-1] != cache["o"][-1]:
            # iovecs of the compressed bodies (what an IOBuf holding them would report); built once —
            # the streams are a pure function of the inputs, so they land at the same offsets every step
            csegs = np.empty((n_msgs, 2), dtype=np.uint64)
            csegs[:, 0] = h_comp.ctypes.data + o[:-1]
            csegs[:, 1] = np.diff(o)
            cache["o"], cache["csegs"] = o, csegs
            cache["cbeg"] = np.arange(n_msgs + 1, dtype=np.uint32)
        if fused:  # receiver: decompress + verify each body against the sender's checksum
            _, status = codec.decompress_verify_batch(cache["csegs"], cache["cbeg"], crc, h_back)
        else:
            _, status = codec.decompress_batch(cache["csegs"], cache["cbeg"], h_back)
        a, b = codec.last_transfer_bytes(); h2d += a; d2h += b
        assert not status.any()
        bytes_h2d, bytes_d2h = h2d, d2h

    def time_e2e(fused=True):
        nonlocal bytes_h2d, bytes_d2h
        e2e_step(fused)  # warm-up: sizes the pinned staging and device arenas
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step(fused)
        dt = time.perf_counter() - t0
        if dist is not None:
            dt = brpc_b200.sharding.max_over_ranks(dt, device="cuda")
        return dt, bytes_h2d, bytes_d2h

    # what bounds the end-to-end path: this box's pinned-memory copy rate, one direction at a time (SURVEY 8(d):
    # measured in the same run, not assumed; tools/pcie_bw.py also times both directions at once)
    def pcie_probe():
        nb = min(256 << 20, len(h_back))
        out = {}
        for name, fn, args in (("h2d_gbs", codec.lib.b2c_memcpy_h2d, (d_back.ptr, h_back.ctypes.data, nb)),
                               ("d2h_gbs", codec.lib.b2c_memcpy_d2h, (h_back.ctypes.data, d_back.ptr, nb))):
            fn(codec.ctx, *args)
            codec.sync()
            a, b = codec.event(), codec.event()
            a.record()
            for _ in range(4):
                fn(codec.ctx, *args)
            b.record()
            codec.sync()
            out[name] = 4 * nb / (b.elapsed_ms_since(a) * 1e-3) / 1e9
        return out

    pcie = pcie_probe() if rank == 0 else None
    # (a) the headline: input blocks in ordinary pageable memory (what IOBuf blocks are today); host threads gather
    #     them into pinned staging and the copy engines move the slabs
    e2e_s, bytes_h2d, bytes_d2h = time_e2e()
    # (a') the same work as three separate calls (crc32c_batch, compress_batch, decompress_batch): the bodies cross
    #      PCIe twice on the sender side and nothing is verified on the receiver side
    e2e_separate_s, _, _ = time_e2e(fused=False)
    # (b) for the record: blocks in page-locked memory + B2C_OPT_INPUT_PINNED (the device gathers them over PCIe
    #     itself).  On this box kernel reads of host memory reach ~35 GB/s against ~55 GB/s for the copy engines,
    #     so (a) stays the default path.
    p_arena = codec.pinned(len(arena))
    p_arena[:] = arena
    segs_pageable = segs.copy()
    segs[:, 0] = segs[:, 0] - np.uint64(arena.ctypes.data) + np.uint64(p_arena.ctypes.data)
    cache.clear()
    codec.set_option(codec.OPT_INPUT_PINNED, 1)
    e2e_zero_copy_s, _, _ = time_e2e()
    codec.set_option(codec.OPT_INPUT_PINNED, 0)
    segs[:] = segs_pageable
    first = b"".join(arena[int(off[i]):int(off[i]) + int(ln[i])].tobytes() for i in range(min(4, n_msgs)))
    assert h_back[:len(first)].tobytes() == first

    if rank == 0:
        hbm_peak, peak_src = read_peaks()
        k_ms, k_n = prof["compress"]
        alg_bytes = payload + comp_bytes  # N read + C written, per launch
        achieved = alg_bytes / (k_ms / max(k_n, 1) * 1e-3) / 1e9 if k_ms > 0 else 0.0
        per_op = {}
        for name, nbytes in (("crc32c", payload), ("compress", payload), ("decompress", payload), ("pack", comp_bytes)):
            t_ms, cnt = prof[name]
            per_op[name] = {"ms_per_step": t_ms / max(cnt, 1),
                            "payload_gbs": nbytes / (t_ms / max(cnt, 1) * 1e-3) / 1e9 if t_ms > 0 else None}
        # CPU baseline on this box's host cores, same buffers, rank 0 only
        cores = os.cpu_count() or 1
        t_cpu = 0.0
        reps = 2
        for _ in range(reps):
            _, dt = cpu.batch_crc32c(arena, off, ln, threads=cores); t_cpu += dt
            o, oo, ol, dt = cpu.batch_compress(arena, off, ln, threads=cores); t_cpu += dt
            _, _, _, dt = cpu.batch_uncompress(o, oo, ol, off, ln, threads=cores); t_cpu += dt
        cpu_gbs = payload * reps / t_cpu / 1e9
        line = {
            "metric": METRIC, "value": payload * world * args.steps / (ms * 1e-3) / 1e9, "unit": "GB/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config_dict(world, args, ln), "clocks": clocks,
            "e2e": {"value": payload * world * e2e_steps / e2e_s / 1e9, "unit": "GB/s",
                    "h2d_bytes_per_step": int(bytes_h2d), "d2h_bytes_per_step": int(bytes_d2h), "steps": e2e_steps,
                    "path": "b2c_snappy_compress_crc_batch (sender: crc32c + compress, one pass) + "
                            "b2c_snappy_decompress_verify_batch (receiver: decompress + crc32c check); inputs = pageable host "
                            "iovecs (1-2 blocks per body) gathered into pinned staging by host threads; outputs = "
                            "page-locked host buffers",
                    "separate_calls_value": payload * world * e2e_steps / e2e_separate_s / 1e9,
                    "pcie_probe": pcie,
                    "pinned_inputs_device_gather_value": payload * world * e2e_steps / e2e_zero_copy_s / 1e9},
            "gpu_launches": int(launches),
            "roofline": {"kernel": "snappy_compress_small2_kernel" if int(ln.max()) <= 1024 else "snappy_compress_*_kernel (all size classes of the batch)", "bound": "hbm", "achieved": achieved,
                         "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                         # dram__bytes_read+write of this kernel from the committed ncu --set full capture of this
                         # very workload (profiles/r1_ncu_full_key_metrics.txt); unknown for other workloads
                         "traffic": 1623731000 if (args.workload, args.corpus, n_msgs) == ("c2", "words", N_MSGS) else None,
                         "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src,
                         "note": "byte-exact greedy LZ77 is a serial dependency chain per fragment: latency/issue bound, "
                                 "not HBM bound (see DESIGN.md)"},
            "ops": per_op, "compression_ratio": comp_bytes / payload,
            "messages_per_s": n_msgs * world * args.steps / (ms * 1e-3),
            **({"decompress_verify": {"ms_per_step": verify_ms, "payload_gbs": payload * world / (verify_ms * 1e-3) / 1e9}}
               if verify_ms else {}),
            "cpu_baseline": {"value": cpu_gbs, "unit": "GB/s", "cores": cores, "kind": cpu.kind,
                             "sample": f"full batch ({n_msgs} bodies, {payload} B) x {reps} passes of crc32c + RawCompress + "
                                       f"RawUncompress, static split over {cores} threads"},
        }
        print(json.dumps(line), flush=True)
    codec.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--messages", type=int, default=N_MSGS, help="bodies per GPU for --workload c2")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS), help="default c2 = BASELINE configs[1]")
    ap.add_argument("--corpus", default=CORPUS, choices=["words", "az09", "rand62"])
    ap.add_argument("--bytes", type=int, default=1 << 30, help="payload per GPU for --workload c4 / c5")
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
