#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 inflate engine (BASELINE.json metric).

Workload (N=1): BASELINE.json configs[1] -- a batch of 100 000 synthetic 64 KiB zlib streams
(`text` payload, level 6, dynamic Huffman), decoded by one launch of the batch kernel per step.
A "step" is one pass of the hot path over the whole batch.  N>1: every rank owns an independent
batch of the same shape (streams shard with no collective) -> weak scaling.

  value      decompressed GB/s, whole job, inputs and outputs resident in HBM (CUDA events)
  e2e        same metric through the C ABI with HOST buffers (infl_buf_batch): H2D of the
             compressed streams and D2H of the decoded bytes inside the timed region
  roofline   algorithmic bytes (compressed in + decompressed out) / kernel time vs measured HBM peak
  cpu_baseline  the reference's own infl_buf (oracle/_ref, compiled unmodified) on the host cores

`--impl reference` times the reference's CPU implementation (oracle/_ref; the oracle port if the
reference could not be compiled) on the same workload shape with all host threads.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

STREAM_BYTES = 65536
METRIC = "decompressed GB/s, batched zlib streams"
UNIT = "GB/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--streams", type=int, default=100_000, help="streams per GPU")
    ap.add_argument("--unique", type=int, default=8192, help="distinct streams generated on the host")
    ap.add_argument("--lanes", type=int, default=0, help="lanes per stream (0 = engine default)")
    ap.add_argument("--shape", type=int, default=-1, help="pipeline shape (engine default when < 0)")
    ap.add_argument("--engine", type=int, default=0, help="3 = warp per stream (default), 2 = lane-decoder pipeline")
    ap.add_argument("--mixed-gib", type=float, default=8.0, help="mixed workload: decompressed GiB per GPU")
    ap.add_argument("--mixed-no-fixed", action="store_true", help="mixed workload without the forced Z_FIXED streams (diagnosis)")
    ap.add_argument("--mixed-unique-mib", type=int, default=1024, help="mixed workload: distinct MiB generated on the host")
    ap.add_argument("--chunk-streams", type=int, default=256, help="chunks workload: 4 MiB streams per step")
    ap.add_argument("--sharded-pool-mib", type=int, default=512, help="mixed_sharded sub-record: distinct MiB generated on the host")
    ap.add_argument("--no-extra", action="store_true", help="headline only: skip the workloads sub-records")
    ap.add_argument("--workload", default="batch", choices=["batch", "single", "mixed", "chunks", "sharded", "wrap"],
                    help="batch = configs[1] (headline); single = configs[3], one large stream, split-point decode")
    ap.add_argument("--single-bytes", type=int, default=1 << 30)
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--e2e-streams", type=int, default=50_000, help="streams per GPU in the end-to-end leg (the same at every N)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--profile", action="store_true", help="short run for ncu: no e2e, no cpu baseline")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons during the timed region (NVML; nvidia-smi semantics)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def build_pool(unique, seed0):
    """Unique pool of config-2 streams: (raws, zlib streams)."""
    from defl_b200 import synth
    t0 = time.time()
    st = synth.make_streams(synth.config2_specs(unique, size=STREAM_BYTES, first_seed=seed0),
                            threads=os.cpu_count() or 1)
    return [r for r, _ in st], [z for _, z in st], time.time() - t0


# ------------------------------------------------------------------------------------------------
def cpu_reference_run(zs, reps_min=3, budget_s=8.0, threads=None, raw_lens=None):
    """The reference's infl_buf (oracle/_ref, else the oracle port) over the pool, all host threads.
    -> dict for `cpu_baseline`."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    from defl_b200 import synth
    threads = threads or (os.cpu_count() or 1)
    n = len(zs)
    src, so, sl = synth.pack(zs, align=1)
    src_off = np.zeros(n + 1, dtype=np.uint64)
    src_off[:n] = so
    src_off[n] = so[-1] + sl[-1]
    assert all(int(src_off[i + 1] - src_off[i]) == int(sl[i]) for i in range(n))
    if raw_lens is None:
        raw_lens = [STREAM_BYTES] * n
    dst_off = np.concatenate([[0], np.cumsum(np.asarray(raw_lens, dtype=np.uint64))]).astype(np.uint64)
    out_total = int(dst_off[-1])
    dst = np.zeros(out_total, dtype=np.uint8)
    codes = np.zeros(n, dtype=np.int32)
    kind = "reference" if ol.have_ref() else "port"
    best, total, reps = None, 0.0, 0
    while reps < reps_min or (total < budget_s and reps < 50):
        if kind == "reference":
            t = ol.ref().ref_infl_buf_batch_mt(src.ctypes.data, src_off.ctypes.data, dst.ctypes.data,
                                               dst_off.ctypes.data, n, 1, threads, None, codes.ctypes.data)
        else:
            t = ol.port().oracle_inflate_batch_mt(src.ctypes.data, src_off.ctypes.data, dst.ctypes.data,
                                                  dst_off.ctypes.data, n, 1, threads, codes.ctypes.data)
        assert (codes == 0).all(), "CPU baseline failed to decode"
        best = t if best is None else min(best, t)
        total += t
        reps += 1
    return {"value": out_total / best / 1e9, "unit": UNIT, "cores": threads, "kind": kind,
            "sample": f"{n} of the workload's streams ({out_total / 2**20:.0f} MiB out), "
                      f"best of {reps} passes, infl_buf per stream on a pthread work queue"}, dst


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    unique = min(args.unique, 4096)
    raws, zs, _ = build_pool(unique, 0)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    from defl_b200 import synth
    threads = os.cpu_count() or 1
    n = len(zs)
    src, so, sl = synth.pack(zs, align=1)
    src_off = np.zeros(n + 1, dtype=np.uint64)
    src_off[:n] = so
    src_off[n] = so[-1] + sl[-1]
    raw_lens = [STREAM_BYTES] * n
    dst_off = np.concatenate([[0], np.cumsum(np.asarray(raw_lens, dtype=np.uint64))]).astype(np.uint64)
    out_total = int(dst_off[-1])
    dst = np.zeros(out_total, dtype=np.uint8)
    codes = np.zeros(n, dtype=np.int32)
    kind = "reference" if ol.have_ref() else "port"

    def step():
        if kind == "reference":
            return ol.ref().ref_infl_buf_batch_mt(src.ctypes.data, src_off.ctypes.data, dst.ctypes.data,
                                                  dst_off.ctypes.data, n, 1, threads, None, codes.ctypes.data)
        return ol.port().oracle_inflate_batch_mt(src.ctypes.data, src_off.ctypes.data, dst.ctypes.data,
                                                 dst_off.ctypes.data, n, 1, threads, codes.ctypes.data)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    assert (codes == 0).all()
    assert dst[:STREAM_BYTES].tobytes() == raws[0]
    val = n * STREAM_BYTES * args.steps / dt / 1e9
    sample = (f"each step = {n} of the workload's 64 KiB streams ({n * STREAM_BYTES / 2**20:.0f} MiB out) "
              f"through infl_buf on {threads} host threads")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "configs[1]: 64 KiB zlib streams (text, level 6, dynamic Huffman), "
                               "reference infl_buf on host cores, bounded sample", "streams_per_step": n,
                   "stream_bytes": STREAM_BYTES},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "same_config_note": "same stream shape as the GPU arm's config (64 KiB text streams, level 6); each step decodes a "
                            f"{n}-stream sample of the 100 000, so GB/s compare per byte, ms_per_step do not",
    }))


def run_single(args, d, dev, stream=None, what=None, steps=None):
    """BASELINE.json configs[3]: ONE zlib stream (default 1 GiB of text/log stripes), device resident, decoded by
    the split-point path.  Timed on the host around the whole call (it contains host-side steps: candidate
    selection, the chunk chain), bracketed by device synchronisation.  -> the record (a JSON-able dict)."""
    import torch
    from defl_b200 import synth
    t0 = time.time()
    raw, z = stream if stream is not None else synth.single_stream(args.single_bytes, seed=7)
    gen_s = time.time() - t0
    steps = steps or args.steps
    t_src = torch.frombuffer(bytearray(z), dtype=torch.uint8).to(dev)
    t_dst = torch.zeros(len(raw), dtype=torch.uint8, device=dev)
    items = np.zeros(1, dtype=d.ITEM_DTYPE)
    items["src"], items["srclen"], items["dst"], items["dstlen"] = t_src.data_ptr(), len(z), t_dst.data_ptr(), len(raw)
    results = np.zeros(1, dtype=d.RESULT_DTYPE)
    from defl_b200 import api
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER | d.INFL_BATCH_DEVICE_PTRS
    for _ in range(max(min(args.warmup, 3), 1)):
        api.infl_buf_batch_raw(items, flags, results)
    torch.cuda.synchronize()
    l0 = d.kernel_launch_count()
    times = []
    for _ in range(steps):
        t_dst.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        api.infl_buf_batch_raw(items, flags, results)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t0)
    launches = d.kernel_launch_count() - l0
    assert int(results["code"][0]) == 0 and int(results["outlen"][0]) == len(raw)
    assert int(results["adler"][0]) == (zlib.adler32(raw) & 0xFFFFFFFF)
    split = bool(int(results["info"][0]) & 4)
    out = t_dst.cpu().numpy()
    assert out.tobytes() == raw, "single-stream output differs"
    del out, t_dst, t_src
    ms = float(np.mean(times)) * 1e3
    peak, peak_src = peaks()
    achieved = (len(z) + len(raw)) / (ms / 1e3) / 1e9
    return {
        "metric": "decompressed GB/s, single zlib stream", "value": len(raw) / (ms / 1e3) / 1e9, "unit": UNIT, "n_gpus": 1,
        "steps": steps, "warmup": max(min(args.warmup, 3), 1), "ms_per_step": ms, "best_ms": min(times) * 1e3,
        "higher_is_better": True, "scaling": "replicas only", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "byte_exact": True,
        "config": {"workload": what or "configs[3]: single deflate stream decoded with speculative intra-stream split points",
                   "raw_bytes": len(raw), "compressed_bytes": len(z), "ratio": len(raw) / len(z), "split_decode": split,
                   "gen_seconds": round(gen_s, 1), "engine": d.engine_info()},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": None, "peak_source": peak_src, "kernel": "split_chunk_kernel (count + decode passes), "
                     "find_block_starts_kernel, marker_resolve_kernel", "algorithmic_bytes_per_launch": len(z) + len(raw)},
        "gpu_launches": int(launches)}


def single_history_stream(total_bytes, seed=11, level=6):
    """ONE deflate() history over `total_bytes` of `text` (no flush points, no history resets)."""
    from defl_b200 import synth
    stripe = 8 << 20
    raw = b"".join(synth.payload("text", min(stripe, total_bytes - o), seed * 7919 + o // stripe)
                   for o in range(0, total_bytes, stripe))
    c = zlib.compressobj(level, zlib.DEFLATED, 15, 8, zlib.Z_DEFAULT_STRATEGY)
    return raw, c.compress(raw) + c.flush()


def run_chunks(args, d, dev, nstreams=None, steps=None):
    """BASELINE.json configs[2]: 4 MiB streams (png and text, level 6) handed over as lists of pieces, never joined
    on the host, through infl_buf_batch_iov (the batched infl_include).  Piece sizes follow SURVEY.md 8(d), driven by
    the LCG of test_fuzz.c:139-142: (ii) half 1-16 B / half 1-32768 B, (iii) 8193-32192 B, and (i) all 1-byte pieces
    on two streams only (4 M descriptors).  Host pieces in, host output out: the timed region holds H2D, the device
    gather, the decode and D2H."""
    import torch
    from defl_b200 import synth, api
    n = nstreams or args.chunk_streams
    Z = zlib.Z_DEFAULT_STRATEGY
    t0 = time.time()
    streams = synth.make_streams([("png" if i & 1 else "text", 4 << 20, 500 + i, 6, Z) for i in range(n)],
                                 threads=os.cpu_count() or 1)
    gen_s = time.time() - t0
    raws, zs = [r for r, _ in streams], [z for _, z in streams]
    src_h, so, sl = synth.pack(zs, align=1)
    h_src = torch.from_numpy(src_h).pin_memory()
    out_len = np.array([len(r) for r in raws], dtype=np.int64)
    out_off = np.concatenate([[0], np.cumsum(out_len)])[:-1]
    h_dst = torch.zeros(int(out_len.sum()), dtype=torch.uint8).pin_memory()
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER

    def piece_sizes(kind, total, state):
        sizes, left = [], total
        if kind == "one":
            return np.ones(total, dtype=np.int64), state
        while left > 0:
            state = (state * 1103515245 + 12345) & 0xFFFFFFFF
            r = state >> 8
            if kind == "mix":
                k = 1 + r % 16 if (state >> 4) & 1 else 1 + r % 32768
            else:
                k = 8193 + r % 24000
            k = min(k, left)
            sizes.append(k)
            left -= k
        return np.array(sizes, dtype=np.int64), state

    def run(kind, idx, steps):
        state, iovs = 0x12345678, []
        items = np.zeros(len(idx), dtype=api.IOV_ITEM_DTYPE)
        npieces = 0
        for j, i in enumerate(idx):
            sizes, state = piece_sizes(kind, int(sl[i]), state)
            iov = np.zeros(len(sizes), dtype=api.IOVEC_DTYPE)
            iov["ptr"] = h_src.data_ptr() + int(so[i]) + np.concatenate([[0], np.cumsum(sizes)])[:-1]
            iov["len"] = sizes
            iovs.append(iov)
            items[j] = (iov.ctypes.data, len(sizes), h_dst.data_ptr() + int(out_off[i]), int(out_len[i]))
            npieces += len(sizes)
        res = np.zeros(len(idx), dtype=d.RESULT_DTYPE)
        call = lambda: d.lib().infl_buf_batch_iov(items.ctypes.data, len(idx), int(flags), res.ctypes.data)
        assert call() == 0, d.last_error()                  # warm-up (allocates staging once)
        h_dst.zero_()
        torch.cuda.synchronize()
        l0 = d.kernel_launch_count()
        t = time.perf_counter()
        for _ in range(steps):
            assert call() == 0, d.last_error()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t) / steps
        launches = (d.kernel_launch_count() - l0) // steps
        assert (res["code"] == 0).all() and (res["adler"] == res["adler_stored"]).all()
        hv = h_dst.numpy()
        for i in idx:
            assert hv[out_off[i]: out_off[i] + out_len[i]].tobytes() == raws[i], f"stream {i} differs"
        outb = int(out_len[idx].sum())
        return {"pieces_per_step": npieces, "streams": len(idx), "ms_per_step": dt * 1e3, "GBps": outb / dt / 1e9,
                "h2d_bytes": int(sl[idx].sum()), "d2h_bytes": outb, "launches": int(launches)}

    allidx = np.arange(n)
    steps = max(steps or args.steps, 1)
    mix = run("mix", allidx, steps)
    large = run("large", allidx, steps)
    one = run("one", allidx[:2], 1)
    peak, peak_src = peaks()
    achieved = (mix["h2d_bytes"] * 2 + mix["h2d_bytes"] + mix["d2h_bytes"]) / (mix["ms_per_step"] / 1e3) / 1e9
    return {
        "metric": "decompressed GB/s, zlib streams given as chunk lists, host to host", "value": mix["GBps"], "unit": UNIT,
        "n_gpus": 1, "steps": steps, "warmup": 1, "ms_per_step": mix["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "byte_exact": True,
        "config": {"workload": "configs[2]: 4 MiB streams (png / text, level 6) as lists of 1 B - 32 KiB pieces through "
                               "infl_buf_batch_iov; value = chunking (ii)", "chunking_ii_mix": mix, "chunking_iii_8k_32k": large,
                   "chunking_i_all_1_byte": one, "engine": d.engine_info(), "gen_seconds": round(gen_s, 1),
                   "timing": "host clock around the call (host pieces in, host bytes out)"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "peak_source": peak_src, "kernel": "whole call (gather_pieces_kernel: 2 x pieces; decode: in + out)",
                     "algorithmic_bytes_per_launch": mix["h2d_bytes"] * 3 + mix["d2h_bytes"]},
        "cpu_baseline": None,
        "e2e": {"value": mix["GBps"], "unit": UNIT, "h2d_bytes_per_step": mix["h2d_bytes"] + 16 * mix["pieces_per_step"],
                "d2h_bytes_per_step": mix["d2h_bytes"]},
        "gpu_launches": mix["launches"] * steps}


def run_mixed_sharded(args, d, dev, rank, world, local, dist, per_gpu_gib=None, pool_mib=None, steps=None):
    """BASELINE.json configs[4] as north_star states it: ONE mixed batch (one descriptor table, the same on every
    rank: same seeds) cut into `world` byte-balanced shards by infl_batch_partition (greedy LPT on srclen + dstlen,
    SURVEY.md 8e); rank r decodes the members with part_of == r from its own device memory, no collective on the
    data path.  The table is a unique pool tiled to world x per_gpu_gib GiB of output (64 GiB at 8 GPUs), every
    member in physically distinct memory.  -> the sub-record (rank 0), None elsewhere."""
    import torch
    from defl_b200 import synth, api
    per_gpu_gib = per_gpu_gib or args.mixed_gib
    pool_mib = pool_mib or args.sharded_pool_mib
    steps = steps or max(2, min(args.steps, 3))
    t0 = time.time()
    specs = synth.config5_specs(pool_mib << 20, first_seed=0)            # every rank: the same pool
    st = synth.make_streams(specs, threads=max(1, (os.cpu_count() or 1) // max(1, world)))
    raws, zs = [r for r, _ in st], [z for _, z in st]
    gen_s = time.time() - t0
    unique = len(zs)
    raw_len = np.array([len(r) for r in raws], dtype=np.int64)
    z_len = np.array([len(z) for z in zs], dtype=np.int64)
    pool_out = int(raw_len.sum())
    reps = max(1, int(round(world * per_gpu_gib * (1 << 30) / pool_out)))
    n = unique * reps
    k = np.arange(n, dtype=np.int64)
    tab_src_len, tab_dst_len = z_len[k % unique], raw_len[k % unique]     # THE table: member i = pool entry i % unique
    part_of, load = api.batch_partition(tab_src_len, tab_dst_len, world)
    mine = np.nonzero(part_of == rank)[0]
    u_of = mine % unique
    # this rank's shard, every member in its own memory
    src_off = np.zeros(len(mine), dtype=np.int64)
    at = 0
    for j, u in enumerate(u_of):
        src_off[j] = at
        at = (at + int(z_len[u]) + 15) // 16 * 16
    h_src = np.zeros(max(at, 16), dtype=np.uint8)
    for j, u in enumerate(u_of):
        h_src[src_off[j]: src_off[j] + z_len[u]] = np.frombuffer(zs[u], dtype=np.uint8)
    t_src = torch.from_numpy(h_src).to(dev)
    del h_src
    dlen = raw_len[u_of]
    dst_off = np.concatenate([[0], np.cumsum(dlen)])[:-1]
    t_dst = torch.zeros(int(dlen.sum()) + 16, dtype=torch.uint8, device=dev)
    items = np.zeros(len(mine), dtype=d.ITEM_DTYPE)
    items["src"] = t_src.data_ptr() + src_off
    items["srclen"] = z_len[u_of]
    items["dst"] = t_dst.data_ptr() + dst_off
    items["dstlen"] = dlen
    results = np.zeros(len(mine), dtype=d.RESULT_DTYPE)
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER | d.INFL_BATCH_DEVICE_PTRS
    in_bytes, out_bytes = int(z_len[u_of].sum()), int(dlen.sum())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(2):
        api.infl_buf_batch_raw(items, flags, results)
    barrier()
    l0 = d.kernel_launch_count()
    times = []
    for _ in range(steps):
        barrier()
        t = time.perf_counter()
        api.infl_buf_batch_raw(items, flags, results)
        torch.cuda.synchronize()
        times.append(time.perf_counter() - t)
    launches = d.kernel_launch_count() - l0
    barrier()
    # byte-exact: code, length and Adler-32 of every member (the stored Adler is zlib's own of the raw bytes); the
    # first copy of every pool entry on this rank against the generator's bytes; every further copy against the first
    assert (results["code"] == 0).all(), f"decode failed: {int((results['code'] != 0).sum())} streams"
    assert (results["outlen"] == dlen.astype(np.uint64)).all() and (results["adler"] == results["adler_stored"]).all()
    first = {}
    for j, u in enumerate(u_of):
        a, b = int(dst_off[j]), int(dst_off[j] + dlen[j])
        if int(u) not in first:
            first[int(u)] = (a, b)
            assert t_dst[a:b].cpu().numpy().tobytes() == raws[int(u)], f"member {mine[j]} differs from the generator's bytes"
        else:
            fa, fb = first[int(u)]
            assert torch.equal(t_dst[a:b], t_dst[fa:fb]), f"member {mine[j]} differs from its twin"
    n_split = int(((results["info"] & 4) != 0).sum())
    mean_ms = float(np.mean(times)) * 1e3
    stat = torch.tensor([mean_ms, float(in_bytes + out_bytes), float(out_bytes), float(len(mine)), float(n_split)],
                        dtype=torch.float64, device=dev)
    if world > 1:
        allst = [torch.zeros_like(stat) for _ in range(world)]
        dist.all_gather(allst, stat)
        allst = torch.stack(allst).cpu().numpy()
    else:
        allst = stat.cpu().numpy()[None, :]
    del t_src, t_dst
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    peak, peak_src = peaks()
    shard_ms = [float(x) for x in allst[:, 0]]
    tmax = max(shard_ms) / 1e3
    alg = float(allst[:, 1].sum())
    tot_out = float(allst[:, 2].sum())
    return {
        "metric": "decompressed GB/s, ONE mixed batch sharded by infl_batch_partition", "value": tot_out / tmax / 1e9,
        "unit": UNIT, "n_gpus": world, "steps": steps, "ms_per_step": max(shard_ms), "byte_exact": True,
        "shard_ms": shard_ms, "shard_ms_min": min(shard_ms), "shard_ms_max": max(shard_ms),
        "shard_algorithmic_bytes": [float(x) for x in allst[:, 1]],
        "byte_imbalance_max_over_mean": float(allst[:, 1].max() / allst[:, 1].mean()),
        "roofline": {"bound": "hbm", "achieved": alg / tmax / 1e9, "peak": peak * world, "unit": "GB/s",
                     "frac": alg / tmax / 1e9 / (peak * world), "peak_source": peak_src + f" x {world} GPUs",
                     "algorithmic_bytes_per_step": alg},
        "config": {"workload": "configs[4]: mixed batch (stored/fixed/dynamic, levels 1/6/9, 64 KiB-4 MiB), "
                               f"{tot_out / 2**30:.1f} GiB out in one table of {n} members, LPT-sharded over {world} GPU(s)",
                   "members": n, "members_per_shard": [int(x) for x in allst[:, 3]], "split_decoded": int(allst[:, 4].sum()),
                   "unique_streams": unique, "unique_out_bytes": pool_out, "gen_seconds": round(gen_s, 1),
                   "timing": "host clock around infl_buf_batch(INFL_BATCH_DEVICE_PTRS) per rank, device-synchronised; "
                             "value = all bytes / slowest shard"},
        "gpu_launches": int(launches)}


def run_wrap(args, d, dev):
    """The rows either side of the hot path (include/defl/infl_wrap.h), device resident: CRC-32 of one
    large buffer, the PNG scanline un-filter on a batch of RGBA images, whole PNG files (container walk + IDAT
    gather + inflate + un-filter) and gzip members.  Every result is compared with the CPU oracle
    (oracle/wrap_oracle.c) or the generator's bytes; the oracle is also the CPU figure beside each number."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import wrap_lib as wl
    from defl_b200 import synth, api
    peak, peak_src = peaks()
    rec = {}
    rng = np.random.default_rng(1)

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            t = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t)
        return float(np.mean(ts))

    # -- CRC-32 of one 1 GiB buffer
    nb = 1 << 30
    t_buf = torch.randint(0, 256, (nb,), dtype=torch.uint8, device=dev)
    items = np.zeros(1, dtype=d.ITEM_DTYPE)
    items[0] = (t_buf.data_ptr(), nb, 0, 0)
    crc = np.zeros(1, dtype=np.uint32)
    call = lambda: d.lib().infl_crc32_batch(items.ctypes.data, 1, int(d.INFL_BATCH_DEVICE_PTRS), crc.ctypes.data)
    dt = timed(call)
    sample = t_buf[: 64 << 20].cpu().numpy().tobytes()
    t0 = time.perf_counter(); want = wl.crc32(sample); cpu_dt = time.perf_counter() - t0
    items[0] = (t_buf.data_ptr(), len(sample), 0, 0)
    call()
    assert int(crc[0]) == want == (zlib.crc32(sample) & 0xFFFFFFFF)
    rec["crc32"] = {"value": nb / dt / 1e9, "unit": "GB/s", "bytes": nb, "ms": dt * 1e3, "byte_exact": True,
                    "roofline": {"bound": "hbm", "achieved": nb / dt / 1e9, "peak": peak, "frac": nb / dt / 1e9 / peak,
                                 "algorithmic_bytes": "1 read per byte"},
                    "cpu_baseline": {"value": len(sample) / cpu_dt / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                                     "sample": "64 MiB, oracle/wrap_oracle.c (bit at a time, RFC 1952 section 8)"}}
    del t_buf

    # -- PNG: 2048 RGBA images of 256 x 256 (smooth content, all five filter types), as scanline data and as files
    w = h = 256
    uniq = 32
    imgs = []
    for k in range(uniq):
        x = np.arange(h * w * 4)
        pix = ((x // 4 * (3 + k) + x % 4 * 31 + rng.integers(0, 3, x.size)) & 255).astype(np.uint8).tobytes()
        f, stride, bpp = wl.png_file(w, h, 8, 6, pix, ftypes=(4, 1, 2, 3, 0), idat=8192)
        imgs.append((f, pix, wl.png_filter_rows(pix, h, stride, bpp, (4, 1, 2, 3, 0))))
    n = 2048
    stride, bpp = w * 4, 4
    rawlen, pixlen = h * (stride + 1), h * stride
    t_raw = torch.cat([torch.frombuffer(bytearray(imgs[i % uniq][2]), dtype=torch.uint8) for i in range(n)]).to(dev)
    t_pix = torch.zeros(n * pixlen, dtype=torch.uint8, device=dev)
    P = ctypes.c_void_p * n
    raws = P(*[t_raw.data_ptr() + i * rawlen for i in range(n)])
    pixs = P(*[t_pix.data_ptr() + i * pixlen for i in range(n)])
    hh = np.full(n, h, dtype=np.uint32); ss = np.full(n, stride, dtype=np.uint32); bb = np.full(n, bpp, dtype=np.uint32)
    bad = np.zeros(n, dtype=np.uint32)
    call = lambda: d.lib().infl_png_unfilter_device(raws, pixs, hh.ctypes.data, ss.ctypes.data, bb.ctypes.data, n,
                                                    bad.ctypes.data, None)
    dt = timed(call)
    host = t_pix.cpu().numpy()
    for i in (0, 1, uniq - 1, n - 1):
        assert host[i * pixlen:(i + 1) * pixlen].tobytes() == imgs[i % uniq][1], "un-filtered pixels differ"
    t0 = time.perf_counter()
    for k in range(uniq):
        assert wl.png_unfilter(imgs[k][2], h, stride, bpp)[1] == imgs[k][1]
    cpu_dt = time.perf_counter() - t0
    alg = n * (rawlen + pixlen)
    rec["png_unfilter"] = {"value": n * pixlen / dt / 1e9, "unit": "GB/s of pixels", "images": n, "image": f"{w}x{h} RGBA8", "ms": dt * 1e3,
                           "byte_exact": True,
                           "roofline": {"bound": "hbm", "achieved": alg / dt / 1e9, "peak": peak, "frac": alg / dt / 1e9 / peak,
                                        "algorithmic_bytes": "scanline data in + pixels out"},
                           "cpu_baseline": {"value": uniq * pixlen / cpu_dt / 1e9, "unit": "GB/s of pixels", "cores": 1, "kind": "port",
                                            "sample": f"{uniq} images, oracle/wrap_oracle.c"}}
    del t_raw
    # whole files, device resident
    blob = b"".join(f for f, _, _ in imgs)
    offs = np.concatenate([[0], np.cumsum([len(f) for f, _, _ in imgs])])
    t_files = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
    t_pix.zero_()
    pitems = np.zeros(n, dtype=d.PNG_ITEM_DTYPE)
    for i in range(n):
        k = i % uniq
        pitems[i] = (t_files.data_ptr() + int(offs[k]), len(imgs[k][0]), t_pix.data_ptr() + i * pixlen, pixlen)
    pres = np.zeros(n, dtype=d.PNG_RESULT_DTYPE)
    call = lambda: d.png_batch_raw(pitems, d.INFL_WRAP_VERIFY | d.INFL_BATCH_DEVICE_PTRS, pres)
    l0 = d.kernel_launch_count()
    dt = timed(call)
    launches = (d.kernel_launch_count() - l0) // 4
    assert (pres["code"] == 0).all()
    host = t_pix.cpu().numpy()
    for i in (0, 5, n - 1):
        assert host[i * pixlen:(i + 1) * pixlen].tobytes() == imgs[i % uniq][1], "decoded PNG differs"
    t0 = time.perf_counter()
    for k in range(uniq):
        assert wl.png_decode(imgs[k][0])[1] == imgs[k][1]
    cpu_dt = time.perf_counter() - t0
    fbytes = sum(len(imgs[i % uniq][0]) for i in range(n))
    rec["png_files"] = {"value": n / dt, "unit": "images/s", "pixel_GBps": n * pixlen / dt / 1e9, "images": n, "ms": dt * 1e3,
                        "file_bytes": fbytes, "idat_per_file": int(pres["nidat"][0]), "launches_per_call": int(launches), "byte_exact": True,
                        "cpu_baseline": {"value": uniq / cpu_dt, "unit": "images/s", "cores": 1, "kind": "port",
                                         "sample": f"{uniq} files: oracle walk + Python zlib + oracle un-filter"}}
    del t_files, t_pix
    # -- gzip members: 20 000 x 64 KiB text
    pool = [synth.payload("text", 65536, 300 + i) for i in range(256)]
    mem = [wl.gzip_member(r, name=b"f%d.txt" % i) for i, r in enumerate(pool)]
    src_h, so, sl = synth.pack(mem, align=16)
    ng = 20000
    t_src = torch.from_numpy(src_h).to(dev)
    t_dst = torch.zeros(ng * 65536, dtype=torch.uint8, device=dev)
    gitems = np.zeros(ng, dtype=d.ITEM_DTYPE)
    k = np.arange(ng)
    gitems["src"] = t_src.data_ptr() + so[k % 256]
    gitems["srclen"] = sl[k % 256]
    gitems["dst"] = t_dst.data_ptr() + k * 65536
    gitems["dstlen"] = 65536
    gres = np.zeros(ng, dtype=d.GZIP_RESULT_DTYPE)
    call = lambda: d.gzip_batch_raw(gitems, d.INFL_WRAP_VERIFY | d.INFL_BATCH_DEVICE_PTRS, gres)
    dt = timed(call)
    assert (gres["code"] == 0).all() and (gres["crc32"] == gres["crc32_stored"]).all()
    assert t_dst[:65536].cpu().numpy().tobytes() == pool[0]
    rec["gzip_members"] = {"value": ng * 65536 / dt / 1e9, "unit": "GB/s decompressed", "members": ng, "ms": dt * 1e3, "byte_exact": True,
                           "note": "header parse kernel + batch inflate + CRC-32 kernel + host combination of the 64 KiB partial CRCs"}
    return rec


def run_mixed(args, d, dev, rank, world, local, dist):
    """BASELINE.json configs[4] shape: stored / fixed / dynamic blocks, levels 1/6/9, sizes log-uniform in
    [64 KiB, 4 MiB], classes text/log/png/rand/zero (synth.config5_specs).  Each GPU owns --mixed-gib GiB of
    output (64 GiB / 8 GPUs by default): a unique pool generated on the host, tiled into physically distinct
    device memory; largest compressed stream first; no collective."""
    import torch
    from defl_b200 import synth
    t0 = time.time()
    specs = synth.config5_specs(args.mixed_unique_mib << 20, first_seed=rank * 1_000_000)
    if args.mixed_no_fixed:
        specs = [(k_, sz, sd, lv, zlib.Z_DEFAULT_STRATEGY) for (k_, sz, sd, lv, _s) in specs]
    st = synth.make_streams(specs, threads=os.cpu_count() or 1)
    raws, zs = [r for r, _ in st], [z for _, z in st]
    gen_s = time.time() - t0
    unique = len(zs)
    raw_len = np.array([len(r) for r in raws], dtype=np.int64)
    pool_out = int(raw_len.sum())
    reps = max(1, int(round(args.mixed_gib * (1 << 30) / pool_out)))
    src_h, so, sl = synth.pack(zs, align=16)
    pool_bytes = int(src_h.size)
    n = unique * reps
    t_src = torch.from_numpy(src_h).to(dev).repeat(reps)
    k = np.arange(n, dtype=np.int64)
    src_off = (k // unique) * pool_bytes + so[k % unique]
    src_len = sl[k % unique]
    dst_len = raw_len[k % unique]
    ro = np.concatenate([[0], np.cumsum(raw_len)])[:-1]
    dst_off = (k // unique) * pool_out + ro[k % unique]
    t_dst = torch.zeros(reps * pool_out, dtype=torch.uint8, device=dev)
    # the product call for device-resident data: infl_buf_batch with INFL_BATCH_DEVICE_PTRS.  It sends the members
    # that would outlast the batch on one warp through the split decoder (all together) and the rest through the
    # warp-per-stream kernel; it returns when everything is decoded, so the host clock around it, bracketed by
    # device synchronisation, is the device time of the whole sequence (the library launches on its own stream,
    # which torch events do not see).
    from defl_b200 import api
    items = np.zeros(n, dtype=d.ITEM_DTYPE)
    items["src"] = t_src.data_ptr() + src_off
    items["srclen"] = src_len
    items["dst"] = t_dst.data_ptr() + dst_off
    items["dstlen"] = dst_len
    results = np.zeros(n, dtype=d.RESULT_DTYPE)
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER | d.INFL_BATCH_DEVICE_PTRS

    class _Call:
        def launch(self, _flags):
            api.infl_buf_batch_raw(items, flags, results)

        def results_host(self):
            return results
    batch = _Call()
    in_bytes, out_bytes = int(src_len.sum()), int(dst_len.sum())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        batch.launch(flags)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = d.kernel_launch_count()
    t_start = time.perf_counter()
    for i in range(args.steps):
        batch.launch(flags)
    barrier()
    total_ms = (time.perf_counter() - t_start) * 1e3
    kern_ms = total_ms / args.steps
    sampler.stop_flag = True
    launches = d.kernel_launch_count() - l0
    sampler.join(timeout=1.0)

    res = batch.results_host()
    assert (res["code"] == 0).all(), f"decode failed: {int((res['code'] != 0).sum())} streams"
    assert (res["outlen"] == dst_len.astype(np.uint64)).all() and (res["adler"] == res["adler_stored"]).all()
    n_split = int(((res["info"] & 4) != 0).sum())
    tile0 = t_dst[:pool_out]
    host0 = tile0.cpu().numpy()
    for i in range(unique):
        assert host0[ro[i]: ro[i] + raw_len[i]].tobytes() == raws[i], f"stream {i} differs"
    for r in range(1, reps):
        assert torch.equal(t_dst[r * pool_out:(r + 1) * pool_out], tile0), f"tile {r}"
    del host0

    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    peak, peak_src = peaks()
    achieved = (in_bytes + out_bytes) / (kern_ms / 1e3) / 1e9
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        pick = list(range(0, unique, max(1, unique // 256)))
        cpu, _ = cpu_reference_run([zs[i] for i in pick], raw_lens=[len(raws[i]) for i in pick])
    if rank == 0:
        kinds = {}
        for sp in specs:
            kinds[sp[0]] = kinds.get(sp[0], 0) + sp[1]
        print(json.dumps({
            "metric": "decompressed GB/s, mixed batch of zlib streams", "value": out_bytes * world * args.steps / (total_ms_max / 1e3) / 1e9,
            "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": total_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[4] shape: mixed batch (stored/fixed/dynamic, levels 1/6/9, 64 KiB-4 MiB), "
                                   f"{out_bytes / 2**30:.1f} GiB out per GPU",
                       "streams_per_gpu": n, "split_decoded_streams": n_split, "unique_streams": unique, "unique_out_bytes": pool_out,
                       "timing": "host clock around infl_buf_batch(INFL_BATCH_DEVICE_PTRS), device-synchronised both sides",
                       "compressed_bytes_per_gpu": in_bytes, "ratio": out_bytes / in_bytes,
                       "class_share_of_bytes": {k_: round(v / pool_out, 3) for k_, v in sorted(kinds.items())},
                       "l2": "inputs+outputs per step exceed L2 (126 MB); no flush needed",
                       "parallelism": f"{world} x independent shards, no collective", "engine": d.engine_info(),
                       "gen_seconds": round(gen_s, 1)},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": None, "peak_source": peak_src,
                         "kernel": "whole sequence: split kernels (find/count/decode/window/marker) + inflate_warp_kernel",
                         "algorithmic_bytes_per_launch": in_bytes + out_bytes, "kernel_ms": kern_ms},
            "cpu_baseline": cpu, "e2e": None, "gpu_launches": int(launches), "clocks": sampler.summary()}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.profile:
        args.no_e2e = args.no_cpu = True

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import defl_b200 as d
    from defl_b200 import synth

    assert torch.cuda.is_available(), "bench.py needs a CUDA device: the engine has no CPU path"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    d.lib()
    if args.lanes:
        d.set_lanes_per_stream(args.lanes)
    if args.shape >= 0:
        d.set_pipeline_shape(args.shape)
    if args.engine:
        d.set_engine(args.engine)

    if args.workload == "single":
        print(json.dumps(run_single(args, d, dev)))
        return
    if args.workload == "mixed":
        run_mixed(args, d, dev, rank, world, local, dist)
        return
    if args.workload == "chunks":
        print(json.dumps(run_chunks(args, d, dev)))
        return
    if args.workload == "wrap":
        print(json.dumps(run_wrap(args, d, dev)))
        return
    if args.workload == "sharded":
        rec = run_mixed_sharded(args, d, dev, rank, world, local, dist)
        if rank == 0:
            print(json.dumps(rec))
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    n = args.streams
    unique = min(args.unique, n)
    raws, zs, gen_s = build_pool(unique, rank * 1_000_000)
    src_h, so, sl = synth.pack(zs, align=16)
    reps = (n + unique - 1) // unique
    pool_bytes = int(src_h.size)

    # device-resident batch: the unique pool tiled `reps` times into physically distinct memory;
    # stream k is pool member k % unique of tile k // unique, so neighbours in launch order differ
    t_pool = torch.from_numpy(src_h).to(dev)
    t_src = t_pool.repeat(reps)
    k = np.arange(n, dtype=np.int64)
    src_off = (k // unique) * pool_bytes + so[k % unique]
    src_len = sl[k % unique]
    dst_off = k * STREAM_BYTES
    dst_len = np.full(n, STREAM_BYTES, dtype=np.int64)
    t_dst = torch.zeros(n * STREAM_BYTES, dtype=torch.uint8, device=dev)
    batch = d.DeviceBatch(t_src, src_off, src_len, t_dst, dst_off, dst_len, order="index")
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER
    in_bytes = int(src_len.sum())
    out_bytes = n * STREAM_BYTES

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3) if not args.profile else args.warmup):
        batch.launch(flags)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = d.kernel_launch_count()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record()
    for i in range(args.steps):
        batch.launch(flags)
        ev[i + 1].record()
    barrier()
    sampler.stop_flag = True
    launches = d.kernel_launch_count() - l0
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    total_ms = ev[0].elapsed_time(ev[-1])
    sampler.join(timeout=1.0)

    # ---- byte-exact check of what was just timed: codes, Adler, tile 0 vs the generator's raw bytes,
    #      every other tile vs tile 0 on the device
    res = batch.results_host()
    assert (res["code"] == 0).all(), f"decode failed: {int((res['code'] != 0).sum())} streams"
    assert (res["outlen"] == STREAM_BYTES).all() and (res["adler"] == res["adler_stored"]).all()
    tile0 = t_dst[: unique * STREAM_BYTES]
    host0 = tile0.cpu().numpy()
    for i in range(0, unique, max(1, unique // 512)):
        assert host0[i * STREAM_BYTES:(i + 1) * STREAM_BYTES].tobytes() == raws[i], f"stream {i} differs"
    want_adler = np.array([zlib.adler32(r) & 0xFFFFFFFF for r in raws], dtype=np.uint32)
    assert (res["adler"][:unique] == want_adler).all()
    for r in range(1, reps):
        m = min(unique, n - r * unique) * STREAM_BYTES
        assert torch.equal(t_dst[r * unique * STREAM_BYTES: r * unique * STREAM_BYTES + m], tile0[:m]), f"tile {r}"
    del host0

    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    value = out_bytes * world * args.steps / (total_ms_max / 1e3) / 1e9

    peak, peak_src = peaks()
    kern_ms = float(np.mean(step_ms))
    achieved = (in_bytes + out_bytes) / (kern_ms / 1e3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src, "kernel": "inflate_batch_kernel" if args.lanes else
                ("inflate_pipeline_kernel" if args.engine == 2 else "inflate_warp_kernel"),
                "algorithmic_bytes_per_launch": in_bytes + out_bytes,
                "kernel_ms": kern_ms}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof) and not args.lanes and args.engine in (0, 3):
        try:
            tj = json.load(open(prof))
            roofline["traffic"] = float(tj["dram_bytes_over_algorithmic_bytes"]) * (in_bytes + out_bytes)
            roofline["traffic_source"] = "ncu dram__bytes_read.sum + dram__bytes_write.sum over algorithmic bytes = %.2f (%s), times this launch's algorithmic bytes" % (
                tj["dram_bytes_over_algorithmic_bytes"], tj.get("source", "profiles/traffic.json"))
        except Exception:
            pass

    # ---- end to end through the C ABI with host buffers ------------------------------------------
    e2e = None
    if not args.no_e2e:
        # the same job on every rank at every N (the curve compares like with like): 50 000 streams per GPU,
        # 3.3 GB out + 1.5 GB in of pinned host memory per rank
        ne = min(n, args.e2e_streams)
        h_src = torch.from_numpy(src_h).repeat((ne + unique - 1) // unique).pin_memory()
        h_dst = torch.zeros(ne * STREAM_BYTES, dtype=torch.uint8).pin_memory()
        items = np.zeros(ne, dtype=d.ITEM_DTYPE)
        items["src"] = h_src.data_ptr() + src_off[:ne]
        items["srclen"] = src_len[:ne]
        items["dst"] = h_dst.data_ptr() + dst_off[:ne]
        items["dstlen"] = STREAM_BYTES
        e_in, e_out = int(src_len[:ne].sum()), ne * STREAM_BYTES
        results = np.zeros(ne, dtype=d.RESULT_DTYPE)
        from defl_b200 import api
        api.infl_buf_batch_raw(items, flags, results)           # warm-up (allocates staging once)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            api.infl_buf_batch_raw(items, flags, results)
            assert int(results["code"][0]) == 0
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        assert (results["code"] == 0).all()
        hv = h_dst.numpy()
        for i in (0, unique - 1, ne - 1):
            assert hv[i * STREAM_BYTES:(i + 1) * STREAM_BYTES].tobytes() == raws[i % unique]
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        # the ceiling of this leg on this box: the same bytes, copies only (H2D of the compressed bytes and D2H of
        # the decoded bytes on two streams at once, pinned memory, all ranks at the same time, nothing decoded)
        c_in = torch.empty(e_in, dtype=torch.uint8, device=dev)
        c_out = torch.empty(e_out, dtype=torch.uint8, device=dev)
        s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
        hs, hd = h_src[:e_in], h_dst[:e_out]

        def copies():
            with torch.cuda.stream(s1):
                c_in.copy_(hs, non_blocking=True)
            with torch.cuda.stream(s2):
                hd.copy_(c_out, non_blocking=True)
        copies()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            copies()
        torch.cuda.synchronize()
        cdt = time.perf_counter() - t0
        ct = torch.tensor([cdt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ct, op=dist.ReduceOp.MAX)
        e2e = {"value": e_out * world * args.e2e_steps / float(tt.item()) / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": e_in + 36 * ne, "d2h_bytes_per_step": e_out + 32 * ne,
               "steps": args.e2e_steps, "streams_per_gpu": ne, "api": "infl_buf_batch (host pointers, pinned)",
               "copy_ceiling": e_out * world * args.e2e_steps / float(ct.item()) / 1e9,
               "copy_ceiling_what": "decoded GB/s if the step were only its pinned H2D + D2H copies (both directions at once, "
                                    "all ranks together), measured right after the timed leg",
               "of_copy_ceiling": float(ct.item()) / float(tt.item())}
        del h_src, h_dst, c_in, c_out, hs, hd

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu, _ = cpu_reference_run(zs[: min(unique, 4096)])

    # ---- the other north_star configs, as sub-records of the same line --------------------------------
    workloads = None
    if not args.profile and not args.no_extra and not args.lanes and args.engine in (0, 3):
        del batch, t_src, t_dst, t_pool, tile0
        torch.cuda.empty_cache()
        workloads = {}
        workloads["mixed_sharded"] = run_mixed_sharded(args, d, dev, rank, world, local, dist)
        if world == 1:
            workloads["single_1gib"] = run_single(args, d, dev, steps=3,
                                                  what="configs[3]: one 1 GiB zlib stream of text/log stripes, built pigz-style "
                                                       "(64 slices, each ending in a full flush: 63 history resets)")
            workloads["single_1gib_one_history"] = run_single(
                args, d, dev, stream=single_history_stream(1 << 30), steps=3,
                what="configs[3]: one 1 GiB zlib stream of text from ONE deflate() history (level 6, no flush points)")
            workloads["chunks"] = run_chunks(args, d, dev, nstreams=256, steps=2)
            workloads["wrap"] = run_wrap(args, d, dev)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3) if not args.profile else args.warmup,
            "ms_per_step": total_ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[1]: batch of 100k synthetic 64 KiB zlib streams (text, level 6, dynamic "
                                   "Huffman), one launch per step" if n == 100_000 else
                                   f"configs[1] shape, {n} streams per GPU",
                       "streams_per_gpu": n, "stream_bytes": STREAM_BYTES, "unique_streams": unique,
                       "compressed_bytes_per_gpu": in_bytes, "ratio": out_bytes / in_bytes,
                       "l2": "inputs+outputs per step (%.1f GB) exceed L2 (126 MB); no flush needed"
                             % ((in_bytes + out_bytes) / 1e9),
                       "parallelism": f"{world} x independent shards, no collective",
                       "engine": d.engine_info(), "gen_seconds": round(gen_s, 1)},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": sampler.summary(), "workloads": workloads,
        }
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
