#!/usr/bin/env python3
"""Per-class decode rate of the batch engine: N streams of one payload class / level / strategy at a time
(device resident, CUDA events).  usage: class_probe.py [size_bytes] [streams]"""
import os, sys, time, zlib
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import defl_b200 as d
from defl_b200 import synth

size = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
nstreams = int(sys.argv[2]) if len(sys.argv) > 2 else 4736
uniq = 64
dev = torch.device("cuda", 0)
cases = [("text", 6, zlib.Z_DEFAULT_STRATEGY), ("text", 1, zlib.Z_DEFAULT_STRATEGY), ("text", 9, zlib.Z_DEFAULT_STRATEGY),
         ("text", 6, zlib.Z_FIXED), ("log", 6, zlib.Z_DEFAULT_STRATEGY), ("png", 6, zlib.Z_DEFAULT_STRATEGY),
         ("png", 1, zlib.Z_DEFAULT_STRATEGY), ("rand", 6, zlib.Z_DEFAULT_STRATEGY), ("zero", 6, zlib.Z_DEFAULT_STRATEGY),
         ("text", 0, zlib.Z_DEFAULT_STRATEGY)]
for kind, level, strat in cases:
    st = synth.make_streams([(kind, size, 100 + i, level, strat) for i in range(uniq)], threads=os.cpu_count())
    zs = [z for _, z in st]
    src_h, so, sl = synth.pack(zs, align=16)
    reps = (nstreams + uniq - 1) // uniq
    t_src = torch.from_numpy(src_h).to(dev).repeat(reps)
    k = np.arange(nstreams, dtype=np.int64)
    batch = d.DeviceBatch(t_src, (k // uniq) * int(src_h.size) + so[k % uniq], sl[k % uniq],
                          torch.zeros(nstreams * size, dtype=torch.uint8, device=dev), k * size, np.full(nstreams, size))
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER
    batch.launch(flags); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); batch.launch(flags); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    res = batch.results_host()
    ok = bool((res["code"] == 0).all())
    print(f"{kind:5s} L{level} strat={strat} ratio {size * uniq / sum(len(z) for z in zs):7.2f}  {nstreams * size / ms / 1e6:8.1f} GB/s  {ms:8.2f} ms ok={ok}", flush=True)
    del batch, t_src
