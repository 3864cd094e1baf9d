"""Host-pointer batch of large members (each >= 4 MiB compressed) through infl_buf_batch: wall time per call.
usage: [DEFL_B200_LIB=...] python profiles/tools/large_host_probe.py [members] [MiB each]"""
import sys, time, zlib
sys.path.insert(0, '/root/repo')
import numpy as np, torch
import defl_b200 as d
from defl_b200 import api, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
mib = int(sys.argv[2]) if len(sys.argv) > 2 else 128
streams = synth.make_streams([("text", mib << 20, 100 + i % 2, 6, zlib.Z_DEFAULT_STRATEGY) for i in range(2)], threads=2)
raws = [streams[i % 2][0] for i in range(n)]; zs = [streams[i % 2][1] for i in range(n)]
src = [torch.frombuffer(bytearray(z), dtype=torch.uint8).pin_memory() for z in zs]
dst = [torch.zeros(len(r), dtype=torch.uint8).pin_memory() for r in raws]
items = np.zeros(n, dtype=d.ITEM_DTYPE)
for i in range(n):
    items[i] = (src[i].data_ptr(), len(zs[i]), dst[i].data_ptr(), len(raws[i]))
res = np.zeros(n, dtype=d.RESULT_DTYPE)
flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER
import os
for grp in (os.environ.get("PROBE_GROUPS", "512").split(",")):
    os.environ["DEFL_BIG_GROUP_MIB"] = grp
    for it in range(3):
        t = time.perf_counter(); api.infl_buf_batch_raw(items, flags, res); dt = time.perf_counter() - t
        assert (res["code"] == 0).all()
    print("group %s MiB, members %d x %d MiB: %.1f ms, %.1f GB/s decoded host to host" % (grp, n, mib, dt * 1e3, sum(map(len, raws)) / dt / 1e9))
assert dst[n - 1].numpy().tobytes() == raws[n - 1]
