#!/usr/bin/env python3
"""Small end-to-end run of every kernel family for compute-sanitizer (memcheck / racecheck / initcheck).
usage: compute-sanitizer --tool memcheck python profiles/tools/sanitize_probe.py [small]"""
import json, os, sys, zlib
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import defl_b200 as d
from defl_b200 import synth
import oracle_lib as ol
small = len(sys.argv) > 1 and sys.argv[1] == "small"
meta = json.load(open(os.path.join(ROOT, "tests", "golden", "corpus.json")))
blob = open(os.path.join(ROOT, "tests", "golden", "corpus.bin"), "rb").read()
ents = meta["files"]
datas, caps = [], []
for e in ents[: (40 if small else len(ents))]:
    datas.append(blob[e["off"]: e["off"] + e["len"]]); caps.append(e["cap"])
Z = zlib.Z_DEFAULT_STRATEGY
specs = [("text", 65536, 1, 6, Z), ("log", 90000, 2, 9, Z), ("png", 50000, 3, 1, Z), ("rand", 40000, 4, 6, Z),
         ("zero", 200000, 5, 6, Z), ("text", 50000, 6, 6, zlib.Z_FIXED)]
streams = synth.make_streams(specs, threads=1)
for engine in (3, 4):
    d.set_engine(engine)
    res, outs = d.infl_buf_batch(datas, caps, 0)
    res, outs = d.infl_buf_batch([z for _, z in streams], [len(r) for r, _ in streams], d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
    assert all(int(g["code"]) == 0 and bytes(o) == r for g, o, (r, _) in zip(res, outs, streams))
    # damaged inputs: truncated, bit flips
    bad = []
    rng = np.random.default_rng(3)
    for r, z in streams[:3]:
        bad.append(z[: len(z) // 2])
        b = bytearray(z); b[int(rng.integers(2, len(b) - 4))] ^= 0x10; bad.append(bytes(b))
    d.infl_buf_batch(bad, [300000] * len(bad), d.INFL_ZLIB)
d.set_engine(3)
# split decode (>= 4 MiB compressed), chunk lists, streaming
raw, z = synth.make_streams([("text", (5 if small else 10) << 20, 7, 1, Z)], threads=1)[0]
if len(z) >= 4 << 20:
    res, outs = d.infl_buf_batch([z], [len(raw)], d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
    assert int(res[0]["code"]) == 0 and bytes(outs[0]) == raw
r0, z0 = streams[0]
pieces = [z0[i:i + 7] for i in range(0, len(z0), 7)]
res, outs = d.infl_buf_batch_iov([pieces], [len(r0)], d.INFL_ZLIB)
assert int(res[0]["code"]) == 0 and bytes(outs[0]) == r0
st = d.InflStream(len(r0) + 8, d.INFL_ZLIB)
rc = 2
for o in range(0, len(z0), 997):
    rc = st.stream(z0[o:o + 997])
while rc == 2:
    rc = st.stream(None)
assert rc == 0 and bytes(st.output()) == r0
st.destroy()
print("sanitize probe ok")
