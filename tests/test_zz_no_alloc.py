"""GPU: infl_batch_decode_device_static -- the batch kernel with its token scratch in a file-scope device array,
no allocation on the call path (include/defl/infl_batch.h; no counterpart in the reference).
Collected last on purpose.

Its output must equal infl_batch_decode_device's on the same device-resident batch, bit for bit (byte work: the
tolerance is zero), on two calls in a row: the array keeps the previous call's contents, so the second call shows
that what must start at zero is cleared.  Every call is made through the C ABI (ctypes)."""
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(d, torch, streams, cap):
    from defl_b200 import synth
    zs = [z for _, z in streams]
    n = len(zs)
    src, so, sl = synth.pack(zs)
    dl = np.full(n, cap, dtype=np.int64)
    do = np.arange(n, dtype=np.int64) * cap
    tsrc = torch.from_numpy(src).cuda()
    tdst = torch.zeros(n * cap, dtype=torch.uint8, device="cuda")
    return d.DeviceBatch(tsrc, so, sl, tdst, do, dl), tdst


def _launch_static(d, torch, b, flags):
    st = torch.cuda.current_stream().cuda_stream
    rc = d.lib().infl_batch_decode_device_static(b.items.data_ptr(),
                                                 b.order.data_ptr() if b.order is not None else None,
                                                 b.n, int(flags), b.results.data_ptr(), st)
    assert rc == d.UNZ_OK, d.last_error()


def test_static_scratch_equals_workspace_path_twice(gpu_lib):
    import torch
    from defl_b200 import synth
    d = gpu_lib
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER
    # two different batches, so that call 2 finds call 1's tokens and ticket in the shared array:
    # (1) 700 text streams of 64 KiB (more streams than the 148 resident warps: the ticket counter matters),
    # (2) mixed classes incl. fixed-Huffman, stored (random) and long-match (zero) data, plus a damaged stream
    b1_streams = synth.make_streams(synth.config2_specs(700, first_seed=77_000))
    specs = []
    for i in range(60):
        kind = ("text", "log", "png", "rand", "zero")[i % 5]
        strat = zlib.Z_FIXED if i % 7 == 3 else zlib.Z_DEFAULT_STRATEGY
        specs.append((kind, 30000 + 977 * i, 88_000 + i, (1, 6, 9)[i % 3], strat))
    b2_streams = synth.make_streams(specs)
    raw, z = b2_streams[5]
    bad = bytearray(z)
    bad[len(bad) // 2] ^= 0x5A
    b2_streams[5] = (raw, bytes(bad))

    for streams, cap in ((b1_streams, 65536), (b2_streams, 131072)):
        want_b, want_dst = _batch(d, torch, streams, cap)
        want_b.launch(flags)                                   # existing entry point, caller's workspace
        torch.cuda.synchronize()
        want_res = want_b.results_host().copy()
        want_out = want_dst.cpu().numpy()

        got_b, got_dst = _batch(d, torch, streams, cap)
        mem0 = torch.cuda.memory_allocated()
        _launch_static(d, torch, got_b, flags)                 # new entry point: no workspace argument
        assert torch.cuda.memory_allocated() == mem0
        torch.cuda.synchronize()
        got_res = got_b.results_host()
        got_out = got_dst.cpu().numpy()

        for f in ("code", "adler", "outlen", "inpos", "adler_stored", "info"):
            assert np.array_equal(got_res[f], want_res[f]), f
        ok = 0
        for i, (raw, _) in enumerate(streams):
            n = int(want_res[i]["outlen"])
            assert np.array_equal(got_out[i * cap: i * cap + n], want_out[i * cap: i * cap + n]), i
            if int(want_res[i]["code"]) == 0:
                assert got_out[i * cap: i * cap + n].tobytes() == raw
                ok += 1
        assert ok >= len(streams) - 1


def test_static_scratch_rejects_oversize(gpu_lib):
    import torch
    d = gpu_lib
    items = torch.zeros((1, 4), dtype=torch.int64, device="cuda")
    res = torch.zeros((1, 32), dtype=torch.uint8, device="cuda")
    L = d.lib()
    rc = L.infl_batch_decode_device_static(items.data_ptr(), None, d.INFL_BATCH_STATIC_MAX_STREAMS + 1, 0,
                                           res.data_ptr(), None)
    assert rc == d.UNZ_ERR and "INFL_BATCH_STATIC_MAX_STREAMS" in d.last_error()
    assert L.infl_batch_decode_device_static(None, None, 3, 0, res.data_ptr(), None) == d.UNZ_ERR
    assert L.infl_batch_decode_device_static(None, None, 0, 0, None, None) == d.UNZ_OK      # empty batch: nothing to do
