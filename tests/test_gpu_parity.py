"""GPU parity tests proper: the sm_100a kernels, called through the C ABI (libdefl_b200.so), against
the oracle (oracle/inflate_oracle.c, pinned by tests/test_oracle.py) and the committed golden
fixtures produced by the compiled reference (tests/golden/make_golden.py).

Bar: bit-exact bytes, equal produced/consumed lengths, equal return codes ({0,-1,1})."""
import hashlib
import zlib

import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu

LANES = (0, 1, 2, 4, 8, 16, 32)   # 0 = the default engine (warp per stream); 1..32 = group per stream


def _check_against_port(d, res, outs, datas, caps, flags, what):
    for i, (data, cap) in enumerate(zip(datas, caps)):
        code, out, r = ol.port_inflate(data, cap, flags)
        g = res[i]
        assert int(g["code"]) == code, (what, i, int(g["info"]) >> 8)
        if code == 0:
            assert int(g["outlen"]) == r.outlen and int(g["inpos"]) == r.inpos, (what, i)
            assert bytes(outs[i]) == out, (what, i)
            assert int(g["adler"]) == r.adler, (what, i)
            assert (int(g["info"]) & 1) == r.has_trailer
            if r.has_trailer:
                assert int(g["adler_stored"]) == r.adler_stored


def test_corpus_raw_every_group_width(gpu_lib, gpu_lib_legacy, corpus):
    """the reference's own corpus (raw deflate, flags 0), test_files.c:174-257: the default engine
    of the product library, then every group width of the retired v1 engine (legacy build) as a cross-check."""
    d = gpu_lib
    res, outs = d.infl_buf_batch([e["data"] for e in corpus], [e["cap"] for e in corpus], 0)
    for e, g, out in zip(corpus, res, outs):
        assert int(g["code"]) == e["ref_code"], e["name"]
        if e["ref_code"] == 0:
            assert hashlib.sha256(bytes(out)).hexdigest() == e["ref_sha256"], e["name"]
    with pytest.raises(ValueError):
        d.set_lanes_per_stream(8)                       # not in the product build
    d = gpu_lib_legacy
    datas = [e["data"] for e in corpus]
    caps = [e["cap"] for e in corpus]
    try:
        for lanes in LANES:
            d.set_lanes_per_stream(lanes)
            res, outs = d.infl_buf_batch(datas, caps, 0)
            for e, g, out in zip(corpus, res, outs):
                assert int(g["code"]) == e["ref_code"], (e["name"], lanes, int(g["info"]) >> 8)
                if e["ref_code"] == 0:
                    assert int(g["outlen"]) == e["ref_outlen"], (e["name"], lanes)
                    assert int(g["inpos"]) == e["ref_inpos"], (e["name"], lanes)
                    assert hashlib.sha256(bytes(out)).hexdigest() == e["ref_sha256"], (e["name"], lanes)
                if "raw_len" in e:
                    assert int(g["adler"]) == e["raw_adler32"], (e["name"], lanes)
    finally:
        d.set_lanes_per_stream(0)


def test_corpus_zlib_twins_and_adler(gpu_lib, corpus):
    d = gpu_lib
    es = [e for e in corpus if "raw_len" in e and e["len"]]
    zs = [b"\x78\x9c" + e["data"] + e["raw_adler32"].to_bytes(4, "big") for e in es]
    res, outs = d.infl_buf_batch(zs, [e["raw_len"] for e in es], d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
    for e, z, g, out in zip(es, zs, res, outs):
        assert int(g["code"]) == 0, e["name"]
        assert hashlib.sha256(bytes(out)).hexdigest() == e["raw_sha256"], e["name"]
        assert int(g["adler"]) == int(g["adler_stored"]) == e["raw_adler32"]
        assert int(g["inpos"]) == len(z) - 4
        assert int(g["info"]) & 3 == 3
    # corrupt trailer: the reference's codes by default (UNZ_OK), UNZ_ERR only when verification is asked for
    bad = [z[:-1] + bytes([z[-1] ^ 0x55]) for z in zs[:20]]
    res, _ = d.infl_buf_batch(bad, [e["raw_len"] for e in es[:20]], d.INFL_ZLIB)
    assert all(int(g["code"]) == 0 and (int(g["info"]) & 3) == 1 for g in res)
    res, _ = d.infl_buf_batch(bad, [e["raw_len"] for e in es[:20]], d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
    assert all(int(g["code"]) == -1 for g in res)


def test_kats_through_infl_buf(gpu_lib, kats):
    """Inline known-answer vectors of test_files.c (:468-476, :627-633) via the one-shot call."""
    d = gpu_lib
    for k in kats:
        data = bytes.fromhex(k["hex"])
        rc, out = d.infl_buf(data, k["cap"], k["flags"])
        assert rc == k["ref_code"], k["name"]
        if rc == 0:
            assert bytes(out[: k["ref_outlen"]]).hex() == k["ref_out_hex"], k["name"]


def test_reference_fuzzer_verdicts(gpu_lib, fuzz_ref):
    """test_fuzz.c:138-246 inputs (seed 0x12345678): same verdict and bytes as the compiled reference."""
    d = gpu_lib
    fz = ol.RefFuzz(fuzz_ref["seed"])
    cases = [fz.case(i) for i in range(fuzz_ref["n"])]
    n = 1500
    # capacity of the reference run was 1 MiB per case; outputs are tiny, so give each what it needs + slack
    caps = [max(fuzz_ref["outlens"][i] + 64, 4096) for i in range(n)]
    res, outs = d.infl_buf_batch(cases[:n], caps, 0)
    for i in range(n):
        code = fuzz_ref["codes"][i]
        if code == 0:
            assert int(res[i]["code"]) == 0, i
            assert int(res[i]["outlen"]) == fuzz_ref["outlens"][i], i
            assert hashlib.sha256(bytes(outs[i])).hexdigest()[:16] == fuzz_ref["sha16"][i], i
        else:
            # a failing case may fail earlier on a smaller dst, never succeed
            assert int(res[i]["code"]) == -1, i


def test_mutants_vs_oracle(gpu_lib, gpu_lib_legacy):
    """Valid, truncated, bit-flipped, over-long streams; short / exact / oversized dst; zlib and raw.  Lanes 0 = the
    product library's default engine; 1, 8, 32 = the retired group-per-stream engine of the legacy build."""
    from test_oracle import _mutants
    by_flag = {0: ([], []), 1: ([], [])}
    for data, cap, flags, raw, m in _mutants(777, 2400):
        by_flag[flags][0].append(data)
        by_flag[flags][1].append(cap)
    try:
        for lanes in (0, 1, 8, 32):
            d = gpu_lib if lanes == 0 else gpu_lib_legacy
            d.set_lanes_per_stream(lanes)
            for flags, (datas, caps) in by_flag.items():
                res, outs = d.infl_buf_batch(datas, caps, flags)
                _check_against_port(d, res, outs, datas, caps, flags, ("mutants", lanes, flags))
    finally:
        gpu_lib_legacy.set_lanes_per_stream(0)


def test_chunk_lists(gpu_lib, corpus):
    """infl_init + N x infl_include + infl, piece sizes 1,2,4,8 cycling (test_files.c:260-331)."""
    d = gpu_lib
    names = {"hello", "hello_world", "json", "html", "text_repeated", "zeros_1k", "repeated_a_258", "ascii",
             "huffman_single_a", "distance_test_1", "length_test_3", "bit_align_7",       # test_files.c:1037-1041
             "png_simulation", "multi_block_1", "uncompressed_medium"}
    picked = [e for e in corpus if e["name"] in names and "raw_len" in e]
    assert len(picked) >= 12
    for e in picked:
        data = e["data"]
        st = d.InflStream(e["cap"], 0)
        pos = k = 0
        while pos < len(data):
            n = 1 << (k % 4)
            st.include(data[pos:pos + n])
            pos += n
            k += 1
        assert st.infl() == d.UNZ_OK, e["name"]
        assert st.output_pos() == e["raw_len"]
        assert hashlib.sha256(bytes(st.output())).hexdigest() == e["raw_sha256"], e["name"]
        assert st.input_pos() == e["ref_inpos"]
        st.destroy()
    st = d.InflStream(16, 0)
    assert st.infl() == d.UNZ_NOOP
    st.include(b"")
    st.include(b"")
    assert st.infl() == d.UNZ_OK and st.output_pos() == 0
    st.destroy()


def test_chunk_lists_batch_iov_large(gpu_lib):
    """BASELINE.json configs[2] shape at reduced size: streams split into 1 B .. 32 KiB pieces."""
    from defl_b200 import synth
    d = gpu_lib
    streams = synth.make_streams([("png", 1 << 20, 1, 6, zlib.Z_DEFAULT_STRATEGY),
                                  ("text", 1 << 20, 2, 6, zlib.Z_DEFAULT_STRATEGY),
                                  ("rand", 300000, 3, 6, zlib.Z_DEFAULT_STRATEGY)])
    s = 0x12345678
    lists = []
    for raw, z in streams:
        pieces, pos = [], 0
        while pos < len(z):
            s = (s * 1103515245 + 12345) & 0xFFFFFFFF
            n = 1 + (s >> 8) % 16 if (s >> 4) & 1 else 1 + (s >> 8) % 32768
            pieces.append(z[pos:pos + n])
            pos += n
        lists.append(pieces)
    res, outs = d.infl_buf_batch_iov(lists, [len(r) for r, _ in streams], d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
    for (raw, z), g, out in zip(streams, res, outs):
        assert int(g["code"]) == 0 and bytes(out) == raw
        assert int(g["adler"]) == (zlib.adler32(raw) & 0xFFFFFFFF)


def test_infl_stream_shim(gpu_lib, corpus):
    """infl_stream(): UNZ_UNFINISHED until the final block is in, then UNZ_OK (test_files.c:688-790);
    UNZ_EFULL + infl_resize_output resumes (stream.c:614-627, mem.c:277-288)."""
    d = gpu_lib
    names = {"hello", "hello_world", "json", "zeros_1k", "huffman_single_a", "multi_block_1",
             "dynamic_huffman_1", "distance_test_1", "length_test_3", "bit_align_7", "png_simulation"}
    for e in [x for x in corpus if x["name"] in names and "raw_len" in x]:      # test_files.c:1043-1048
        data = e["data"]
        st = d.InflStream(e["cap"], 0)
        rc, pos, k = d.UNZ_UNFINISHED, 0, 0
        while pos < len(data):
            n = 64 << (k % 5)
            rc = st.stream(data[pos:pos + n])
            pos += n
            k += 1
            if rc != d.UNZ_UNFINISHED:
                break
        polls = 0
        while rc == d.UNZ_UNFINISHED and polls < 5:          # test_files.c:761-767: a few empty calls
            rc = st.stream(None)
            polls += 1
        assert rc == d.UNZ_OK, e["name"]
        assert hashlib.sha256(bytes(st.output())).hexdigest() == e["raw_sha256"], e["name"]
        st.destroy()
    # byte-wise (test_files.c:792-869): one-byte pieces, then at most 32 empty calls; the library must not pay a
    # device round trip per byte
    e = next(x for x in corpus if x["name"] == "png_simulation")
    st = d.InflStream(e["cap"], 0)
    l0 = d.kernel_launch_count()
    rc = d.UNZ_UNFINISHED
    for i in range(len(e["data"])):
        rc = st.stream(e["data"][i:i + 1])
        if rc != d.UNZ_UNFINISHED:
            break
    polls = 0
    while rc == d.UNZ_UNFINISHED and polls < 32:
        rc = st.stream(None)
        polls += 1
    assert rc == d.UNZ_OK and hashlib.sha256(bytes(st.output())).hexdigest() == e["raw_sha256"]
    assert d.kernel_launch_count() - l0 < 2000, "a launch per byte"
    st.destroy()
    e = next(x for x in corpus if x["name"] == "png_simulation")
    data = e["data"]
    st = d.InflStream(e["raw_len"] // 2, 0)
    assert st.stream(data) == d.UNZ_EFULL
    assert st.resize_output(e["raw_len"]) == d.UNZ_OK
    assert st.stream(None) == d.UNZ_OK
    assert hashlib.sha256(bytes(st.output())).hexdigest() == e["raw_sha256"]
    st.destroy()


def test_edge_shapes(gpu_lib, gpu_lib_legacy):
    """Longest match, farthest distance, overlapping short periods, empty/ragged batch members."""
    d = gpu_lib
    rnd = np.random.default_rng(5)
    raws = [
        b"",                                                    # empty payload, valid stream
        b"a",
        b"ab" * 40000,                                           # period 2, len 258 matches
        b"abc" * 30000, b"abcde" * 20000, b"abcdefg" * 12000,    # short periods < 8
        bytes(rnd.integers(0, 256, 32768, dtype=np.uint8)) * 3,  # distance 32768
        bytes(rnd.integers(0, 4, 70000, dtype=np.uint8)),        # short codes, dense matches
        b"\0" * 500000,
    ]
    datas, caps = [], []
    for raw in raws:
        for level, strat in ((6, zlib.Z_DEFAULT_STRATEGY), (9, zlib.Z_DEFAULT_STRATEGY), (1, zlib.Z_FIXED),
                             (0, zlib.Z_DEFAULT_STRATEGY), (6, zlib.Z_HUFFMAN_ONLY), (6, zlib.Z_RLE)):
            c = zlib.compressobj(level, zlib.DEFLATED, 15, 8, strat)
            datas.append(c.compress(raw) + c.flush())
            caps.append(len(raw))
    try:
        for lanes in LANES:
            dd = gpu_lib if lanes == 0 else gpu_lib_legacy
            dd.set_lanes_per_stream(lanes)
            res, outs = dd.infl_buf_batch(datas, caps, dd.INFL_ZLIB)
            _check_against_port(dd, res, outs, datas, caps, 1, ("edges", lanes))
            for i, out in enumerate(outs):
                assert bytes(out) == raws[i // 6]
    finally:
        gpu_lib_legacy.set_lanes_per_stream(0)
    # an empty batch and NULL-sized members
    res, outs = d.infl_buf_batch([b"", zlib.compress(b"xyz")], [4, 3], d.INFL_ZLIB)
    assert int(res[0]["code"]) == 0 and int(res[0]["outlen"]) == 0 and bytes(outs[1]) == b"xyz"


def test_config2_reduced_and_device_resident(gpu_lib):
    """BASELINE.json configs[1] at a size the oracle finishes in seconds: 64 KiB text streams, level 6,
    device-resident batch through infl_batch_decode_device, compared with the oracle byte for byte."""
    import torch
    from defl_b200 import synth
    d = gpu_lib
    n = 600
    streams = synth.make_streams(synth.config2_specs(n, first_seed=10_000))
    zs = [z for _, z in streams]
    src, so, sl = synth.pack(zs)
    dl = np.full(n, 65536, dtype=np.int64)
    do = np.arange(n, dtype=np.int64) * 65536
    tsrc = torch.from_numpy(src).cuda()
    tdst = torch.zeros(n * 65536, dtype=torch.uint8, device="cuda")
    b = d.DeviceBatch(tsrc, so, sl, tdst, do, dl)
    b.launch(d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
    torch.cuda.synchronize()
    res = b.results_host()
    out = tdst.cpu().numpy()
    for i, (raw, z) in enumerate(streams):
        code, o, r = ol.port_inflate(z, 65536, 1)
        assert code == 0 and int(res[i]["code"]) == 0
        assert int(res[i]["adler"]) == r.adler == int(res[i]["adler_stored"])
        assert out[i * 65536:(i + 1) * 65536].tobytes() == o == raw


def test_gather_kernel(gpu_lib):
    """K2: descriptor-driven gather, every alignment and tiny lengths (reference mem.c:121-155 joins on the host)."""
    import torch
    d = gpu_lib
    rnd = np.random.default_rng(11)
    blob = rnd.integers(0, 256, 1 << 20, dtype=np.uint8)
    tb = torch.from_numpy(blob).cuda()
    lens = np.concatenate([rnd.integers(0, 40, 3000), rnd.integers(0, 70000, 40)]).astype(np.uint64)
    srcoff = rnd.integers(0, (1 << 20) - 70000, lens.size).astype(np.uint64)
    dstoff = np.zeros(lens.size, dtype=np.uint64)
    dstoff[1:] = np.cumsum(lens)[:-1]
    total = int(lens.sum())
    out = torch.zeros(total + 16, dtype=torch.uint8, device="cuda")
    t_src = torch.from_numpy((srcoff + np.uint64(tb.data_ptr())).astype(np.int64)).cuda()
    t_len = torch.from_numpy(lens.astype(np.int64)).cuda()
    t_off = torch.from_numpy(dstoff.astype(np.int64)).cuda()
    rc = d.lib().infl_gather_device(t_src.data_ptr(), t_len.data_ptr(), t_off.data_ptr(), lens.size,
                                    out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    want = np.concatenate([blob[int(o):int(o) + int(n)] for o, n in zip(srcoff, lens)])
    assert np.array_equal(out.cpu().numpy()[:total], want)


def test_pipeline_shapes(gpu_lib_legacy):
    """Every compiled (decoder warps, slots, worker warps) variant of the retired lane-decoder pipeline (v2, legacy
    build), mixed payloads."""
    from defl_b200 import synth
    d = gpu_lib_legacy
    specs = []
    for i in range(160):
        kind = ("text", "log", "png", "rand", "zero")[i % 5]
        size = (1000, 65536, 200000, 33000, 5)[(i // 5) % 5]
        level, strat = ((6, zlib.Z_DEFAULT_STRATEGY), (1, zlib.Z_DEFAULT_STRATEGY), (9, zlib.Z_DEFAULT_STRATEGY),
                        (6, zlib.Z_FIXED), (0, zlib.Z_DEFAULT_STRATEGY))[(i // 25) % 5]
        specs.append((kind, size, 500 + i, level, strat))
    streams = synth.make_streams(specs)
    zs = [z for _, z in streams]
    caps = [len(r) for r, _ in streams]
    try:
        d.set_engine(2)
        for shape in range(8):
            d.set_pipeline_shape(shape)
            res, outs = d.infl_buf_batch(zs, caps, d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
            for i, (raw, z) in enumerate(streams):
                assert int(res[i]["code"]) == 0, (shape, i, int(res[i]["info"]) >> 8)
                assert bytes(outs[i]) == raw, (shape, i)
                assert int(res[i]["adler"]) == (zlib.adler32(raw) & 0xFFFFFFFF)
    finally:
        d.set_pipeline_shape(0)
        d.set_engine(3)


def test_split_single_large_stream(gpu_lib):
    """BASELINE.json configs[3] at reduced size: one large stream decoded by many warps from speculative block
    starts (split_decode.cuh).  Bytes, lengths and Adler-32 equal the oracle's; streams the block search cannot
    split (stored-only, fixed-Huffman) and damaged streams come out of the one-warp decoder with the oracle's code."""
    import torch
    from defl_b200 import synth
    d = gpu_lib
    raw, z = synth.single_stream(40 << 20, seed=3, slice_bytes=8 << 20)
    assert zlib.decompress(z) == raw
    # host pointers
    res, outs = d.infl_buf_batch([z], [len(raw)], d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
    g = res[0]
    assert int(g["code"]) == 0 and int(g["info"]) & 4, (int(g["code"]), int(g["info"]))
    assert bytes(outs[0]) == raw
    assert int(g["outlen"]) == len(raw) and int(g["inpos"]) == len(z) - 4
    assert int(g["adler"]) == int(g["adler_stored"]) == (zlib.adler32(raw) & 0xFFFFFFFF)
    # one continuous deflate history (no flush points), other payloads, with a stored stretch in the middle
    parts = [synth.payload("text", 9 << 20, 1), synth.payload("rand", 1 << 20, 2), synth.payload("png", 6 << 20, 3),
             synth.payload("log", 8 << 20, 4)]
    raw2 = b"".join(parts)
    z2 = zlib.compress(raw2, 6)
    code, want, r = ol.port_inflate(z2, len(raw2), 1)
    assert code == 0 and want == raw2
    res, outs = d.infl_buf_batch([z2, z], [len(raw2), len(raw)], d.INFL_ZLIB)
    assert int(res[0]["code"]) == 0 and int(res[0]["info"]) & 4
    assert bytes(outs[0]) == raw2 and bytes(outs[1]) == raw
    assert int(res[0]["adler"]) == r.adler and int(res[0]["inpos"]) == r.inpos
    # the same stream with splitting switched off: identical result record
    res1, outs1 = d.infl_buf_batch([z2], [len(raw2)], d.INFL_ZLIB | d.INFL_BATCH_NO_SPLIT)
    assert bytes(outs1[0]) == raw2 and not (int(res1[0]["info"]) & 4)
    for f in ("code", "adler", "outlen", "inpos", "adler_stored"):
        assert int(res1[0][f]) == int(res[0][f]), f
    # cannot be split: stored only / fixed Huffman; damaged: truncated, one byte short of output, flipped bit
    rnd = synth.payload("rand", 6 << 20, 9)
    c = zlib.compressobj(6, zlib.DEFLATED, 15, 8, zlib.Z_FIXED)
    txt = synth.payload("text", 12 << 20, 5)
    zfix = c.compress(txt) + c.flush()
    bad = bytearray(z2)
    bad[len(bad) // 2] ^= 0x10
    cases = [(zlib.compress(rnd, 6), len(rnd)), (zfix, len(txt)), (z2[: len(z2) * 3 // 4], len(raw2)),
             (z2, len(raw2) - 1), (bytes(bad), len(raw2))]
    res, outs = d.infl_buf_batch([c_[0] for c_ in cases], [c_[1] for c_ in cases], d.INFL_ZLIB)
    for (data, cap), g, out in zip(cases, res, outs):
        code, want, r = ol.port_inflate(data, cap, 1)
        assert int(g["code"]) == code
        if code == 0:
            assert bytes(out) == want and int(g["adler"]) == r.adler


@pytest.mark.gpu
def test_split_many_streams_in_one_batch(gpu_lib):
    """BASELINE.json configs[4] in small: a device-resident batch whose large members would outlast the rest on
    one warp each.  They go through the split decoder TOGETHER (one launch sequence for all of them), the small
    ones and the unsplittable ones (stored, fixed Huffman) through the warp-per-stream kernel; every record
    equals the oracle's."""
    import torch
    from defl_b200 import synth, api
    d = gpu_lib
    Z = zlib.Z_DEFAULT_STRATEGY
    specs = [("text", 3 << 20, 1, 6, Z), ("log", 4 << 20, 2, 9, Z), ("png", 2 << 20, 3, 1, Z), ("text", 1500000, 4, 1, Z),
             ("rand", 1 << 20, 5, 6, Z), ("text", 2 << 20, 6, 6, zlib.Z_FIXED), ("zero", 3 << 20, 7, 6, Z),
             ("text", 700000, 8, 9, Z), ("log", 70000, 9, 6, Z), ("text", 65536, 10, 6, Z), ("png", 900000, 11, 6, Z),
             ("text", 4000000, 12, 6, Z), ("png", 3 << 20, 13, 6, zlib.Z_FIXED), ("log", 3 << 20, 14, 1, zlib.Z_FIXED)]
    streams = synth.make_streams(specs, threads=4)
    dev = torch.device("cuda", 0)
    srcs = [torch.frombuffer(bytearray(z), dtype=torch.uint8).to(dev) for _, z in streams]
    dsts = [torch.zeros(len(r) + 8, dtype=torch.uint8, device=dev) for r, _ in streams]
    items = np.zeros(len(streams), dtype=d.ITEM_DTYPE)
    for i, ((raw, z), s, t) in enumerate(zip(streams, srcs, dsts)):
        items[i] = (s.data_ptr(), len(z), t.data_ptr() + (i % 3), len(raw))      # outputs at odd alignments too
    results = np.zeros(len(streams), dtype=d.RESULT_DTYPE)
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER | d.INFL_BATCH_DEVICE_PTRS
    api.infl_buf_batch_raw(items, flags, results)
    torch.cuda.synchronize()
    nsplit = 0
    for i, ((raw, z), t, g) in enumerate(zip(streams, dsts, results)):
        code, want, r = ol.port_inflate(z, len(raw), 1)
        assert code == 0 and int(g["code"]) == 0, (i, int(g["code"]), int(g["info"]))
        assert int(g["outlen"]) == len(raw) and int(g["inpos"]) == r.inpos
        assert int(g["adler"]) == r.adler == int(g["adler_stored"])
        assert t[(i % 3): (i % 3) + len(raw)].cpu().numpy().tobytes() == raw, i
        nsplit += bool(int(g["info"]) & 4)
        if specs[i][4] == zlib.Z_FIXED:
            # fixed-Huffman streams have no block headers to search for: their chunks start on token boundaries
            # inside the blocks (split_decode.cuh, MODE_FIXED_MID)
            assert int(g["info"]) & 4, ("fixed-Huffman stream was not split", i)
    assert nsplit >= 9, nsplit
    # a damaged member: same verdict as the oracle, the others unharmed
    bad = bytearray(streams[0][1]); bad[len(bad) // 2] ^= 0x40
    tb = torch.frombuffer(bad, dtype=torch.uint8).to(dev)
    items[0]["src"] = tb.data_ptr()
    api.infl_buf_batch_raw(items, flags, results)
    code, _, _ = ol.port_inflate(bytes(bad), len(streams[0][0]), 1)
    exp = code if code != 0 else (0 if zlib.adler32(ol.port_inflate(bytes(bad), len(streams[0][0]), 1)[1]) else 0)
    assert (int(results[0]["code"]) == 0) == (code == 0 and int(results[0]["adler"]) == int(results[0]["adler_stored"]))
    assert all(int(g["code"]) == 0 for g in results[1:])


@pytest.mark.gpu
def test_split_paths_on_damaged_streams(gpu_lib):
    """Mutants of large dynamic and fixed-Huffman streams through the device-resident batch call (they take the
    split decoder, which hands anything it cannot finish to the one-warp decoder): per member the verdict, and for
    accepted members the bytes / length / Adler-32, equal the oracle's."""
    import torch
    from defl_b200 import synth, api
    d = gpu_lib
    Z = zlib.Z_DEFAULT_STRATEGY
    base = synth.make_streams([("text", 2 << 20, 31, 6, Z), ("png", 2 << 20, 32, 6, zlib.Z_FIXED),
                               ("log", 3 << 20, 33, 6, zlib.Z_FIXED), ("text", 1500000, 34, 9, Z)], threads=4)
    rng = np.random.default_rng(99)
    cases = []
    for raw, z in base:
        cases.append((z, len(raw)))                                   # intact
        cases.append((z[: len(z) * 2 // 3], len(raw)))               # truncated
        cases.append((z, len(raw) - 1))                               # one byte short of output
        for _ in range(4):                                            # bit flips anywhere in the body
            b = bytearray(z)
            pos = int(rng.integers(2, len(b) - 4))
            b[pos] ^= 1 << int(rng.integers(0, 8))
            cases.append((bytes(b), len(raw)))
        b = bytearray(z); b[-1] ^= 0xFF                               # damaged Adler trailer
        cases.append((bytes(b), len(raw)))
    dev = torch.device("cuda", 0)
    srcs = [torch.frombuffer(bytearray(z), dtype=torch.uint8).to(dev) for z, _ in cases]
    dsts = [torch.zeros(cap + 8, dtype=torch.uint8, device=dev) for _, cap in cases]
    items = np.zeros(len(cases), dtype=d.ITEM_DTYPE)
    for i, ((z, cap), s, t) in enumerate(zip(cases, srcs, dsts)):
        items[i] = (s.data_ptr(), len(z), t.data_ptr(), cap)
    results = np.zeros(len(cases), dtype=d.RESULT_DTYPE)
    api.infl_buf_batch_raw(items, d.INFL_ZLIB | d.INFL_BATCH_DEVICE_PTRS, results)     # reference semantics: no Adler veto
    torch.cuda.synchronize()
    for i, ((z, cap), t, g) in enumerate(zip(cases, dsts, results)):
        code, want, r = ol.port_inflate(z, cap, 1)
        assert int(g["code"]) == code, (i, int(g["code"]), code, int(g["info"]))
        if code == 0:
            assert int(g["outlen"]) == len(want) and int(g["adler"]) == r.adler, i
            assert t[: len(want)].cpu().numpy().tobytes() == want, i


@pytest.mark.gpu
def test_all_gpus_in_one_process(gpu_lib):
    """INFL_BATCH_ALL_GPUS: one host batch sharded over every visible device inside one process (byte-balanced, one
    host thread and one copy/compute pipeline per device, no exchange).  With one GPU it is the same path with one
    shard.  Results equal the oracle's and do not depend on how many devices took part."""
    import torch
    from defl_b200 import synth
    d = gpu_lib
    specs = synth.config5_specs(24 << 20, first_seed=4242, lo=4096, hi=1 << 20)
    streams = synth.make_streams(specs, threads=4)
    zs = [z for _, z in streams]
    caps = [len(r) for r, _ in streams]
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER
    res1, outs1 = d.infl_buf_batch(zs, caps, flags)
    resn, outsn = d.infl_buf_batch(zs, caps, flags | d.INFL_BATCH_ALL_GPUS)
    for (raw, z), a, b, oa, ob in zip(streams, res1, resn, outs1, outsn):
        assert int(a["code"]) == 0 and int(b["code"]) == 0
        assert bytes(oa) == raw and bytes(ob) == raw
        for f in ("adler", "outlen", "inpos", "adler_stored"):
            assert int(a[f]) == int(b[f]), f
    assert torch.cuda.device_count() >= 1


@pytest.mark.gpu
def test_engine4_parity(gpu_lib, corpus, fuzz_ref):
    """The v4 warp decoder (warp_decode4.cuh: literals+match tokens, shared-memory rings and stage; selectable,
    not the default) against the same pins: the reference's corpus, its fuzzer verdicts, mutants vs the oracle
    and every payload class at sizes with full 32-lane windows."""
    from defl_b200 import synth
    from test_oracle import _mutants
    d = gpu_lib
    try:
        d.set_engine(4)
        assert "engine=4" in d.engine_info()
        datas = [e["data"] for e in corpus]
        caps = [e["cap"] for e in corpus]
        res, outs = d.infl_buf_batch(datas, caps, 0)
        for e, g, out in zip(corpus, res, outs):
            assert int(g["code"]) == e["ref_code"], (e["name"], int(g["info"]) >> 8)
            if e["ref_code"] == 0:
                assert int(g["outlen"]) == e["ref_outlen"] and int(g["inpos"]) == e["ref_inpos"], e["name"]
                assert hashlib.sha256(bytes(out)).hexdigest() == e["ref_sha256"], e["name"]
        fz = ol.RefFuzz(fuzz_ref["seed"])
        n = 1000
        cases = [fz.case(i) for i in range(n)]
        fcaps = [max(fuzz_ref["outlens"][i] + 64, 4096) for i in range(n)]
        res, outs = d.infl_buf_batch(cases, fcaps, 0)
        for i in range(n):
            if fuzz_ref["codes"][i] == 0:
                assert int(res[i]["code"]) == 0 and int(res[i]["outlen"]) == fuzz_ref["outlens"][i], i
                assert hashlib.sha256(bytes(outs[i])).hexdigest()[:16] == fuzz_ref["sha16"][i], i
            else:
                assert int(res[i]["code"]) != 0, i
        for flags in (0, 1):
            ms = [(dd, cap) for dd, cap, fl, raw, m in _mutants(2024, 1200) if fl == flags]
            res, outs = d.infl_buf_batch([m[0] for m in ms], [m[1] for m in ms], flags)
            _check_against_port(d, res, outs, [m[0] for m in ms], [m[1] for m in ms], flags, "engine4 mutants")
        Z = zlib.Z_DEFAULT_STRATEGY
        specs = [("text", 65536, 1, 6, Z), ("text", 900000, 2, 9, Z), ("log", 700000, 3, 6, Z), ("png", 400000, 4, 1, Z),
                 ("rand", 300000, 5, 6, Z), ("zero", 2000000, 6, 6, Z), ("text", 500000, 7, 6, zlib.Z_FIXED),
                 ("png", 300000, 8, 6, zlib.Z_HUFFMAN_ONLY), ("log", 400000, 9, 1, zlib.Z_RLE), ("text", 3, 10, 6, Z)]
        streams = synth.make_streams(specs, threads=4)
        res, outs = d.infl_buf_batch([z for _, z in streams], [len(r) for r, _ in streams],
                                     d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
        for (raw, z), g, out in zip(streams, res, outs):
            assert int(g["code"]) == 0, int(g["info"]) >> 8
            assert bytes(out) == raw and int(g["adler"]) == (zlib.adler32(raw) & 0xFFFFFFFF)
            assert int(g["inpos"]) == len(z) - 4
    finally:
        d.set_engine(3)


@pytest.mark.gpu
def test_chunk_lists_scattered_host_pieces_and_device_pieces(gpu_lib):
    """infl_buf_batch_iov with pieces that are NOT neighbours in memory: (a) host pieces in separate allocations,
    handed over in shuffled allocation order (small ones ride the pinned bounce buffer, large ones one DMA each);
    (b) INFL_BATCH_DEVICE_PTRS with about a million device-resident pieces (1-byte pieces included), so
    gather_pieces_kernel sees real descriptor counts.  Reference semantics: the stream is the concatenation of the
    pieces in call order (mem.c:114-216)."""
    import torch
    from defl_b200 import synth, api
    d = gpu_lib
    Z = zlib.Z_DEFAULT_STRATEGY
    streams = synth.make_streams([("text", 300000, 61, 6, Z), ("png", 200000, 62, 6, Z), ("log", 250000, 63, 9, Z),
                                  ("rand", 50000, 64, 6, Z)], threads=2)
    rng = np.random.default_rng(12)
    # (a) host: every piece its own numpy allocation, created in random order so that neighbours in the stream are
    # not neighbours on the heap
    piece_lists = []
    for raw, z in streams:
        cuts, at = [], 0
        while at < len(z):
            n = int(rng.choice([1, 3, 17, 200, 5000, 9000, 40000]))
            cuts.append((at, min(n, len(z) - at)))
            at += n
        order = rng.permutation(len(cuts))
        bufs = [None] * len(cuts)
        for j in order:
            a, n = cuts[j]
            bufs[j] = np.frombuffer(z[a:a + n], dtype=np.uint8).copy()
        piece_lists.append(bufs)
    res, outs = d.infl_buf_batch_iov(piece_lists, [len(r) for r, _ in streams], d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER)
    for (raw, z), g, out in zip(streams, res, outs):
        assert int(g["code"]) == 0 and bytes(out) == raw, int(g["info"]) >> 8
        assert int(g["inpos"]) == len(z) - 4
    # (b) device pieces: stream 0 and 2 as 1-byte pieces (about 215 000), stream 1 and 3 as 1..64-byte pieces
    dev = torch.device("cuda", 0)
    keep, items, caps = [], np.zeros(len(streams), dtype=api.IOV_ITEM_DTYPE), []
    total_pieces = 0
    reps = 7                                                      # the same four streams seven times: > 10^6 pieces
    items = np.zeros(len(streams) * reps, dtype=api.IOV_ITEM_DTYPE)
    dsts = []
    for r in range(reps):
        for i, (raw, z) in enumerate(streams):
            t = torch.frombuffer(bytearray(z), dtype=torch.uint8).to(dev)
            if i % 2 == 0:
                sizes = np.ones(len(z), dtype=np.int64)
            else:
                sizes = []
                left = len(z)
                while left:
                    n = min(left, int(rng.integers(1, 65)))
                    sizes.append(n)
                    left -= n
                sizes = np.array(sizes, dtype=np.int64)
            iov = np.zeros(len(sizes), dtype=api.IOVEC_DTYPE)
            iov["ptr"] = t.data_ptr() + np.concatenate([[0], np.cumsum(sizes)])[:-1]
            iov["len"] = sizes
            dst = torch.zeros(len(raw) + 8, dtype=torch.uint8, device=dev)
            keep += [t, iov]
            dsts.append(dst)
            items[r * len(streams) + i] = (iov.ctypes.data, len(sizes), dst.data_ptr(), len(raw))
            total_pieces += len(sizes)
    assert total_pieces > 1_000_000
    results = np.zeros(len(items), dtype=d.RESULT_DTYPE)
    rc = d.lib().infl_buf_batch_iov(items.ctypes.data, len(items), int(d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER | d.INFL_BATCH_DEVICE_PTRS),
                                    results.ctypes.data)
    assert rc == 0, d.last_error()
    torch.cuda.synchronize()
    for k, g in enumerate(results):
        raw = streams[k % len(streams)][0]
        assert int(g["code"]) == 0 and int(g["outlen"]) == len(raw), (k, int(g["info"]) >> 8)
        assert dsts[k][: len(raw)].cpu().numpy().tobytes() == raw, k


@pytest.mark.gpu
def test_degenerate_codes_and_window_shapes(gpu_lib):
    """The shapes that stress the speculative pass's bookkeeping: codes of one or two bits (a lane's token segment fills
    long before its sub-sequence ends, warm-up tokens included), Huffman-only / RLE / fixed-code streams, blocks of a
    few hundred bytes (sub-sequences shorter than the warm-up), every size around the window and checkpoint grains,
    tight and exact output capacities.  Against the oracle port: bytes, lengths, codes, Adler-32."""
    d = gpu_lib
    rng = np.random.default_rng(77)
    payloads = []
    for n in (1, 2, 3, 7, 8, 9, 63, 64, 65, 255, 256, 257, 1023, 1024, 1025, 4095, 4096, 4097, 65535, 65536, 1 << 20):
        payloads.append(bytes(n))                                               # one symbol: 1-bit codes
        payloads.append(bytes(rng.integers(0, 2, n, dtype=np.uint8)))           # two symbols
        payloads.append(bytes(rng.integers(0, 4, n, dtype=np.uint8) * 85))      # four symbols: 2-bit codes
        payloads.append(bytes(rng.integers(0, 256, n, dtype=np.uint8)))         # incompressible
    datas, caps = [], []
    for p in payloads:
        for strategy in (zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE, zlib.Z_FIXED, zlib.Z_DEFAULT_STRATEGY):
            c = zlib.compressobj(6, zlib.DEFLATED, -15, 8, strategy)
            datas.append(c.compress(p) + c.flush()); caps.append(len(p))
    # many small blocks: a full flush every 100 .. 700 bytes of text-like data
    words = [bytes(rng.integers(97, 123, int(k), dtype=np.uint8)) for k in rng.integers(2, 9, 400)]
    text = b" ".join(words[int(i)] for i in rng.integers(0, 400, 60000))
    c = zlib.compressobj(6, zlib.DEFLATED, -15)
    parts, at = [], 0
    while at < len(text):
        k = int(rng.integers(100, 700))
        parts.append(c.compress(text[at:at + k]) + c.flush(zlib.Z_FULL_FLUSH)); at += k
    parts.append(c.flush())
    datas.append(b"".join(parts)); caps.append(len(text))
    # capacities one short, exact, and with slack
    more_d, more_c = [], []
    for dd, cc in list(zip(datas, caps))[::7]:
        if cc > 1:
            more_d += [dd, dd]; more_c += [cc - 1, cc + 100]
    datas += more_d; caps += more_c
    res, outs = d.infl_buf_batch(datas, caps, 0)
    _check_against_port(d, res, outs, datas, caps, 0, "degenerate")


@pytest.mark.gpu
def test_split_replay_pieces_on_and_off(gpu_lib):
    """The C pass replays a chunk's kept tokens whole or in pieces cut at decode-window ends (DEFL_SPLIT_PIECES = 0 / 1;
    2 = by the fill rule): the same bytes, Adler-32 and positions either way, with stored blocks inside the streams
    (their bytes are kept as literal tokens), on streams with and without flush points and at odd output alignments."""
    import os
    import torch
    from defl_b200 import synth, api
    d = gpu_lib
    rng = np.random.default_rng(5)
    raws = []
    for seed, n in ((21, 9 << 20), (22, 14 << 20), (23, 6 << 20)):
        raws.append(synth.make_streams([("text", n, seed, 6, zlib.Z_DEFAULT_STRATEGY)], threads=2)[0][0])
    streams = []
    # (a) one deflate history; (b) full flushes every ~1 MiB (empty stored blocks); (c) incompressible stretches inside
    # compressible text: real stored blocks between dynamic ones
    streams.append((raws[0], zlib.compress(raws[0], 6)))
    c = zlib.compressobj(6)
    z = b"".join(c.compress(raws[1][i:i + (1 << 20)]) + c.flush(zlib.Z_FULL_FLUSH) for i in range(0, len(raws[1]), 1 << 20)) + c.flush()
    streams.append((raws[1], z))
    mix = bytearray(raws[2])
    for at in range(1 << 20, len(mix) - (300 << 10), 1 << 20):
        mix[at:at + (200 << 10)] = rng.integers(0, 256, 200 << 10, dtype=np.uint8).tobytes()
    mix = bytes(mix)
    streams.append((mix, zlib.compress(mix, 6)))
    dev = torch.device("cuda", 0)
    srcs = [torch.frombuffer(bytearray(z), dtype=torch.uint8).to(dev) for _, z in streams]
    flags = d.INFL_ZLIB | d.INFL_BATCH_VERIFY_ADLER | d.INFL_BATCH_DEVICE_PTRS
    old = os.environ.get("DEFL_SPLIT_PIECES")
    try:
        for mode in ("0", "1", "2"):
            os.environ["DEFL_SPLIT_PIECES"] = mode
            dsts = [torch.zeros(len(r) + 8, dtype=torch.uint8, device=dev) for r, _ in streams]
            items = np.zeros(len(streams), dtype=d.ITEM_DTYPE)
            for i, ((raw, z), s, t) in enumerate(zip(streams, srcs, dsts)):
                items[i] = (s.data_ptr(), len(z), t.data_ptr() + (i % 3), len(raw))
            results = np.zeros(len(streams), dtype=d.RESULT_DTYPE)
            api.infl_buf_batch_raw(items, flags, results)
            torch.cuda.synchronize()
            for i, ((raw, z), t, g) in enumerate(zip(streams, dsts, results)):
                assert int(g["code"]) == 0 and int(g["info"]) & 4, (mode, i, int(g["code"]), hex(int(g["info"])))
                assert int(g["outlen"]) == len(raw) and int(g["inpos"]) == len(z) - 4, (mode, i)   # the trailer is not deflate data
                assert int(g["adler"]) == (zlib.adler32(raw) & 0xFFFFFFFF) == int(g["adler_stored"]), (mode, i)
                assert t[(i % 3): (i % 3) + len(raw)].cpu().numpy().tobytes() == raw, (mode, i)
    finally:
        if old is None:
            os.environ.pop("DEFL_SPLIT_PIECES", None)
        else:
            os.environ["DEFL_SPLIT_PIECES"] = old


@pytest.mark.gpu
def test_large_host_members_pipelined(gpu_lib):
    """Host-pointer batches whose members are >= 4 MiB compressed go through the split decoder one after another,
    three slots deep (copy in / decode / copy home overlap): five such members between small ones, one of them damaged,
    outputs at their own host buffers -- every record and byte as the oracle has it."""
    from defl_b200 import synth
    d = gpu_lib
    specs = [("text", 15 << 20, 31, 6, zlib.Z_DEFAULT_STRATEGY), ("text", 70000, 32, 6, zlib.Z_DEFAULT_STRATEGY),
             ("text", 17 << 20, 33, 6, zlib.Z_DEFAULT_STRATEGY), ("png", 9 << 20, 34, 6, zlib.Z_DEFAULT_STRATEGY),
             ("log", 3000, 35, 9, zlib.Z_DEFAULT_STRATEGY), ("text", 16 << 20, 36, 1, zlib.Z_DEFAULT_STRATEGY),
             ("text", 15 << 20, 37, 6, zlib.Z_DEFAULT_STRATEGY)]
    streams = synth.make_streams(specs, threads=4)
    datas = [z for _, z in streams]
    big = [i for i, z in enumerate(datas) if len(z) >= (4 << 20)]
    assert len(big) >= 4, [len(z) for z in datas]
    bad = bytearray(datas[big[1]]); bad[len(bad) // 3] ^= 0x10
    datas[big[1]] = bytes(bad)
    caps = [len(r) for r, _ in streams]
    for _ in range(2):                                   # the second call reuses the slots
        res, outs = d.infl_buf_batch(datas, caps, d.INFL_ZLIB)
        _check_against_port(d, res, outs, datas, caps, 1, "large host members")
        assert sum(bool(int(g["info"]) & 4) for g in res) >= len(big) - 1
