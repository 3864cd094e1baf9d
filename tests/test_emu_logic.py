"""CPU: the decode core (defl_b200/csrc/inflate_core.cuh) run by a lock-step SIMT emulator
(tests/emu/) against the oracle.  A debug aid for the container without a GPU -- it checks the
kernel's *logic* for every group width; the parity tests proper are the -m gpu ones, which run
the real sm_100a code through the C ABI."""
import ctypes
import os
import random
import subprocess
import zlib

import pytest

import oracle_lib as ol

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_DIR = os.path.join(HERE, "emu")
EMU_SO = os.path.join(EMU_DIR, "libemu.so")
CSRC = os.path.join(os.path.dirname(HERE), "defl_b200", "csrc")
CORE = os.path.join(CSRC, "inflate_core.cuh")


class StreamOut(ctypes.Structure):
    _fields_ = [("code", ctypes.c_int32), ("adler", ctypes.c_uint32), ("outlen", ctypes.c_uint32),
                ("inpos", ctypes.c_uint32), ("adler_stored", ctypes.c_uint32), ("info", ctypes.c_uint32)]


def load_emu():
    srcs = [os.path.join(EMU_DIR, "emu_main.cpp"), os.path.join(EMU_DIR, "simt_emu.h"), CORE,
            os.path.join(CSRC, "pipeline.cuh"), os.path.join(CSRC, "warp_decode.cuh"),
            os.path.join(CSRC, "warp_decode4.cuh"), os.path.join(CSRC, "split_decode.cuh")]
    if not os.path.exists(EMU_SO) or any(os.path.getmtime(s) > os.path.getmtime(EMU_SO) for s in srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-fPIC", "-shared", "-pthread", "-w",
                               "-o", EMU_SO, srcs[0]])
    return ctypes.CDLL(EMU_SO)


@pytest.fixture(scope="module")
def emu():
    return load_emu()


def run(emu, G, data, cap, flags, mis=0):
    buf = ctypes.create_string_buffer(len(data) + 16)
    base = ctypes.addressof(buf)
    off = (-base) % 4 + mis
    ctypes.memmove(base + off, data, len(data))
    dst = ctypes.create_string_buffer(cap + 16)
    o = StreamOut()
    rc = emu.emu_inflate(G, ctypes.c_void_p(base + off), len(data), dst, cap, flags, ctypes.byref(o))
    return rc, dst.raw[: o.outlen], o


def same(rc, out, o, rc0, out0, r0):
    if rc != rc0:
        return False
    return rc != 0 or (out == out0 and o.inpos == r0.inpos and o.adler == r0.adler and
                       o.adler_stored == r0.adler_stored and (o.info & 1) == r0.has_trailer)


def test_corpus_logic(emu, corpus):
    for i, e in enumerate(corpus):
        rc0, out0, r0 = ol.port_inflate(e["data"], e["cap"], 0)
        for G in (1, 8) if i % 3 else (2, 32):
            rc, out, o = run(emu, G, e["data"], e["cap"], 0, mis=i % 4)
            assert same(rc, out, o, rc0, out0, r0), (e["name"], G, o.info >> 8)


def test_mutants_logic(emu):
    from test_oracle import _mutants
    for it, (d, cap, flags, raw, m) in enumerate(_mutants(4242, 500)):
        rc0, out0, r0 = ol.port_inflate(d, cap, flags)
        G = (1, 2, 4, 8, 16, 32)[it % 6]
        rc, out, o = run(emu, G, d, cap, flags, mis=it % 4)
        assert same(rc, out, o, rc0, out0, r0), (it, G, m, o.info >> 8)


# ---- the lane-decoder pipeline (defl_b200/csrc/pipeline.cuh): one emulated CTA ------------------------
import numpy as np

ITEM = np.dtype([("src", "<u8"), ("srclen", "<u8"), ("dst", "<u8"), ("dstlen", "<u8")])
RESULT = np.dtype([("code", "<i4"), ("adler", "<u4"), ("outlen", "<u8"), ("inpos", "<u8"),
                   ("adler_stored", "<u4"), ("info", "<u4")])


def run_pipeline(emu, shape, datas, caps, flags):
    D, SL, K = shape
    n = len(datas)
    srcs = [np.frombuffer(bytes(d) + b"\0" * 8, dtype=np.uint8).copy() for d in datas]
    dsts = [np.zeros(c + 16, dtype=np.uint8) for c in caps]
    items = np.zeros(n, dtype=ITEM)
    for i in range(n):
        items[i] = (srcs[i].ctypes.data + (i % 4 if len(datas[i]) else 0) * 0, len(datas[i]), dsts[i].ctypes.data, caps[i])
    res = np.full(n, -7, dtype=np.int32).repeat(8).view(RESULT).copy()
    emu.emu_pipeline(D, SL, K, ctypes.c_void_p(items.ctypes.data), n, flags, ctypes.c_void_p(res.ctypes.data))
    return res, [bytes(dsts[i][: int(res[i]["outlen"])]) for i in range(n)]


def check_pipeline(emu, shape, datas, caps, flags, what):
    res, outs = run_pipeline(emu, shape, datas, caps, flags)
    for i, (d, cap) in enumerate(zip(datas, caps)):
        rc0, out0, r0 = ol.port_inflate(d, cap, flags)
        g = res[i]
        assert int(g["code"]) == rc0, (what, i, int(g["info"]) >> 8, rc0)
        if rc0 == 0:
            assert outs[i] == out0, (what, i)
            assert int(g["inpos"]) == r0.inpos and int(g["adler"]) == r0.adler, (what, i)
            assert (int(g["info"]) & 1) == r0.has_trailer


def test_pipeline_corpus_logic(emu, corpus):
    datas = [e["data"] for e in corpus]
    caps = [e["cap"] for e in corpus]
    check_pipeline(emu, (1, 32, 3), datas, caps, 0, "corpus")


def test_pipeline_mutants_logic(emu):
    from test_oracle import _mutants
    by = {0: ([], []), 1: ([], [])}
    for d, cap, flags, raw, m in _mutants(31337, 400):
        by[flags][0].append(d)
        by[flags][1].append(cap)
    check_pipeline(emu, (2, 5, 3), by[0][0], by[0][1], 0, "mut0")
    check_pipeline(emu, (1, 32, 2), by[1][0], by[1][1], 1, "mut1")


# ---- the warp-per-stream decoder (defl_b200/csrc/warp_decode.cuh): one emulated warp ------------------
def run_warp(emu, data, cap, flags, mis=0):
    buf = ctypes.create_string_buffer(len(data) + 16)
    base = ctypes.addressof(buf)
    off = (-base) % 4 + mis
    ctypes.memmove(base + off, data, len(data))
    dst = ctypes.create_string_buffer(cap + 16)
    o = StreamOut()
    rc = emu.emu_warp_inflate(ctypes.c_void_p(base + off), len(data), dst, cap, flags, ctypes.byref(o))
    return rc, dst.raw[: o.outlen], o


def test_warp_corpus_logic(emu, corpus):
    for i, e in enumerate(corpus):
        rc0, out0, r0 = ol.port_inflate(e["data"], e["cap"], 0)
        rc, out, o = run_warp(emu, e["data"], e["cap"], 0, mis=i % 4)
        assert same(rc, out, o, rc0, out0, r0), (e["name"], o.info >> 8)


def test_warp_mutants_logic(emu):
    from test_oracle import _mutants
    for it, (d, cap, flags, raw, m) in enumerate(_mutants(9001, 500)):
        rc0, out0, r0 = ol.port_inflate(d, cap, flags)
        rc, out, o = run_warp(emu, d, cap, flags, mis=it % 4)
        assert same(rc, out, o, rc0, out0, r0), (it, m, o.info >> 8)


def test_warp_large_streams_logic(emu):
    """Streams long enough for full 32-lane windows, several blocks, every payload class."""
    from defl_b200 import synth
    specs = [("text", 300000, 1, 6, zlib.Z_DEFAULT_STRATEGY), ("log", 200000, 2, 9, zlib.Z_DEFAULT_STRATEGY),
             ("png", 150000, 3, 1, zlib.Z_DEFAULT_STRATEGY), ("rand", 70000, 4, 6, zlib.Z_DEFAULT_STRATEGY),
             ("zero", 600000, 5, 6, zlib.Z_DEFAULT_STRATEGY), ("text", 120000, 6, 6, zlib.Z_FIXED),
             ("text", 65536, 7, 6, zlib.Z_DEFAULT_STRATEGY), ("png", 65536, 8, 6, zlib.Z_HUFFMAN_ONLY)]
    for k, (raw, z) in enumerate(synth.make_streams(specs, threads=1)):
        rc0, out0, r0 = ol.port_inflate(z, len(raw), 1)
        assert rc0 == 0 and out0 == raw
        rc, out, o = run_warp(emu, z, len(raw), 1, mis=k % 4)
        assert same(rc, out, o, rc0, out0, r0), (k, o.info >> 8)
        # one byte short of output space, and a truncated input: same verdict as the oracle
        for d, cap in ((z, len(raw) - 1), (z[: len(z) * 2 // 3], len(raw))):
            rc0, out0, r0 = ol.port_inflate(d, cap, 1)
            rc, out, o = run_warp(emu, d, cap, 1)
            assert rc == rc0 == -1, (k, cap)


# ---- split decode of one large stream (defl_b200/csrc/split_decode.cuh): chunks, markers ---------------
def test_split_chunks_and_markers_logic(emu):
    """Chunks that start at true block boundaries (from the oracle's block index), decoded in isolation into
    16-bit symbols with markers for the unknown window, then marker resolution: bytes equal the oracle's."""
    from defl_b200 import synth
    for kind, size, level in (("text", 700000, 6), ("log", 900000, 6), ("png", 400000, 1)):
        raw, z = synth.make_streams([(kind, size, 21, level, zlib.Z_DEFAULT_STRATEGY)], threads=1)[0]
        bits, outs_at = ol.port_block_index(z, 1)
        assert len(bits) >= 6
        # every other block start is a candidate (stream-relative bits: the index counts from the body)
        cand = [bits[0] + 16] + [b + 16 for b in bits[2::2]]
        arr = (ctypes.c_uint64 * len(cand))(*cand)
        buf = ctypes.create_string_buffer(len(z) + 16)
        base = ctypes.addressof(buf)
        off = (-base) % 4
        ctypes.memmove(base + off, z, len(z))
        dst = ctypes.create_string_buffer(len(raw) + 16)
        end_bit = ctypes.c_uint64(0)
        emu.emu_split_inflate.restype = ctypes.c_long
        n = emu.emu_split_inflate(ctypes.c_void_p(base + off), len(z), arr, len(cand), dst, len(raw), ctypes.byref(end_bit))
        assert n == len(raw), (kind, n)
        assert dst.raw[:n] == raw, kind
        assert (end_bit.value + 7) // 8 == len(z) - 4
    emu.emu_split_replayed.restype = ctypes.c_long
    assert emu.emu_split_replayed() > 0      # some chunks took the kept-token replay path


def test_split_mid_block_fixed_logic(emu):
    """Fixed-Huffman streams (zlib's Z_FIXED: one block, or a few) have no block starts to split at: chunks start
    at token boundaries found by the probe, link where a chunk's trajectory meets a later candidate, and the chain
    walk carries whether the block they start in is the stream's last.  Bytes equal the oracle's."""
    from defl_b200 import synth
    emu.emu_split_inflate_mid.restype = ctypes.c_long
    before = emu.emu_split_replayed()
    for kind, size, granule, probe in (("text", 260000, 16384, 16384), ("log", 200000, 8192, 4096),
                                       ("png", 150000, 16384, 16384)):
        raw, z = synth.make_streams([(kind, size, 33, 6, zlib.Z_FIXED)], threads=1)[0]
        rc0, out0, r0 = ol.port_inflate(z, len(raw), 1)
        assert rc0 == 0 and out0 == raw
        buf = ctypes.create_string_buffer(len(z) + 16)
        base = ctypes.addressof(buf)
        off = (-base) % 4 + 1
        ctypes.memmove(base + off, z, len(z))
        dst = ctypes.create_string_buffer(len(raw) + 16)
        end_bit = ctypes.c_uint64(0)
        ncand = ctypes.c_uint32(0)
        nchain = ctypes.c_uint32(0)
        n = emu.emu_split_inflate_mid(ctypes.c_void_p(base + off), len(z), 2, granule, probe, dst, len(raw),
                                      ctypes.byref(end_bit), ctypes.byref(ncand), ctypes.byref(nchain))
        assert n == len(raw), (kind, n, ncand.value, nchain.value)
        assert dst.raw[:n] == raw, kind
        assert (end_bit.value + 7) // 8 == len(z) - 4
        assert ncand.value >= 4 and nchain.value >= ncand.value // 2, (kind, ncand.value, nchain.value)
    assert emu.emu_split_replayed() > before


# ---- the v4 warp decoder (defl_b200/csrc/warp_decode4.cuh): one emulated warp ------------------------------
def run_warp4(emu, data, cap, flags, mis=0, dmis=0):
    buf = ctypes.create_string_buffer(len(data) + 32)
    base = ctypes.addressof(buf)
    off = (-base) % 16 + mis
    ctypes.memmove(base + off, data, len(data))
    dst = ctypes.create_string_buffer(cap + 64)
    dbase = ctypes.addressof(dst)
    doff = (-dbase) % 16 + dmis
    o = StreamOut()
    rc = emu.emu_warp4_inflate(ctypes.c_void_p(base + off), len(data), ctypes.c_void_p(dbase + doff), cap, flags,
                               ctypes.byref(o))
    return rc, dst.raw[doff: doff + o.outlen], o


def test_warp4_corpus_and_mutants_logic(emu, corpus):
    from test_oracle import _mutants
    for i, e in enumerate(corpus):
        rc0, out0, r0 = ol.port_inflate(e["data"], e["cap"], 0)
        rc, out, o = run_warp4(emu, e["data"], e["cap"], 0, mis=i % 16, dmis=(i * 5) % 16)
        assert same(rc, out, o, rc0, out0, r0), (e["name"], o.info >> 8)
    for it, (d, cap, flags, raw, m) in enumerate(_mutants(9004, 250)):
        rc0, out0, r0 = ol.port_inflate(d, cap, flags)
        rc, out, o = run_warp4(emu, d, cap, flags, mis=it % 16, dmis=(it * 3) % 16)
        assert same(rc, out, o, rc0, out0, r0), (it, m, o.info >> 8)


def test_warp4_large_streams_logic(emu):
    from defl_b200 import synth
    specs = [("text", 200000, 1, 6, zlib.Z_DEFAULT_STRATEGY), ("log", 120000, 2, 9, zlib.Z_DEFAULT_STRATEGY),
             ("png", 60000, 3, 1, zlib.Z_DEFAULT_STRATEGY), ("rand", 70000, 4, 6, zlib.Z_DEFAULT_STRATEGY),
             ("zero", 300000, 5, 6, zlib.Z_DEFAULT_STRATEGY), ("text", 90000, 6, 6, zlib.Z_FIXED),
             ("png", 40000, 8, 6, zlib.Z_HUFFMAN_ONLY)]
    for k, (raw, z) in enumerate(synth.make_streams(specs, threads=1)):
        rc0, out0, r0 = ol.port_inflate(z, len(raw), 1)
        assert rc0 == 0 and out0 == raw
        rc, out, o = run_warp4(emu, z, len(raw), 1, mis=(k * 5) % 16, dmis=(k * 7) % 16)
        assert same(rc, out, o, rc0, out0, r0), (k, o.info >> 8)
