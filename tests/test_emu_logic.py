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
CORE = os.path.join(os.path.dirname(HERE), "defl_b200", "csrc", "inflate_core.cuh")


class StreamOut(ctypes.Structure):
    _fields_ = [("code", ctypes.c_int32), ("adler", ctypes.c_uint32), ("outlen", ctypes.c_uint32),
                ("inpos", ctypes.c_uint32), ("adler_stored", ctypes.c_uint32), ("info", ctypes.c_uint32)]


@pytest.fixture(scope="module")
def emu():
    srcs = [os.path.join(EMU_DIR, "emu_main.cpp"), os.path.join(EMU_DIR, "simt_emu.h"), CORE]
    if not os.path.exists(EMU_SO) or any(os.path.getmtime(s) > os.path.getmtime(EMU_SO) for s in srcs):
        subprocess.check_call(["g++", "-std=c++17", "-O1", "-fPIC", "-shared", "-pthread", "-w",
                               "-o", EMU_SO, srcs[0]])
    return ctypes.CDLL(EMU_SO)


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
