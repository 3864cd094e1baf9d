#!/usr/bin/env python3
"""The kernels' .cuh files compiled against the lock-step SIMT emulator (tests/emu) under AddressSanitizer /
ThreadSanitizer: the stand-in for compute-sanitizer memcheck / racecheck, which is closed on this GPU pool.

  g++ -std=c++17 -O1 -g -fsanitize=address -fPIC -shared -pthread -w -o /tmp/libemu_asan.so tests/emu/emu_main.cpp
  LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0 EMU=/tmp/libemu_asan.so \
      python profiles/tools/emu_sanitize.py /tmp/libemu_asan.so 150
  (likewise -fsanitize=thread / libtsan.so)
Every lane is a host thread and every shuffle / ballot / __syncwarp a pthread barrier, so a missing __syncwarp
between lanes shows up as a ThreadSanitizer data race and an out-of-bounds access to src / dst / the token
scratch / the shared-memory structs as an AddressSanitizer report."""
import os, sys
os.environ['EMU'] = sys.argv[1]
import ctypes, zlib, time
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import oracle_lib as ol
from defl_b200 import synth
class StreamOut(ctypes.Structure):
    _fields_ = [("code", ctypes.c_int32), ("adler", ctypes.c_uint32), ("outlen", ctypes.c_uint32),
                ("inpos", ctypes.c_uint32), ("adler_stored", ctypes.c_uint32), ("info", ctypes.c_uint32)]
emu = ctypes.CDLL(os.environ.get('EMU','/tmp/w/libemu4.so'))
def run4(data, cap, flags, mis=0, dmis=0):
    buf = ctypes.create_string_buffer(len(data) + 32)
    base = ctypes.addressof(buf)
    off = (-base) % 16 + mis
    ctypes.memmove(base + off, data, len(data))
    dst = ctypes.create_string_buffer(cap + 64)
    dbase = ctypes.addressof(dst); doff = (-dbase) % 16 + dmis
    o = StreamOut()
    rc = emu.emu_warp4_inflate(ctypes.c_void_p(base + off), len(data), ctypes.c_void_p(dbase+doff), cap, flags, ctypes.byref(o))
    return rc, dst.raw[doff: doff+o.outlen], o
def same(rc, out, o, rc0, out0, r0):
    if rc != rc0: return False
    return rc != 0 or (out == out0 and o.inpos == r0.inpos and o.adler == r0.adler and
                       o.adler_stored == r0.adler_stored and (o.info & 1) == r0.has_trailer)

def run3(data, cap, flags, mis=0):
    buf = ctypes.create_string_buffer(len(data) + 16); base = ctypes.addressof(buf); off = (-base) % 4 + mis
    ctypes.memmove(base + off, data, len(data)); dst = ctypes.create_string_buffer(cap + 16); o = StreamOut()
    rc = emu.emu_warp_inflate(ctypes.c_void_p(base + off), len(data), dst, cap, flags, ctypes.byref(o))
    return rc, dst.raw[: o.outlen], o
from test_oracle import _mutants
n=int(sys.argv[2]); bad=0
for it,(d,cap,flags,raw,m) in enumerate(_mutants(555, n)):
    rc0,out0,r0=ol.port_inflate(d,cap,flags)
    for f in (run3, run4):
        rc,out,o=f(d,cap,flags,mis=it%4)
        if not same(rc,out,o,rc0,out0,r0): bad+=1; print('FAIL',it,f.__name__)
Z=zlib.Z_DEFAULT_STRATEGY
specs=[("text",90000,1,6,Z),("log",60000,2,9,Z),("png",30000,3,1,Z),("zero",100000,5,6,Z),("text",40000,6,6,zlib.Z_FIXED),("rand",30000,4,6,Z)]
for k,(raw,z) in enumerate(synth.make_streams(specs,threads=1)):
    rc0,out0,r0=ol.port_inflate(z,len(raw),1)
    for f in (run3, run4):
        rc,out,o=f(z,len(raw),1,mis=k%4)
        if not same(rc,out,o,rc0,out0,r0): bad+=1; print('FAIL synth',k,f.__name__)
print('done bad',bad)
