"""ctypes access to the CHECKERS under oracle/ (test infrastructure only).

* ``port``  -- oracle/liboracle.so, the CPU restatement (oracle/inflate_oracle.c)
* ``ref``   -- oracle/_ref/libdefl_ref.so, the reference's own sources compiled unmodified
               (present when it was built in the container that has /root/reference)

Nothing under defl_b200/ imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
PORT_SO = os.path.join(ORACLE_DIR, "liboracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libdefl_ref.so")


class OracleResult(ctypes.Structure):
    _fields_ = [
        ("code", ctypes.c_int32),
        ("adler", ctypes.c_uint32),
        ("outlen", ctypes.c_uint64),
        ("inpos", ctypes.c_uint64),
        ("adler_stored", ctypes.c_uint32),
        ("has_trailer", ctypes.c_uint32),
        ("nblocks", ctypes.c_uint32),
        ("reserved", ctypes.c_uint32),
    ]


def build_port():
    src = os.path.join(ORACLE_DIR, "inflate_oracle.c")
    if (not os.path.exists(PORT_SO)) or os.path.getmtime(PORT_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "port"], stdout=subprocess.DEVNULL)
    return PORT_SO


_port = None
_ref = None


def port():
    global _port
    if _port is None:
        lib = ctypes.CDLL(build_port())
        lib.oracle_inflate.restype = ctypes.c_int
        lib.oracle_inflate.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
                                       ctypes.c_size_t, ctypes.c_int, ctypes.POINTER(OracleResult)]
        lib.oracle_inflate_chunks.restype = ctypes.c_int
        lib.oracle_inflate_chunks.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                              ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int,
                                              ctypes.POINTER(OracleResult)]
        lib.oracle_adler32.restype = ctypes.c_uint32
        lib.oracle_adler32.argtypes = [ctypes.c_uint32, ctypes.c_void_p, ctypes.c_size_t]
        lib.oracle_lens_ok.restype = ctypes.c_int
        lib.oracle_lens_ok.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.oracle_block_index.restype = ctypes.c_long
        lib.oracle_block_index.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int,
                                           ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t]
        lib.oracle_inflate_batch_mt.restype = ctypes.c_double
        lib.oracle_inflate_batch_mt.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int,
                                                ctypes.c_int, ctypes.c_void_p]
        _port = lib
    return _port


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    global _ref
    if _ref is None:
        lib = ctypes.CDLL(REF_SO)
        lib.ref_infl_buf.restype = ctypes.c_int
        lib.ref_infl_buf.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p,
                                     ctypes.c_uint32, ctypes.c_int]
        lib.ref_infl_buf_ex.restype = ctypes.c_int
        lib.ref_infl_buf_ex.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p,
                                        ctypes.c_uint32, ctypes.c_int,
                                        ctypes.POINTER(ctypes.c_uint32), ctypes.POINTER(ctypes.c_uint32)]
        lib.ref_infl_chunks.restype = ctypes.c_int
        lib.ref_infl_chunks.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                        ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int,
                                        ctypes.POINTER(ctypes.c_uint32)]
        lib.ref_infl_stream_pieces.restype = ctypes.c_int
        lib.ref_infl_stream_pieces.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32,
                                               ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int,
                                               ctypes.POINTER(ctypes.c_uint32)]
        lib.ref_infl_buf_batch_mt.restype = ctypes.c_double
        lib.ref_infl_buf_batch_mt.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                              ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int,
                                              ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        _ref = lib
    return _ref


def _buf(data):
    return (ctypes.c_char * max(len(data), 1)).from_buffer_copy(bytes(data) or b"\0")


def port_inflate(data, cap, flags=0):
    """-> (code, out_bytes, OracleResult)"""
    data = bytes(data)
    dst = ctypes.create_string_buffer(max(cap, 1))
    res = OracleResult()
    src = _buf(data)
    rc = port().oracle_inflate(ctypes.addressof(src), len(data), ctypes.addressof(dst), cap, flags,
                               ctypes.byref(res))
    return rc, dst.raw[: res.outlen], res


def port_block_index(data, flags=0, maxblocks=1 << 20):
    """-> (bit offsets of the block headers counted from the deflate body, output offsets)"""
    data = bytes(data)
    bits = np.zeros(maxblocks, dtype=np.uint64)
    outs = np.zeros(maxblocks, dtype=np.uint64)
    src = _buf(data)
    n = port().oracle_block_index(ctypes.addressof(src), len(data), flags, bits.ctypes.data, outs.ctypes.data, maxblocks)
    assert n >= 0
    return [int(b) for b in bits[:n]], [int(o) for o in outs[:n]]


def port_inflate_chunks(chunks, cap, flags=0):
    data = b"".join(bytes(c) for c in chunks)
    offs = np.zeros(len(chunks) + 1, dtype=np.uint64)
    if chunks:
        offs[1:] = np.cumsum([len(c) for c in chunks])
    dst = ctypes.create_string_buffer(max(cap, 1))
    res = OracleResult()
    src = _buf(data)
    rc = port().oracle_inflate_chunks(ctypes.addressof(src), offs.ctypes.data, len(chunks),
                                      ctypes.addressof(dst), cap, flags, ctypes.byref(res))
    return rc, dst.raw[: res.outlen], res


def ref_inflate(data, cap, flags=0):
    """-> (code, out_bytes, outlen, inpos) from the compiled reference"""
    data = bytes(data)
    dst = ctypes.create_string_buffer(max(cap, 1))
    ol, ip = ctypes.c_uint32(), ctypes.c_uint32()
    src = _buf(data)
    rc = ref().ref_infl_buf_ex(ctypes.addressof(src), len(data), ctypes.addressof(dst), cap, flags,
                               ctypes.byref(ol), ctypes.byref(ip))
    return rc, dst.raw[: ol.value], ol.value, ip.value


def ref_inflate_chunks(chunks, cap, flags=0):
    data = b"".join(bytes(c) for c in chunks)
    offs = np.zeros(len(chunks) + 1, dtype=np.uint64)
    if chunks:
        offs[1:] = np.cumsum([len(c) for c in chunks])
    dst = ctypes.create_string_buffer(max(cap, 1))
    ol = ctypes.c_uint32()
    src = _buf(data)
    rc = ref().ref_infl_chunks(ctypes.addressof(src), offs.ctypes.data, len(chunks),
                               ctypes.addressof(dst), cap, flags, ctypes.byref(ol))
    return rc, dst.raw[: ol.value], ol.value


def ref_inflate_stream(data, piece, cap, flags=0):
    data = bytes(data)
    dst = ctypes.create_string_buffer(max(cap, 1))
    ol = ctypes.c_uint32()
    src = _buf(data)
    rc = ref().ref_infl_stream_pieces(ctypes.addressof(src), len(data), piece,
                                      ctypes.addressof(dst), cap, flags, ctypes.byref(ol))
    return rc, dst.raw[: ol.value], ol.value


def takes_ft_path(data, flags=0):
    """True when the reference decodes `data` entirely on its contiguous fast path
    (infl_stored_direct / infl_ft_full), i.e. NOT via the stored-prefix hand-off to the
    huff-based general path (infl.c:256-258).  Hand-off happens iff the first block is
    stored, every stored block up to some point is well-formed, and a later block header
    has BTYPE != 0."""
    data = bytes(data)
    p = 0
    if flags == 1:
        if len(data) < 3:
            return True
        p = 2
    elif flags != 0:
        return False
    if p >= len(data):
        return True
    if (data[p] >> 1) & 3:
        return True  # first block not stored -> infl_ft_full does everything
    cap_unknown = True  # caller decides overflow separately; walk well-formed stored blocks
    while True:
        if p >= len(data):
            return True  # ran out -> UNZ_ERR on the stored path itself
        hdr = data[p]
        if (hdr >> 1) & 3:
            return False  # hand-off
        if len(data) - p < 5:
            return True
        ln = data[p + 1] | (data[p + 2] << 8)
        nl = data[p + 3] | (data[p + 4] << 8)
        if (ln ^ (~nl & 0xFFFF)) or len(data) - p - 5 < ln:
            return True
        p += 5 + ln
        if hdr & 1:
            return True
    return cap_unknown


# --- the reference fuzzer's generator, test/test_fuzz.c:138-246 (LCG + mutations) -------------
class RefFuzz:
    def __init__(self, seed=0x12345678):
        self.s = seed

    def rand(self):
        self.s = (self.s * 1103515245 + 12345) & 0xFFFFFFFF
        return self.s

    def generate(self, max_size=10000):
        size = (self.rand() % max_size) + 1
        fmt = self.rand() % 4
        out = bytearray()
        if fmt == 0:
            if size >= 10:
                out.append(0x01)
                ln = self.rand() % 100
                out += bytes([ln & 0xFF, (ln >> 8) & 0xFF, (~ln) & 0xFF, (~(ln >> 8)) & 0xFF])
                i = 0
                while i < ln and len(out) < size:
                    out.append(self.rand() & 0xFF)
                    i += 1
        elif fmt == 1:
            out.append(0x03)
            while len(out) < size - 1:
                if self.rand() % 10 == 0:
                    out.append(0x00)
                    break
                out.append(0x30 + (self.rand() % 144))
        else:
            if fmt == 2:
                out += b"\x78\x9c"
            while len(out) < size:
                out.append(self.rand() & 0xFF)
        return bytes(out)

    def mutated(self):
        size = (self.rand() % 10000) + 1
        data = bytearray(self.rand() & 0xFF for _ in range(size))
        strategy = self.rand() % 5
        if strategy == 0:
            idx = self.rand() % size
            data[idx] ^= 1 << (self.rand() % 8)
        elif strategy == 1:
            idx = self.rand() % size
            data[idx] = self.rand() & 0xFF
        elif strategy == 2:
            if size > 1:
                idx = self.rand() % (size - 1)
                data[idx + 1:size] = data[idx:size - 1]
                data[idx] = self.rand() & 0xFF
        elif strategy == 3:
            if size > 2:
                idx = self.rand() % (size - 1)
                data[idx:size - 1] = data[idx + 1:size]
        else:
            if size > 10:
                start = self.rand() % (size - 10)
                ln = (self.rand() % 10) + 1
                for i in range(ln // 2):
                    a, b = start + i, start + ln - 1 - i
                    data[a], data[b] = data[b], data[a]
        return bytes(data)

    def case(self, i):
        return self.generate() if i % 2 == 0 else self.mutated()
