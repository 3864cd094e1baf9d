"""Test helpers for include/defl/infl_wrap.h: the CPU oracle (oracle/wrap_oracle.c) and generators of gzip members
and PNG files (test infrastructure; never used by the product)."""
import ctypes
import struct
import zlib

import numpy as np

import oracle_lib as ol

_o = None


def oracle():
    global _o
    if _o is None:
        lib = ol.port()
        lib.wrap_oracle_crc32.restype = ctypes.c_uint32
        lib.wrap_oracle_crc32.argtypes = [ctypes.c_char_p, ctypes.c_uint64]
        lib.wrap_oracle_gzip_body.restype = ctypes.c_uint32
        lib.wrap_oracle_gzip_body.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p]
        lib.wrap_oracle_png_unfilter.restype = ctypes.c_int
        lib.wrap_oracle_png_unfilter.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32]
        lib.wrap_oracle_png_idat.restype = ctypes.c_int64
        lib.wrap_oracle_png_idat.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
        lib.wrap_oracle_zip_list.restype = ctypes.c_int64
        lib.wrap_oracle_zip_list.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint32]
        _o = lib
    return _o


def crc32(b):
    return int(oracle().wrap_oracle_crc32(bytes(b), len(b)))


def gzip_body(b):
    crc, isz = ctypes.c_uint32(), ctypes.c_uint32()
    off = oracle().wrap_oracle_gzip_body(bytes(b), len(b), ctypes.byref(crc), ctypes.byref(isz))
    return int(off), int(crc.value), int(isz.value)


def png_unfilter(raw, height, stride, bpp):
    out = ctypes.create_string_buffer(max(height * stride, 1))
    rc = oracle().wrap_oracle_png_unfilter(bytes(raw), out, height, stride, bpp)
    return rc, out.raw[: height * stride]


def png_idat(f):
    hdr = (ctypes.c_uint32 * 5)()
    z = ctypes.create_string_buffer(max(len(f), 1))
    nidat = ctypes.c_uint32()
    n = oracle().wrap_oracle_png_idat(bytes(f), len(f), hdr, z, len(f), ctypes.byref(nidat))
    return int(n), list(hdr), z.raw[: max(n, 0)], int(nidat.value)


def png_decode(f):
    """The whole PNG flow on the CPU: container walk, zlib (Python's), un-filter.  -> (code, pixels, hdr)"""
    n, hdr, z, nidat = png_idat(f)
    if n < 0:
        return -1, b"", hdr
    w, h, depth, ctype, interlace = hdr
    chan = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}.get(ctype)
    if chan is None or interlace or depth not in (1, 2, 4, 8, 16) or not w or not h:
        return -1, b"", hdr
    stride = (w * chan * depth + 7) // 8
    bpp = max(1, chan * depth // 8)
    try:
        raw = zlib.decompress(z)
    except zlib.error:
        return -1, b"", hdr
    if len(raw) != h * (stride + 1):
        return -1, b"", hdr
    rc, pix = png_unfilter(raw, h, stride, bpp)
    return (-1 if rc else 0), pix, hdr


def zip_list(z, cap=4096):
    """-> list of dicts (the oracle's walk of the central directory), or None."""
    out = (ctypes.c_uint64 * (8 * cap))()
    n = oracle().wrap_oracle_zip_list(bytes(z), len(z), out, cap)
    if n < 0:
        return None
    ents = []
    for k in range(min(n, cap)):
        r = out[8 * k: 8 * k + 8]
        ents.append({"local_off": r[0], "data_off": r[1], "csize": r[2], "usize": r[3], "crc32": r[4], "method": r[5] & 0xFFFF,
                     "gpflags": r[5] >> 16, "name_off": r[6], "name_len": r[7] & 0xFFFF, "ok": r[7] >> 16})
    return ents


# ---- generators -------------------------------------------------------------------------------------------------
def gzip_member(raw, level=6, name=None, comment=None, extra=None, hcrc=False, mtime=0):
    flg = (4 if extra is not None else 0) | (8 if name is not None else 0) | (16 if comment is not None else 0) | (2 if hcrc else 0)
    head = bytes([0x1F, 0x8B, 8, flg]) + struct.pack("<I", mtime) + bytes([0, 3])
    if extra is not None:
        head += struct.pack("<H", len(extra)) + extra
    if name is not None:
        head += name + b"\0"
    if comment is not None:
        head += comment + b"\0"
    if hcrc:
        head += struct.pack("<H", zlib.crc32(head) & 0xFFFF)
    c = zlib.compressobj(level, zlib.DEFLATED, -15)
    body = c.compress(raw) + c.flush()
    return head + body + struct.pack("<II", zlib.crc32(raw) & 0xFFFFFFFF, len(raw) & 0xFFFFFFFF)


def _chunk(tag, data):
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def _paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)


def png_filter_rows(pix, height, stride, bpp, ftypes):
    """Forward filtering (PNG spec 9.2) of packed rows with the filter type given per row -> raw scanline data."""
    pix = np.frombuffer(bytes(pix), dtype=np.uint8).reshape(height, stride).astype(np.int32)
    out = bytearray()
    zero = np.zeros(stride, dtype=np.int32)
    for y in range(height):
        row, up = pix[y], (pix[y - 1] if y else zero)
        a = np.concatenate([np.zeros(bpp, dtype=np.int32), row[:-bpp]]) if stride > bpp else np.zeros(stride, dtype=np.int32)
        c = np.concatenate([np.zeros(bpp, dtype=np.int32), up[:-bpp]]) if stride > bpp else np.zeros(stride, dtype=np.int32)
        ft = ftypes[y % len(ftypes)]
        if ft == 0:
            f = row
        elif ft == 1:
            f = row - a
        elif ft == 2:
            f = row - up
        elif ft == 3:
            f = row - ((a + up) >> 1)
        else:
            pa, pb, pc = np.abs(up - c), np.abs(a - c), np.abs(a + up - 2 * c)
            pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, up, c))
            f = row - pred
        out.append(ft)
        out += (f & 255).astype(np.uint8).tobytes()
    return bytes(out)


def png_file(width, height, depth, ctype, pix, ftypes=(0, 1, 2, 3, 4), idat=8192, level=6, interlace=0, extra_chunks=True):
    chan = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
    stride = (width * chan * depth + 7) // 8
    bpp = max(1, chan * depth // 8)
    raw = png_filter_rows(pix, height, stride, bpp, ftypes)
    z = zlib.compress(raw, level)
    out = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", width, height, depth, ctype, 0, 0, interlace))
    if extra_chunks:
        out += _chunk(b"gAMA", struct.pack(">I", 45455))
        if ctype == 3:
            out += _chunk(b"PLTE", bytes(range(256)) * 3)
        out += _chunk(b"tEXt", b"Comment\0defl_b200 test image")
    for o in range(0, len(z), idat):
        out += _chunk(b"IDAT", z[o:o + idat])
    return out + _chunk(b"IEND", b""), stride, bpp
