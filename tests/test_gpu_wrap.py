"""GPU: include/defl/infl_wrap.h through the C ABI against the wrapper oracle (oracle/wrap_oracle.c) and Python's
zlib / gzip as a second opinion: CRC-32, gzip members, whole PNG files -- container walk on the
device, IDATs gathered by descriptors, inflate, scanline un-filter (8f-1, 8f-2; the reference's README.md:120-155)."""
import gzip
import zlib

import numpy as np
import pytest

import wrap_lib as wl

pytestmark = pytest.mark.gpu


def test_crc32_batch(gpu_lib):
    d = gpu_lib
    rng = np.random.default_rng(3)
    bufs = [rng.integers(0, 256, n, dtype=np.uint8).tobytes()
            for n in (0, 1, 2, 15, 16, 17, 255, 256, 257, 4095, 65535, 65536, 65537, 131072, 1000003, 5 << 20)]
    bufs.append(b"123456789")
    crc = d.crc32_batch(bufs)
    for b, c in zip(bufs, crc):
        assert int(c) == wl.crc32(b) == (zlib.crc32(b) & 0xFFFFFFFF), len(b)


def test_gzip_members(gpu_lib):
    from defl_b200 import synth
    d = gpu_lib
    Z = zlib.Z_DEFAULT_STRATEGY
    raws = [synth.payload(k, n, 70 + i) for i, (k, n) in enumerate(
        (("text", 65536), ("log", 300000), ("png", 100000), ("rand", 50000), ("zero", 400000), ("text", 0), ("text", 1)))]
    members = [wl.gzip_member(raws[0]), wl.gzip_member(raws[1], level=9, name=b"access.log"),
               wl.gzip_member(raws[2], level=1, comment=b"a comment", hcrc=True),
               wl.gzip_member(raws[3], extra=b"\x41\x42\x04\x00abcd"), wl.gzip_member(raws[4], name=b"z", comment=b"c", extra=b"", hcrc=True),
               wl.gzip_member(raws[5]), gzip.compress(raws[6], mtime=0)]
    res, outs = d.gzip_batch(members, [len(r) + 16 for r in raws], d.INFL_WRAP_VERIFY)
    for m, raw, g, out in zip(members, raws, res, outs):
        off, crc, isz = wl.gzip_body(m)
        assert int(g["code"]) == 0 and bytes(out) == raw, int(g["info"])
        assert int(g["crc32"]) == wl.crc32(raw) == crc == int(g["crc32_stored"]) and int(g["isize_stored"]) == isz
        assert int(g["inpos"]) == len(m) and int(g["info"]) & d.INFL_WRAP_INFO_CRC_MATCH
    # damaged members: wrong CRC, wrong ISIZE, truncated body, not gzip at all, a flipped bit in the body
    bad = []
    b = bytearray(members[0]); b[-8] ^= 1; bad.append(bytes(b))
    b = bytearray(members[1]); b[-1] ^= 0x40; bad.append(bytes(b))
    bad.append(members[2][: len(members[2]) // 2])
    bad.append(b"\x1f\x8b\x09" + members[0][3:])
    b = bytearray(members[0]); b[len(b) // 2] ^= 0x08; bad.append(bytes(b))
    res, outs = d.gzip_batch(bad, [400000] * len(bad), d.INFL_WRAP_VERIFY)
    assert all(int(g["code"]) != 0 for g in res), [int(g["code"]) for g in res]
    assert int(res[3]["info"]) & d.INFL_WRAP_INFO_BAD_HEADER
    # without verification a wrong CRC is reported, not refused (the reference never checks a trailer either)
    res, outs = d.gzip_batch(bad[:1], [400000], 0)
    assert int(res[0]["code"]) == 0 and not (int(res[0]["info"]) & d.INFL_WRAP_INFO_CRC_MATCH) and bytes(outs[0]) == raws[0]


def test_gzip_large_member_takes_the_split_decoder(gpu_lib):
    import torch
    from defl_b200 import synth, api
    d = gpu_lib
    raw = synth.payload("text", 12 << 20, 91)
    m = wl.gzip_member(raw, level=1, name=b"big.txt")
    assert len(m) > 4 << 20
    dev = torch.device("cuda", 0)
    t_src = torch.frombuffer(bytearray(m), dtype=torch.uint8).to(dev)
    t_dst = torch.zeros(len(raw), dtype=torch.uint8, device=dev)
    items = np.zeros(1, dtype=d.ITEM_DTYPE)
    items[0] = (t_src.data_ptr(), len(m), t_dst.data_ptr(), len(raw))
    res = np.zeros(1, dtype=d.GZIP_RESULT_DTYPE)
    d.gzip_batch_raw(items, d.INFL_WRAP_VERIFY | d.INFL_BATCH_DEVICE_PTRS, res)
    assert int(res[0]["code"]) == 0 and int(res[0]["info"]) & 4 and int(res[0]["crc32"]) == (zlib.crc32(raw) & 0xFFFFFFFF)
    assert t_dst.cpu().numpy().tobytes() == raw


def _png_cases():
    rng = np.random.default_rng(4)
    cases = []
    shapes = ((7, 5, 8, 6), (33, 40, 8, 2), (64, 3, 8, 0), (10, 9, 16, 6), (13, 7, 16, 2), (17, 6, 8, 4), (19, 4, 4, 0),
              (21, 8, 1, 0), (9, 5, 2, 3), (5, 70, 8, 3), (1, 1, 8, 0), (256, 200, 8, 6), (640, 97, 8, 2), (100, 333, 16, 4))
    for k, (w, h, depth, ctype) in enumerate(shapes):
        chan = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
        stride = (w * chan * depth + 7) // 8
        if k % 2:
            pix = rng.integers(0, 256, h * stride, dtype=np.uint8).tobytes()
        else:                                             # smooth content: the filters really predict something
            x = (np.arange(h * stride) // max(1, chan * depth // 8))
            pix = ((x * 3 + (x // stride) * 5) & 255).astype(np.uint8).tobytes()
        for ft, idat in (((0,), 8192), ((1,), 1), ((2,), 64), ((3,), 7), ((4,), 8192), ((4, 3, 2, 1, 0), 100)):
            if idat == 1 and h * stride > 20000:
                idat = 3
            cases.append((wl.png_file(w, h, depth, ctype, pix, ftypes=ft, idat=idat)[0], pix, (w, h, depth, ctype)))
    return cases


def test_png_files(gpu_lib):
    d = gpu_lib
    cases = _png_cases()
    res, outs = d.png_batch([c[0] for c in cases], [len(c[1]) + 8 for c in cases], d.INFL_WRAP_VERIFY)
    for (f, pix, shape), g, out in zip(cases, res, outs):
        code, want, hdr = wl.png_decode(f)
        assert code == 0 and want == pix
        assert int(g["code"]) == 0, (shape, int(g["info"]))
        assert bytes(out) == pix, shape
        assert (int(g["width"]), int(g["height"]), int(g["bit_depth"]), int(g["color_type"])) == shape
        assert int(g["nidat"]) == wl.png_idat(f)[3]


def test_png_refusals_match_the_oracle(gpu_lib):
    d = gpu_lib
    rng = np.random.default_rng(5)
    pix = rng.integers(0, 256, 20 * 30 * 3, dtype=np.uint8).tobytes()
    good, stride, bpp = wl.png_file(30, 20, 8, 2, pix, idat=500)
    bad = [good[:8] + b"\0" + good[9:],                                   # IHDR length field damaged
           b"\x89PNX" + good[4:],                                        # signature
           wl.png_file(30, 20, 8, 2, pix, interlace=1)[0],               # Adam7: not handled here
           good[:-12],                                                   # no IEND
           good[: len(good) // 2]]                                       # cut inside an IDAT
    # a filter byte of 5 in the third row, and scanline data one byte short
    import struct
    raw = bytearray(wl.png_filter_rows(pix, 20, stride, bpp, (1,)))
    raw[2 * (stride + 1)] = 5
    head = b"\x89PNG\r\n\x1a\n" + wl._chunk(b"IHDR", struct.pack(">IIBBBBB", 30, 20, 8, 2, 0, 0, 0))
    bad.append(head + wl._chunk(b"IDAT", zlib.compress(bytes(raw))) + wl._chunk(b"IEND", b""))
    bad.append(head + wl._chunk(b"IDAT", zlib.compress(bytes(raw[:-1]))) + wl._chunk(b"IEND", b""))
    files = bad + [good]
    res, outs = d.png_batch(files, [len(pix) + 8] * len(files), 0)
    for f, g in zip(files, res):
        code, want, hdr = wl.png_decode(f)
        assert (int(g["code"]) == 0) == (code == 0), (int(g["code"]), code, int(g["info"]))
    assert int(res[2]["info"]) & d.INFL_WRAP_INFO_INTERLACED and int(res[5]["info"]) & d.INFL_WRAP_INFO_BAD_FILTER
    assert int(res[-1]["code"]) == 0 and bytes(outs[-1]) == pix


def test_png_batch_device_resident_many_files(gpu_lib):
    """Device-resident files, a few thousand of them: the container walk, the gather and the un-filter at batch size."""
    import torch
    d = gpu_lib
    rng = np.random.default_rng(6)
    base = []
    for k in range(24):
        w, h = int(rng.integers(8, 120)), int(rng.integers(8, 120))
        ctype = (0, 2, 4, 6)[k % 4]
        chan = {0: 1, 2: 3, 4: 2, 6: 4}[ctype]
        x = np.arange(h * w * chan)
        pix = ((x // chan * 7 + x % chan * 31 + rng.integers(0, 3, x.size)) & 255).astype(np.uint8).tobytes()
        f = wl.png_file(w, h, 8, ctype, pix, ftypes=(4, 1, 2, 3, 0), idat=int(rng.choice([50, 700, 8192])))[0]
        base.append((f, pix))
    n = 3000
    dev = torch.device("cuda", 0)
    blob = b"".join(f for f, _ in base)
    offs = np.concatenate([[0], np.cumsum([len(f) for f, _ in base])])
    t_files = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
    pixoff = np.concatenate([[0], np.cumsum([len(base[i % 24][1]) + 16 for i in range(n)])])
    t_pix = torch.zeros(int(pixoff[-1]), dtype=torch.uint8, device=dev)
    items = np.zeros(n, dtype=d.PNG_ITEM_DTYPE)
    for i in range(n):
        k = i % 24
        items[i] = (t_files.data_ptr() + int(offs[k]), len(base[k][0]), t_pix.data_ptr() + int(pixoff[i]), len(base[k][1]))
    res = np.zeros(n, dtype=d.PNG_RESULT_DTYPE)
    d.png_batch_raw(items, d.INFL_WRAP_VERIFY | d.INFL_BATCH_DEVICE_PTRS, res)
    torch.cuda.synchronize()
    assert (res["code"] == 0).all()
    host = t_pix.cpu().numpy()
    for i in range(0, n, 7):
        want = base[i % 24][1]
        assert host[pixoff[i]: pixoff[i] + len(want)].tobytes() == want, i


def test_zip_archives(gpu_lib):
    import io
    import zipfile
    from test_wrap_oracle import _archive
    d = gpu_lib
    z, files = _archive()
    ent, names = d.zip_list(z)
    want = wl.zip_list(z)
    assert len(ent) == len(want) == len(files)
    for e, w_, nm in zip(ent, want, names):
        for f in ("local_off", "data_off", "csize", "usize", "crc32", "method", "gpflags", "name_off", "name_len", "ok"):
            assert int(e[f]) == w_[f], f
    res, outs = d.zip_extract(z, ent, d.INFL_WRAP_VERIFY)
    zf = zipfile.ZipFile(io.BytesIO(z))
    for e, nm, g, out in zip(ent, names, res, outs):
        assert int(g["code"]) == 0 and bytes(out) == zf.read(nm.decode()) == files[nm.decode()], nm
        assert int(g["crc32"]) == int(e["crc32"]) and int(g["info"]) & d.INFL_WRAP_INFO_CRC_MATCH
    # a damaged member: its CRC no longer matches (or the body no longer decodes)
    bad = bytearray(z)
    bad[int(ent[3]["data_off"]) + 40] ^= 0x20
    res, outs = d.zip_extract(bytes(bad), ent, d.INFL_WRAP_VERIFY)
    assert int(res[3]["code"]) != 0 and all(int(res[i]["code"]) == 0 for i in (0, 1, 2, 4))
    with pytest.raises(RuntimeError):
        d.zip_list(z[:-30])
