"""CPU: the wrapper oracle (oracle/wrap_oracle.c: CRC-32, gzip member header, PNG container walk and un-filter)
pinned against independent implementations -- Python's zlib.crc32 and gzip module, and forward filtering."""
import gzip
import io
import zlib

import numpy as np

import wrap_lib as wl


def test_crc32_against_zlib():
    rng = np.random.default_rng(1)
    for n in (0, 1, 2, 3, 255, 256, 257, 65535, 65536, 65537, 300001):
        b = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert wl.crc32(b) == (zlib.crc32(b) & 0xFFFFFFFF), n
    assert wl.crc32(b"123456789") == 0xCBF43926          # the check value of CRC-32/ISO-HDLC


def test_gzip_header_against_the_gzip_module():
    raw = b"hello gzip " * 1000
    buf = io.BytesIO()
    with gzip.GzipFile(filename="a-name.txt", mode="wb", fileobj=buf, mtime=123) as g:
        g.write(raw)
    member = buf.getvalue()
    off, crc, isz = wl.gzip_body(member)
    assert off > 10 and crc == (zlib.crc32(raw) & 0xFFFFFFFF) and isz == len(raw)
    assert zlib.decompress(member[off:-8], -15) == raw
    for kw in ({}, {"name": b"n"}, {"comment": b"c c"}, {"extra": b"\x01\x02\x03\x04\x05"}, {"hcrc": True},
               {"name": b"nm", "comment": b"cm", "extra": b"xx", "hcrc": True}):
        m = wl.gzip_member(raw, **kw)
        assert gzip.decompress(m) == raw                  # the generator writes what the gzip module reads
        off, crc, isz = wl.gzip_body(m)
        assert zlib.decompress(m[off:-8], -15) == raw and crc == (zlib.crc32(raw) & 0xFFFFFFFF), kw
    assert wl.gzip_body(b"\x1f\x8b\x07" + bytes(30))[0] == 0     # CM != 8
    assert wl.gzip_body(member[:12])[0] == 0                      # too short


def test_png_unfilter_round_trip_and_walk():
    rng = np.random.default_rng(2)
    for (w, h, depth, ctype) in ((7, 5, 8, 6), (33, 40, 8, 2), (64, 3, 8, 0), (10, 9, 16, 6), (13, 7, 16, 2), (17, 6, 8, 4),
                                 (19, 4, 4, 0), (21, 8, 1, 0), (9, 5, 2, 3), (5, 70, 8, 3), (1, 1, 8, 0)):
        chan = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
        stride = (w * chan * depth + 7) // 8
        pix = rng.integers(0, 256, h * stride, dtype=np.uint8).tobytes()
        for ft in ((0,), (1,), (2,), (3,), (4,), (4, 3, 2, 1, 0)):
            f, stride2, bpp = wl.png_file(w, h, depth, ctype, pix, ftypes=ft, idat=37)
            assert stride2 == stride
            code, got, hdr = wl.png_decode(f)
            assert code == 0 and got == pix and hdr[:4] == [w, h, depth, ctype], (w, h, depth, ctype, ft)
    n, hdr, z, nidat = wl.png_idat(f)
    assert nidat == (len(z) + 36) // 37
    assert wl.png_idat(b"\x89PNG\r\n\x1a\n" + bytes(40))[0] < 0
    assert wl.png_unfilter(bytes([5, 1, 2, 3]), 1, 3, 1)[0] == 1


def _archive():
    import io
    import zipfile
    rng = np.random.default_rng(8)
    files = {"a.txt": b"hello zip\n" * 500, "dir/b.bin": rng.integers(0, 256, 70000, dtype=np.uint8).tobytes(),
             "empty": b"", "c.log": b"2024-01-01 INFO something happened\n" * 4000, "one": b"x"}
    buf = io.BytesIO()
    with zipfile.ZipFile(buf, "w") as z:
        for i, (name, data) in enumerate(files.items()):
            z.writestr(zipfile.ZipInfo(name), data, compress_type=zipfile.ZIP_STORED if i == 1 else zipfile.ZIP_DEFLATED)
        z.comment = b"archive comment"
    return buf.getvalue(), files


def test_zip_directory_against_zipfile():
    import io
    import zipfile
    z, files = _archive()
    ents = wl.zip_list(z)
    infos = zipfile.ZipFile(io.BytesIO(z)).infolist()
    assert len(ents) == len(infos) == len(files)
    for e, i in zip(ents, infos):
        assert z[e["name_off"]: e["name_off"] + e["name_len"]].decode() == i.filename
        assert (e["csize"], e["usize"], e["crc32"], e["method"], e["local_off"]) == (i.compress_size, i.file_size, i.CRC, i.compress_type, i.header_offset)
        assert e["ok"] == 1
        data = z[e["data_off"]: e["data_off"] + e["csize"]]
        assert (zlib.decompress(data, -15) if e["method"] == 8 else data) == files[i.filename]
    assert wl.zip_list(z[:-5]) is None and wl.zip_list(b"PK" + bytes(40)) is None
