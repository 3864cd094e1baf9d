#!/usr/bin/env python
"""Regenerate tests/golden/* from the reference tree (run in the container that has
/root/reference; the fixtures it writes are committed so GPU boxes need neither).

  corpus.bin / corpus.json  the reference's own test corpus, test/data/compressed/* (raw DEFLATE,
                            gendata.py:71-73) packed back to back, with for every file: the paired
                            raw length / sha256 / Adler-32 (test/data/raw/*), and what the COMPILED
                            REFERENCE (oracle/_ref, flags=0) returns: code, produced and consumed bytes
  kats.json                 inline known-answer vectors of test/test_files.c (:468-476, :627-633,
                            :885-888) with the compiled reference's verdict
  fuzz_ref.json             verdict of the compiled reference on the first N inputs of its own
                            fuzzer (test/test_fuzz.c:138-246, seed 0x12345678)
"""
import glob
import hashlib
import json
import os
import sys
import zlib

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402

REF = "/root/reference"
NFUZZ = 3000


def main():
    assert ol.have_ref(), "build oracle/_ref first: make -C oracle ref"
    files = sorted(glob.glob(os.path.join(REF, "test/data/compressed/*")))
    blob = bytearray()
    entries = []
    for f in files:
        name = os.path.basename(f)
        comp = open(f, "rb").read()
        rawp = os.path.join(REF, "test/data/raw", name)
        raw = open(rawp, "rb").read() if os.path.exists(rawp) else None
        cap = (len(raw) if raw is not None else 1 << 20) + 1000   # test_files.c:213 uses raw + 1000
        code, out, outlen, inpos = ol.ref_inflate(comp, cap, 0)
        e = {"name": name, "off": len(blob), "len": len(comp), "cap": cap,
             "ref_code": code, "ref_outlen": outlen, "ref_inpos": inpos,
             "ref_sha256": hashlib.sha256(out).hexdigest() if code == 0 else None}
        if raw is not None:
            assert code == 0 and out == raw, name
            e.update(raw_len=len(raw), raw_sha256=hashlib.sha256(raw).hexdigest(),
                     raw_adler32=zlib.adler32(raw) & 0xFFFFFFFF)
        entries.append(e)
        blob += comp
    open(os.path.join(HERE, "corpus.bin"), "wb").write(bytes(blob))
    json.dump({"source": "reference test/data/compressed (commit def57dfa)", "files": entries},
              open(os.path.join(HERE, "corpus.json"), "w"), indent=1)

    kats = [
        # name, hex, flags, dst capacity   (test/test_files.c)
        ("invalid_block_type", "07", 0, 100),                                       # :469
        ("buffer_overflow_protection", "010a00f5ff4142434445464748494a", 0, 5),      # :473-476
        ("truncated_stream", "789c", 1, 100),                                       # :468
        ("invalid_zlib_method", "00000300", 1, 100),                                # :470
        ("invalid_zlib_checksum", "78000300", 1, 100),                              # :471
        ("unsupported_zlib_dictionary", "78bb000000000300", 1, 100),                # :472
        ("regression_case_1", "010100feff41", 0, 10),                               # :627-629
        ("stored_prefix_then_fixed", "000300fcff6162634b494d0300", 0, 10),          # :630-633
        ("empty_input_raw", "", 0, 16),
        ("empty_input_zlib", "", 1, 16),
        ("fixed_a", "4b0400", 0, 8),
        ("zlib_hello", zlib.compress(b"hello, hello, hello").hex(), 1, 64),
        ("zlib_hello_bad_adler", (zlib.compress(b"hello, hello, hello")[:-1] + b"\x00").hex(), 1, 64),
        ("zlib_hello_no_trailer", zlib.compress(b"hello, hello, hello")[:-4].hex(), 1, 64),
        ("zlib_stream_with_flags0", zlib.compress(b"hello, hello, hello").hex(), 0, 64),
        ("raw_with_flags3_is_raw", zlib.compress(b"abcabcabc", 6)[2:-4].hex(), 3, 64),
        ("dst_one_short", zlib.compress(b"0123456789" * 10).hex(), 1, 99),
        ("dst_exact", zlib.compress(b"0123456789" * 10).hex(), 1, 100),
    ]
    out = []
    for name, hx, flags, cap in kats:
        data = bytes.fromhex(hx)
        code, outb, outlen, inpos = ol.ref_inflate(data, cap, flags)
        out.append({"name": name, "hex": hx, "flags": flags, "cap": cap, "ref_code": code,
                    "ref_outlen": outlen, "ref_out_hex": outb.hex() if code == 0 else None})
    json.dump(out, open(os.path.join(HERE, "kats.json"), "w"), indent=1)

    fz = ol.RefFuzz()
    codes, lens, shas = [], [], []
    for i in range(NFUZZ):
        d = fz.case(i)
        code, outb, outlen, inpos = ol.ref_inflate(d, 1 << 20, 0)   # MAX_OUTPUT_SIZE, test_fuzz.c
        codes.append(code)
        lens.append(outlen if code == 0 else 0)
        shas.append(hashlib.sha256(outb).hexdigest()[:16] if code == 0 else "")
    json.dump({"seed": 0x12345678, "n": NFUZZ, "cap": 1 << 20, "codes": codes, "outlens": lens, "sha16": shas},
              open(os.path.join(HERE, "fuzz_ref.json"), "w"))
    print("corpus files", len(entries), "bytes", len(blob), "kats", len(out), "fuzz", NFUZZ,
          "fuzz ok", sum(1 for c in codes if c == 0))


if __name__ == "__main__":
    main()
