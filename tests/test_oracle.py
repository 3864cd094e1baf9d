"""CPU: pin oracle/inflate_oracle.c (the restatement) against the reference's golden vectors,
against the compiled reference (oracle/_ref, when present) and against zlib."""
import hashlib
import random
import zlib

import pytest

import oracle_lib as ol


def test_corpus_against_golden(corpus):
    assert len(corpus) == 144
    paired = 0
    for e in corpus:
        code, out, res = ol.port_inflate(e["data"], e["cap"], 0)
        assert code == e["ref_code"], e["name"]
        if code == 0:
            assert res.outlen == e["ref_outlen"], e["name"]
            assert res.inpos == e["ref_inpos"], e["name"]
            assert hashlib.sha256(out).hexdigest() == e["ref_sha256"], e["name"]
        if "raw_len" in e:
            paired += 1
            assert code == 0 and len(out) == e["raw_len"]
            assert hashlib.sha256(out).hexdigest() == e["raw_sha256"]
            assert res.adler == e["raw_adler32"]            # Adler-32 KAT (zlib.adler32 of raw/)
            if e["len"]:
                assert out == zlib.decompress(e["data"], -15)   # independent byte oracle
    assert paired == 139


def test_corpus_zlib_wrapped_twins(corpus):
    for e in corpus:
        if "raw_len" not in e or e["len"] == 0:
            continue
        z = b"\x78\x9c" + e["data"] + e["raw_adler32"].to_bytes(4, "big")
        code, out, res = ol.port_inflate(z, e["raw_len"], 1)
        assert code == 0 and hashlib.sha256(out).hexdigest() == e["raw_sha256"], e["name"]
        assert res.has_trailer == 1 and res.adler_stored == res.adler == e["raw_adler32"]
        assert res.inpos == len(z) - 4     # stops right before the trailer


def test_kats(kats):
    for k in kats:
        data = bytes.fromhex(k["hex"])
        flags = k["flags"]
        code, out, res = ol.port_inflate(data, k["cap"], 1 if flags == 1 else 0)
        assert code == k["ref_code"], k["name"]
        if code == 0:
            assert out.hex() == k["ref_out_hex"], k["name"]


def test_reference_fuzzer_verdicts(fuzz_ref):
    fz = ol.RefFuzz(fuzz_ref["seed"])
    for i in range(fuzz_ref["n"]):
        d = fz.case(i)
        code, out, res = ol.port_inflate(d, fuzz_ref["cap"], 0)
        assert code == fuzz_ref["codes"][i], i
        if code == 0:
            assert res.outlen == fuzz_ref["outlens"][i], i
            assert hashlib.sha256(out).hexdigest()[:16] == fuzz_ref["sha16"][i], i


def _mutants(seed, count):
    rnd = random.Random(seed)
    for it in range(count):
        n = rnd.choice([0, 1, 5, 40, 300, 2000, 9000])
        kind = it % 4
        if kind == 0:
            raw = bytes(rnd.choice(b"abcdefgh ") for _ in range(n))
        elif kind == 1:
            raw = bytes(rnd.getrandbits(8) for _ in range(n))
        elif kind == 2:
            raw = (b"hello world, " * (n // 13 + 1))[:n]
        else:
            raw = bytes((i * 7 + (i >> 3)) & 0xFF for i in range(n))
        lvl = rnd.choice([0, 1, 6, 9])
        strat = rnd.choice([zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE])
        wb = rnd.choice([-15, 15])
        c = zlib.compressobj(lvl, zlib.DEFLATED, wb, 8, strat)
        d = bytearray(c.compress(raw) + c.flush())
        m = rnd.randrange(6)
        if m == 1 and d:
            d = d[: rnd.randrange(len(d) + 1)]
        elif m == 2 and d:
            for _ in range(rnd.randrange(1, 4)):
                d[rnd.randrange(len(d))] ^= 1 << rnd.randrange(8)
        elif m == 3 and d:
            d[rnd.randrange(len(d))] = rnd.getrandbits(8)
        elif m == 4:
            d += bytes(rnd.getrandbits(8) for _ in range(rnd.randrange(8)))
        cap = rnd.choice([n, n + 1000, max(n - 1, 0), n // 2])
        yield bytes(d), cap, (1 if wb == 15 else 0), raw, m


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_port_equals_compiled_reference_on_mutants():
    """Return code, produced bytes and consumed length vs the reference itself, on valid,
    truncated, bit-flipped and over-long streams and on short/exact/oversized dst."""
    checked = 0
    for d, cap, flags, raw, m in _mutants(1234, 3000):
        if not ol.takes_ft_path(d, flags):
            continue   # stored-prefix hand-off: decided by the absent huff library (parity unpinned)
        code, out, res = ol.port_inflate(d, cap, flags)
        rcode, rout, rlen, rinpos = ol.ref_inflate(d, cap, flags)
        assert code == rcode
        if code == 0:
            assert out == rout and res.inpos == rinpos
        checked += 1
    assert checked > 2900


def test_port_equals_zlib_on_valid_streams():
    for d, cap, flags, raw, m in _mutants(99, 600):
        if m != 0 or cap < len(raw):
            continue
        code, out, res = ol.port_inflate(d, cap, flags)
        assert code == 0 and out == raw
        assert res.adler == (zlib.adler32(raw) & 0xFFFFFFFF)


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_chunk_lists_match_reference(corpus):
    """infl_include semantics: any split of the input decodes to the same bytes
    (reference test_file_chunked, test_files.c:260-331: piece sizes 1,2,4,8 cycling)."""
    for e in corpus[:40]:
        if "raw_len" not in e:
            continue
        d, pieces, pos, k = e["data"], [], 0, 0
        while pos < len(d):
            n = 1 << (k % 4)
            pieces.append(d[pos:pos + n])
            pos += n
            k += 1
        code, out, res = ol.port_inflate_chunks(pieces, e["cap"], 0)
        rcode, rout, rlen = ol.ref_inflate_chunks(pieces, e["cap"], 0)
        assert code == rcode == 0 and out == rout, e["name"]
    code, out, res = ol.port_inflate_chunks([], 16, 0)
    rcode, rout, rlen = ol.ref_inflate_chunks([], 16, 0)
    assert code == rcode == 1      # UNZ_NOOP: nothing was included
    code, out, res = ol.port_inflate_chunks([b"", b""], 16, 0)
    rcode, rout, rlen = ol.ref_inflate_chunks([b"", b""], 16, 0)
    assert code == rcode == 0      # only empty pieces => UNZ_OK


def test_lens_acceptance_rule():
    import numpy as np
    lib = ol.port()

    def ok(lens):
        a = np.array(lens, dtype=np.uint8)
        return lib.oracle_lens_ok(a.ctypes.data, len(lens))

    assert ok([1, 1]) and ok([1]) and ok([0, 0, 0]) and ok([2, 2, 2])     # incomplete accepted
    assert not ok([1, 1, 1])                                              # over-subscribed (test_files.c:557)
    assert not ok([16])                                                   # too long (test_files.c:556)
    assert ok([15] * 288)
