"""Deterministic synthetic zlib streams for tests and bench.py.

Payload classes (each a seeded generator): ``text`` (words from a 4096-word lexicon),
``log`` (templated timestamp/id lines, cf. reference test/data/gendata.py:411-414), ``png``
(filter byte + RGBA gradient rows, cf. gendata.py:451-457), ``rand`` (uniform bytes -> zlib
emits stored blocks), ``zero`` (constant bytes).  Streams are produced by the system zlib
through Python's ``zlib`` module (the *compressor* only; decoding is never done here).
"""
import zlib
from concurrent.futures import ThreadPoolExecutor
import os

import numpy as np

SEED_BASE = 0xDEF10000
_LEX = None


def _lexicon():
    global _LEX
    if _LEX is None:
        rng = np.random.Generator(np.random.PCG64(0xDEF1))
        lens = rng.integers(2, 10, size=4096)
        rows = np.full((4096, 10), ord(" "), dtype=np.uint8)
        letters = rng.integers(ord("a"), ord("z") + 1, size=(4096, 10)).astype(np.uint8)
        col = np.arange(10)[None, :]
        rows = np.where(col < lens[:, None], letters, rows)
        _LEX = (rows, lens + 1)  # word + one space
    return _LEX


def payload(kind, nbytes, seed):
    rng = np.random.Generator(np.random.PCG64(SEED_BASE + int(seed)))
    nbytes = int(nbytes)
    if nbytes == 0:
        return b""
    if kind == "text":
        rows, wl = _lexicon()
        k = nbytes // 3 + 16
        idx = rng.integers(0, 4096, size=k)
        # zipf-ish skew so some words are common, as in natural text
        idx = (idx * rng.integers(0, 4096, size=k)) >> 12
        mask = np.arange(10)[None, :] < wl[idx][:, None]
        out = rows[idx][mask]
        while out.size < nbytes:
            out = np.concatenate([out, out])
        return out[:nbytes].tobytes()
    if kind == "log":
        tmpl = b"2024-01-15 10:23:45.123 INFO  [worker-07] RequestHandler id=00012345 status=200 latency=0012ms path=/api/v1/items/0042\n"
        w = len(tmpl)
        lines = nbytes // w + 1
        a = np.tile(np.frombuffer(tmpl, dtype=np.uint8), (lines, 1)).copy()

        def digits(cols, values):
            v = values.copy()
            for c in reversed(cols):
                a[:, c] = (v % 10 + ord("0")).astype(np.uint8)
                v //= 10

        def span(token):
            at = tmpl.index(token)
            return list(range(at, at + len(token)))

        t = np.cumsum(rng.integers(1, 900, size=lines))
        digits(span(b"123"), t % 1000)
        digits(span(b"45"), (t // 1000) % 60)
        digits(span(b"23"), (t // 60000) % 60)
        digits(span(b"07"), rng.integers(0, 16, size=lines))
        digits(span(b"00012345"), rng.integers(0, 10 ** 8, size=lines))
        status = rng.choice(np.array([200, 200, 200, 200, 404, 500, 302]), size=lines)
        digits(span(b"200"), status)
        digits(span(b"=0012")[1:], rng.integers(1, 3000, size=lines))
        digits(span(b"0042"), rng.integers(0, 100, size=lines))
        lvl = rng.integers(0, 20, size=lines)
        at = tmpl.index(b"INFO ")
        a[lvl == 0, at:at + 5] = np.frombuffer(b"WARN ", dtype=np.uint8)
        a[lvl == 1, at:at + 5] = np.frombuffer(b"ERROR", dtype=np.uint8)
        return a.reshape(-1)[:nbytes].tobytes()
    if kind == "png":
        width = 256
        stride = 1 + width * 4
        rows = nbytes // stride + 1
        x = np.arange(width, dtype=np.int64)[None, :]
        y = np.arange(rows, dtype=np.int64)[:, None]
        s = int(rng.integers(0, 256))
        noise = rng.integers(0, 3, size=(rows, width))
        img = np.empty((rows, stride), dtype=np.uint8)
        img[:, 0] = 0
        px = img[:, 1:].reshape(rows, width, 4)
        px[:, :, 0] = ((x + s + noise) & 255)
        px[:, :, 1] = ((y * 2 + s) & 255)
        px[:, :, 2] = (((x + y) >> 1) & 255)
        px[:, :, 3] = 255
        return img.reshape(-1)[:nbytes].tobytes()
    if kind == "rand":
        return rng.integers(0, 256, size=nbytes, dtype=np.uint8).tobytes()
    if kind == "zero":
        return bytes([int(rng.integers(0, 256))]) * nbytes
    raise ValueError(kind)


def deflate(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, wbits=15):
    c = zlib.compressobj(level, zlib.DEFLATED, wbits, 8, strategy)
    return c.compress(data) + c.flush()


def make_streams(specs, threads=None):
    """specs: list of (kind, nbytes, seed, level, strategy) -> list of (raw, zlib stream)."""
    threads = threads or min(32, os.cpu_count() or 1)

    def one(sp):
        kind, nbytes, seed, level, strategy = sp
        raw = payload(kind, nbytes, seed)
        return raw, deflate(raw, level, strategy)

    if threads <= 1 or len(specs) < 4:
        return [one(s) for s in specs]
    with ThreadPoolExecutor(threads) as ex:
        return list(ex.map(one, specs, chunksize=max(1, len(specs) // (threads * 8))))


def config2_specs(nstreams, size=65536, first_seed=0):
    """BASELINE.json configs[1]: 64 KiB `text` streams, level 6, dynamic Huffman."""
    return [("text", size, first_seed + i, 6, zlib.Z_DEFAULT_STRATEGY) for i in range(nstreams)]


def config5_specs(total_bytes, first_seed=0, lo=65536, hi=4 << 20):
    """BASELINE.json configs[4]: sizes log-uniform in [64 KiB, 4 MiB]; 50% text / 20% log / 15% png /
    10% rand / 5% zero; levels cycle 1/6/9; 10% forced Z_FIXED; 5% level 0 (stored)."""
    rng = np.random.Generator(np.random.PCG64(SEED_BASE ^ 0x5555 ^ first_seed))
    specs, tot, i = [], 0, 0
    kinds = ["text"] * 10 + ["log"] * 4 + ["png"] * 3 + ["rand"] * 2 + ["zero"]
    while tot < total_bytes:
        size = int(np.exp(rng.uniform(np.log(lo), np.log(hi))))
        kind = kinds[int(rng.integers(0, len(kinds)))]
        level = (1, 6, 9)[i % 3]
        strategy = zlib.Z_DEFAULT_STRATEGY
        u = rng.uniform()
        if u < 0.10:
            strategy = zlib.Z_FIXED
        elif u < 0.15:
            level = 0
        specs.append((kind, size, first_seed + i, level, strategy))
        tot += size
        i += 1
    return specs


def pack(blobs, align=16):
    """Concatenate byte strings at `align`-byte aligned offsets -> (uint8 array, offsets, lengths)."""
    lens = np.array([len(b) for b in blobs], dtype=np.int64)
    offs = np.zeros(len(blobs), dtype=np.int64)
    at = 0
    for i, n in enumerate(lens):
        offs[i] = at
        at = (at + int(n) + align - 1) // align * align
    buf = np.zeros(max(at, 1), dtype=np.uint8)
    for b, o in zip(blobs, offs):
        buf[o:o + len(b)] = np.frombuffer(b, dtype=np.uint8)
    return buf, offs, lens


def single_stream(total_bytes, seed=0, level=6, stripe=1 << 20, slice_bytes=16 << 20, threads=None):
    """BASELINE.json configs[3]: ONE zlib stream over `total_bytes` of text/log stripes.  Built pigz-style so the
    host needs seconds, not minutes: slices are deflated in parallel as raw deflate ending in a full flush
    (byte-aligned, non-final) and concatenated; the last slice ends the stream.  -> (raw, zlib stream)"""
    threads = threads or min(32, os.cpu_count() or 1)
    nstripes = (total_bytes + stripe - 1) // stripe

    def stripe_bytes(i):
        n = min(stripe, total_bytes - i * stripe)
        return payload("text" if i % 2 == 0 else "log", n, seed * 100003 + i)

    with ThreadPoolExecutor(threads) as ex:
        raw = b"".join(ex.map(stripe_bytes, range(nstripes)))
    nslices = max(1, (total_bytes + slice_bytes - 1) // slice_bytes)

    def one(k):
        c = zlib.compressobj(level, zlib.DEFLATED, -15, 8, zlib.Z_DEFAULT_STRATEGY)
        part = raw[k * slice_bytes:(k + 1) * slice_bytes]
        return c.compress(part) + c.flush(zlib.Z_FINISH if k == nslices - 1 else zlib.Z_FULL_FLUSH)

    with ThreadPoolExecutor(threads) as ex:
        body = b"".join(ex.map(one, range(nslices)))
    return raw, b"\x78\x9c" + body + (zlib.adler32(raw) & 0xFFFFFFFF).to_bytes(4, "big")
