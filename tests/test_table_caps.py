"""CPU: can a code-length set overflow a decode table, and is the verdict then the reference's?

The reference gives up with UNZ_ERR when a table would need more than 2048 (literal/length, 10-bit root) or 512
(distance, 8-bit root) entries (infl.c:524-525); the kernels lay their tables out the same way (root + one
sub-table per root prefix, sized by the longest code under that prefix, inflate_core.cuh build_table) with caps
kLitTab = 1536 and kDistTab = 512.  Canonical codes are ordered by length, so the layout depends only on how many
codes each length > root has; a small dynamic programme over those counts gives the EXACT worst case over every
non-over-subscribed set of <= 288 / <= 32 symbols:
  literal/length: 1024 + 344 entries at most -- below both caps, never an error;
  distance:        256 + 278 = 534 at most -- ABOVE both caps (the same number, 512), so some distance sets are an
                   error in the reference and must be one here; where exactly is pinned below with sets around 512.
The extremal sets and sets right at the cap are encoded as real dynamic blocks and decoded by the compiled
reference, the oracle port and the emulated kernels (v3 and v4): same verdict everywhere."""
import ctypes
import functools

import pytest

import oracle_lib as ol

LIT_ROOT, LIT_SYMS, LIT_CAP_REF, LIT_CAP_OURS = 10, 288, 2048, 1536
DIST_ROOT, DIST_SYMS, DIST_CAP_REF, DIST_CAP_OURS = 8, 32, 512, 512
CL_ORDER = [16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15]


def table_size(lens, root):
    """Entries the layout needs for these code lengths (mirror of build_table's closed form)."""
    cnt = [0] * 16
    for l in lens:
        cnt[l] += 1
    cnt[0] = 0
    code = prev = 0
    cur_base, lo_prev, l_prev, hi_last = 1 << root, 0, 0, 0
    left = 1
    for l in range(1, 16):
        left = (left << 1) - cnt[l]
        assert left >= 0, "over-subscribed"
        code = (code + prev) << 1
        prev = cnt[l]
        if l > root and cnt[l]:
            lo = code >> (l - root)
            if l_prev:
                cur_base += (lo - lo_prev) << (l_prev - root)
            lo_prev, l_prev = lo, l
            hi_last = (code + cnt[l] - 1) >> (l - root)
    return cur_base + ((hi_last - lo_prev + 1) << (l_prev - root)) if l_prev else 1 << root


def worst_case(root, nsyms):
    """max over all count vectors (c_{root+1} .. c_15), sum <= nsyms, of the sub-table entries; with a witness.
    State after the codes of length < l: f = codes' worth of the current root prefix already taken, in units of
    2^-l (the long codes start on a prefix boundary: every shorter code fills whole prefixes)."""
    @functools.lru_cache(maxsize=None)
    def best(l, f, left, started):
        # -> (entries from length l on, counts tuple); `started`: a longer-than-root length is already present
        if l > 15:
            return 0, ()
        P = 1 << (l - root)
        # this length absent
        val, tail = best(l + 1, f * 2, left, started)
        out = (val, (0,) + tail)
        for c in range(1, left + 1):
            if f + c > (1 << l):           # (cannot happen for these symbol counts; keeps Kraft honest)
                break
            full = (f + c) // P            # root prefixes that end inside this run: sized by this length
            rest = (f + c) % P
            v2, t2 = best(l + 1, rest * 2, left - c, True)
            here = full * P
            if not any(t2):                # the last present length also owns the prefix it ends in
                here += P if rest else 0
                v2 = 0
            if here + v2 > out[0]:
                out = (here + v2, (c,) + t2)
        return out
    val, counts = best(root + 1, 0, nsyms, False)
    return val, {root + 1 + i: c for i, c in enumerate(counts) if c}


class BitWriter:
    def __init__(self):
        self.acc, self.n, self.out = 0, 0, bytearray()

    def put(self, v, nbits):
        self.acc |= v << self.n
        self.n += nbits
        while self.n >= 8:
            self.out.append(self.acc & 255)
            self.acc >>= 8
            self.n -= 8

    def done(self):
        if self.n:
            self.out.append(self.acc & 255)
        return bytes(self.out)


def canon(lens):
    cnt = [0] * 16
    for l in lens:
        cnt[l] += 1
    cnt[0] = 0
    nxt, code = [0] * 16, 0
    for l in range(1, 16):
        code = (code + cnt[l - 1]) << 1
        nxt[l] = code
    codes = []
    for l in lens:
        codes.append(nxt[l] if l else 0)
        if l:
            nxt[l] += 1
    return codes


def dynamic_block(lit_lens, dist_lens, payload_syms=()):
    """One final dynamic block with exactly these code lengths; it holds `payload_syms` literals and end-of-block."""
    assert len(lit_lens) >= 257 and lit_lens[256]
    w = BitWriter()
    w.put(1, 1); w.put(2, 2)
    w.put(len(lit_lens) - 257, 5); w.put(len(dist_lens) - 1, 5); w.put(19 - 4, 4)
    for s in CL_ORDER:                       # code-length code: every symbol 5 bits (incomplete, accepted: infl.c:485-490)
        w.put(5, 3)
    rev = lambda v, n: int(format(v, "0%db" % n)[::-1], 2)
    for l in list(lit_lens) + list(dist_lens):
        w.put(rev(l, 5), 5)                  # symbol l of a flat 5-bit code is code l
    codes = canon(lit_lens)
    for s in list(payload_syms) + [256]:
        w.put(rev(codes[s], lit_lens[s]), lit_lens[s])
    return w.done()


def lens_from_counts(counts, nsyms, root, must_have=None):
    """A length vector with these counts of long codes (and nothing else but what `must_have` needs)."""
    lens = []
    for l in sorted(counts):
        lens += [l] * counts[l]
    lens += [0] * (nsyms - len(lens))
    return lens


def test_worst_case_tables_fit_both_caps():
    lit_max, lit_w = worst_case(LIT_ROOT, LIT_SYMS)
    dist_max, dist_w = worst_case(DIST_ROOT, DIST_SYMS)
    # the dynamic programme agrees with the layout formula on its own witness
    assert table_size(lens_from_counts(lit_w, LIT_SYMS, LIT_ROOT), LIT_ROOT) == (1 << LIT_ROOT) + lit_max
    assert table_size(lens_from_counts(dist_w, DIST_SYMS, DIST_ROOT), DIST_ROOT) == (1 << DIST_ROOT) + dist_max
    # literal/length: no accepted set reaches either cap; distance: the worst case is beyond both caps, which are equal
    assert (1 << LIT_ROOT) + lit_max <= LIT_CAP_OURS <= LIT_CAP_REF, (lit_max, lit_w)
    assert (1 << DIST_ROOT) + dist_max == 534 and DIST_CAP_OURS == DIST_CAP_REF == 512, (dist_max, dist_w)


def test_layout_formula_against_random_sets():
    import random
    rnd = random.Random(7)
    for root, nsyms, cap in ((LIT_ROOT, LIT_SYMS, LIT_CAP_OURS), (DIST_ROOT, DIST_SYMS, DIST_CAP_OURS)):
        worst = worst_case(root, nsyms)[0]
        for _ in range(3000):
            lens, left = [], 1 << 15
            for _s in range(rnd.randint(1, nsyms)):
                l = rnd.choice([0, 0] + list(range(1, 16)) + [15, 14, 13, 12, 11])
                if l and (1 << (15 - l)) <= left:
                    left -= 1 << (15 - l)
                    lens.append(l)
                else:
                    lens.append(0)
            size = table_size(lens, root)
            assert size - (1 << root) <= worst and (size <= cap or root == DIST_ROOT)


def _dist_sets_around_the_cap():
    """Distance length sets whose tables need 480 .. 534 entries: both sides of the 512-entry cap."""
    import random
    rnd = random.Random(11)
    out = {}
    _, w = worst_case(DIST_ROOT, DIST_SYMS)
    base = lens_from_counts(w, DIST_SYMS, DIST_ROOT)
    for it in range(60000):
        if it % 2:
            lens = list(base)
            for _k in range(rnd.randint(1, 10)):              # shorten / drop some of the witness's long codes
                i = rnd.randrange(len(lens))
                lens[i] = rnd.choice([0, lens[i], max(9, lens[i] - 1), min(15, lens[i] + 1)]) if lens[i] else 0
        else:
            lens = [rnd.choice([9, 10, 11, 12, 12, 13, 14, 15, 15]) for _k in range(rnd.randint(8, 32))]
            lens += [0] * (DIST_SYMS - len(lens))
        lens.sort(key=lambda l: (l == 0, l))
        cnt, left, ok = [0] * 16, 1, True
        for l in lens:
            cnt[l] += 1
        for l in range(1, 16):
            left = (left << 1) - cnt[l]
            ok = ok and left >= 0
        if not ok:
            continue
        size = table_size(lens, DIST_ROOT)
        if 480 <= size <= 534 and size not in out:
            out[size] = lens
    return out


def _streams():
    lit_max, lit_w = worst_case(LIT_ROOT, LIT_SYMS)
    dist_max, dist_w = worst_case(DIST_ROOT, DIST_SYMS)
    lit = lens_from_counts(lit_w, LIT_SYMS, LIT_ROOT)
    assert lit[256]
    dist = lens_from_counts(dist_w, DIST_SYMS, DIST_ROOT)
    plain_lit = [8] * 144 + [9] * 112 + [7] * 24 + [8] * 8
    cases = [("worst literal/length table (1368 entries)", dynamic_block(lit, [5] * 30, payload_syms=(65, 66, 200)), 0, 3),
             ("worst distance table (534 entries)", dynamic_block(plain_lit, dist, payload_syms=(120, 121)), -1, 0),
             ("both worst", dynamic_block(lit, dist, payload_syms=(1,)), -1, 0)]
    sets = _dist_sets_around_the_cap()
    assert any(sz <= 512 for sz in sets) and any(sz > 512 for sz in sets), sorted(sets)
    for size, lens in sorted(sets.items()):
        cases.append((f"distance table of {size} entries", dynamic_block(plain_lit, lens, payload_syms=(7,)),
                      0 if size <= 512 else -1, 1 if size <= 512 else 0))
    return cases


def test_verdicts_at_the_cap_reference_and_port():
    for name, z, want, nout in _streams():
        code, out, r = ol.port_inflate(z, 64, 0)
        assert code == want and (code != 0 or len(out) == nout), (name, code)
        if ol.have_ref():
            rcode, rout, routlen, rinpos = ol.ref_inflate(z, 64, 0)
            assert rcode == want, (name, rcode)
            if want == 0:
                assert routlen == nout and bytes(rout[:nout]) == out, name


def test_verdicts_at_the_cap_emulated_kernels():
    import test_emu_logic as te
    emu = te.load_emu()
    for name, z, want, nout in _streams():
        code0, out0, r0 = ol.port_inflate(z, 64, 0)
        assert code0 == want
        rc, out, o = te.run_warp(emu, z, 64, 0)
        assert te.same(rc, out, o, code0, out0, r0), (name, "v3", rc, o.info >> 8)
        rc, out, o = te.run_warp4(emu, z, 64, 0)
        assert te.same(rc, out, o, code0, out0, r0), (name, "v4", rc, o.info >> 8)
