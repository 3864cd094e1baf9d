// inflate_core.cuh -- the DEFLATE decode core: one *group* of G lanes (G = 1..32, a slice of a
// warp) decodes one stream.
//
// Replaces, from scratch, the reference's contiguous decode path
//   infl_ft_full / infl_ft_stored / infl_ft_dynamic / infl_ft_build / infl_ft_block
//   (/root/reference/src/infl/infl.c:452-1104) and the zlib header check (zlib.h:45-96),
// and adds Adler-32 (absent from the reference).  Results (bytes, produced length, consumed
// length, UNZ_OK / UNZ_ERR) are defined by oracle/inflate_oracle.c.
//
// Shape of the computation (B200-first, not the reference's):
//  * All G lanes of a group run the bit reader and the symbol decode REDUNDANTLY and in
//    lock-step: every lane holds the same bit window, looks up the same shared-memory table
//    slot (a broadcast, no bank conflict) and therefore already knows (length, distance)
//    when a match arrives -- no shuffles, no producer/consumer queue.  A warp-wide
//    instruction costs the same whether 1 or 32 lanes are active, so redundancy is free and
//    32/G streams advance per issued instruction.
//  * The lanes split only where there is real data parallelism: LZ77 copies, stored-block
//    copies, Huffman table construction (ranks by match.any), table zeroing, Adler-32.
//  * Tables live in shared memory as 16-bit entries (10-bit root + sub-tables for
//    literal/length, 8-bit root for distance): 4 KiB per stream instead of the reference's
//    10 KiB of uint32 entries (infl.c:382-383,406-414), so ~48 streams fit per SM.
//  * The bit reader is a 64-bit window in two registers plus one prefetched word; a peek is
//    a single funnel shift (SHF.R) and refills are aligned 32-bit loads on the read-only path.
#pragma once
#include <stdint.h>

namespace defl {

constexpr int kLitRootBits  = 10;
constexpr int kLitRoot      = 1 << kLitRootBits;
constexpr int kLitTab       = 1536;  // root + worst-case sub-tables (<= 448)
constexpr int kDistRootBits = 8;
constexpr int kDistRoot     = 1 << kDistRootBits;
constexpr int kDistTab      = 512;   // root + worst-case sub-tables (<= 256)
constexpr uint32_t kAdlerMod = 65521u;
constexpr uint32_t kFull     = 0xFFFFFFFFu;   // every lane of a warp

// failure reasons (diagnostics only; every one of them is UNZ_ERR at the API)
enum Reason : uint32_t {
  R_NONE = 0, R_ZLIB_HEADER = 1, R_TRUNCATED = 2, R_BLOCK_TYPE = 3, R_STORED_LEN = 4,
  R_OUTPUT_FULL = 5, R_CODE_LENGTHS = 6, R_TABLE = 7, R_SYMBOL = 8, R_DISTANCE = 9,
  R_ADLER = 10
};

// ---- table entry formats (16 bit) -------------------------------------------------------
// literal/length table:
//   0                         : no code maps here (incomplete set, or symbols 286/287)
//   bit15 = 1                 : length symbol   [14:11] code bits  [10:8] extra bits  [7:0] base-3
//   bit15 = 0, [10:9] = 00    : literal         [14:11] code bits  [7:0] byte
//   bit15 = 0, [10:9] = 01    : end of block    [14:11] code bits
//   bit15 = 0, [10:9] = 11    : sub-table link  [14:11] sub-table index bits  [8:0] offset - root
// distance table:
//   0                         : no code (or symbols 30/31)
//   bit15 = 1                 : sub-table link  [14:11] sub-table index bits  [8:0] offset - root
//   bit15 = 0                 : distance symbol [14:11] code bits  [10:7] extra bits  [1:0] m,
//                               distance = (m << extra_bits) + 1 + extra
constexpr uint32_t kLitIsLen   = 0x8000u;
constexpr uint32_t kLitKind    = 0x0600u;
constexpr uint32_t kLitEob     = 0x0200u;
constexpr uint32_t kLitSub     = 0x0600u;
constexpr uint32_t kDistSub    = 0x8000u;

// order in which the code-length-code lengths arrive (RFC 1951 3.2.7; reference table `ord`,
// apicommon.h:51-52), packed 5 bits each so no constant-memory lookup is needed
constexpr uint64_t pack5(const int* v, int n) {
  uint64_t r = 0;
  for (int i = 0; i < n; i++) r |= (uint64_t)v[i] << (5 * i);
  return r;
}
constexpr int kClOrdA[12] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4};
constexpr int kClOrdB[7]  = {12, 3, 13, 2, 14, 1, 15};
constexpr uint64_t kClOrderLo = pack5(kClOrdA, 12);
constexpr uint64_t kClOrderHi = pack5(kClOrdB, 7);
DEFL_DEV uint32_t cl_order(uint32_t i) {
  return i < 12u ? (uint32_t)(kClOrderLo >> (5u * i)) & 31u : (uint32_t)(kClOrderHi >> (5u * (i - 12u))) & 31u;
}

// scratch needed only while a block header is parsed and its tables are built
struct BuildScratch {
  uint8_t  lens[320];     // code lengths of the current dynamic block (HLIT + HDIST)
  uint8_t  cltab[128];    // 7-bit lookup for the code-length code: sym << 3 | bits
  uint32_t cnt[16];       // codes per length
  uint32_t next[16];      // next canonical code per length
  uint16_t sub_lo[16];    // per length > root: first root prefix that holds such codes
  uint16_t sub_base[16];  // per length > root: table offset of that prefix's sub-table
  uint8_t  cl[32];        // the 19 code-length-code lengths
};

struct GroupSmem {
  uint16_t     lit[kLitTab];
  uint16_t     dist[kDistTab];
  BuildScratch s;
};

struct StreamOut {
  int32_t  code;
  uint32_t adler;
  uint32_t outlen;
  uint32_t inpos;
  uint32_t adler_stored;
  uint32_t info;
};

template <int G>
struct Lane {
  unsigned gmask;   // lanes of this group inside the warp
  int      lane;    // 0..G-1 inside the group
  int      wlane;   // lane inside the warp
};

// ---- bit reader ---------------------------------------------------------------------------
// Bit positions are counted from the 4-byte-aligned word that holds src[0].
struct BitReader {
  const uint32_t* w;
  uint32_t        wend;   // number of words that hold input bytes
  uint32_t        widx;   // index of the word in `lo`
  uint32_t        lo, hi, nx;
  uint32_t        bo;     // consumed bits of `lo` (may reach 63 between norm() calls)

  DEFL_DEV uint32_t ld(uint32_t i) const { return i < wend ? simt::ld_in32(w + i) : 0u; }
  DEFL_DEV void seek(uint64_t bitpos) {
    widx = (uint32_t)(bitpos >> 5);
    bo   = (uint32_t)(bitpos & 31u);
    lo = ld(widx); hi = ld(widx + 1); nx = ld(widx + 2);
  }
  DEFL_DEV uint32_t peek() const { return simt::funnel_r(lo, hi, bo); }
  DEFL_DEV void skip(uint32_t n) { bo += n; }
  DEFL_DEV void norm() {
    if (bo >= 32u) {
      bo -= 32u; lo = hi; hi = nx; widx++;
      nx = ld(widx + 2);
    }
  }
  DEFL_DEV uint64_t bitpos() const { return ((uint64_t)widx << 5) + bo; }
};

// ---- Adler-32 over dst[from, to), order-free partial sums ------------------------------------
// a = 1 + sum x_i ; b = n + sum (n - i) x_i  (i = 0-based position)  =>  keep A = sum x_i and
// W = sum i*x_i per lane (mod 65521) and combine once at the end of the stream.
template <int G>
DEFL_DEV void adler_range(const Lane<G>& L, const uint8_t* dst, uint32_t from, uint32_t to,
                          uint32_t& A, uint32_t& W) {
  while (from < to) {
    const uint32_t n = (to - from) < 32768u ? (to - from) : 32768u;
    const uint8_t* p = dst + from;
    uint32_t ta = 0;
    uint64_t tw = 0;
    uint32_t head = (4u - (uint32_t)((uintptr_t)p & 3u)) & 3u;
    if (head > n) head = n;
    const uint32_t nwords = (n - head) >> 2;
    const uint32_t tail0  = head + (nwords << 2);
    if (L.lane == 0) {
      for (uint32_t j = 0; j < head; j++) { uint32_t b = p[j]; ta += b; tw += j * b; }
      for (uint32_t j = tail0; j < n; j++) { uint32_t b = p[j]; ta += b; tw += j * b; }
    }
    const uint32_t* pw = reinterpret_cast<const uint32_t*>(p + head);
#pragma unroll 8
    for (uint32_t k = L.lane; k < nwords; k += G) {
      const uint32_t v = pw[k];
      const uint32_t s = simt::dp4a(v, 0x01010101u, 0u);
      const uint32_t q = simt::dp4a(v, 0x03020100u, 0u);
      ta += s;
      tw += (uint64_t)((head + (k << 2)) * s + q);
    }
    A = (A + ta) % kAdlerMod;
    W = (uint32_t)((W + (uint64_t)(from % kAdlerMod) * (ta % kAdlerMod) + tw % kAdlerMod) % kAdlerMod);
    from += n;
  }
}

template <int G>
DEFL_DEV uint32_t group_sum(const Lane<G>& L, uint32_t v) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += simt::shfl(L.gmask, v, L.wlane ^ o);
  return v;
}

// ---- copies -----------------------------------------------------------------------------------
// Non-overlapping copy from the (read-only) input to the output: stored blocks.
template <int G>
DEFL_DEV void copy_from_input(const Lane<G>& L, uint8_t* d, const uint8_t* s, uint32_t n) {
  uint32_t head = (4u - (uint32_t)((uintptr_t)d & 3u)) & 3u;
  if (head > n) head = n;
  for (uint32_t j = L.lane; j < head; j += G) d[j] = simt::ld_in8(s + j);
  d += head; s += head; n -= head;
  const uint32_t  nw = n >> 2;
  const uint32_t  sh = (uint32_t)((uintptr_t)s & 3u) * 8u;
  const uint32_t* sw = reinterpret_cast<const uint32_t*>((uintptr_t)s & ~(uintptr_t)3);
  uint32_t*       dw = reinterpret_cast<uint32_t*>(d);
  // eight independent loads per lane before the first store: one warp alone on a multi-megabyte stored
  // stream is bound by load latency, not by issue
  constexpr uint32_t U = 8;
  uint32_t k = L.lane;
  if (sh == 0) {
#pragma unroll 1
    for (; k + (U - 1u) * G < nw; k += U * G) {
      uint32_t v[U];
#pragma unroll
      for (uint32_t u = 0; u < U; u++) v[u] = simt::ld_in32(sw + k + u * G);
#pragma unroll
      for (uint32_t u = 0; u < U; u++) dw[k + u * G] = v[u];
    }
    for (; k < nw; k += G) dw[k] = simt::ld_in32(sw + k);
  } else {
#pragma unroll 1
    for (; k + (U - 1u) * G < nw; k += U * G) {
      uint32_t a[U], b[U];
#pragma unroll
      for (uint32_t u = 0; u < U; u++) { a[u] = simt::ld_in32(sw + k + u * G); b[u] = simt::ld_in32(sw + k + u * G + 1); }
#pragma unroll
      for (uint32_t u = 0; u < U; u++) dw[k + u * G] = simt::funnel_r(a[u], b[u], sh);
    }
    for (; k < nw; k += G) dw[k] = simt::funnel_r(simt::ld_in32(sw + k), simt::ld_in32(sw + k + 1), sh);
  }
  for (uint32_t j = (nw << 2) + L.lane; j < n; j += G) d[j] = simt::ld_in8(s + j);
}

// LZ77 match: d[i] = d[i - dist] for i in [0, len), byte i depends on byte i - dist.
template <int G>
DEFL_DEV void copy_match(const Lane<G>& L, uint8_t* d, uint32_t dist, uint32_t len) {
  const uint8_t* s = d - dist;
  if (G == 1) {
    for (uint32_t i = 0; i < len; i++) d[i] = s[i];
    return;
  }
  simt::syncwarp(L.gmask);  // literals stored by lane 0 must be visible to the other lanes
  if (dist >= len) {
    for (uint32_t i = L.lane; i < len; i += G) d[i] = s[i];
  } else if (dist >= (uint32_t)G) {
    // overlapping, period >= G: a round of G bytes only reads bytes of earlier rounds
    for (uint32_t base = 0; base < len; base += G) {
      const uint32_t i = base + L.lane;
      if (i < len) d[i] = s[i];
      simt::syncwarp(L.gmask);
    }
  } else if (dist == 1) {
    const uint8_t b = s[0];
    for (uint32_t i = L.lane; i < len; i += G) d[i] = b;
  } else {
    // short period: every byte comes from the dist bytes before the match
    for (uint32_t i = L.lane; i < len; i += G) d[i] = s[i % dist];
  }
  simt::syncwarp(L.gmask);
}

// ---- Huffman table construction ------------------------------------------------------------------
DEFL_DEV uint32_t lit_entry(uint32_t sym, uint32_t len) {
  if (sym < 256u) return (len << 11) | sym;
  if (sym == 256u) return (len << 11) | kLitEob;
  if (sym > 285u) return 0u;  // 286/287 take code space but never decode (infl.c:433-438)
  const uint32_t idx = sym - 257u;
  uint32_t xb, base;
  if (idx < 8u)        { xb = 0; base = idx; }                       // 3..10
  else if (idx == 28u) { xb = 0; base = 255u; }                      // 258
  else                 { xb = (idx >> 2) - 1u; base = (4u + (idx & 3u)) << xb; }
  return kLitIsLen | (len << 11) | (xb << 8) | base;                 // base holds length - 3
}

DEFL_DEV uint32_t dist_entry(uint32_t sym, uint32_t len) {
  if (sym > 29u) return 0u;   // 30/31 never decode (infl.c:445-446)
  const uint32_t xb = sym < 4u ? 0u : (sym >> 1) - 1u;
  const uint32_t m  = sym < 4u ? sym : (2u | (sym & 1u));
  return (len << 11) | (xb << 7) | m;
}

// Build one decode table from code lengths.  Accept/reject exactly as infl_ft_build
// (infl.c:478-490): over-subscribed => false; incomplete or empty => true.
// EMPTY: what a slot no code maps to holds (0; the v4 decoder asks for kLitEob with 0 code bits in the
// literal/length table, so that "not a plain literal" is one test)
template <int G, int ROOT_BITS, int TAB, bool IS_DIST, uint32_t EMPTY = 0u>
DEFL_DEV bool build_table(const Lane<G>& L, BuildScratch& sm, uint16_t* tab, const uint8_t* lens,
                          uint32_t n) {
  constexpr uint32_t ROOT = 1u << ROOT_BITS;

  if (L.lane < 16) sm.cnt[L.lane] = 0;
  if (G < 16) for (int i = L.lane + G; i < 16; i += G) sm.cnt[i] = 0;
  simt::syncwarp(L.gmask);
  for (uint32_t s = L.lane; s < n; s += G) {
    const uint32_t l = lens[s];
    if (l) simt::atomic_add_shared(&sm.cnt[l], 1u);
  }
  simt::syncwarp(L.gmask);

  // canonical first codes, Kraft check and sub-table layout: uniform, every lane the same
  bool     ok = true;
  int32_t  left = 1;
  uint32_t code = 0, prevcnt = 0;
  uint32_t cur_base = ROOT, lo_prev = 0, l_prev = 0, hi_last = 0;
#pragma unroll 1
  for (uint32_t l = 1; l <= 15; l++) {
    const uint32_t c = sm.cnt[l];
    left = (left << 1) - (int32_t)c;
    if (left < 0) ok = false;
    code = (code + prevcnt) << 1;
    prevcnt = c;
    if (L.lane == 0) sm.next[l] = code;
    if (l > (uint32_t)ROOT_BITS) {
      uint32_t lo = 0xFFFFu, base = 0;
      if (c) {
        lo = code >> (l - ROOT_BITS);
        if (l_prev) cur_base += (lo - lo_prev) << (l_prev - ROOT_BITS);
        base = cur_base;
        lo_prev = lo; l_prev = l;
        hi_last = (code + c - 1u) >> (l - ROOT_BITS);
      }
      if (L.lane == 0) { sm.sub_lo[l] = (uint16_t)lo; sm.sub_base[l] = (uint16_t)base; }
    }
  }
  if (!ok) return false;
  const uint32_t total = l_prev ? cur_base + ((hi_last - lo_prev + 1u) << (l_prev - ROOT_BITS)) : ROOT;
  if (total > (uint32_t)TAB) return false;  // cannot happen for <= 288 / 32 symbols

  {
    uint32_t* t32 = reinterpret_cast<uint32_t*>(tab);
    for (uint32_t i = L.lane; i < (total + 1u) / 2u; i += G) t32[i] = EMPTY | (EMPTY << 16);
  }
  simt::syncwarp(L.gmask);

#pragma unroll 1
  for (uint32_t base = 0; base < n; base += G) {
    const uint32_t s = base + L.lane;
    const uint32_t l = s < n ? lens[s] : 0u;
    uint32_t rank = 0, same = 1;
    if (G > 1) {
      const unsigned m = simt::match_any(L.gmask, l);
      rank = simt::popc(m & simt::lanemask_lt());
      same = simt::popc(m);
    }
    const uint32_t c0 = l ? sm.next[l] : 0u;
    simt::syncwarp(L.gmask);
    if (l && rank == 0) sm.next[l] = c0 + same;
    simt::syncwarp(L.gmask);
    if (!l) continue;

    const uint32_t cd    = c0 + rank;
    uint32_t entry = IS_DIST ? dist_entry(s, l) : lit_entry(s, l);
    if (EMPTY != 0u && entry == 0u) entry = EMPTY;
    if (l <= (uint32_t)ROOT_BITS) {
      const uint32_t r = simt::brev(cd) >> (32u - l);
      for (uint32_t i = r; i < ROOT; i += 1u << l) tab[i] = (uint16_t)entry;
    } else {
      const uint32_t p = cd >> (l - ROOT_BITS);
      // longest code length living under root prefix p decides that sub-table's size
      uint32_t lm = l;
      for (uint32_t k = 15; k > l; k--)
        if (sm.sub_lo[k] <= p) { lm = k; break; }
      const uint32_t sb   = lm - ROOT_BITS;
      const uint32_t off  = sm.sub_base[lm] + ((p - sm.sub_lo[lm]) << sb);
      const uint32_t slot = simt::brev(p) >> (32u - ROOT_BITS);
      tab[slot] = (uint16_t)(IS_DIST ? (kDistSub | (sb << 11) | (off - ROOT))
                                     : (kLitSub | (sb << 11) | (off - ROOT)));
      const uint32_t rb  = l - ROOT_BITS;                       // bits of this code below the root
      const uint32_t low = cd & ((1u << rb) - 1u);
      const uint32_t r   = simt::brev(low) >> (32u - rb);
      for (uint32_t i = r; i < (1u << sb); i += 1u << rb) tab[off + i] = (uint16_t)entry;
    }
  }
  simt::syncwarp(L.gmask);
  return true;
}

// 7-bit lookup for the code-length code (19 symbols, lengths 0..7).  Over-subscribed => false.
template <int G>
DEFL_DEV bool build_cl_table(const Lane<G>& L, BuildScratch& sm) {
  uint64_t cnt = 0;
  for (int i = 0; i < 19; i++) cnt += 1ull << (8 * sm.cl[i]);
  cnt &= ~0xFFull;
  int32_t  left = 1;
  uint64_t next = 0;
  uint32_t code = 0, prevcnt = 0;
  bool     ok = true;
  for (int l = 1; l <= 7; l++) {
    const uint32_t c = (uint32_t)(cnt >> (8 * l)) & 0xFFu;
    left = (left << 1) - (int32_t)c;
    if (left < 0) ok = false;
    code = (code + prevcnt) << 1;
    prevcnt = c;
    next |= (uint64_t)code << (8 * l);
  }
  if (!ok) return false;
  for (int i = L.lane; i < 32; i += G) reinterpret_cast<uint32_t*>(sm.cltab)[i] = 0u;
  simt::syncwarp(L.gmask);
  for (uint32_t s = 0; s < 19; s++) {
    const uint32_t l = sm.cl[s];
    if (!l) continue;
    const uint32_t cd = (uint32_t)(next >> (8 * l)) & 0xFFu;
    next += 1ull << (8 * l);
    const uint32_t r = simt::brev(cd) >> (32u - l);
    for (uint32_t i = r + ((uint32_t)L.lane << l); i < 128u; i += (uint32_t)G << l)
      sm.cltab[i] = (uint8_t)((s << 3) | l);
  }
  simt::syncwarp(L.gmask);
  return true;
}


// ---- block pieces shared by the group-per-stream kernel and the lane-decoder pipeline ---------------

// Stored block (infl_ft_stored, infl.c:659-715).  `br` stands right after the 3 header bits.
// Copies into dst[pos..), advances pos and re-seeks br.  Returns a Reason.
template <int G>
DEFL_DEV uint32_t stored_block(const Lane<G>& L, BitReader& br, const uint8_t* src, uint32_t srclen,
                               uint64_t base_bit, uint8_t* dst, uint32_t& pos, uint32_t cap) {
  const uint64_t bytepos = ((br.bitpos() - base_bit) + 7u) >> 3;  // from src[0]
  if (bytepos + 4u > srclen) return R_TRUNCATED;
  const uint8_t* p = src + bytepos;
  const uint32_t len  = simt::ld_in8(p) | ((uint32_t)simt::ld_in8(p + 1) << 8);
  const uint32_t nlen = simt::ld_in8(p + 2) | ((uint32_t)simt::ld_in8(p + 3) << 8);
  if ((len ^ (~nlen & 0xFFFFu)) != 0u) return R_STORED_LEN;
  if (len > cap - pos) return R_OUTPUT_FULL;
  if ((uint64_t)len > srclen - bytepos - 4u) return R_TRUNCATED;
  simt::syncwarp(L.gmask);
  copy_from_input<G>(L, dst + pos, p + 4, len);
  simt::syncwarp(L.gmask);
  pos += len;
  br.seek(base_bit + (bytepos + 4u + len) * 8u);
  return R_NONE;
}

// Fixed (btype 1) or dynamic (btype 2) block: leave decode tables for it in lit/dist.
// `br` stands right after the 3 header bits.  Returns a Reason.
template <int G, uint32_t EMPTY = 0u>
DEFL_DEV uint32_t huffman_block_setup(const Lane<G>& L, uint16_t* lit, uint16_t* dist, BuildScratch& s,
                                      BitReader& br, uint64_t end_bit, uint32_t btype, bool& fixed_ready) {
  if (btype == 1) {
    if (!fixed_ready) {
      // fixed code lengths (RFC 1951 3.2.6; reference table fxd, apicommon.h:53-64)
      simt::syncwarp(L.gmask);
      for (uint32_t i = L.lane; i < 320u; i += G)
        s.lens[i] = i < 144u ? 8 : i < 256u ? 9 : i < 280u ? 7 : i < 288u ? 8 : 5;
      simt::syncwarp(L.gmask);
      build_table<G, kLitRootBits, kLitTab, false, EMPTY>(L, s, lit, s.lens, 288u);
      build_table<G, kDistRootBits, kDistTab, true>(L, s, dist, s.lens + 288, 32u);
      fixed_ready = true;
    }
    return R_NONE;
  }
  // ---- dynamic header (infl_ft_dynamic, infl.c:903-1000)
  fixed_ready = false;
  int64_t rem = (int64_t)(end_bit - br.bitpos());
  if (rem < 14) return R_TRUNCATED;
  uint32_t v = br.peek();
  const uint32_t hlit = (v & 31u) + 257u, hdist = ((v >> 5) & 31u) + 1u, hclen = ((v >> 10) & 15u) + 4u;
  const uint32_t n = hlit + hdist;
  br.skip(14); br.norm();
  rem -= 14;
  if (rem < (int64_t)(3u * hclen)) return R_TRUNCATED;
  simt::syncwarp(L.gmask);
  if (L.lane == 0) {
    for (int i = 0; i < 20; i++) s.cl[i] = 0;
  }
  for (uint32_t i = L.lane; i < 80u; i += G) reinterpret_cast<uint32_t*>(s.lens)[i] = 0u;
  simt::syncwarp(L.gmask);
  for (uint32_t i = 0; i < hclen; i++) {
    const uint32_t v3 = br.peek() & 7u;
    br.skip(3); br.norm();
    if (L.lane == 0) s.cl[cl_order(i)] = (uint8_t)v3;
  }
  rem -= 3 * (int64_t)hclen;
  simt::syncwarp(L.gmask);
  if (!build_cl_table<G>(L, s)) return R_CODE_LENGTHS;

  uint32_t i = 0, prev = 0, reason = R_NONE;
#pragma unroll 1
  while (i < n) {
    if (rem < 7) { reason = R_TRUNCATED; break; }       // infl.c:946-948
    v = br.peek();
    const uint32_t e = s.cltab[v & 127u];
    if (!e) { reason = R_CODE_LENGTHS; break; }         // infl.c:951-952
    const uint32_t l = e & 7u, sym = e >> 3;
    uint32_t used = l;
    v >>= l;
    if (sym < 16u) {
      if (L.lane == 0) s.lens[i] = (uint8_t)sym;
      prev = sym;
      i++;
    } else {
      uint32_t rep, xb;
      if (sym == 16u)      { xb = 2; rep = 3u + (v & 3u); }
      else if (sym == 17u) { xb = 3; rep = 3u + (v & 7u); }
      else                 { xb = 7; rep = 11u + (v & 127u); }
      if (rem - (int64_t)l < (int64_t)xb || (sym == 16u && i == 0u) || i + rep > n) {
        reason = (rem - (int64_t)l < (int64_t)xb) ? R_TRUNCATED : R_CODE_LENGTHS;  // infl.c:960-985
        break;
      }
      used += xb;
      if (sym == 16u) {
        for (uint32_t k = L.lane; k < rep; k += G) s.lens[i + k] = (uint8_t)prev;
      } else {
        prev = 0;
      }
      i += rep;
    }
    br.skip(used); br.norm();
    rem -= used;
  }
  if (reason) return reason;
  simt::syncwarp(L.gmask);
  if (!build_table<G, kLitRootBits, kLitTab, false, EMPTY>(L, s, lit, s.lens, hlit) ||
      !build_table<G, kDistRootBits, kDistTab, true>(L, s, dist, s.lens + hlit, hdist))
    return R_TABLE;
  return R_NONE;
}

// zlib header (infl.c:1027-1039, zlib.h:45-96).  Returns a Reason; body = bytes to skip.
DEFL_DEV uint32_t zlib_header(const uint8_t* src, uint32_t srclen, uint32_t& body) {
  body = 0;
  if (srclen < 2) return R_TRUNCATED;
  const uint32_t cmf = simt::ld_in8(src), flg = simt::ld_in8(src + 1);
  if ((cmf & 15u) != 8u || (cmf >> 4) > 7u || ((cmf << 8) + flg) % 31u != 0u || (flg & 0x20u)) return R_ZLIB_HEADER;
  body = 2;
  return R_NONE;
}

// A symbol-level failure met at or beyond the end of the input is a truncation first: the symbol
// that ran off the end precedes it in stream order (matters to infl_stream() only).
DEFL_DEV uint32_t settle_reason(uint32_t reason, int64_t avail_bits) {
  if (avail_bits < 0 || ((reason == R_SYMBOL || reason == R_DISTANCE) && avail_bits < 15)) return R_TRUNCATED;
  return reason;
}

// Result record of a stream that decoded to its final block: Adler-32 from the partial sums,
// consumed bytes (mem.c:296-329), the stored trailer and the optional verification.
template <int G>
DEFL_DEV void finish_stream(const Lane<G>& L, const uint8_t* src, uint32_t srclen, uint32_t pos,
                            uint64_t consumed_bits, uint32_t A, uint32_t W, bool zlib, int flags, StreamOut& out) {
  const uint32_t sa = group_sum<G>(L, A) % kAdlerMod;
  const uint32_t sw = group_sum<G>(L, W) % kAdlerMod;
  const uint32_t nm = pos % kAdlerMod;
  const uint32_t a  = (1u + sa) % kAdlerMod;
  const uint32_t b  = (uint32_t)((nm + (uint64_t)nm * sa + kAdlerMod - sw) % kAdlerMod);
  out.adler  = (b << 16) | a;
  out.outlen = pos;
  out.inpos  = (uint32_t)((consumed_bits + 7u) >> 3);
  out.code   = 0;
  out.adler_stored = 0;
  out.info = 0;
  if (zlib && srclen - out.inpos >= 4u) {
    const uint8_t* t = src + out.inpos;
    out.adler_stored = ((uint32_t)simt::ld_in8(t) << 24) | ((uint32_t)simt::ld_in8(t + 1) << 16) |
                       ((uint32_t)simt::ld_in8(t + 2) << 8) | simt::ld_in8(t + 3);
    out.info |= 1u;                                   // INFL_INFO_HAS_TRAILER
    if (out.adler_stored == out.adler) out.info |= 2u; // INFL_INFO_ADLER_MATCH
  }
  if (zlib && (flags & 0x0100) && (out.info & 3u) != 3u) {  // INFL_BATCH_VERIFY_ADLER
    out.code = -1;
    out.info |= R_ADLER << 8;
  }
}

// ---- one stream, decoded start to finish by one group of G lanes -------------------------------------
template <int G>
DEFL_DEV void inflate_stream(const Lane<G>& L, GroupSmem& sm, const uint8_t* src, uint32_t srclen,
                             uint8_t* dst, uint32_t cap, int flags, StreamOut& out) {
  out.code = -1; out.adler = 1; out.outlen = 0; out.inpos = 0; out.adler_stored = 0; out.info = 0;

  // all pieces empty => UNZ_OK, nothing produced (infl.c:1547-1550)
  if (srclen == 0) { out.code = 0; return; }

  const bool zlib = (flags & 0xFF) == 1;
  uint32_t   body = 0;
  if (zlib) {
    const uint32_t hr = zlib_header(src, srclen, body);
    if (hr) { out.info = hr << 8; return; }
  }

  const uint32_t mis      = (uint32_t)((uintptr_t)src & 3u);
  const uint64_t base_bit = (uint64_t)mis * 8u;
  const uint64_t end_bit  = base_bit + (uint64_t)srclen * 8u;
  BitReader br;
  br.w    = reinterpret_cast<const uint32_t*>((uintptr_t)src - mis);
  br.wend = (uint32_t)(((uint64_t)mis + srclen + 3u) >> 2);
  br.seek(base_bit + body * 8u);

  uint32_t pos = 0, adler_pos = 0, A = 0, W = 0;
  uint32_t reason = R_NONE;
  bool     fixed_ready = false;

#pragma unroll 1
  for (;;) {
    // ---- block header (infl.c:1063-1070)
    if ((int64_t)(end_bit - br.bitpos()) < 3) { reason = R_TRUNCATED; break; }
    const uint32_t hdr    = br.peek();
    const uint32_t bfinal = hdr & 1u, btype = (hdr >> 1) & 3u;
    br.skip(3); br.norm();

    if (btype == 0) {
      reason = stored_block<G>(L, br, src, srclen, base_bit, dst, pos, cap);
      if (reason) break;
    } else if (btype == 3) {
      reason = R_BLOCK_TYPE;
      break;
    } else {
      reason = huffman_block_setup<G>(L, sm.lit, sm.dist, sm.s, br, end_bit, btype, fixed_ready);
      if (reason) break;

      // ---- symbols (infl_ft_block, infl.c:717-901)
#pragma unroll 1
      for (;;) {
        if (br.widx > br.wend) { reason = R_TRUNCATED; break; }  // far past the input: stop early
        uint32_t v = br.peek();
        uint32_t e = sm.lit[v & (kLitRoot - 1)];
        if ((e & (kLitIsLen | kLitKind)) == kLitSub)
          e = sm.lit[kLitRoot + (e & 0x1FFu) + ((v >> kLitRootBits) & ((1u << ((e >> 11) & 15u)) - 1u))];
        if (e == 0u) { reason = R_SYMBOL; break; }
        uint32_t cl = (e >> 11) & 15u;

        if (!(e & kLitIsLen)) {
          if (e & kLitKind) {            // end of block
            br.skip(cl); br.norm();
            break;
          }
          if (pos >= cap) { reason = R_OUTPUT_FULL; break; }
          if (L.lane == 0) dst[pos] = (uint8_t)e;
          pos++;
          // a second literal from the same 32-bit window (>= 17 bits are still valid)
          v >>= cl;
          uint32_t e2 = sm.lit[v & (kLitRoot - 1)];
          if ((e2 & (kLitIsLen | kLitKind)) == 0u && e2 != 0u && pos < cap) {
            if (L.lane == 0) dst[pos] = (uint8_t)e2;
            pos++;
            cl += (e2 >> 11) & 15u;
          }
          br.skip(cl); br.norm();
          continue;
        }

        // length + distance
        const uint32_t lxb = (e >> 8) & 7u;
        const uint32_t len = (e & 255u) + 3u + simt::bfe(v, cl, lxb);
        br.skip(cl + lxb); br.norm();
        v = br.peek();
        uint32_t de = sm.dist[v & (kDistRoot - 1)];
        if (de & kDistSub)
          de = sm.dist[kDistRoot + (de & 0x1FFu) + ((v >> kDistRootBits) & ((1u << ((de >> 11) & 15u)) - 1u))];
        if (de == 0u) { reason = R_DISTANCE; break; }
        const uint32_t dcl = (de >> 11) & 15u, dxb = (de >> 7) & 15u;
        const uint32_t dist = ((de & 3u) << dxb) + 1u + simt::bfe(v, dcl, dxb);
        br.skip(dcl + dxb); br.norm();

        if (dist > pos) { reason = R_DISTANCE; break; }          // infl.c:816-817
        if (len > cap - pos) { reason = R_OUTPUT_FULL; break; }  // infl.c:819-821
        copy_match<G>(L, dst + pos, dist, len);
        pos += len;
      }
      if (reason) {
        reason = settle_reason(reason, (int64_t)(end_bit - br.bitpos()));
        break;
      }
      // the reference fails a symbol whose bits are not all there (infl.c:741-745,806-810);
      // here over-reads see zeros/garbage and are caught once, at the end of the block
      if (br.bitpos() > end_bit) { reason = R_TRUNCATED; break; }
    }

    simt::syncwarp(L.gmask);
    adler_range<G>(L, dst, adler_pos, pos, A, W);
    adler_pos = pos;
    if (bfinal) break;
  }

  if (reason) { out.info = reason << 8; out.outlen = pos; return; }
  finish_stream<G>(L, src, srclen, pos, br.bitpos() - base_bit, A, W, zlib, flags, out);
}

}  // namespace defl
