// warp_decode.cuh -- batch inflate v3: ONE WARP PER STREAM, 32 lanes decode 32 sub-sequences of the
// current Huffman block at once.
//
// Why (profiles/r01_v2_*): with one decoding *thread* per stream (v2) an SM holds ~48 streams -- the
// decode tables of a stream take 4 KiB of shared memory -- and every symbol is a ~150-cycle dependent
// chain, so the SM idles on latency.  Here the 32 lanes of a warp share ONE table set and each decodes
// its own slice of the block's bits, so 32x more symbol chains are in flight per table.
//
// How (replaces the reference's serial loop infl_ft_block, infl.c:717-901, result-identical):
//   1. window  : the next 32*S bits of the block are cut into 32 sub-sequences of S bits.
//   2. spec    : lane i decodes from bit i*S *assuming* a symbol starts there, writes tokens to its own
//                segment of a global scratch and notes, every 128 bits, where a token started
//                ("checkpoints").  Lane 0 starts at a known boundary.
//   3. fix-up  : Huffman codes self-synchronise: lane i re-decodes from the position where lane i-1
//                really ended until it meets its own speculative trajectory at a checkpoint; from there
//                on the speculative tokens are the true ones.  Repeated until no lane's end moves
//                (a fixed point; at most 31 rounds, in practice one).
//   4. resolve : the warp walks the tokens in stream order, 32 at a time, and produces the bytes
//                one byte per lane (coalesced), which is also where distance / capacity are checked.
// Stored blocks, block headers, table construction and Adler-32 are the warp-cooperative pieces of
// inflate_core.cuh.
#pragma once
#include "inflate_core.cuh"
#include "../../include/defl/infl_batch.h"

namespace defl {

constexpr uint32_t kSubMax    = 8192;                 // bits per lane per window, at most
constexpr uint32_t kSubMin    = 512;
constexpr uint32_t kSubGrain  = 128;                  // a sub-sequence is a multiple of this many bits
constexpr uint32_t kCpTokens  = 8;                    // checkpoint spacing: where every 8th speculative token starts
constexpr uint32_t kSpecCap   = 1024;                 // tokens per lane per window (speculative pass, warm-up included)
constexpr uint32_t kFixCap    = 1024;                 // tokens per lane per window (fix-up pass)
constexpr uint32_t kCpPerLane = kSpecCap / kCpTokens;
constexpr size_t   kWarpScratchBytes = 32u * (kSpecCap + kFixCap) * 4u + 32u * kCpPerLane * 2u;
static_assert(kWarpScratchBytes % 16 == 0, "scratch slices stay 16-byte aligned");

#ifndef DEFL_WARM_DYNAMIC
#define DEFL_WARM_DYNAMIC 512
#endif
#ifndef DEFL_WARM_FIXED
#define DEFL_WARM_FIXED 4096
#endif
constexpr uint32_t kWarmDynamic = DEFL_WARM_DYNAMIC;  // bits a lane decodes ahead of its sub-sequence to get in step (dynamic codes)
constexpr uint32_t kWarmFixed   = DEFL_WARM_FIXED;    // fixed-Huffman codes are nearly fixed-length and re-synchronise slowly
#ifndef DEFL_STAGE_V3
#define DEFL_STAGE_V3 512
#endif
constexpr int      kStageV3   = DEFL_STAGE_V3;                  // output bytes resolved per token batch
constexpr uint32_t kTokIsMatch = 0x80000000u;         // [8:0] length, [23:9] distance - 1
constexpr uint32_t kTokIsDep   = 0x40000000u;         // resolver-side mark
                                                      // literal token: [26:24] count (1..2), [23:0] bytes

// lane status after a pass: 0 = ran to the end of its sub-sequence; else why it stopped
enum : uint32_t { LS_OK = 0, LS_EOB = 1, LS_FULL = 2, LS_FAIL = 3, LS_NONE = 4 };   // LS_FAIL | reason << 8

struct WarpSmem {
  uint16_t lit[kLitTab];
  uint16_t dist[kDistTab];
  uint32_t stop_lane;            // speculative pass: the lowest lane that met the end of the block (or no code)
  uint32_t pad_[3];
  union {
    BuildScratch s;              // while a block header is parsed
    uint16_t     stage[kStageV3];  // while tokens are resolved (bytes, or 16-bit symbols in split decode)
  };
};
static_assert(sizeof(WarpSmem) % 16 == 0, "WarpSmem alignment");

// ---- per-lane bit reader: a 96-bit window in registers plus one word already on its way (w3 is only
// ever moved, so its load has a whole refill period to land); positions relative to the window base word ----
struct SubReader {
  const uint32_t* w;
  uint32_t        wend, widx;
  uint32_t        w0, w1, w2, w3;
  uint32_t        bo;

  DEFL_DEV uint32_t ld(uint32_t i) const { return i < wend ? simt::ld_in32(w + i) : 0u; }
  // The lanes of an SM stream through more 128-byte lines than its L1 holds, and a line's first word comes
  // from HBM: ask for the line two ahead (into L2) whenever the reader enters a new one.
  DEFL_DEV void ahead(uint32_t i) const {
    if ((i & 31u) == 0u && i + 64u < wend) simt::prefetch_l2(w + i + 64u);
  }
  DEFL_DEV void seek(uint32_t word, uint32_t bit) {
    widx = word + (bit >> 5);
    bo   = bit & 31u;
    w0 = ld(widx); w1 = ld(widx + 1); w2 = ld(widx + 2); w3 = ld(widx + 3);
    if (widx + 32u < wend) simt::prefetch_l2(w + widx + 32u);
    if (widx + 64u < wend) simt::prefetch_l2(w + widx + 64u);
  }
  DEFL_DEV void consume(uint32_t n) {     // n <= 48, bo < 32 on entry
    bo += n;
    // a loop, not two ifs: if-converted, the second step's moves wait for the word the first step has only
    // just asked for, taken or not (13 % of all stall samples in profiles/r01_v3b_*); a real branch does not
#pragma unroll 1
    while (bo >= 32u) { w0 = w1; w1 = w2; w2 = w3; widx++; bo -= 32u; w3 = ld(widx + 3); ahead(widx + 3u); }
  }
};

// The reader of the speculative pass.  Same window as SubReader, but the words ahead travel global -> shared by
// cp.async, 16 bytes at a time, into two slots per lane (ring[slot * 128 + lane * 4 + word]) instead of one word at a
// time into a register: a quarter of the load instructions and of the L2 sectors touched (the lanes of an SM stream
// through more lines than its L1 holds, so a 4-byte load mostly paid for a whole sector), and nothing waits for a
// quad until four to eight words after it was requested.  drain() before the shared memory is used for anything else.
// Words are counted from the 16-byte boundary at or below the stream's first word, so a refill is a shift, a mask and
// one LDS; what lies past the stream's last word is zero-filled by the copy itself (its source size ends there).
struct RingReader {
  uint32_t        w0, w1, w2;
  uint32_t        bo;
  uint32_t        ai;              // the word that enters the window next (16-byte-boundary numbering)
  uint32_t        aend;            // one past the stream's last word, same numbering
  uint32_t*       ring;            // this lane's 4 words of slot 0
  const uint8_t*  quads;           // that boundary

  // quad q holds words 4q .. 4q + 3
  DEFL_DEV void request(uint32_t q) {
    const uint32_t first = 4u * q;
    const uint32_t nw = first < aend ? (aend - first < 4u ? aend - first : 4u) : 0u;
    simt::cp_async16_n(ring + (q & 1u) * 128u, quads + (nw ? (size_t)q * 16u : 0u), nw * 4u);
    simt::cp_async_commit();
  }
  DEFL_DEV void seek(const uint32_t* w, uint32_t wend, uint32_t word, uint32_t bit, uint32_t* ring_lane) {
    const uint32_t widx = word + (bit >> 5);
    const uint32_t skew = (uint32_t)(((uintptr_t)w >> 2) & 3u);
    bo = bit & 31u;
    w0 = widx < wend ? simt::ld_in32(w + widx) : 0u;
    w1 = widx + 1u < wend ? simt::ld_in32(w + widx + 1u) : 0u;
    w2 = widx + 2u < wend ? simt::ld_in32(w + widx + 2u) : 0u;
    ring  = ring_lane;
    quads = reinterpret_cast<const uint8_t*>((uintptr_t)w & ~(uintptr_t)15);
    aend  = wend + skew;
    ai    = widx + 3u + skew;
    request(ai >> 2); request((ai >> 2) + 1u);
  }
  DEFL_DEV void consume(uint32_t n) {     // n <= 48, bo < 32 on entry
    bo += n;
    // a loop, not two ifs: if-converted, the second step's moves wait for the word the first step has only just
    // asked for, taken or not
#pragma unroll 1
    while (bo >= 32u) {
      simt::cp_async_wait<1>();                          // the quad of word ai has landed (the next may be on its way)
      w0 = w1; w1 = w2;
      w2 = ring[(ai & 4u) * 32u + (ai & 3u)];
      if ((ai & 3u) == 3u) request((ai >> 2) + 2u);      // its slot is free now
      ai++; bo -= 32u;
    }
  }
  DEFL_DEV void drain() { simt::cp_async_wait<0>(); }
};

DEFL_DEV uint32_t fail_stat(uint32_t reason) { return LS_FAIL | (reason << 8); }
DEFL_DEV uint32_t low_bits(uint32_t v, uint32_t n) { return v & ~(0xFFFFFFFFu << n); }    // n < 32

// Decode the token that starts at the reader's position.  `avail` = input bits from here to the end
// of the stream; a symbol whose bits are not all there fails (infl.c:741-745,806-810).
// A second literal joins the first in one token when it starts less than `pack_before` bits from here (0: never).
// Returns LS_OK with (tok, used), LS_EOB with used = the code's bits, or LS_FAIL | reason << 8.
template <typename READER>
DEFL_DEV uint32_t decode_token(const uint16_t* lit, const uint16_t* dist, const READER& br, uint32_t avail,
                               uint32_t pack_before, uint32_t& tok, uint32_t& used) {
  const uint32_t v0 = simt::funnel_r(br.w0, br.w1, br.bo);
  const uint32_t v1 = simt::funnel_r(br.w1, br.w2, br.bo);
  uint32_t       e  = lit[v0 & (kLitRoot - 1)];
  if ((e & (kLitIsLen | kLitKind)) == kLitSub)
    e = lit[kLitRoot + (e & 0x1FFu) + ((v0 >> kLitRootBits) & ((1u << ((e >> 11) & 15u)) - 1u))];
  const uint32_t cl = (e >> 11) & 15u;
  used = cl;
  tok = 0;
  if (e == 0u) return fail_stat(avail < 15u ? R_TRUNCATED : R_SYMBOL);
  if ((e & (kLitIsLen | kLitKind)) == kLitEob) return cl > avail ? fail_stat(R_TRUNCATED) : (uint32_t)LS_EOB;

  const bool     is_len = (e & kLitIsLen) != 0u;
  const uint32_t xb   = (e >> 8) & 7u;                        // 0 for a literal
  const uint32_t val  = (e & 255u) + simt::bfe(v0, cl, xb);   // byte, or length - 3
  const uint32_t use1 = cl + xb;                              // <= 20
  const uint32_t v2   = simt::funnel_r(v0, v1, use1);         // the 32 bits after this symbol
  uint32_t       use2;
  if (is_len) {
    uint32_t x = dist[v2 & (kDistRoot - 1)];
    if (x & kDistSub)
      x = dist[kDistRoot + (x & 0x1FFu) + ((v2 >> kDistRootBits) & ((1u << ((x >> 11) & 15u)) - 1u))];
    if (x == 0u) return fail_stat((use1 > avail || avail - use1 < 15u) ? R_TRUNCATED : R_DISTANCE);
    const uint32_t dcl = (x >> 11) & 15u, dxb = (x >> 7) & 15u;
    const uint32_t d1  = ((x & 3u) << dxb) + simt::bfe(v2, dcl, dxb);   // distance - 1
    use2 = dcl + dxb;
    tok  = kTokIsMatch | (d1 << 9) | (val + 3u);
  } else {
    const uint32_t x   = lit[v2 & (kLitRoot - 1)];
    const uint32_t xcl = (x >> 11) & 15u;
    const bool two = use1 < pack_before && (x & (kLitIsLen | kLitKind)) == 0u && x != 0u && use1 + xcl <= avail;
    use2 = two ? xcl : 0u;
    tok  = two ? ((2u << 24) | ((x & 255u) << 8) | val) : ((1u << 24) | val);
  }
  used = use1 + use2;
  if (used > avail) return fail_stat(R_TRUNCATED);
  return LS_OK;
}

// what one lane knows about its sub-sequence after the passes
struct LaneTraj {
  uint32_t land;   // window-relative bit where its trajectory ended (first token start >= its limit, or the stop)
  uint32_t stat;   // LS_*
  uint32_t n;      // speculative tokens written (warm-up tokens in front included)
  uint32_t m;      // first speculative token that belongs to the true trajectory
  uint32_t nf;     // fix-up tokens in front of them
  uint32_t T;      // where its true trajectory starts
  uint32_t org;    // checkpoint c says: speculative token kCpTokens * c starts at bit org + cps[c]
  uint32_t ncp;    // checkpoints recorded (the v4 engine's own scheme, warp_decode4.cuh)
  uint32_t s_land, s_stat, s_n;   // how the speculative trajectory ended (valid again after any later merge)
};

// speculative pass: decode from `start` (window-relative bits) to `limit`.
// The loop is the hot loop of the whole engine, so it is written for the instruction count:
//  * every trip yields exactly one token, so the lanes' token counts move in step and the checkpoint store (where
//    every 8th token starts) is a uniform branch, taken by all lanes in the same trip;
//  * the warm-up runs in the same loop: the lane starts `warm` bits early and simply remembers which token was the
//    first to start at or after `start` (tj.m; the tokens before it stay in the segment, unused).  Codes
//    self-synchronise, so by then the lane is most likely ON the true trajectory and its first kept token starts
//    exactly where its predecessor will end: no fix-up needed.  If the warm-up runs into anything that is not a
//    token (end of block, no code) the lane starts again, blind, at `start`;
//  * a second literal joins a token only if it STARTS before the end of the phase (start for the warm-up, limit
//    after it), so that a trajectory leaves a phase on the first SYMBOL boundary at or after its end -- the same
//    one for a lane and for its warmed-up successor, however their literal pairs fell;
//  * no end-of-input arithmetic (it stops 64 bits short of the end; decode_token() takes over there, and for
//    anything unusual), one code path for "length + distance" and "literal + maybe a second literal" (the second
//    lookup goes to lit[] or dist[] by index, not by branch), a 32-bit token index.
DEFL_DEV void lane_spec(const uint16_t* lit, const uint16_t* dist, const uint32_t* words, uint32_t wend,
                        uint32_t bw, uint32_t start, uint32_t limit, uint32_t endr, uint32_t* seg,
                        uint16_t* cps, LaneTraj& tj, uint32_t warm, uint32_t* ring_lane, bool active) {
  // ALL 32 lanes come here (an inactive one only keeps the others company at the two barriers below).
  RingReader rr;
  uint32_t   org = start - warm;
  uint32_t   r = org, r0 = start, m0 = 0, ti = 0, stat = LS_OK;
  if (active) rr.seek(words, wend, bw, r, ring_lane);
  const uint32_t  endr64  = endr >= 64u ? endr - 64u : 0u;
  const uint32_t  dist_at = (uint32_t)(dist - lit);            // dist[] follows lit[] in shared memory
  // Two phases, the warm-up and the sub-sequence itself, through the same loop body.  The lanes start each phase
  // TOGETHER (lane 0 has no warm-up and waits): a lane that went on alone would run the loop on its own, and every
  // trip of it would cost the warp a full pass.
#pragma unroll 1
  for (uint32_t phase = 0; phase < 2u; phase++) {
    simt::syncwarp(kFull);
    const uint32_t lim = phase == 0u ? start : limit;          // the end of this phase
    const uint32_t fast_limit = lim < endr64 ? lim : endr64;
    if (active && (phase == 1u || warm != 0u)) {
      for (;;) {
        // ---- fast loop: a checkpoint, then up to kCpTokens tokens
        while (r < fast_limit && ti < kSpecCap) {
          if ((ti & (kCpTokens - 1u)) == 0u) cps[ti / kCpTokens] = (uint16_t)(r - org);
          bool odd = false;                                           // something that is not a token
#pragma unroll 1
          do {
            const uint32_t v0 = simt::funnel_r(rr.w0, rr.w1, rr.bo);
            uint32_t       e  = lit[v0 & (kLitRoot - 1)];
            if ((e & (kLitIsLen | kLitKind)) == kLitSub)
              e = lit[kLitRoot + (e & 0x1FFu) + low_bits(v0 >> kLitRootBits, (e >> 11) & 15u)];
            if (e < 0x0800u || (e & (kLitIsLen | kLitKind)) == kLitEob) { odd = true; break; }   // no code here / end of block
            const uint32_t v1 = simt::funnel_r(rr.w1, rr.w2, rr.bo);
            const bool     is_len = (e & kLitIsLen) != 0u;
            const uint32_t cl   = (e >> 11) & 15u;
            const uint32_t xb   = (e >> 8) & 7u;                        // 0 for a literal
            const uint32_t val  = (e & 255u) + low_bits(v0 >> cl, xb);  // byte, or length - 3
            const uint32_t use1 = cl + xb;
            const uint32_t v2   = simt::funnel_r(v0, v1, use1);         // the 32 bits after this symbol
            uint32_t       x    = lit[is_len ? dist_at + (v2 & (kDistRoot - 1)) : (v2 & (kLitRoot - 1))];
            if (is_len && (x & kDistSub))
              x = dist[kDistRoot + (x & 0x1FFu) + low_bits(v2 >> kDistRootBits, (x >> 11) & 15u)];
            if (is_len && x == 0u) { odd = true; break; }               // no distance code here
            const uint32_t xcl = (x >> 11) & 15u;
            const uint32_t dxb = (x >> 7) & 15u;
            const uint32_t d1  = ((x & 3u) << dxb) + low_bits(v2 >> xcl, dxb);          // distance - 1
            const bool     two = (x & (kLitIsLen | kLitKind)) == 0u && x >= 0x0800u && r + use1 < lim;   // a second literal joins
            // selects, not branches: literal lanes and match lanes sit side by side in every warp
            const uint32_t tok_m = kTokIsMatch | (d1 << 9) | (val + 3u);
            const uint32_t tok_l = simt::byte_perm(val, x, 0x1140u) + (two ? (2u << 24) : (1u << 24));   // the second byte only counts when two
            const uint32_t use2  = is_len ? xcl + dxb : (two ? xcl : 0u);
            seg[ti++] = is_len ? tok_m : tok_l;
            r += use1 + use2;
            rr.consume(use1 + use2);
          } while ((ti & (kCpTokens - 1u)) != 0u && r < fast_limit);
          if (odd) break;
        }
        // ---- one careful step: the ends of the phase, of the input and of the block, failures
        if (r >= lim) break;
        uint32_t s = LS_FULL, tok = 0, used = 0;
        if (ti != kSpecCap) {
          if ((ti & (kCpTokens - 1u)) == 0u) cps[ti / kCpTokens] = (uint16_t)(r - org);
          s = decode_token(lit, dist, rr, endr > r ? endr - r : 0u, lim - r, tok, used);
        }
        if (s == LS_OK) {
          seg[ti++] = tok;
          r += used;
          rr.consume(used);
          continue;
        }
        if (phase == 0u) {                                          // start again, blind, where the sub-sequence begins
          org = start; r = start; ti = 0;
          rr.drain();
          rr.seek(words, wend, bw, r, ring_lane);
        } else {
          stat = s;
          if (s == LS_EOB) r += used;
        }
        break;
      }
    }
    if (phase == 0u) { m0 = ti; r0 = r; }                            // the first kept token
  }
  if (active) rr.drain();
  if (!active) return;
  tj.land = r; tj.stat = stat; tj.n = ti; tj.m = m0; tj.nf = 0; tj.T = r0; tj.org = org;
  tj.s_land = r; tj.s_stat = stat; tj.s_n = ti;
}

// fix-up pass: decode from the true start `T` until the speculative trajectory is met at a checkpoint
DEFL_DEV void lane_redo(const uint16_t* lit, const uint16_t* dist, const uint32_t* words, uint32_t wend,
                        uint32_t bw, uint32_t limit, uint32_t endr, uint32_t* fix,
                        const uint16_t* cps, uint32_t T, LaneTraj& tj) {
  SubReader br;
  br.w = words; br.wend = wend;
  br.seek(bw, T);
  uint32_t r = T, f = 0, k = 0;
  const uint32_t ncp = (tj.s_n + kCpTokens - 1u) / kCpTokens;    // checkpoints 0 .. ncp - 1 were written
  uint32_t target = k < ncp ? tj.org + cps[k] : 0xFFFFFFFFu;
  tj.T = T;
  for (;;) {
    while (r > target) {
      k++;
      target = k < ncp ? tj.org + cps[k] : 0xFFFFFFFFu;
    }
    if (r == target) {            // met the speculative trajectory: its tokens from kCpTokens * k on are true
      tj.m = kCpTokens * k; tj.nf = f;
      tj.land = tj.s_land; tj.stat = tj.s_stat; tj.n = tj.s_n;
      return;
    }
    uint32_t stat = LS_OK;
    if (r >= limit) {
      // ran through the whole sub-sequence without meeting it
    } else if (f == kFixCap) {
      stat = LS_FULL;
    } else {
      uint32_t tok, used;
      const uint32_t s = decode_token(lit, dist, br, endr > r ? endr - r : 0u, limit - r, tok, used);
      if (s == LS_OK) {
        fix[f++] = tok;
        r += used;
        br.consume(used);
        continue;
      }
      stat = s;
      if (s == LS_EOB) r += used;
    }
    tj.land = r; tj.stat = stat; tj.n = 0; tj.m = 0; tj.nf = f;
    return;
  }
}

// ---- resolve: tokens -> bytes, in stream order ------------------------------------------------------------
// toks[0..count) are consecutive tokens of the stream; the bytes they produce start at dst[P].
// Checks what needs the absolute position (infl.c:816-821): distance > produced so far, output full.
// Returns a Reason; P advances over everything produced before a failing token.
// SYM = uint8_t: plain bytes.  SYM = uint16_t (split decode of one large stream, split_decode.cuh): the
// chunk does not know the 32 KiB before its start, so a byte copied from there becomes the marker
// 0x8000 | k (k = index into that window) and markers are copied around like bytes; `hist` = 32768 then.
//
// Up to 32 tokens per batch, ONE TOKEN PER LANE (profiles/r01_v3_*: the earlier one-byte-per-lane walk cost
// 2.7 warp instructions per byte, half of the kernel).  The batch is assembled in a shared-memory stage laid
// out with the same 4-byte phase as its place in dst, then stored as aligned words:
//   1. lanes with a literal token, or a short match whose source lies wholly before the batch, fill their
//      own symbols: the source is read as aligned 32-bit words and funnel-shifted into place;
//   2. long matches that source before the batch are copied by the whole warp, one after another;
//   3. matches that read symbols of this very batch follow in token order, by the whole warp, from the stage;
//   4. the stage goes to dst as aligned words (symbols at the ragged ends singly).
template <typename SYM> struct SymWord;
template <> struct SymWord<uint8_t>  { static constexpr uint32_t kPer = 4, kShift = 2, kBits = 8,  kMask = 0xFFu; };
template <> struct SymWord<uint16_t> { static constexpr uint32_t kPer = 2, kShift = 1, kBits = 16, kMask = 0xFFFFu; };
#ifndef DEFL_LANE_WORDS
#define DEFL_LANE_WORDS 3   // 12 bytes: 189 GB/s on config 2 against 185 (16 bytes) and 154 (8 bytes), profiles/README.md round 2
#endif
#ifndef DEFL_LANE_WORDS16
#define DEFL_LANE_WORDS16 6   // 12 symbols: mixed batch 80.5 -> 83.6 GB/s against 4 (72 registers; at 64 it spilled and lost)
#endif
// a lane copies matches of up to this many words' worth of symbols itself (bytes / 16-bit symbols)
template <typename SYM> struct LaneWords { static constexpr uint32_t k = sizeof(SYM) == 2 ? DEFL_LANE_WORDS16 : DEFL_LANE_WORDS; };

// Adler-32 partial sums of the bytes a lane has sent to dst (order-free form: sum x and sum (position - base) * x),
// taken from the very words the resolver stores -- no second pass over the output.
struct AdlerAcc {
  uint32_t ta;
  uint64_t tw;
  uint32_t base;
};

template <typename SYM>
DEFL_DEV uint32_t resolve_tokens(uint16_t* stage_mem, int lane, const uint32_t* toks, uint32_t count, SYM* dst0,
                                 uint32_t cap, uint32_t& P, uint32_t hist, AdlerAcc* ad = nullptr) {
  using SW = SymWord<SYM>;
  constexpr uint32_t kPer = SW::kPer;
  constexpr uint32_t kLaneWords = LaneWords<SYM>::k;
  constexpr uint32_t kRoom = (uint32_t)(kStageV3 * sizeof(uint16_t) / sizeof(SYM)) - kPer;   // symbols per batch
  SYM* const      st  = reinterpret_cast<SYM*>(stage_mem);
  uint32_t* const stw = reinterpret_cast<uint32_t*>(stage_mem);
  uint32_t h = 0;
  uint32_t nxt = (uint32_t)lane < count ? simt::ld_tok(toks + lane) : 0u;
#pragma unroll 1
  while (h < count) {
    uint32_t n = count - h < 32u ? count - h : 32u;
    uint32_t tok = nxt;
    nxt = h + 32u + (uint32_t)lane < count ? simt::ld_tok(toks + h + 32u + lane) : 0u;    // lands during this batch
    uint32_t olen = (tok >> 31) ? (tok & 511u) : ((tok >> 24) & 7u);

    uint32_t inc = olen;                                   // inclusive prefix sum of output lengths
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = simt::shfl_up(kFull, inc, d);
      if (lane >= d) inc += t;
    }
    const uint32_t o  = inc - olen;                        // this token's first symbol inside the batch
    const uint32_t d1 = (tok >> 9) & 0x7FFFu;              // distance - 1
    const bool bad_dist = (tok >> 31) && d1 >= P + o + hist;
    const bool bad_full = olen != 0u && inc > cap - P;
    const bool nofit    = inc > kRoom;
    // all three are monotone or first-hit: keep the tokens before the first stop
    const unsigned stop = simt::ballot(kFull, bad_dist || bad_full || nofit || (uint32_t)lane >= n);
    const uint32_t keep = stop ? (uint32_t)simt::ffs(stop) - 1u : 32u;
    if (keep == 0u) {
      // token 0 of the batch fails (one token always fits the stage: <= 258 symbols)
      const uint32_t bd = simt::shfl(kFull, (uint32_t)bad_dist, 0);
      return bd ? R_DISTANCE : R_OUTPUT_FULL;
    }
    const bool short_batch = keep != 32u && h + keep < count;
    n = keep;
    if ((uint32_t)lane >= n) { tok = 0u; olen = 0u; }
    const uint32_t B  = simt::shfl(kFull, inc, (int)n - 1);
    SYM* const     dst = dst0 + P;
    const uint32_t a0 = (uint32_t)(((uintptr_t)dst & 3u) / sizeof(SYM));      // stage index of the batch's symbol 0
    const bool     is_match = (tok >> 31) != 0u;
    const bool     dep  = is_match && (d1 + 1u < o + olen);                   // reads symbols of this batch
    const bool     wide = is_match && !dep && olen > kLaneWords * kPer;       // phase 2

    simt::syncwarp(kFull);            // the previous batch's stage has been stored
    // ---- 1. every lane fills its own short token
    if (!is_match) {
      if (olen) st[a0 + o] = (SYM)(tok & 255u);
      if (olen > 1u) st[a0 + o + 1u] = (SYM)((tok >> 8) & 255u);
    } else if (!dep && !wide) {
      const int32_t   g0 = (int32_t)o - (int32_t)(d1 + 1u);                   // first source symbol, relative to dst (< 0)
      const uintptr_t sa = (uintptr_t)(dst + g0);
      const uint32_t* wp = reinterpret_cast<const uint32_t*>(sa & ~(uintptr_t)3);
      const uint32_t  sh = (uint32_t)(sa & 3u) * 8u;
      const uint32_t  last_word = (uint32_t)(((sa & 3u) + olen * sizeof(SYM) - 1u) >> 2);   // words the source spans - 1
      // 16-bit symbols: what lies before the chunk is never read (those symbols become markers below)
      const uint32_t* const wmin = sizeof(SYM) == 2 ? reinterpret_cast<const uint32_t*>((uintptr_t)dst0 & ~(uintptr_t)3) : nullptr;
      // all the words first (independent loads in flight together), then shift and store
      uint32_t w[kLaneWords + 1];
#pragma unroll
      for (uint32_t i = 0; i <= kLaneWords; i++)
        w[i] = (i <= last_word && (sizeof(SYM) == 1 || wp + i >= wmin)) ? simt::ld_outw(wp + i) : 0u;
#pragma unroll
      for (uint32_t wi = 0; wi < kLaneWords; wi++) {
        const uint32_t k = wi * kPer;
        const uint32_t v = simt::funnel_r(w[wi], w[wi + 1], sh);
#pragma unroll
        for (uint32_t i = 0; i < kPer; i++) {
          if (k + i < olen) {
            uint32_t x = (v >> (i * SW::kBits)) & SW::kMask;
            if (sizeof(SYM) == 2) {
              const int32_t gi = (int32_t)P + g0 + (int32_t)(k + i);
              if (gi < 0) x = 0x8000u | (uint32_t)(gi + 32768);
            }
            st[a0 + o + k + i] = (SYM)x;
          }
        }
      }
    }
    // ---- 2. long matches from before the batch: one symbol per lane across all of them, 128 per step
    if (simt::any(kFull, wide)) {
      const unsigned le = simt::lanemask_le();
      const uint32_t p_lo = simt::reduce_min(kFull, wide ? o : 0xFFFFFFFFu);
      const uint32_t p_hi = simt::reduce_max(kFull, wide ? o + olen : 0u);
      const bool     valid = olen != 0u;
      const uint32_t ks = o >> 5, kb = 1u << (o & 31u);
#pragma unroll 1
      for (uint32_t base = p_lo & ~31u; base < p_hi; base += 128u) {
        uint32_t val[4];
        bool     have[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const uint32_t rb = base + 32u * (uint32_t)u;
          have[u] = false;
          val[u] = 0;
          if (rb < p_hi) {                                   // uniform
            const uint32_t k = rb >> 5;
            const uint32_t before = (uint32_t)simt::popc(simt::ballot(kFull, valid && ks < k));
            const uint32_t starts = simt::reduce_or(kFull, (valid && ks == k) ? kb : 0u);
            const int      i  = (int)((before + (uint32_t)simt::popc(starts & le) - 1u) & 31u);
            const uint32_t t  = simt::shfl(kFull, tok, i);
            const uint32_t wd = simt::shfl(kFull, (uint32_t)wide, i);
            const uint32_t j  = rb + (uint32_t)lane;
            if (j < B && wd) {
              const int32_t back = (int32_t)j - (int32_t)(((t >> 9) & 0x7FFFu) + 1u);
              if (sizeof(SYM) == 2 && (int32_t)P + back < 0) val[u] = 0x8000u | (uint32_t)((int32_t)P + back + 32768);
              else val[u] = simt::ld_out(dst + back);
              have[u] = true;
            }
          }
        }
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (have[u]) st[a0 + base + 32u * (uint32_t)u + (uint32_t)lane] = (SYM)val[u];
      }
    }
    // ---- 3. matches that read symbols of this very batch: in token order, each by the whole warp
    unsigned depmask = simt::ballot(kFull, dep);
#pragma unroll 1
    while (depmask) {
      const int i = simt::ffs(depmask) - 1;
      depmask &= depmask - 1u;
      const uint32_t t  = simt::shfl(kFull, tok, i);
      const uint32_t oi = simt::shfl(kFull, o, i);
      const uint32_t len = t & 511u, dist = ((t >> 9) & 0x7FFFu) + 1u;
      simt::syncwarp(kFull);
      if (dist >= 32u || dist >= len) {
        // no lane reads what another writes in the same step; a later step may (dist < len)
        for (uint32_t r0 = 0; r0 < len; r0 += 32u) {
          const uint32_t r = r0 + (uint32_t)lane;
          if (r < len) {
            const int32_t so = (int32_t)oi - (int32_t)dist + (int32_t)r;
            uint32_t v;
            if (so >= 0) v = st[a0 + so];
            else if (sizeof(SYM) == 2 && (int32_t)P + so < 0) v = 0x8000u | (uint32_t)((int32_t)P + so + 32768);
            else v = simt::ld_out(dst + so);
            st[a0 + oi + r] = (SYM)v;
          }
          if (dist < len) simt::syncwarp(kFull);
        }
      } else {
        // a period shorter than a step: every symbol is one of the `dist` symbols before the match
        for (uint32_t r = (uint32_t)lane; r < len; r += 32u) {
          const int32_t so = (int32_t)oi - (int32_t)dist + (int32_t)(r % dist);
          uint32_t v;
          if (so >= 0) v = st[a0 + so];
          else if (sizeof(SYM) == 2 && (int32_t)P + so < 0) v = 0x8000u | (uint32_t)((int32_t)P + so + 32768);
          else v = simt::ld_out(dst + so);
          st[a0 + oi + r] = (SYM)v;
        }
      }
    }
    simt::syncwarp(kFull);
    // ---- 4. stage -> dst: aligned words, the ragged ends symbol by symbol
    {
      const uint32_t s_end = a0 + B;
      const uint32_t w_lo  = (a0 + kPer - 1u) >> SW::kShift;
      const uint32_t w_hi  = s_end >> SW::kShift;
      uint32_t* const gw = reinterpret_cast<uint32_t*>((uintptr_t)dst & ~(uintptr_t)3);     // word of stage word 0
      uint32_t bs = 0, bq = 0;               // this lane's bytes of the batch: sum x, sum (offset in the batch) * x
      if (w_hi > w_lo) {
        for (uint32_t w = w_lo + (uint32_t)lane; w < w_hi; w += 32u) {
          const uint32_t v = stw[w];
          gw[w] = v;
          if (sizeof(SYM) == 1 && ad) {
            const uint32_t sv = simt::dp4a(v, 0x01010101u, 0u);
            bs += sv;
            bq += ((w << 2) - a0) * sv + simt::dp4a(v, 0x03020100u, 0u);
          }
        }
        const uint32_t head = (w_lo << SW::kShift) - a0;                    // < kPer
        const uint32_t tail = s_end - (w_hi << SW::kShift);                 // < kPer
        if ((uint32_t)lane < head) {
          const SYM b = st[a0 + lane];
          dst[lane] = b;
          if (sizeof(SYM) == 1 && ad) { bs += b; bq += (uint32_t)lane * b; }
        }
        if ((uint32_t)lane < tail) {
          const SYM b = st[s_end - tail + lane];
          dst[B - tail + lane] = b;
          if (sizeof(SYM) == 1 && ad) { bs += b; bq += (B - tail + lane) * b; }
        }
      } else {
        if ((uint32_t)lane < B) {                                           // B < 2 * kPer here
          const SYM b = st[a0 + lane];
          dst[lane] = b;
          if (sizeof(SYM) == 1 && ad) { bs += b; bq += (uint32_t)lane * b; }
        }
      }
      if (sizeof(SYM) == 1 && ad) {
        ad->ta += bs;
        ad->tw += (uint64_t)(P - ad->base) * bs + bq;
      }
    }
    P += B;
    h += n;
    if (short_batch) nxt = h + (uint32_t)lane < count ? simt::ld_tok(toks + h + lane) : 0u;
  }
  simt::syncwarp(kFull);
  return R_NONE;
}

// ---- one window of a Huffman block -------------------------------------------------------------------------
struct WarpScratch {
  uint32_t* spec;   // [32][kSpecCap]
  uint32_t* fix;    // [32][kFixCap]
  uint16_t* cps;    // [32][kCpPerLane]
};

DEFL_DEV WarpScratch scratch_of(void* base) {
  WarpScratch s;
  s.spec = static_cast<uint32_t*>(base);
  s.fix  = s.spec + 32u * kSpecCap;
  s.cps  = reinterpret_cast<uint16_t*>(s.fix + 32u * kFixCap);
  return s;
}

// Where the COUNT pass of the split decode keeps a chunk's tokens for the DECODE pass (split_decode.cuh).
struct TokSink {
  uint32_t* buf;      // nullptr: nothing is kept
  uint32_t  cap, n;   // budget and fill, in tokens
  bool      ok;       // false once the budget ran out (n is meaningless then)
};

// Decode from absolute bit `cur` (a token boundary inside a Huffman block whose tables are in sm) as far
// as one window reaches.  On return `cur` is where the stream goes on; the result is LS_OK (window done,
// block goes on), LS_EOB (block ended at cur), or LS_FAIL | reason << 8.
// COUNT: only add up what the window's tokens would produce (first pass of the split decode).
template <typename SYM, bool COUNT>
DEFL_DEV uint32_t decode_window(WarpSmem& sm, const WarpScratch& sc, int lane, const uint32_t* words,
                                uint32_t wend, uint64_t end_bit, uint64_t& cur, SYM* dst, uint32_t cap,
                                uint32_t& pos, uint32_t hist, TokSink* sink = nullptr,
                                uint64_t soft_end = ~0ull, uint32_t warm_bits = kWarmDynamic, AdlerAcc* ad = nullptr) {
  // soft_end (> cur): no token that STARTS at or after this bit is decoded in this window, so the window ends on
  // the first token boundary >= soft_end (the caller compares `cur` with it: split decode links chunks there)
  const uint32_t bw   = (uint32_t)(cur >> 5);
  const uint32_t rel0 = (uint32_t)(cur & 31u);
  const uint64_t left64 = end_bit - ((uint64_t)bw << 5);
  const uint32_t endr = left64 > 0x40000000ull ? 0x40000000u : (uint32_t)left64;   // window-relative end of input
  const uint64_t soft64 = soft_end == ~0ull ? ~0ull : soft_end - ((uint64_t)bw << 5);
  const uint32_t softr = soft64 > (uint64_t)endr ? endr : (uint32_t)soft64;        // window-relative soft end
  const uint32_t remaining = softr - rel0;
  uint32_t S = (remaining + 31u) / 32u;
  S = (S + kSubGrain - 1u) / kSubGrain * kSubGrain;
  S = S < kSubMin ? kSubMin : (S > kSubMax ? kSubMax : S);

  const uint32_t start = rel0 + (uint32_t)lane * S;
  const uint32_t limit = (soft_end != ~0ull && start + S > softr) ? softr : start + S;
  const bool     active = start < softr || lane == 0;
  uint32_t* seg = sc.spec + (uint32_t)lane * kSpecCap;
  uint32_t* fix = sc.fix + (uint32_t)lane * kFixCap;
  uint16_t* cps = sc.cps + (uint32_t)lane * kCpPerLane;

  LaneTraj tj;
  tj.land = start; tj.stat = LS_NONE; tj.n = 0; tj.m = 0; tj.nf = 0; tj.T = start; tj.org = start;
  tj.s_land = start; tj.s_stat = LS_NONE; tj.s_n = 0;
  const uint32_t warm = (lane >= 1 && active) ? (warm_bits < S ? warm_bits : S) : 0u;      // lane 0 starts on a known boundary
  // the resolver's stage is idle during the speculative pass: it holds the lanes' input rings
  simt::syncwarp(kFull);
  lane_spec(sm.lit, sm.dist, words, wend, bw, start, limit, endr, seg, cps, tj, warm,
            reinterpret_cast<uint32_t*>(sm.stage) + lane * 4, active);
  simt::syncwarp(kFull);

  // fix-up rounds until every lane starts where its predecessor ended
#pragma unroll 1
  for (int round = 0; round < 32; round++) {
    const uint32_t pland = simt::shfl_up(kFull, tj.land, 1);
    const uint32_t pstat = simt::shfl_up(kFull, tj.stat, 1);
    const bool need = lane >= 1 && active && pstat == LS_OK && pland != tj.T;
    if (!simt::any(kFull, need)) break;
    if (need) lane_redo(sm.lit, sm.dist, words, wend, bw, limit, endr, fix, cps, pland, tj);
    simt::syncwarp(kFull);
  }

  // a lane's fix-up tokens move in front of its true speculative tokens when the discarded ones left room:
  // one run of tokens per lane for the resolver instead of two
  if (active && tj.nf != 0u && tj.nf <= tj.m && tj.n > tj.m) {
    uint32_t* to = seg + (tj.m - tj.nf);
    for (uint32_t i = 0; i < tj.nf; i++) to[i] = simt::ld_tok(fix + i);
    tj.m -= tj.nf;
    tj.nf = 0;
  }

  // the first lane whose true trajectory stopped ends the window; without one, the last active lane does
  const unsigned act  = simt::ballot(kFull, active);
  const unsigned stop = simt::ballot(kFull, active && tj.stat != LS_OK);
  const int      last = stop ? simt::ffs(stop) - 1 : 31 - simt::clz(act);
  simt::syncwarp(kFull);

  uint32_t reason = R_NONE;
  if (COUNT) {
    uint32_t bytes = 0;
    if (sink && sink->buf && sink->ok) {
      // the warp walks the true tokens in stream order: adds up their bytes and keeps them (coalesced)
      uint32_t w = sink->n;
      bool     ok = true;
#pragma unroll 1
      for (int L = 0; L <= last; L++) {
        const uint32_t nf = simt::shfl(kFull, tj.nf, L);
        const uint32_t m  = simt::shfl(kFull, tj.m, L);
        const uint32_t n  = simt::shfl(kFull, tj.n, L);
#pragma unroll 1
        for (int part = 0; part < 2; part++) {
          const uint32_t* p = part == 0 ? sc.fix + (uint32_t)L * kFixCap : sc.spec + (uint32_t)L * kSpecCap + m;
          const uint32_t  c = part == 0 ? nf : (n > m ? n - m : 0u);
          if (c > sink->cap - w) ok = false;
          for (uint32_t i = (uint32_t)lane; i < c; i += 32u) {
            const uint32_t t = simt::ld_tok(p + i);
            bytes += (t >> 31) ? (t & 511u) : ((t >> 24) & 7u);
            if (ok) sink->buf[w + i] = t;
          }
          if (ok) w += c;
        }
      }
      sink->n = w;
      sink->ok = ok;
    } else if (lane <= last) {   // every lane adds up its own true tokens
      for (uint32_t i = 0; i < tj.nf; i++) {
        const uint32_t t = simt::ld_tok(fix + i);
        bytes += (t >> 31) ? (t & 511u) : ((t >> 24) & 7u);
      }
      for (uint32_t i = tj.m; i < tj.n; i++) {
        const uint32_t t = simt::ld_tok(seg + i);
        bytes += (t >> 31) ? (t & 511u) : ((t >> 24) & 7u);
      }
    }
    simt::syncwarp(kFull);
    pos += simt::reduce_add(kFull, bytes);
  } else {
#pragma unroll 1
    for (int L = 0; L <= last && reason == R_NONE; L++) {
      const uint32_t nf = simt::shfl(kFull, tj.nf, L);
      const uint32_t m  = simt::shfl(kFull, tj.m, L);
      const uint32_t n  = simt::shfl(kFull, tj.n, L);
      if (nf) reason = resolve_tokens<SYM>(sm.stage, lane, sc.fix + (uint32_t)L * kFixCap, nf, dst, cap, pos, hist, ad);
      if (reason == R_NONE && n > m)
        reason = resolve_tokens<SYM>(sm.stage, lane, sc.spec + (uint32_t)L * kSpecCap + m, n - m, dst, cap, pos, hist, ad);
    }
  }
  if (reason) return fail_stat(reason);

  const uint32_t land = simt::shfl(kFull, tj.land, last);
  const uint32_t stat = simt::shfl(kFull, tj.stat, last);
  cur = ((uint64_t)bw << 5) + land;
  if (stat == LS_FULL) return LS_OK;        // a token segment filled up: the next window starts there
  return stat;
}

// ---- one stream, start to finish, by one warp ------------------------------------------------------------------
// Where a decode that ran out of input (or output room) can be taken up again: the last block boundary it passed.
// infl_stream() keeps one per handle on the device (reference: the saved state of stream.c:468-794).
struct StreamResume {
  uint64_t bit;      // bits consumed from src[0] up to that boundary
  uint32_t pos;      // bytes produced up to it
  uint32_t valid;    // 0: start at the stream's first byte
};

DEFL_DEV void inflate_stream_warp(WarpSmem& sm, const WarpScratch& sc, int lane, const uint8_t* src,
                                  uint32_t srclen, uint8_t* dst, uint32_t cap, int flags, StreamOut& out,
                                  StreamResume* rs = nullptr) {
  Lane<32> L;
  L.gmask = kFull; L.lane = lane; L.wlane = lane;
  out.code = -1; out.adler = 1; out.outlen = 0; out.inpos = 0; out.adler_stored = 0; out.info = 0;
  if (srclen == 0) { out.code = 0; return; }       // all pieces empty => UNZ_OK (infl.c:1547-1550)

  const bool zlib = (flags & 0xFF) == 1;
  const bool resume = rs != nullptr && rs->valid != 0u;
  const uint64_t rbit = resume ? rs->bit : 0u;
  uint32_t   pos = resume ? rs->pos : 0u;
  simt::syncwarp(kFull);                             // every lane has read *rs before lane 0 writes it again
  uint32_t   body = 0;
  if (zlib && !resume) {
    const uint32_t hr = zlib_header(src, srclen, body);
    if (hr) { out.info = hr << 8; return; }
  }
  const uint32_t mis      = (uint32_t)((uintptr_t)src & 3u);
  const uint64_t base_bit = (uint64_t)mis * 8u;
  const uint64_t end_bit  = base_bit + (uint64_t)srclen * 8u;
  BitReader br;
  br.w    = reinterpret_cast<const uint32_t*>((uintptr_t)src - mis);
  br.wend = (uint32_t)(((uint64_t)mis + srclen + 3u) >> 2);
  br.seek(resume ? base_bit + rbit : base_bit + body * 8u);

  uint32_t reason = R_NONE;
  bool     fixed_ready = false;
  // Adler-32: from the resolver's own stores (and over a stored block right after its copy); a stream that is
  // taken up at a block boundary (infl_stream) sums its whole output once it is complete instead
  const bool fused = rs == nullptr;
  uint32_t A = 0, W = 0;
  AdlerAcc ad;
  ad.ta = 0; ad.tw = 0; ad.base = 0;
#pragma unroll 1
  for (;;) {
    if ((int64_t)(end_bit - br.bitpos()) < 3) { reason = R_TRUNCATED; break; }
    const uint32_t hdr    = br.peek();
    const uint32_t bfinal = hdr & 1u, btype = (hdr >> 1) & 3u;
    br.skip(3); br.norm();
    if (btype == 0) {
      const uint32_t pos0 = pos;
      reason = stored_block<32>(L, br, src, srclen, base_bit, dst, pos, cap);
      if (reason) break;
      if (fused) { simt::syncwarp(kFull); adler_range<32>(L, dst, pos0, pos, A, W); }
    } else if (btype == 3) {
      reason = R_BLOCK_TYPE;
      break;
    } else {
      simt::syncwarp(kFull);
      reason = huffman_block_setup<32>(L, sm.lit, sm.dist, sm.s, br, end_bit, btype, fixed_ready);
      if (reason) break;
      simt::syncwarp(kFull);
      uint64_t cur = br.bitpos();
      uint32_t st;
#pragma unroll 1
      do {
        ad.base = pos;
        st = decode_window<uint8_t, false>(sm, sc, lane, br.w, br.wend, end_bit, cur, dst, cap, pos, 0u, nullptr, ~0ull,
                                           btype == 1u ? kWarmFixed : kWarmDynamic, fused ? &ad : nullptr);
        if (fused) {                                   // fold this window's sums (positions relative to its start)
          const uint32_t a = ad.ta % kAdlerMod;
          A = (A + a) % kAdlerMod;
          W = (uint32_t)((W + (uint64_t)(ad.base % kAdlerMod) * a + ad.tw % kAdlerMod) % kAdlerMod);
          ad.ta = 0; ad.tw = 0;
        }
      } while (st == LS_OK);
      if (st != LS_EOB) { reason = st >> 8; break; }
      br.seek(cur);
    }
    if (bfinal) break;
    if (rs != nullptr && lane == 0) {                // a block boundary: a later call can go on from here
      rs->bit = br.bitpos() - base_bit; rs->pos = pos; rs->valid = 1u;
    }
  }
  if (reason) { out.info = reason << 8; out.outlen = pos; return; }

  simt::syncwarp(kFull);
  if (!fused) adler_range<32>(L, dst, 0, pos, A, W);
  finish_stream<32>(L, src, srclen, pos, br.bitpos() - base_bit, A, W, zlib, flags, out);
}

}  // namespace defl
