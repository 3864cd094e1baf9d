// warp_decode4.cuh -- batch inflate v4: one warp per stream, built for the instruction count.
//
// Same plan as v3 (warp_decode.cuh: 32 lanes decode 32 sub-sequences of a Huffman block speculatively, fix
// up where a lane started off a token boundary, then resolve the tokens in stream order), re-cut after the
// round-1 profile (profiles/r01_v3e_*: 4.4 warp instructions per output byte, issue-bound):
//   * a TOKEN is now "up to 4 literals, then a match" in 64 bits: 6.4 output bytes per token on text
//     instead of 4.1, and one loop iteration of a lane yields a whole token (the match part runs with all
//     lanes in step, the literal part is a short inner loop);
//   * the lanes read their input from a shared-memory ring (36 words per lane) that is topped up with
//     16-byte cp.async every fourth iteration, all lanes together: no per-token refill branch;
//   * the resolver keeps the last kilobytes of output in a shared-memory stage: matches that reach back
//     less than that never touch global memory, the stage goes to dst as aligned 16-byte stores, and
//     Adler-32 is taken from those very registers (no second pass over dst);
//   * tokens of a batch whose source is final are copied in one round, the rest in further rounds
//     (multi-round resolution) instead of one dependent token after another.
// Replaces infl_ft_block + the copy kernels (reference src/infl/infl.c:717-901, :269-376); result-identical
// (oracle/inflate_oracle.c).  Block headers, tables, stored blocks: inflate_core.cuh.
#pragma once
#include "inflate_core.cuh"
#include "warp_decode.cuh"

namespace defl {
namespace v4 {

constexpr uint32_t kSubMax    = 8192;                 // bits per lane per window, at most
constexpr uint32_t kSubMin    = 512;
constexpr uint32_t kCpBits    = 256;                  // checkpoint spacing (a token is at most 94 bits)
constexpr uint32_t kCpPerLane = kSubMax / kCpBits;
constexpr uint32_t kSpecCap   = 1024;                 // tokens per lane per window (speculative pass)
constexpr uint32_t kFixCap    = 256;                  // tokens per lane per window (fix-up pass)
constexpr size_t   kWarpScratchBytes = 32u * (kSpecCap + kFixCap) * 8u + 32u * 96u * 4u;
static_assert(kWarpScratchBytes % 16 == 0, "scratch slices stay 16-byte aligned");

constexpr uint32_t kEmptyLit  = kLitEob;              // literal/length slots no code maps to: "end of block, 0 bits"
constexpr uint32_t kNotLit    = kLitIsLen | kLitKind; // any of these bits: not a plain literal

// token: x = [8:0] match length (0: none)  [23:9] distance - 1  [26:24] literals in front (0..4)
//        y = the literal bytes, the first one in bits 7:0
DEFL_DEV uint32_t tok_len(uint32_t x)  { return x & 511u; }
DEFL_DEV uint32_t tok_d1(uint32_t x)   { return (x >> 9) & 0x7FFFu; }
DEFL_DEV uint32_t tok_nlit(uint32_t x) { return (x >> 24) & 7u; }

constexpr uint32_t kRingWords = 36;                   // per lane: 32 ring words + a copy of the first 4 (no wrap inside a read)
constexpr uint32_t kStageBytes = 4096;                // resolver stage
constexpr uint32_t kBatchMax   = 1024;                // output bytes of one resolver batch, at most (+4: a trimmed token's literals)
constexpr uint32_t kKeepBytes  = 1024;                // history that survives a shift of the stage

struct WarpSmem4 {
  uint16_t lit[kLitTab];
  uint16_t dist[kDistTab];
  uint32_t stop_lane;                                 // speculative pass: the lowest lane that met the end of the block
  uint32_t pad_[3];
  union {
    BuildScratch s;                                   // while a block header is parsed
    uint32_t     ring[32 * kRingWords];               // speculative pass: the lanes' input rings
    uint8_t      stage[kStageBytes + 64];             // resolver: output not yet in dst + recent history
  };
};
static_assert(sizeof(WarpSmem4) % 16 == 0, "WarpSmem4 alignment");

// ---- where a stream's bits are ---------------------------------------------------------------------------
// "A-space": bit positions counted from the 16-byte boundary at or below the 4-byte word that holds src[0].
struct Input4 {
  const uint32_t* words;     // 4-byte aligned word that holds src[0]
  uint32_t        wend;      // words that hold input bytes
  const uint8_t*  q0;        // the 16-byte boundary
  uint32_t        skew;      // words between q0 and words (0..3)
  uint32_t        abytes;    // bytes from q0 to the end of the input
};

// ---- the speculative pass's reader: a ring of 32 words per lane in shared memory ----------------------------
// Positions r are relative to a window base that is a multiple of 1024 bits in A-space, so the ring slot of
// the word that holds bit r is (r >> 5) & 31 and quad (r >> 7) lives in slot quad & 7.
struct Ring4 {
  uint32_t*      ring;       // this lane's 36 words
  const uint8_t* qbase;      // address of quad 0 of the window
  uint32_t       qoff;       // that quad's index counted from q0
  uint32_t       abytes, skew;
  uint32_t       have;       // quads [.., have) have been asked for

  uint32_t       qlo, qhi;   // quads [qlo, qhi) of the window lie wholly inside the input: one plain 16-byte copy each

  DEFL_DEV void request(uint32_t q) {
    uint32_t* const slot = ring + (q & 7u) * 4u;
    const uint8_t*  g    = qbase + (size_t)q * 16u;
    if (q >= qlo && q < qhi) {
      simt::cp_async16(slot, g, true);
      if ((q & 7u) == 0u) simt::cp_async16(ring + 32u, g, true);
      return;
    }
    request_edge(q, slot, g);
  }
  // the quads at the ends of the input: nothing outside [first word, last byte] is read
  DEFL_DEV void request_edge(uint32_t q, uint32_t* slot, const uint8_t* g) {
    const uint32_t aq   = qoff + q;                               // quad index from q0
    const uint32_t left = abytes > aq * 16u ? abytes - aq * 16u : 0u;
    if (aq == 0u && skew != 0u) {
      // the first quad starts before the stream's first word: word by word, and nothing below that word
#pragma unroll 1
      for (uint32_t i = 0; i < 4; i++) {
        const bool in = i >= skew && left > i * 4u;
        simt::cp_async4(slot + i, in ? g + i * 4u : g + skew * 4u, in);
        if ((q & 7u) == 0u) simt::cp_async4(ring + 32u + i, in ? g + i * 4u : g + skew * 4u, in);
      }
      return;
    }
    const uint32_t nb = left < 16u ? left : 16u;
    simt::cp_async16_n(slot, nb ? g : qbase, nb);
    if ((q & 7u) == 0u) simt::cp_async16_n(ring + 32u, nb ? g : qbase, nb);
  }
  DEFL_DEV void init(uint32_t* ring_lane, const Input4& in, uint32_t base_quad, uint32_t r) {
    ring = ring_lane; abytes = in.abytes; skew = in.skew; qoff = base_quad;
    qbase = in.q0 + (size_t)base_quad * 16u;
    const uint32_t full = in.abytes >> 4;                         // whole quads of input, counted from q0
    qlo = (base_quad == 0u && in.skew != 0u) ? 1u : 0u;
    qhi = full > base_quad ? full - base_quad : 0u;
    have = r >> 7;
    const uint32_t lim = have + 8u;
#pragma unroll 1
    for (; have < lim; have++) request(have);
    simt::cp_async_commit();
  }
  // every fourth token: whatever was asked for before has landed by now; ask for up to three more quads
  // (a token is at most 94 bits, so four of them use less than three quads)
  DEFL_DEV void topup(uint32_t r) {
    simt::cp_async_wait<0>();
    const uint32_t lim = (r >> 7) + 8u;
#pragma unroll
    for (int j = 0; j < 3; j++)
      if (have < lim) { request(have); have++; }
    simt::cp_async_commit();
  }
  DEFL_DEV void drain() { simt::cp_async_wait<0>(); }
};

// ---- one token, every check made (end of input, codes that do not exist): the slow path -----------------------
// Decodes from the reader's position: up to 4 literals and a match.  `avail` = input bits from here to the end
// of the stream; a symbol whose bits are not all there fails (infl.c:741-745,806-810).
// Returns LS_OK  (tok complete, used = its bits),
//         LS_EOB (tok = the literals in front of the end-of-block code, used = their bits, extra = the code's bits),
//         LS_FAIL | reason << 8 (tok / used likewise: what was decoded before the failing symbol).
DEFL_DEV uint32_t careful_token(const uint16_t* lit, const uint16_t* dist, SubReader& br, uint32_t avail,
                                uint2& tok, uint32_t& used, uint32_t& extra) {
  uint32_t lits = 0, nlit = 0;
  used = 0; extra = 0;
  for (;;) {
    const uint32_t v0 = simt::funnel_r(br.w0, br.w1, br.bo);
    const uint32_t v1 = simt::funnel_r(br.w1, br.w2, br.bo);
    uint32_t       e  = lit[v0 & (kLitRoot - 1)];
    if ((e & kNotLit) == kLitSub)
      e = lit[kLitRoot + (e & 0x1FFu) + ((v0 >> kLitRootBits) & ((1u << ((e >> 11) & 15u)) - 1u))];
    const uint32_t cl = (e >> 11) & 15u;
    tok = make_uint2(nlit << 24, lits);
    if ((e & kNotLit) == kLitEob) {
      if (cl == 0u) return fail_stat(avail < 15u ? R_TRUNCATED : R_SYMBOL);     // no code maps here
      if (cl > avail) return fail_stat(R_TRUNCATED);
      extra = cl;
      return LS_EOB;
    }
    if (!(e & kLitIsLen)) {
      if (cl > avail) return fail_stat(R_TRUNCATED);
      lits |= (e & 255u) << (8u * nlit);
      nlit++;
      used += cl; avail -= cl;
      br.consume(cl);
      // the same rule as the fast loop (lane_spec4), so both cut a run of literals into the same tokens
      if (nlit == 4u || used >= 32u) { tok = make_uint2(nlit << 24, lits); return LS_OK; }
      continue;
    }
    const uint32_t xb   = (e >> 8) & 7u;
    const uint32_t len  = (e & 255u) + simt::bfe(v0, cl, xb) + 3u;
    const uint32_t use1 = cl + xb;                                  // <= 20
    const uint32_t v2   = simt::funnel_r(v0, v1, use1);
    uint32_t x = dist[v2 & (kDistRoot - 1)];
    if (x & kDistSub)
      x = dist[kDistRoot + (x & 0x1FFu) + ((v2 >> kDistRootBits) & ((1u << ((x >> 11) & 15u)) - 1u))];
    if (x == 0u) return fail_stat((use1 > avail || avail - use1 < 15u) ? R_TRUNCATED : R_DISTANCE);
    const uint32_t dcl = (x >> 11) & 15u, dxb = (x >> 7) & 15u;
    const uint32_t d1  = ((x & 3u) << dxb) + simt::bfe(v2, dcl, dxb);
    const uint32_t u   = use1 + dcl + dxb;
    if (u > avail) return fail_stat(R_TRUNCATED);
    tok = make_uint2((nlit << 24) | (d1 << 9) | len, lits);
    used += u;
    br.consume(u);
    return LS_OK;
  }
}


// ---- speculative pass: one lane, one sub-sequence -------------------------------------------------------------
// Decodes from `start` (after a warm-up of `warm` bits in front of it whose tokens are thrown away) to the first
// token boundary >= limit.  Positions are window-relative bits; `bw` = index in in.words of the window's word 0.
// Every lane of the warp yields one token per trip, so the trips are counted once for all: every fourth trip the
// rings are topped up and the lane looks whether it still has to go on, every sixteenth it notes where it stands
// (a checkpoint: position | tokens written << 20) for the fix-up pass.  Nothing of this is done per token.
constexpr uint32_t kCpTokens = 16;
constexpr uint32_t kCpMax    = 96;        // checkpoints per lane: (warm-up tokens + kSpecCap) / 16, rounded up

// literal k of a token goes to byte k of the token's second word
#define DEFL_LIT_SEL(k) ((k) == 0 ? 0x3214u : (k) == 1 ? 0x3240u : (k) == 2 ? 0x3410u : 0x4210u)

DEFL_DEV void lane_spec4(const uint16_t* lit, const uint16_t* dist, const Input4& in, uint32_t base_quad,
                         uint32_t bw, uint32_t start, uint32_t limit, uint32_t endr, uint2* seg, uint32_t* cps,
                         LaneTraj& tj, uint32_t warm, uint32_t* ring_lane, uint32_t* stop_lane, uint32_t lane) {
  // stop_lane: nullptr inside fixed-Huffman blocks (a lane that starts blind there meets the 7-bit end-of-block code
  // within a few symbols: the hint would be wrong most of the time)
  uint32_t r = start - warm, r0 = start;
  Ring4 rg;
  rg.init(ring_lane, in, base_quad, r);
  uint32_t       stat = LS_OK, trips4 = 0, ncp = 0;
  uint32_t       ti = 0;                       // tokens kept
  uint32_t       keep = warm == 0u ? 1u : 0u;  // 0 while warming up
  const uint32_t fast_limit = endr >= 128u ? (limit < endr - 128u ? limit : endr - 128u) : 0u;
  for (;;) {
    // ---- fast loop: no end-of-input arithmetic (it stops 128 bits short of the end), one token per trip
    while (r < fast_limit) {
      // -- every fourth trip
      rg.topup(r);
      // a lane below met the end of the block (or something that is no code): most likely everything from
      // here on is not part of the block.  Stop as if the token segment was full; if that lane turns out to
      // have been off the true trajectory, the next window simply starts here.
      if ((stop_lane && lane > *reinterpret_cast<volatile uint32_t*>(stop_lane)) || ti + 4u > kSpecCap) {
        if (!keep) { r0 = start; r = start; }
        stat = LS_FULL;
        break;
      }
      if ((trips4 & 3u) == 0u && ncp < kCpMax) cps[ncp++] = r | (ti << 20);
      trips4++;
      bool odd = false;
#pragma unroll 1
      for (int j = 0; j < 4 && r < fast_limit; j++) {
        if (!keep && r >= start) { keep = 1u; r0 = r; }       // the warm-up ends on this token boundary
        const uint32_t* p = rg.ring + ((r >> 5) & 31u);
        const uint32_t a0 = p[0], a1 = p[1], a2 = p[2], a3 = p[3];
        uint32_t V0 = simt::funnel_r(a0, a1, r), V1 = simt::funnel_r(a1, a2, r), V2 = simt::funnel_r(a2, a3, r);
        uint32_t u = 0, nlit = 0, lits = 0, e = 0, x = V0;
        // literals: up to 4, as long as the next one starts in the first 32 bits.  Four steps without a branch:
        // every lane looks up all four, a lane that has met its first non-literal (kept in e) only stops taking
        // them.  (Lanes differ in how many literals they have; a loop would run as long as the longest.)
        bool go = true, more = false;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const uint32_t ek  = lit[x & (kLitRoot - 1)];
          const bool     isl = (ek & kNotLit) == 0u;
          if (go && !isl) { e = ek; more = (ek & kNotLit) == kLitSub; }
          const bool take = go && isl;
          lits = take ? simt::byte_perm(lits, ek, DEFL_LIT_SEL(k)) : lits;
          nlit = take ? (uint32_t)k + 1u : nlit;
          u   += take ? ek >> 11 : 0u;
          go   = take && u < 32u;
          x = simt::funnel_r(V0, V1, u);
        }
        if (more) {
          // a code longer than the root table's index: the same steps, rolled, with the second-level lookup
#pragma unroll 1
          for (;;) {
            if ((e & kNotLit) == kLitSub)
              e = lit[kLitRoot + (e & 0x1FFu) + ((x >> kLitRootBits) & ((1u << ((e >> 11) & 15u)) - 1u))];
            if (e & kNotLit) break;
            lits |= (e & 255u) << (8u * nlit);
            nlit++;
            u += e >> 11;
            if (nlit == 4u || u >= 32u) { e = 0u; break; }
            x = simt::funnel_r(V0, V1, u);
            e = lit[x & (kLitRoot - 1)];
          }
        }
        uint32_t w0 = nlit << 24, used = u;
        if (e & kLitIsLen) {
          if (u >= 32u) { V0 = V1; V1 = V2; V2 = 0u; }       // the funnel shifts below take u mod 32
          const uint32_t M0 = simt::funnel_r(V0, V1, u), M1 = simt::funnel_r(V1, V2, u);
          const uint32_t cl = (e >> 11) & 15u, xb = (e >> 8) & 7u;
          const uint32_t len  = (e & 255u) + simt::bfe(M0, cl, xb) + 3u;
          const uint32_t use1 = cl + xb;
          const uint32_t D    = simt::funnel_r(M0, M1, use1);
          uint32_t xd = dist[D & (kDistRoot - 1)];
          if (xd & kDistSub)
            xd = dist[kDistRoot + (xd & 0x1FFu) + ((D >> kDistRootBits) & ((1u << ((xd >> 11) & 15u)) - 1u))];
          const uint32_t dcl = (xd >> 11) & 15u, dxb = (xd >> 7) & 15u;
          const uint32_t d1  = ((xd & 3u) << dxb) + simt::bfe(D, dcl, dxb);
          odd  = xd == 0u;
          w0  |= (d1 << 9) | len;
          used = u + use1 + dcl + dxb;
        } else if (e != 0u) {
          odd = true;                                        // end of block, or no code here
        }
        if (odd) {
          // the literals in front of the odd symbol are good; the careful step below deals with the symbol
          if (nlit) { if (keep) seg[ti] = make_uint2(nlit << 24, lits); ti += keep; r += u; }
          break;
        }
        if (keep) seg[ti] = make_uint2(w0, lits);
        ti += keep;
        r += used;
      }
      if (odd) break;
    }
    // ---- one careful step: the ends of the sub-sequence, the input and the block, failures
    if (r >= limit || stat != LS_OK) break;
    if (!keep && r >= start) { keep = 1u; r0 = r; }
    if (ti == kSpecCap) { stat = LS_FULL; break; }
    SubReader br;
    br.w = in.words; br.wend = in.wend;
    br.seek(bw, r);
    uint2    tok;
    uint32_t used, extra;
    const uint32_t s = careful_token(lit, dist, br, endr > r ? endr - r : 0u, tok, used, extra);
    if (s != LS_OK && !keep) {
      // the warm-up ran into something that is not a token: start blind at `start`, as without a warm-up
      r = start;
      rg.drain();
      rg.init(ring_lane, in, base_quad, r);
      continue;
    }
    if (s == LS_OK || tok_nlit(tok.x) != 0u) { if (keep) seg[ti] = tok; ti += keep; r += used; }
    if (s != LS_OK) {
      stat = s;
      if (s == LS_EOB) r += extra;
      break;
    }
  }
  rg.drain();
  if (stop_lane && (stat == LS_EOB || (stat & 0xFFu) == LS_FAIL)) simt::atomic_min_shared(stop_lane, lane);
  tj.land = r; tj.stat = stat; tj.n = ti; tj.m = 0; tj.nf = 0; tj.T = r0; tj.ncp = ncp;
  tj.s_land = r; tj.s_stat = stat; tj.s_n = ti;
}

// fix-up pass: decode from the true start `T` until the speculative trajectory is met at a checkpoint
DEFL_DEV void lane_redo4(const uint16_t* lit, const uint16_t* dist, const Input4& in, uint32_t bw, uint32_t start,
                         uint32_t limit, uint32_t endr, uint2* fix, const uint32_t* cps, uint32_t T, LaneTraj& tj) {
  SubReader br;
  br.w = in.words; br.wend = in.wend;
  br.seek(bw, T);
  uint32_t r = T, f = 0, k = 0;
  const uint32_t ncp = tj.ncp;
  uint32_t c = k < ncp ? cps[k] : 0xFFFFFFFFu;
  tj.T = T;
  for (;;) {
    // (checkpoints taken during the warm-up lie below `start` <= T and are skipped like any other)
    while (k < ncp && (c & 0xFFFFFu) < r) {
      k++;
      c = k < ncp ? cps[k] : 0xFFFFFFFFu;
    }
    if (k < ncp && (c & 0xFFFFFu) == r) {    // met the speculative trajectory: its tokens from c >> 20 on are true
      tj.m = c >> 20; tj.nf = f;
      tj.land = tj.s_land; tj.stat = tj.s_stat; tj.n = tj.s_n;
      return;
    }
    uint32_t stat = LS_OK;
    if (r >= limit) {
      // ran through the whole sub-sequence without meeting it
    } else if (f == kFixCap) {
      stat = LS_FULL;
    } else {
      uint2    tok;
      uint32_t used, extra;
      const uint32_t s = careful_token(lit, dist, br, endr > r ? endr - r : 0u, tok, used, extra);
      if (s == LS_OK || tok_nlit(tok.x) != 0u) { fix[f++] = tok; r += used; }
      if (s == LS_OK) continue;
      stat = s;
      if (s == LS_EOB) r += extra;
    }
    tj.land = r; tj.stat = stat; tj.n = 0; tj.m = 0; tj.nf = f;
    return;
  }
}

// ---- output side: the stage, its way to dst, Adler-32 ---------------------------------------------------------
// The stage holds output positions [S0, P): what is not in dst yet (F..P, less than 16 bytes after every
// batch) and what went there recently (V..F) -- a match that starts at or after V is copied inside shared
// memory.  (dst + S0) is a 16-byte boundary, so stage -> dst is aligned 16-byte loads and stores.
struct Out4 {
  uint8_t* dst;
  uint32_t cap;
  uint32_t P, F, V, S0;
  uint32_t ta;          // this lane's sum of the bytes it sent to dst since the last fold
  uint64_t tw;          // ... and of (position - abase) * byte
  uint32_t abase;
  uint32_t A, W;        // folded partial sums mod 65521 (the order-free form finish_stream() takes)
};

DEFL_DEV void out_reset_stage(Out4& o, uint32_t pos) {
  o.P = o.F = o.V = pos;
  o.S0 = pos - (uint32_t)((uintptr_t)(o.dst + pos) & 15u);
}
DEFL_DEV void out_begin(Out4& o, uint8_t* dst, uint32_t cap) {
  o.dst = dst; o.cap = cap; o.ta = 0; o.tw = 0; o.abase = 0; o.A = 0; o.W = 0;
  out_reset_stage(o, 0);
}
DEFL_DEV void fold_adler(Out4& o) {
  const uint32_t a = o.ta % kAdlerMod;
  o.A = (o.A + a) % kAdlerMod;
  o.W = (uint32_t)((o.W + (uint64_t)(o.abase % kAdlerMod) * a + o.tw % kAdlerMod) % kAdlerMod);
  o.ta = 0; o.tw = 0;
  o.abase = o.F;
}

// stage -> dst: every whole 16-byte group below `upto`
DEFL_DEV void flush_groups(uint8_t* st, int lane, Out4& o, uint32_t upto) {
  const uint32_t mis = (uint32_t)((uintptr_t)(o.dst + o.F) & 15u);
  if (mis) {                                   // after a forced flush F is not on a 16-byte boundary
    uint32_t nh = 16u - mis;
    if (nh > upto - o.F) nh = upto - o.F;
    if ((uint32_t)lane < nh) {
      const uint32_t b = st[o.F - o.S0 + lane];
      o.dst[o.F + lane] = (uint8_t)b;
      o.ta += b;
      o.tw += (uint64_t)((o.F + lane - o.abase) * b);
    }
    o.F += nh;
  }
  const uint32_t ng = (upto - o.F) >> 4;
  const uint4*   sp = reinterpret_cast<const uint4*>(st + (o.F - o.S0));
  uint4*         dp = reinterpret_cast<uint4*>(o.dst + o.F);
  const uint32_t j0 = o.F - o.abase;
  for (uint32_t g = (uint32_t)lane; g < ng; g += 32u) {
    const uint4 v = sp[g];
    dp[g] = v;
    const uint32_t s0 = simt::dp4a(v.x, 0x01010101u, 0u), s1 = simt::dp4a(v.y, 0x01010101u, 0u);
    const uint32_t s2 = simt::dp4a(v.z, 0x01010101u, 0u), s3 = simt::dp4a(v.w, 0x01010101u, 0u);
    uint32_t q = simt::dp4a(v.x, 0x03020100u, 0u);
    q = simt::dp4a(v.y, 0x03020100u, q);
    q = simt::dp4a(v.z, 0x03020100u, q);
    q = simt::dp4a(v.w, 0x03020100u, q);
    const uint32_t s = s0 + s1 + s2 + s3;
    o.ta += s;
    o.tw += (uint64_t)((j0 + 16u * g) * s + q + 4u * s1 + 8u * s2 + 12u * s3);
  }
  o.F += ng * 16u;
}
// ... and the bytes after the last whole group (before the stage is used for anything else)
DEFL_DEV void flush_all(uint8_t* st, int lane, Out4& o) {
  simt::syncwarp(kFull);
  flush_groups(st, lane, o, o.P);
  const uint32_t nt = o.P - o.F;               // < 16
  if ((uint32_t)lane < nt) {
    const uint32_t b = st[o.F - o.S0 + lane];
    o.dst[o.F + lane] = (uint8_t)b;
    o.ta += b;
    o.tw += (uint64_t)((o.F + lane - o.abase) * b);
  }
  o.F = o.P;
  simt::syncwarp(kFull);
}
// make room for one more batch: move the newest kilobyte to the front of the stage
DEFL_DEV void stage_room(uint8_t* st, int lane, Out4& o) {
  if (o.P - o.S0 + kBatchMax + 4u <= kStageBytes) return;
  const uint32_t have = o.F - o.S0;
  const uint32_t delta = have > kKeepBytes ? (have - kKeepBytes) & ~15u : 0u;
  simt::syncwarp(kFull);
  if (delta < 512u) {                          // cannot happen while F follows P within 16 bytes; be safe
    flush_all(st, lane, o);
    out_reset_stage(o, o.P);
    return;
  }
  const uint32_t n = o.P - o.S0 - delta;       // bytes that stay; delta >= 512: a round never reads what it wrote
  for (uint32_t i = (uint32_t)lane * 16u; i < n; i += 512u) {
    const uint4 v = *reinterpret_cast<const uint4*>(st + delta + i);
    *reinterpret_cast<uint4*>(st + i) = v;
  }
  o.S0 += delta;
  if ((int32_t)(o.V - o.S0) < 0) o.V = o.S0;
  simt::syncwarp(kFull);
}

// ---- resolve: tokens -> bytes, in stream order -------------------------------------------------------------------
// toks[0..count) are consecutive tokens of the stream; what they produce starts at position o.P.
// Checks what needs the absolute position, in the serial decoder's order (infl.c:816-821): a literal that does
// not fit, distance > produced so far, a match that does not fit.  Returns a Reason; everything in front of the
// failing symbol has been produced.
DEFL_DEV uint32_t resolve4(WarpSmem4& sm, int lane, Out4& o, const uint2* toks, uint32_t count) {
  uint8_t* const st = sm.stage;
  uint32_t h = 0;
  uint2    nxt = (uint32_t)lane < count ? simt::ld_tok2(toks + lane) : make_uint2(0u, 0u);
#pragma unroll 1
  while (h < count) {
    uint32_t n = count - h < 32u ? count - h : 32u;
    uint2    tok = nxt;
    nxt = h + 32u + (uint32_t)lane < count ? simt::ld_tok2(toks + h + 32u + lane) : make_uint2(0u, 0u);
    if ((uint32_t)lane >= n) tok = make_uint2(0u, 0u);
    stage_room(st, lane, o);
    if (o.P - o.abase > (1u << 19)) fold_adler(o);

    uint32_t       len = tok_len(tok.x), nlit = tok_nlit(tok.x);
    const uint32_t d1 = tok_d1(tok.x);
    const uint32_t lw = tok.y;                                         // literal j in bits 8j..8j+7
    uint32_t       olen = len + nlit;
    uint32_t       inc = olen;                                         // inclusive prefix sum of output lengths
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = simt::shfl_up(kFull, inc, d);
      if (lane >= d) inc += t;
    }
    const uint32_t ob   = inc - olen;                                  // this token's first byte inside the batch
    const uint32_t room = o.cap - o.P;
    uint32_t bad = R_NONE, lits_fit = nlit;
    if (ob + nlit > room) {
      lits_fit = room > ob ? room - ob : 0u;
      bad = R_OUTPUT_FULL;
    } else if (len) {
      if (d1 >= o.P + ob + nlit) bad = R_DISTANCE;
      else if (len > room - ob - nlit) bad = R_OUTPUT_FULL;
    }
    const bool     nofit = inc > kBatchMax;
    const unsigned stop  = simt::ballot(kFull, bad != R_NONE || nofit || (uint32_t)lane >= n);
    uint32_t       reason = R_NONE;
    uint32_t       keep = 32u;
    if (stop) {
      const int fs = simt::ffs(stop) - 1;
      keep = (uint32_t)fs;
      reason = simt::shfl(kFull, nofit ? (uint32_t)R_NONE : bad, fs);
      if (reason != R_NONE) {
        // the failing token: its literals that fit are still produced, its match is not
        keep = (uint32_t)fs + 1u;
        if (lane == fs) { len = 0u; nlit = lits_fit; olen = nlit; inc = ob + nlit; }
      }
    }
    const bool short_batch = reason == R_NONE && keep != 32u && h + keep < count;
    n = keep;
    if ((uint32_t)lane >= n) { len = 0u; nlit = 0u; olen = 0u; }
    const uint32_t B = n ? simt::shfl(kFull, inc, (int)n - 1) : 0u;

    // ---- the batch's bytes -> stage.  A token's bytes (its literals, then its match) are assembled in
    // registers as the aligned words they occupy in the stage and stored as whole words.  A word that two
    // tokens share is written by the earlier one, which takes the later one's head bytes by shuffle when both
    // are copied in the same round, and from the stage otherwise (then nobody else writes it in this round).
    // Rounds: a token is copied once its source is final (most batches take one round).
    const uint32_t ts = o.P + ob;                                      // where the token's bytes go
    const uint32_t mp = ts + nlit;                                     // ... its match
    const uint32_t sp = mp - d1 - 1u;                                  // where the match comes from
    const uint32_t q  = ts - o.S0;                                     // stage index of the token's first byte
    const uint32_t a  = q & 3u;
    const uint32_t ms = a + nlit;                                      // the match's first byte in the token's word frame
    const uint32_t e_ = (a + olen) & 3u;
    const uint32_t klast = (a + olen + 3u) / 4u - 1u;                  // last word of the token (olen > 0)
    const bool from_stage = sp >= o.V && sp - o.S0 >= 8u;
    const bool from_dst   = sp + len <= o.F && sp >= 16u;
    const bool fast = len == 0u || (len <= 16u && d1 + 1u >= len && (from_stage || from_dst));
    unsigned   pend = simt::ballot(kFull, olen != 0u);
    uint32_t* const stw = reinterpret_cast<uint32_t*>(st);
    simt::syncwarp(kFull);
#pragma unroll 1
    while (pend) {
      const int      first = simt::ffs(pend) - 1;
      const uint32_t done  = simt::shfl(kFull, ts, first);            // everything below is final
      const bool     ready = ((pend >> lane) & 1u) && fast && (len == 0u || !from_stage || sp + len <= done);
      const unsigned rm    = simt::ballot(kFull, ready);
      if (!((rm >> first) & 1u)) {
        // the first pending token is overlapping, long, straddles the stage's lower end or reads its own
        // literals: the whole warp copies it, one byte per lane and step
        const uint32_t tx   = simt::shfl(kFull, tok.x, first);
        const uint32_t lw0  = simt::shfl(kFull, lw, first);
        const uint32_t nl0  = simt::shfl(kFull, nlit, first);
        const uint32_t ts0  = done;
        const uint32_t len0 = simt::shfl(kFull, len, first), dist0 = tok_d1(tx) + 1u;
        if ((uint32_t)lane < nl0) st[ts0 - o.S0 + lane] = (uint8_t)(lw0 >> (8u * lane));
        simt::syncwarp(kFull);
        const uint32_t mp0 = ts0 + nl0;
        for (uint32_t j = (uint32_t)lane; j < len0; j += 32u) {
          const uint32_t s = mp0 - dist0 + (dist0 < len0 ? j % dist0 : j);
          const uint32_t b = s >= o.V ? (uint32_t)st[s - o.S0] : simt::ld_out8(o.dst + s);
          st[mp0 - o.S0 + j] = (uint8_t)b;
        }
        pend &= pend - 1u;
        simt::syncwarp(kFull);
        continue;
      }
      uint32_t D[6];
#pragma unroll
      for (int k = 0; k < 6; k++) D[k] = 0u;
      if (ready) {
        const uint64_t L64 = (uint64_t)lw << (8u * a);
        D[0] = (uint32_t)L64; D[1] = (uint32_t)(L64 >> 32);
        if (len) {
          uint32_t w[7];
          uint32_t sh, wl;
          if (from_stage) {
            const uint32_t idx = sp - o.S0 - ms;                       // >= 1
            const uint32_t* wp = stw + (idx >> 2);
            sh = (idx & 3u) * 8u;
            wl = ((idx & 3u) + ms + len - 1u) >> 2;                    // the word that holds the last source byte
#pragma unroll
            for (uint32_t k = 0; k < 7; k++) w[k] = k <= wl ? wp[k] : 0u;
          } else {
            const uintptr_t ad = (uintptr_t)(o.dst + sp) - ms;
            const uint32_t* wp = reinterpret_cast<const uint32_t*>(ad & ~(uintptr_t)3);
            sh = (uint32_t)(ad & 3u) * 8u;
            wl = ((uint32_t)(ad & 3u) + ms + len - 1u) >> 2;
#pragma unroll
            for (uint32_t k = 0; k < 7; k++) w[k] = k <= wl ? simt::ld_outw(wp + k) : 0u;
          }
          const uint64_t M64 = ~0ull << (8u * ms);                     // bytes of words 0, 1 that belong to the match
          const uint32_t m0 = (uint32_t)M64, m1 = (uint32_t)(M64 >> 32);
          D[0] = (simt::funnel_r(w[0], w[1], sh) & m0) | (D[0] & ~m0);
          D[1] = (simt::funnel_r(w[1], w[2], sh) & m1) | (D[1] & ~m1);
#pragma unroll
          for (int k = 2; k < 6; k++) D[k] = simt::funnel_r(w[k], w[k + 1], sh);
        }
      }
      // the later neighbour's head bytes (its word 0 is this token's last word when e_ != 0)
      const uint32_t nextD0 = simt::shfl(kFull, D[0], (lane + 1) & 31);
      if (ready) {
        uint32_t* const qw = stw + (q >> 2);
        const bool prev_act = lane > 0 && ((rm >> (lane - 1)) & 1u);
        const bool next_act = lane < 31 && ((rm >> (lane + 1)) & 1u);
        uint32_t k0 = 0u;
        if (a != 0u) {
          if (prev_act) k0 = 1u;                                       // the earlier token writes that word
          else { const uint32_t lo = (1u << (8u * a)) - 1u; D[0] = (D[0] & ~lo) | (qw[0] & lo); }
        }
        uint32_t tail = 0u;
        const uint32_t keepm = e_ ? (1u << (8u * e_)) - 1u : 0xFFFFFFFFu;   // this token's bytes in its last word
        if (e_ != 0u) tail = next_act ? nextD0 : qw[klast];
#pragma unroll
        for (uint32_t k = 0; k < 6; k++) {
          if (k >= k0 && k <= klast) {
            uint32_t v = D[k];
            if (k == klast) v = (v & keepm) | (tail & ~keepm);
            qw[k] = v;
          }
        }
      }
      pend &= ~rm;
      simt::syncwarp(kFull);
    }
    // ---- the stage's whole 16-byte groups go to dst
    o.P += B;
    if (o.P - o.F >= 512u) flush_groups(st, lane, o, o.P);           // a full trip of 16-byte stores for every lane
    h += n;
    if (reason != R_NONE) return reason;
    if (short_batch) nxt = h + (uint32_t)lane < count ? simt::ld_tok2(toks + h + lane) : make_uint2(0u, 0u);
  }
  return R_NONE;
}


// ---- one window of a Huffman block -------------------------------------------------------------------------
struct WarpScratch4 {
  uint2*    spec;   // [32][kSpecCap]
  uint2*    fix;    // [32][kFixCap]
  uint32_t* cps;    // [32][kCpPerLane]
};
DEFL_DEV WarpScratch4 scratch4_of(void* base) {
  WarpScratch4 s;
  s.spec = static_cast<uint2*>(base);
  s.fix  = s.spec + 32u * kSpecCap;
  s.cps  = reinterpret_cast<uint32_t*>(s.fix + 32u * kFixCap);
  return s;
}

// Decode from A-space bit `cur` (a token boundary inside a Huffman block whose tables are in sm) as far as one
// window reaches.  On return `cur` is where the stream goes on; the result is LS_OK (window done, block goes
// on), LS_EOB (block ended at cur), or LS_FAIL | reason << 8.
DEFL_DEV uint32_t decode_window4(WarpSmem4& sm, const WarpScratch4& sc, int lane, const Input4& in,
                                 uint64_t end_bit, uint64_t& cur, Out4& o, uint32_t warm_bits) {
  const uint32_t base_quad = (uint32_t)(cur >> 10) * 8u;               // the window's base: a multiple of 1024 bits
  const uint32_t rel0   = (uint32_t)(cur & 1023u);
  const uint64_t left64 = end_bit - (cur - rel0);
  const uint32_t endr   = left64 > 0x40000000ull ? 0x40000000u : (uint32_t)left64;   // window-relative end of input
  const uint32_t bw     = base_quad * 4u - in.skew;                     // index in in.words of the window's word 0
  uint32_t S = (endr - rel0 + 31u) / 32u;
  S = (S + kCpBits - 1u) / kCpBits * kCpBits;
  S = S < kSubMin ? kSubMin : (S > kSubMax ? kSubMax : S);

  const uint32_t start  = rel0 + (uint32_t)lane * S;
  const uint32_t limit  = start + S;
  const bool     active = start < endr || lane == 0;
  uint2*    seg = sc.spec + (uint32_t)lane * kSpecCap;
  uint2*    fix = sc.fix + (uint32_t)lane * kFixCap;
  uint32_t* cps = sc.cps + (uint32_t)lane * kCpMax;

  LaneTraj tj;
  tj.land = start; tj.stat = LS_NONE; tj.n = 0; tj.m = 0; tj.nf = 0; tj.T = start; tj.ncp = 0;
  tj.s_land = start; tj.s_stat = LS_NONE; tj.s_n = 0;
  const uint32_t warm = (lane >= 1 && active) ? (warm_bits < S ? warm_bits : S) : 0u;   // lane 0 starts on a known boundary
  if (lane == 0) sm.stop_lane = 32u;
  simt::syncwarp(kFull);                                                // the stage is the lanes' input rings now
  if (active) lane_spec4(sm.lit, sm.dist, in, base_quad, bw, start, limit, endr, seg, cps, tj, warm,
                         sm.ring + lane * kRingWords, warm_bits == kWarmDynamic ? &sm.stop_lane : nullptr, (uint32_t)lane);
  simt::syncwarp(kFull);

  // fix-up rounds until every lane starts where its predecessor ended
#pragma unroll 1
  for (int round = 0; round < 32; round++) {
    const uint32_t pland = simt::shfl_up(kFull, tj.land, 1);
    const uint32_t pstat = simt::shfl_up(kFull, tj.stat, 1);
    const bool need = lane >= 1 && active && pstat == LS_OK && pland != tj.T;
    if (!simt::any(kFull, need)) break;
    if (need) lane_redo4(sm.lit, sm.dist, in, bw, start, limit, endr, fix, cps, pland, tj);
    simt::syncwarp(kFull);
  }

  // a lane's fix-up tokens move in front of its true speculative tokens when the discarded ones left room
  if (active && tj.nf != 0u && tj.nf <= tj.m && tj.n > tj.m) {
    uint2* to = seg + (tj.m - tj.nf);
    for (uint32_t i = 0; i < tj.nf; i++) to[i] = simt::ld_tok2(fix + i);
    tj.m -= tj.nf;
    tj.nf = 0;
  }

  // the first lane whose true trajectory stopped ends the window; without one, the last active lane does
  const unsigned act  = simt::ballot(kFull, active);
  const unsigned stop = simt::ballot(kFull, active && tj.stat != LS_OK);
  const int      last = stop ? simt::ffs(stop) - 1 : 31 - simt::clz(act);
  simt::syncwarp(kFull);

  out_reset_stage(o, o.P);                                              // the rings overwrote the stage
  uint32_t reason = R_NONE;
#pragma unroll 1
  for (int L = 0; L <= last && reason == R_NONE; L++) {
    const uint32_t nf = simt::shfl(kFull, tj.nf, L);
    const uint32_t m  = simt::shfl(kFull, tj.m, L);
    const uint32_t n  = simt::shfl(kFull, tj.n, L);
    if (nf) reason = resolve4(sm, lane, o, sc.fix + (uint32_t)L * kFixCap, nf);
    if (reason == R_NONE && n > m) reason = resolve4(sm, lane, o, sc.spec + (uint32_t)L * kSpecCap + m, n - m);
  }
  flush_all(sm.stage, lane, o);
  if (reason) return fail_stat(reason);

  const uint32_t land = simt::shfl(kFull, tj.land, last);
  const uint32_t stat = simt::shfl(kFull, tj.stat, last);
  cur = (cur - rel0) + land;
  if (stat == LS_FULL) return LS_OK;        // a token segment filled up: the next window starts there
  return stat;
}

// ---- one stream, start to finish, by one warp ------------------------------------------------------------------
DEFL_DEV void inflate_stream_warp4(WarpSmem4& sm, const WarpScratch4& sc, int lane, const uint8_t* src,
                                   uint32_t srclen, uint8_t* dst, uint32_t cap, int flags, StreamOut& out) {
  Lane<32> L;
  L.gmask = kFull; L.lane = lane; L.wlane = lane;
  out.code = -1; out.adler = 1; out.outlen = 0; out.inpos = 0; out.adler_stored = 0; out.info = 0;
  if (srclen == 0) { out.code = 0; return; }       // all pieces empty => UNZ_OK (infl.c:1547-1550)

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
  Input4 in;
  in.words  = br.w;
  in.wend   = br.wend;
  in.skew   = (uint32_t)(((uintptr_t)br.w >> 2) & 3u);
  in.q0     = reinterpret_cast<const uint8_t*>((uintptr_t)br.w & ~(uintptr_t)15);
  in.abytes = in.skew * 4u + mis + srclen;
  const uint64_t abias = (uint64_t)in.skew * 32u;   // A-space = bit position from br.w + this

  Out4 o;
  out_begin(o, dst, cap);
  uint32_t reason = R_NONE;
  bool     fixed_ready = false;
#pragma unroll 1
  for (;;) {
    if ((int64_t)(end_bit - br.bitpos()) < 3) { reason = R_TRUNCATED; break; }
    const uint32_t hdr    = br.peek();
    const uint32_t bfinal = hdr & 1u, btype = (hdr >> 1) & 3u;
    br.skip(3); br.norm();
    if (btype == 0) {
      // stored block: input -> dst directly; Adler-32 over what it wrote (still in L2)
      uint32_t pos = o.P;
      reason = stored_block<32>(L, br, src, srclen, base_bit, dst, pos, cap);
      if (reason) break;
      adler_range<32>(L, dst, o.P, pos, o.A, o.W);
      out_reset_stage(o, pos);
    } else if (btype == 3) {
      reason = R_BLOCK_TYPE;
      break;
    } else {
      simt::syncwarp(kFull);
      reason = huffman_block_setup<32, kEmptyLit>(L, sm.lit, sm.dist, sm.s, br, end_bit, btype, fixed_ready);
      if (reason) break;
      simt::syncwarp(kFull);
      uint64_t cur = br.bitpos() + abias;
      uint32_t st;
#pragma unroll 1
      do {
        st = decode_window4(sm, sc, lane, in, end_bit + abias, cur, o, btype == 1u ? kWarmFixed : kWarmDynamic);
      } while (st == LS_OK);
      if (st != LS_EOB) { reason = st >> 8; break; }
      br.seek(cur - abias);
    }
    if (bfinal) break;
  }
  if (reason) { out.info = reason << 8; out.outlen = o.P; return; }
  fold_adler(o);
  finish_stream<32>(L, src, srclen, o.P, br.bitpos() - base_bit, o.A, o.W, zlib, flags, out);
}

}  // namespace v4
}  // namespace defl
