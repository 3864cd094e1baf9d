// pipeline.cuh -- the lane-decoder pipeline (batch inflate v2).
//
// Why: profiles/r01_v1_ncu_full_summary.json -- with one warp per stream every symbol costs ~54 warp
// issue slots (all 32 lanes run the bit reader redundantly) and the SM is issue-bound at ~40 GB/s.
// Here the serial part of DEFLATE (bit reader + Huffman lookups, reference infl_ft_block
// infl.c:717-901) runs as ONE THREAD per stream, 32 streams per warp instruction, and emits
// 32-bit tokens into a per-stream shared-memory ring.  The parallel part (LZ77 copies, reference
// infl.c:269-376,822-896; stored blocks; table construction; Adler-32) is done by worker warps:
// a worker pops up to 32 tokens of one stream, prefix-sums their output lengths and resolves the
// produced bytes one byte per lane (coalesced stores), falling back to in-order handling only for
// matches that read bytes produced by the same batch.
//
// Roles inside a CTA:   D decoder warps (SL active lanes each = SL stream slots)
//                       K worker warps  (slot s is served by worker s % K)
// Protocol per slot (shared memory, volatile words + block fences):
//   lane  -> worker : tokens q[tail..), then `req` (START / NEXT block header / FINISH)
//   worker-> lane   : `head` (tokens consumed), `ack` (RUN at bit position / DONE / FAIL)
// A request is served only when the slot's ring is empty, so everything a worker does for a
// stream (copies, stored blocks, result record) happens in stream order.
#pragma once
#include "inflate_core.cuh"
#include "../../include/defl/infl_batch.h"

namespace defl {

constexpr int      kQCap     = 64;           // tokens per ring
constexpr int      kSteps    = 8;            // decode steps between two publications of `tail`
constexpr int      kStage    = 512;          // output bytes resolved per batch
constexpr uint32_t kTokMatch = 0x80000000u;  // [8:0] length, [23:9] distance - 1
constexpr uint32_t kTokDep   = 0x40000000u;  // worker-side mark: match reads bytes of its own batch
                                             // literal token: [26:24] count (1..3), [23:0] bytes
#ifndef DEFL_WATCHDOG_SPINS
#define DEFL_WATCHDOG_SPINS (1u << 24)
#endif

enum Req : uint32_t { REQ_NONE = 0, REQ_START = 1, REQ_NEXT = 2, REQ_FINISH = 3 };
enum Ack : uint32_t { ACK_NONE = 0, ACK_RUN = 1, ACK_DONE = 2 };

struct Slot {
  uint16_t lit[kLitTab];
  uint16_t dist[kDistTab];
  uint32_t q[kQCap];
  // decoder lane -> worker
  uint32_t tail;
  uint32_t req;
  uint32_t req_bit_lo, req_bit_hi;   // BitReader coordinates
  uint32_t req_reason;               // REQ_FINISH: 0 = final block decoded, else a Reason
  uint32_t stream;                   // index into items[]
  // worker -> decoder lane
  uint32_t head;
  uint32_t ack;
  uint32_t ack_bit_lo, ack_bit_hi;
  uint32_t ack_pos;
  uint32_t ack_bfinal;
  // worker-private
  uint32_t wpos;
  uint32_t fixed_ready;
  uint32_t dst_lo, dst_hi;
};
static_assert(sizeof(Slot) % 16 == 0, "Slot must keep 16-byte alignment in an array");

// a worker either resolves a token batch or serves a request, never both at once
struct WorkerSmem {
  union {
    BuildScratch s;
    struct {
      uint8_t stage[kStage];
      uint8_t own[kStage];
    };
  };
};
static_assert(sizeof(WorkerSmem) % 16 == 0, "WorkerSmem alignment");

DEFL_DEV uint32_t uniform_vload(const uint32_t* p, int lane) {
  uint32_t v = 0;
  if (lane == 0) v = simt::vload(p);
  return simt::shfl(kFull, v, 0);
}

struct StreamView {
  const uint8_t* src;
  uint8_t*       dst;
  uint32_t       srclen, cap;
  uint64_t       base_bit, end_bit;
  bool           oversize;
};

DEFL_DEV StreamView view_of(const infl_batch_item_t& it) {
  StreamView v;
  v.src = static_cast<const uint8_t*>(it.src);
  v.dst = static_cast<uint8_t*>(it.dst);
  v.oversize = it.srclen > 0xFFFFFFFFull || it.dstlen > 0xFFFFFFFFull;
  v.srclen = (uint32_t)it.srclen;
  v.cap = (uint32_t)it.dstlen;
  const uint32_t mis = (uint32_t)((uintptr_t)v.src & 3u);
  v.base_bit = (uint64_t)mis * 8u;
  v.end_bit = v.base_bit + (uint64_t)v.srclen * 8u;
  return v;
}

DEFL_DEV void reader_for(BitReader& br, const StreamView& v) {
  const uint32_t mis = (uint32_t)((uintptr_t)v.src & 3u);
  br.w = reinterpret_cast<const uint32_t*>((uintptr_t)v.src - mis);
  br.wend = (uint32_t)(((uint64_t)mis + v.srclen + 3u) >> 2);
}

// =====================================================================================================
//  decoder warp: one thread per stream
// =====================================================================================================
enum LaneMode : int { M_IDLE = 0, M_WAIT = 1, M_RUN = 2, M_EXIT = 3 };

// Bit reader of a decoder lane: a 128-bit window in four registers.  A decode step consumes at
// most 48 bits (15 + 5 length, 15 + 13 distance) and starts with bo < 32, so both table lookups
// of a step come out of the window with funnel shifts and the refill happens once, at the end;
// the word loaded there is first needed >= 64 bits later, which hides the L1/L2 latency.
struct LaneReader {
  const uint32_t* w;
  uint32_t        wend, widx;
  uint32_t        w0, w1, w2, w3;
  uint32_t        bo;

  DEFL_DEV uint32_t ld(uint32_t i) const { return i < wend ? simt::ld_in32(w + i) : 0u; }
  DEFL_DEV void seek(uint64_t bitpos) {
    widx = (uint32_t)(bitpos >> 5);
    bo   = (uint32_t)(bitpos & 31u);
    w0 = ld(widx); w1 = ld(widx + 1); w2 = ld(widx + 2); w3 = ld(widx + 3);
  }
  DEFL_DEV void advance() {
    w0 = w1; w1 = w2; w2 = w3; widx++; bo -= 32u;
    w3 = ld(widx + 3);
  }
  DEFL_DEV uint64_t bitpos() const { return ((uint64_t)widx << 5) + bo; }
};

struct LaneState {
  LaneReader br;
  uint64_t   end_bit;
  uint32_t   pos, cap, tail, bfinal;
  int        mode;
};

// tokens first, then the request word
DEFL_DEV void lane_post(Slot& sl, LaneState& st, uint32_t req, uint32_t reason) {
  const uint64_t bp = st.br.bitpos();
  simt::fence_block();
  simt::vstore(&sl.tail, st.tail);
  sl.req_bit_lo = (uint32_t)bp;
  sl.req_bit_hi = (uint32_t)(bp >> 32);
  sl.req_reason = reason;
  simt::fence_block();
  simt::vstore(&sl.req, req);
  st.mode = M_WAIT;
}

DEFL_DEV void lane_fail(Slot& sl, LaneState& st, uint32_t reason) {
  lane_post(sl, st, REQ_FINISH, settle_reason(reason, (int64_t)(st.end_bit - st.br.bitpos())));
}

// One decode step of one stream: a match, or up to two literals, or the end of the block.
// Literal and match lanes run the same instruction sequence (one literal/length lookup, one
// second lookup in the distance or the literal table, one token store, one refill); only
// sub-table links, the end of a block and failures branch away.
DEFL_DEV void lane_step(Slot& sl, LaneState& st) {
  LaneReader&    br = st.br;
  const uint32_t v0 = simt::funnel_r(br.w0, br.w1, br.bo);
  const uint32_t v1 = simt::funnel_r(br.w1, br.w2, br.bo);
  uint32_t       e  = sl.lit[v0 & (kLitRoot - 1)];
  if ((e & (kLitIsLen | kLitKind)) == kLitSub)
    e = sl.lit[kLitRoot + (e & 0x1FFu) + ((v0 >> kLitRootBits) & ((1u << ((e >> 11) & 15u)) - 1u))];
  const uint32_t cl = (e >> 11) & 15u;
  if (e == 0u || (e & (kLitIsLen | kLitKind)) == kLitEob) {
    if (e == 0u) { lane_fail(sl, st, R_SYMBOL); return; }
    br.bo += cl;                                         // end of block
    if (br.bo >= 32u) br.advance();
    if (br.bitpos() > st.end_bit) { lane_post(sl, st, REQ_FINISH, R_TRUNCATED); return; }
    lane_post(sl, st, st.bfinal ? REQ_FINISH : REQ_NEXT, R_NONE);
    return;
  }
  const bool     is_len = (e & kLitIsLen) != 0u;
  const uint32_t xb   = (e >> 8) & 7u;                   // 0 for a literal
  const uint32_t val  = (e & 255u) + simt::bfe(v0, cl, xb);   // byte, or length - 3
  const uint32_t use1 = cl + xb;                         // <= 20
  const uint32_t v2   = simt::funnel_r(v0, v1, use1);    // the 32 bits after this symbol
  const uint16_t* tab2 = is_len ? sl.dist : sl.lit;
  uint32_t        x    = tab2[v2 & (is_len ? (uint32_t)(kDistRoot - 1) : (uint32_t)(kLitRoot - 1))];
  uint32_t        tok, use2, adv;
  if (is_len) {
    if (x & kDistSub)
      x = sl.dist[kDistRoot + (x & 0x1FFu) + ((v2 >> kDistRootBits) & ((1u << ((x >> 11) & 15u)) - 1u))];
    const uint32_t dcl = (x >> 11) & 15u, dxb = (x >> 7) & 15u;
    const uint32_t dist = ((x & 3u) << dxb) + 1u + simt::bfe(v2, dcl, dxb);
    const uint32_t len = val + 3u;
    use2 = dcl + dxb;
    if (x == 0u || dist > st.pos || len > st.cap - st.pos) {
      // the failing symbol is not consumed beyond its length code, as in the in-order decoder
      br.bo += use1;
      if (x != 0u) br.bo += use2;
      while (br.bo >= 32u) br.advance();
      lane_fail(sl, st, x == 0u ? R_DISTANCE : (dist > st.pos ? R_DISTANCE : R_OUTPUT_FULL));
      return;
    }
    tok = kTokMatch | ((dist - 1u) << 9) | len;
    adv = len;
  } else {
    if (st.pos >= st.cap) { lane_fail(sl, st, R_OUTPUT_FULL); return; }
    const bool two = (x & (kLitIsLen | kLitKind)) == 0u && x != 0u && st.pos + 1u < st.cap;
    use2 = two ? (x >> 11) & 15u : 0u;
    tok  = two ? ((2u << 24) | ((x & 255u) << 8) | val) : ((1u << 24) | val);
    adv  = two ? 2u : 1u;
  }
  sl.q[st.tail & (kQCap - 1)] = tok;
  st.tail++;
  st.pos += adv;
  br.bo += use1 + use2;
  if (br.bo >= 32u) br.advance();
  if (br.bo >= 32u) br.advance();
}

// `slots`: this warp's SL slots.  Lanes >= SL only take part in the exit vote.
DEFL_DEV void decoder_warp(Slot* slots, int SL, int lane, const infl_batch_item_t* items,
                           const uint32_t* order, uint32_t n, unsigned int* ticket) {
  Slot&     sl = slots[lane < SL ? lane : 0];
  LaneState st;
  st.mode = lane < SL ? M_IDLE : M_EXIT;
  st.tail = 0; st.pos = 0; st.cap = 0; st.bfinal = 0; st.end_bit = 0;
  st.br.w = nullptr; st.br.wend = 0; st.br.widx = 0; st.br.w0 = st.br.w1 = st.br.w2 = st.br.w3 = 0; st.br.bo = 0;
  uint32_t idle = 0;

  for (;;) {
    if (st.mode == M_IDLE) {
      const uint32_t k = simt::atomic_add_global(ticket, 1u);
      if (k >= n) {
        st.mode = M_EXIT;
      } else {
        const uint32_t   idx = order ? order[k] : k;
        const StreamView sv  = view_of(items[idx]);
        const uint32_t   mis = (uint32_t)((uintptr_t)sv.src & 3u);
        st.br.w = reinterpret_cast<const uint32_t*>((uintptr_t)sv.src - mis);
        st.br.wend = (uint32_t)(((uint64_t)mis + sv.srclen + 3u) >> 2);
        st.br.widx = 0; st.br.bo = 0;
        st.end_bit = sv.end_bit;
        st.cap = sv.cap;
        sl.stream = idx;
        lane_post(sl, st, REQ_START, R_NONE);
      }
    } else if (st.mode == M_WAIT) {
      const uint32_t a = simt::vload(&sl.ack);
      if (a) {
        simt::fence_block();
        const uint64_t bit = ((uint64_t)sl.ack_bit_hi << 32) | sl.ack_bit_lo;
        st.pos = sl.ack_pos;
        st.bfinal = sl.ack_bfinal;
        simt::vstore(&sl.ack, ACK_NONE);
        if (a == ACK_RUN) { st.br.seek(bit); st.mode = M_RUN; }
        else st.mode = M_IDLE;                       // the worker wrote the stream's result record
      }
    }
    bool ran = false;
    if (st.mode == M_RUN) {
      const uint32_t head = simt::vload(&sl.head);
      if ((uint32_t)kQCap - (st.tail - head) >= (uint32_t)kSteps) {
        if (st.br.widx > st.br.wend) {               // far past the input: stop early
          lane_fail(sl, st, R_TRUNCATED);
        } else {
          const uint32_t t0 = st.tail;
#pragma unroll 1
          for (int it = 0; it < kSteps && st.mode == M_RUN; it++) lane_step(sl, st);
          if (st.mode == M_RUN && st.tail != t0) {
            simt::fence_block();
            simt::vstore(&sl.tail, st.tail);
          }
        }
        ran = true;
      }
    }
    if (simt::all(kFull, st.mode == M_EXIT)) break;
    if (simt::ballot(kFull, ran) == 0u) {
      simt::backoff(64);
      if (++idle > DEFL_WATCHDOG_SPINS) break;          // watchdog: never spin forever
    } else {
      idle = 0;
    }
  }
}

// =====================================================================================================
//  worker warps: byte-parallel token resolution, block headers, stored blocks, results
// =====================================================================================================
DEFL_DEV uint8_t* slot_dst(const Slot& sl) {
  return reinterpret_cast<uint8_t*>(((uint64_t)sl.dst_hi << 32) | sl.dst_lo);
}

// Resolve up to 32 tokens of one stream into dst[wpos..).  Returns the number of tokens consumed.
DEFL_DEV uint32_t process_batch(Slot& sl, WorkerSmem& ws, int lane, uint32_t head, uint32_t avail) {
  uint32_t n = avail < 32u ? avail : 32u;
  uint32_t tok = (uint32_t)lane < n ? sl.q[(head + lane) & (kQCap - 1)] : 0u;
  uint32_t olen = (tok >> 31) ? (tok & 511u) : ((tok >> 24) & 7u);

  uint32_t inc = olen;                                   // inclusive prefix sum of output lengths
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t t = simt::shfl_up(kFull, inc, d);
    if (lane >= d) inc += t;
  }
  const unsigned fit = simt::ballot(kFull, inc <= (uint32_t)kStage);   // monotone: a prefix of ones
  const uint32_t nfit = fit == kFull ? 32u : (uint32_t)simt::ffs(~fit) - 1u;
  if (nfit < n) n = nfit;                                // n >= 1: one token is at most 258 bytes
  if ((uint32_t)lane >= n) { tok = 0u; olen = 0u; }
  const uint32_t B = simt::shfl(kFull, inc, (int)n - 1);
  const uint32_t o = inc - olen;                         // exclusive prefix: this token's first byte
  const uint32_t P = sl.wpos;
  uint8_t*       dst = slot_dst(sl) + P;

  const bool dep = (tok >> 31) && (((tok >> 9) & 0x7FFFu) + 1u < o + olen);
  if (dep) tok |= kTokDep;

  // Byte-parallel resolution, 32 bytes per round.  The token that owns byte j is found without
  // shared memory: tokens sit in lane order, so it is (tokens starting before the round) +
  // (token starts at or before j inside the round) - 1; the start bits of a round come from one
  // warp-wide OR.  Four rounds are addressed and loaded before any of them is stored, so a batch
  // pays the L2 latency of its match sources once.
  const bool     valid = olen != 0u;
  const uint32_t ks = o >> 5, kb = 1u << (o & 31u);
  const unsigned le = simt::lanemask_le();
  simt::syncwarp(kFull);
#pragma unroll 1
  for (uint32_t base = 0; base < B; base += 128u) {
    uint32_t val[4];
    bool     have[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const uint32_t rb = base + 32u * (uint32_t)u;
      have[u] = false;
      val[u] = 0;
      if (rb < B) {                                      // uniform
        const uint32_t k = rb >> 5;
        const uint32_t before = (uint32_t)simt::popc(simt::ballot(kFull, valid && ks < k));
        const uint32_t starts = simt::reduce_or(kFull, (valid && ks == k) ? kb : 0u);
        const int      i  = (int)((before + (uint32_t)simt::popc(starts & le) - 1u) & 31u);
        const uint32_t t  = simt::shfl(kFull, tok, i);
        const uint32_t oi = simt::shfl(kFull, o, i);
        const uint32_t j  = rb + (uint32_t)lane;
        if (j < B) {
          if (!(t >> 31)) {
            val[u] = (t >> (8u * (j - oi))) & 255u;
            have[u] = true;
          } else if (!(t & kTokDep)) {
            val[u] = simt::ld_out8(dst + j - (((t >> 9) & 0x7FFFu) + 1u));
            have[u] = true;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const uint32_t j = base + 32u * (uint32_t)u + (uint32_t)lane;
      if (have[u]) {
        ws.stage[j] = (uint8_t)val[u];
        dst[j] = (uint8_t)val[u];
      }
    }
  }

  // matches that read bytes of this very batch: in token order, each by the whole warp
  unsigned depmask = simt::ballot(kFull, dep);
#pragma unroll 1
  while (depmask) {
    const int i = simt::ffs(depmask) - 1;
    depmask &= depmask - 1u;
    const uint32_t t  = simt::shfl(kFull, tok, i);
    const uint32_t oi = simt::shfl(kFull, o, i);
    const uint32_t len = t & 511u, dist = ((t >> 9) & 0x7FFFu) + 1u;
    simt::syncwarp(kFull);
    for (uint32_t r = (uint32_t)lane; r < len; r += 32u) {
      const uint32_t sr = dist < len ? r % dist : r;     // a period shorter than the match repeats
      const int32_t  so = (int32_t)oi - (int32_t)dist + (int32_t)sr;
      const uint8_t  val = so < 0 ? (uint8_t)simt::ld_out8(dst + so) : ws.stage[so];
      ws.stage[oi + r] = val;
      dst[oi + r] = val;
    }
  }
  simt::syncwarp(kFull);
  if (lane == 0) sl.wpos = P + B;
  return n;
}

DEFL_DEV void write_result(infl_batch_result_t* results, uint32_t idx, const StreamOut& o, int lane) {
  if (lane == 0) {
    infl_batch_result_t r;
    r.code = o.code; r.adler = o.adler; r.outlen = o.outlen; r.inpos = o.inpos;
    r.adler_stored = o.adler_stored; r.info = o.info;
    results[idx] = r;
  }
}

DEFL_DEV void worker_ack(Slot& sl, int lane, uint32_t ack, uint64_t bit, uint32_t pos, uint32_t bfinal) {
  simt::syncwarp(kFull);
  if (lane == 0) {
    sl.ack_bit_lo = (uint32_t)bit;
    sl.ack_bit_hi = (uint32_t)(bit >> 32);
    sl.ack_pos = pos;
    sl.ack_bfinal = bfinal;
    sl.wpos = pos;
    simt::vstore(&sl.req, REQ_NONE);
    simt::fence_block();
    simt::vstore(&sl.ack, ack);
  }
  simt::syncwarp(kFull);
}

// Serve one request of a slot whose ring is empty.  All 32 lanes walk the header bits together
// (the cooperative table construction of inflate_core.cuh with G = 32).
DEFL_DEV void serve_request(Slot& sl, WorkerSmem& ws, int lane, uint32_t req, const infl_batch_item_t* items,
                            int flags, infl_batch_result_t* results) {
  Lane<32> L;
  L.gmask = kFull; L.lane = lane; L.wlane = lane;
  simt::fence_block();
  const uint32_t   idx = sl.stream;
  const StreamView sv  = view_of(items[idx]);
  const bool       zlib = (flags & 0xFF) == 1;
  const uint64_t   req_bit = ((uint64_t)sl.req_bit_hi << 32) | sl.req_bit_lo;
  const uint32_t   req_reason = sl.req_reason;
  uint32_t         pos = sl.wpos;
  bool             fixed_ready = sl.fixed_ready != 0u;
  BitReader        br;
  reader_for(br, sv);

  StreamOut out;
  out.code = -1; out.adler = 1; out.outlen = 0; out.inpos = 0; out.adler_stored = 0; out.info = 0;
  uint32_t reason = R_NONE;
  bool     finished = false;        // final block decoded

  if (req == REQ_START) {
    pos = 0; fixed_ready = false;
    simt::syncwarp(kFull);
    if (lane == 0) {
      sl.dst_lo = (uint32_t)(uintptr_t)sv.dst;
      sl.dst_hi = (uint32_t)((uint64_t)(uintptr_t)sv.dst >> 32);
    }
    if (sv.oversize) {
      out.info = R_TRUNCATED << 8;
      write_result(results, idx, out, lane);
      worker_ack(sl, lane, ACK_DONE, 0, 0, 0);
      return;
    }
    if (sv.srclen == 0) {           // all pieces empty => UNZ_OK, nothing produced (infl.c:1547-1550)
      out.code = 0;
      write_result(results, idx, out, lane);
      worker_ack(sl, lane, ACK_DONE, 0, 0, 0);
      return;
    }
    uint32_t body = 0;
    if (zlib) reason = zlib_header(sv.src, sv.srclen, body);
    if (reason) {
      out.info = reason << 8;
      write_result(results, idx, out, lane);
      worker_ack(sl, lane, ACK_DONE, 0, 0, 0);
      return;
    }
    br.seek(sv.base_bit + body * 8u);
  } else {
    br.seek(req_bit);
    if (req == REQ_FINISH) {
      reason = req_reason;
      finished = reason == R_NONE;
    }
  }

  if (req != REQ_FINISH) {
#pragma unroll 1
    for (;;) {
      if ((int64_t)(sv.end_bit - br.bitpos()) < 3) { reason = R_TRUNCATED; break; }
      const uint32_t hdr    = br.peek();
      const uint32_t bfinal = hdr & 1u, btype = (hdr >> 1) & 3u;
      br.skip(3); br.norm();
      if (btype == 0) {
        reason = stored_block<32>(L, br, sv.src, sv.srclen, sv.base_bit, sv.dst, pos, sv.cap);
        if (reason) break;
        if (bfinal) { finished = true; break; }
      } else if (btype == 3) {
        reason = R_BLOCK_TYPE;
        break;
      } else {
        reason = huffman_block_setup<32>(L, sl.lit, sl.dist, ws.s, br, sv.end_bit, btype, fixed_ready);
        if (reason) break;
        simt::syncwarp(kFull);
        if (lane == 0) sl.fixed_ready = fixed_ready ? 1u : 0u;
        worker_ack(sl, lane, ACK_RUN, br.bitpos(), pos, bfinal);
        return;
      }
    }
  }

  if (finished) {
    uint32_t A = 0, W = 0;
    simt::syncwarp(kFull);
    adler_range<32>(L, sv.dst, 0, pos, A, W);
    finish_stream<32>(L, sv.src, sv.srclen, pos, br.bitpos() - sv.base_bit, A, W, zlib, flags, out);
  } else {
    out.info = reason << 8;
    out.outlen = pos;
  }
  write_result(results, idx, out, lane);
  worker_ack(sl, lane, ACK_DONE, 0, pos, 0);
}

// worker `w` of K serves slots w, w+K, ... < S.  One poll looks at all of them at once: lane i
// reads the control words of the i-th owned slot, a ballot finds the slots with work, and the
// pick rotates so no slot starves.
DEFL_DEV void worker_warp(Slot* slots, int S, int w, int K, WorkerSmem& ws, int lane, const uint32_t* live,
                          const infl_batch_item_t* items, int flags, infl_batch_result_t* results) {
  const int  mine  = w + lane * K;
  const bool valid = mine < S;
  uint32_t   idle = 0, start = 0;
  for (;;) {
    uint32_t req = 0, tail = 0, head = 0;
    if (valid) {
      req  = simt::vload(&slots[mine].req);      // read before tail: a posted request implies
      tail = simt::vload(&slots[mine].tail);     // that the final tail is visible
      head = simt::vload(&slots[mine].head);
    }
    const unsigned m = simt::ballot(kFull, valid && (tail - head >= 32u || req != 0u));
    if (m == 0u) {
      if (uniform_vload(live, lane) == 0u) break;
      simt::backoff(200);
      if (++idle > DEFL_WATCHDOG_SPINS) break;              // watchdog
      continue;
    }
    idle = 0;
    const unsigned later = m & ~((start ? (1u << start) : 1u) - 1u);
    const int      pick  = simt::ffs(later ? later : m) - 1;
    start = (uint32_t)(pick + 1) & 31u;
    Slot&          sl = slots[w + pick * K];
    const uint32_t r  = simt::shfl(kFull, req, pick);
    const uint32_t h  = simt::shfl(kFull, head, pick);
    const uint32_t av = simt::shfl(kFull, tail, pick) - h;
    if (av) {
      simt::fence_block();
      const uint32_t n = process_batch(sl, ws, lane, h, av);
      simt::syncwarp(kFull);
      if (lane == 0) { simt::fence_block(); simt::vstore(&sl.head, h + n); }
      simt::syncwarp(kFull);
    } else {
      serve_request(sl, ws, lane, r, items, flags, results);
    }
  }
}

}  // namespace defl
