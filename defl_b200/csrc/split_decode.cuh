// split_decode.cuh -- ONE large stream decoded by the whole GPU (BASELINE.json configs[3]).
//
// The reference decodes a stream serially (infl_ft_full, infl.c:1002-1104).  Here the stream is cut at
// deflate block starts found by a speculative search, and the pieces ("chunks") are decoded in parallel,
// one warp each, with the same lane-speculative decoder as the batch path (warp_decode.cuh):
//   F  find_block_starts_kernel : per 32 KiB granule of the input, the first bit position that holds a
//                                 complete, plausible dynamic-block header (rapidgzip-style filters:
//                                 BTYPE, HLIT/HDIST range, Kraft-complete code-length code, Kraft-complete
//                                 literal/length and distance codes, an end-of-block code).
//   A  split_chunk_kernel<COUNT>: every candidate decodes (tokens only) until a block boundary falls on a
//                                 later candidate, counting its output bytes.  The host then walks the
//                                 links from the stream's first block: every chunk reached that way starts
//                                 at a TRUE block boundary, so false candidates can never corrupt anything.
//   C  split_chunk_kernel<DECODE>: the chunks on the chain decode again, now resolving tokens into 16-bit
//                                 symbols at their final offsets; bytes copied from the unknown 32 KiB
//                                 before a chunk become markers 0x8000 | k.
//   D  marker_resolve_kernel    : every marker is followed back through the previous chunks' windows until
//                                 it lands on a byte; the bytes go to dst and into Adler-32 partial sums.
// Anything unusual (no candidates, a chain that does not reach the final block, any decode failure) makes
// the host fall back to the one-warp decoder, which defines the exact error behaviour.
#pragma once
#include "warp_decode.cuh"

namespace defl {

enum : uint32_t { CH_LINKED = 0, CH_FINAL = 1, CH_FAIL = 2, CH_REPLAYED = 3 };   // CH_FAIL | reason << 8
constexpr uint32_t kNoTokens = 0xFFFFFFFFu;

constexpr uint32_t kMaxPieces  = 7;         // a chunk is replayed as up to 8 pieces
#ifndef DEFL_PIECE_BYTES
#define DEFL_PIECE_BYTES 98304
#endif
constexpr uint32_t kPieceBytes = DEFL_PIECE_BYTES;
struct SplitChunkOut {
  uint64_t end_bit;     // stream-relative bit where the chunk stopped (a block boundary unless it failed)
  uint32_t status;      // CH_*
  uint32_t link;        // CH_LINKED: index of the candidate at end_bit
  uint32_t out_bytes;   // bytes the chunk produces
  uint32_t ntok;        // COUNT: tokens kept for the DECODE pass, or kNoTokens
  // chunks that start INSIDE a block (mid-block mode) do not know whether that block is the stream's last one:
  uint32_t link_final;  // CH_LINKED inside a block: 0 / 1 = that block is not / is final, 2 = as this chunk's first block
  uint32_t has_eob;     // the chunk's first block (header unseen) ended inside it; if it was final, the chunk is:
  uint32_t eob_bytes;   //   bytes produced up to there
  uint32_t eob_ntok;    //   tokens kept up to there (or kNoTokens)
  uint64_t eob_bit;     //   stream-relative bit after its end-of-block code
  // COUNT, tokens kept: places where the chunk can be cut for the replay (ends of decode windows about
  // kPieceBytes of output apart): the first piece_tok[m] tokens produce exactly piece_pos[m] bytes
  uint32_t npiece;
  uint32_t piece_tok[kMaxPieces];
  uint32_t piece_pos[kMaxPieces];
};
enum : uint32_t { FIN_NO = 0, FIN_YES = 1, FIN_UNKNOWN = 2 };
enum : uint32_t { MODE_BLOCK_STARTS = 0, MODE_FIXED_MID = 1 };   // SplitStreamDesc::mode: what the candidates are

// One stream of a split decode.  All the kernels below work on MANY streams at once: a table of these plus
// cumulative bases (granules, candidates, chunks, windows, D2 blocks) says which stream a unit belongs to.
struct SplitStreamDesc {
  const uint8_t* src;
  uint8_t*       dst;
  uint64_t srclen;
  uint64_t total;        // output bytes (known after the count pass)
  uint64_t sym_base;     // first 16-bit symbol of the stream in sym[] (a multiple of 8)
  uint64_t tok_base;     // first kept token of the stream in tokbuf[]
  uint64_t blk_base;     // first D2 block of the stream (also its slot in partial[])
  uint32_t body;         // bytes of zlib header in front of the deflate data
  uint32_t gran_base, ngran;   // F: granules searched
  uint32_t cand_base, ncand;   // candidates in cand[], stream-relative bit positions, increasing
  uint32_t off_base, nchunk;   // chain chunks: off[off_base .. off_base + nchunk] = output offsets
  uint32_t win_base;           // first of the stream's nchunk - 1 resolved windows in wbuf
  uint32_t nblk;               // D2 blocks
  uint32_t mode;               // MODE_*: candidates are block starts (found by F) or token boundaries inside
                               // fixed-Huffman blocks (found by P; every candidate but the first)
};

// the stream that owns unit x: the last one whose base is <= x (streams without units share their successor's base)
template <typename F>
DEFL_DEV uint32_t stream_of(uint32_t nstreams, uint64_t x, F base_of) {
  uint32_t lo = 0, hi = nstreams;
  while (hi - lo > 1u) { const uint32_t mid = (lo + hi) >> 1; if (base_of(mid) <= x) lo = mid; else hi = mid; }
  return lo;
}

struct SplitJob {
  uint32_t stream;      // index into the stream table
  uint32_t start_final; // mid-block chunks: FIN_* of the block the chunk starts in (COUNT: FIN_UNKNOWN)
  uint32_t cand;        // index into the stream's candidates: where the chunk starts
  uint32_t cap;         // DECODE: exactly the bytes the COUNT pass reported
  uint64_t out_off;     // DECODE: offset of the chunk's first symbol
  uint64_t tok_off;     // where the chunk's tokens are kept (COUNT writes, DECODE reads)
  uint32_t tok_cap;     // COUNT: token budget; DECODE (replay): 0 = the job starts the stream, 32768 = unknown history before it
  uint32_t ntok;        // DECODE: tokens to replay, or kNoTokens = decode again
};

// stored block inside a chunk (infl_ft_stored, infl.c:659-715): count it, or copy it as symbols
template <typename SYM, bool COUNT>
DEFL_DEV uint32_t stored_block_sym(int lane, BitReader& br, const uint8_t* src, uint32_t srclen, uint64_t base_bit,
                                   SYM* dst, uint32_t& pos, uint32_t cap) {
  const uint64_t bytepos = ((br.bitpos() - base_bit) + 7u) >> 3;
  if (bytepos + 4u > srclen) return R_TRUNCATED;
  const uint8_t* p = src + bytepos;
  const uint32_t len  = simt::ld_in8(p) | ((uint32_t)simt::ld_in8(p + 1) << 8);
  const uint32_t nlen = simt::ld_in8(p + 2) | ((uint32_t)simt::ld_in8(p + 3) << 8);
  if ((len ^ (~nlen & 0xFFFFu)) != 0u) return R_STORED_LEN;
  if (len > cap - pos) return R_OUTPUT_FULL;
  if ((uint64_t)len > srclen - bytepos - 4u) return R_TRUNCATED;
  if (!COUNT) {
    simt::syncwarp(kFull);
    for (uint32_t i = (uint32_t)lane; i < len; i += 32u) dst[pos + i] = (SYM)simt::ld_in8(p + 4 + i);
    simt::syncwarp(kFull);
  }
  pos += len;
  br.seek(base_bit + (bytepos + 4u + len) * 8u);
  return R_NONE;
}

// One chunk: from candidate `me` to the first block boundary that is a later candidate, or through the
// final block.  cand[] holds stream-relative bit positions in increasing order.
template <typename SYM, bool COUNT>
DEFL_DEV void inflate_chunk_warp(WarpSmem& sm, const WarpScratch& sc, int lane, const uint8_t* src, uint32_t srclen,
                                 const uint64_t* cand, uint32_t ncand, uint32_t me, SYM* dst, uint32_t cap,
                                 uint32_t hist, SplitChunkOut& out, TokSink* sink = nullptr,
                                 const uint32_t* saved = nullptr, uint32_t nsaved = kNoTokens,
                                 uint32_t mode = MODE_BLOCK_STARTS, uint32_t start_final = FIN_UNKNOWN) {
  out.link_final = FIN_NO; out.has_eob = 0; out.eob_bytes = 0; out.eob_ntok = kNoTokens; out.eob_bit = 0;
  out.npiece = 0;
  uint32_t mark_pos = 0;                      // output position of the last piece mark
  if (!COUNT && saved && nsaved != kNoTokens) {
    // the COUNT pass kept this chunk's tokens: only the byte-producing half of the work is left
    uint32_t p = 0;
    simt::syncwarp(kFull);
    const uint32_t r = resolve_tokens<SYM>(sm.stage, lane, saved, nsaved, dst, cap, p, hist);
    out.end_bit = 0; out.link = 0; out.ntok = nsaved; out.out_bytes = p;
    out.status = r ? (CH_FAIL | (r << 8)) : (uint32_t)CH_REPLAYED;
    return;
  }
  Lane<32> L;
  L.gmask = kFull; L.lane = lane; L.wlane = lane;
  const uint32_t mis      = (uint32_t)((uintptr_t)src & 3u);
  const uint64_t base_bit = (uint64_t)mis * 8u;
  const uint64_t end_bit  = base_bit + (uint64_t)srclen * 8u;
  BitReader br;
  br.w    = reinterpret_cast<const uint32_t*>((uintptr_t)src - mis);
  br.wend = (uint32_t)(((uint64_t)mis + srclen + 3u) >> 2);
  const uint64_t start = cand[me];
  br.seek(base_bit + start);
  const bool mid = mode == MODE_FIXED_MID;          // candidates (but the first) are token boundaries inside fixed blocks
  bool       in_first_unseen = mid && me != 0u;     // the chunk starts inside a block whose header it never saw

  uint32_t pos = 0, reason = R_NONE, next = me + 1u, status = CH_FAIL;
  bool     fixed_ready = false;
#pragma unroll 1
  for (;;) {
    uint32_t bfinal, btype;
    if (in_first_unseen) {
      bfinal = start_final; btype = 1u;
    } else {
      const uint64_t q = br.bitpos() - base_bit;
      if (!mid && q != start) {
        while (next < ncand && cand[next] < q) next++;
        if (next < ncand && cand[next] == q) { status = CH_LINKED; break; }
      }
      if ((int64_t)(end_bit - br.bitpos()) < 3) { reason = R_TRUNCATED; break; }
      const uint32_t hdr = br.peek();
      bfinal = hdr & 1u; btype = (hdr >> 1) & 3u;
      br.skip(3); br.norm();
    }
    if (btype == 0) {
      const uint32_t pos0 = pos;
      const uint64_t bit0 = br.bitpos();
      reason = stored_block_sym<SYM, COUNT>(lane, br, src, srclen, base_bit, dst, pos, cap);
      if (reason) break;
      // kept tokens (COUNT pass): stored bytes go in as literal tokens, two to a token, so that such a chunk is only
      // replayed in the DECODE pass like any other.  (An empty stored block -- what a flush leaves, 63 of them in a
      // pigz-style stream -- adds nothing; before, every chunk holding one was decoded a second time, a whole chunk
      // on one warp, and those 63 set the time of the pass.)
      if (COUNT && sink && sink->buf && sink->ok && pos != pos0) {
        const uint32_t len = pos - pos0, nt = (len + 1u) >> 1;
        if (nt > sink->cap - sink->n) {
          sink->ok = false;
        } else {
          const uint8_t* p = src + ((((bit0 - base_bit) + 7u) >> 3) + 4u);
          for (uint32_t i = (uint32_t)lane; i < nt; i += 32u) {
            const uint32_t b0 = simt::ld_in8(p + 2u * i);
            const bool     two = 2u * i + 1u < len;
            const uint32_t b1 = two ? simt::ld_in8(p + 2u * i + 1u) : 0u;
            sink->buf[sink->n + i] = ((two ? 2u : 1u) << 24) | (b1 << 8) | b0;
          }
          sink->n += nt;
        }
      }
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
      bool     linked = false;
#pragma unroll 1
      do {
        uint64_t soft = ~0ull;
        if (mid && btype == 1u) {
          // inside a fixed block a later candidate may lie on this chunk's trajectory: stop exactly there
          const uint64_t q = cur - base_bit;
          while (next < ncand && cand[next] < q) next++;
          if (next < ncand && cand[next] == q && q != start) { linked = true; break; }
          if (next < ncand && cand[next] == q) next++;        // the chunk's own start
          if (next < ncand) soft = base_bit + cand[next];
        }
        st = decode_window<SYM, COUNT>(sm, sc, lane, br.w, br.wend, end_bit, cur, dst, cap, pos, hist, sink, soft,
                                       btype == 1u ? kWarmFixed : kWarmDynamic);
        if (COUNT && sink && sink->buf && sink->ok && out.npiece < kMaxPieces && pos - mark_pos >= kPieceBytes &&
            (st == LS_OK || st == LS_EOB)) {
          out.piece_tok[out.npiece] = sink->n; out.piece_pos[out.npiece] = pos; out.npiece++;
          mark_pos = pos;
        }
      } while (st == LS_OK);
      if (linked) {
        br.seek(cur);
        status = CH_LINKED;
        out.link_final = in_first_unseen ? start_final : bfinal;
        break;
      }
      if (st != LS_EOB) { reason = st >> 8; break; }
      br.seek(cur);
    }
    if (in_first_unseen) {
      in_first_unseen = false;
      if (bfinal == FIN_UNKNOWN) {
        // remember where the chunk ends if that block was the last one, then go on as if it was not
        out.has_eob = 1; out.eob_bytes = pos; out.eob_bit = br.bitpos() - base_bit;
        out.eob_ntok = (COUNT && sink && sink->buf && sink->ok) ? sink->n : kNoTokens;
        bfinal = 0;
      }
    }
    if (bfinal) { status = CH_FINAL; break; }
  }
  out.end_bit   = br.bitpos() - base_bit;
  out.status    = reason ? (CH_FAIL | (reason << 8)) : status;
  out.link      = next;
  out.out_bytes = pos;
  out.ntok      = (COUNT && sink && sink->buf && sink->ok) ? sink->n : kNoTokens;
}

// P: a probable token boundary inside a fixed-Huffman block, for one nominal chunk start (one LANE per chunk).
// Decodes serially from bit p (an arbitrary position) for probe_bits: fixed codes re-synchronise, so where this
// trajectory stands then is very likely a boundary of the true one.  A wrong guess costs nothing: no chunk's true
// trajectory will land on it.  Stops (no candidate) at a final block, at any block that is not fixed, on any error.
// lit / dist: the fixed-code tables.  Returns the absolute bit position or kNoCandidate.
constexpr uint64_t kNoCandidate = ~0ull;
DEFL_DEV uint64_t fixed_probe_lane(const uint16_t* lit, const uint16_t* dist, const uint32_t* words, uint32_t wend,
                                   uint64_t p, uint64_t end_bit, uint32_t probe_bits) {
  if (p + probe_bits + 64u >= end_bit) return kNoCandidate;
  SubReader br;
  br.w = words; br.wend = wend;
  br.seek((uint32_t)(p >> 5), (uint32_t)(p & 31u));
  uint64_t r = p;
  const uint64_t target = p + probe_bits;
#pragma unroll 1
  for (;;) {
    if (r >= target) return r;
    uint32_t tok, used;
    const uint64_t left = end_bit - r;
    const uint32_t s = decode_token(lit, dist, br, left > 0x7FFFFFFFull ? 0x7FFFFFFFu : (uint32_t)left, 0u, tok, used);
    if (s == LS_OK) { r += used; br.consume(used); continue; }
    // Not yet on the true trajectory, the bits may spell a code that does not exist or an end-of-block that is
    // none: step one bit and go on, the trajectory is still to be met.  (A true end of block that is followed by
    // the final block or by one that is not fixed also lands here; what the probe reports after that is a wrong
    // guess, which costs one chunk nobody links to.)
    if (s != LS_EOB) { r += 1u; br.consume(1u); continue; }
    r += used; br.consume(used);
    const uint32_t hdr = simt::funnel_r(br.w0, br.w1, br.bo) & 7u;       // BFINAL, BTYPE of the next block
    if ((hdr & 1u) || ((hdr >> 1) & 3u) != 1u) continue;
    r += 3u; br.consume(3u);
  }
}

#ifdef __CUDACC__
#ifndef DEFL_SPLIT_CTAS_PER_SM
#define DEFL_SPLIT_CTAS_PER_SM 7     // x 4 warps, as the batch kernel (kernels.cuh)
#endif
template <bool COUNT, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, DEFL_SPLIT_CTAS_PER_SM)
split_chunk_kernel(const SplitStreamDesc* __restrict__ streams, const uint64_t* __restrict__ cand,
                   const SplitJob* __restrict__ jobs, uint32_t njobs, uint16_t* __restrict__ sym,
                   SplitChunkOut* __restrict__ outs, unsigned int* __restrict__ ticket,
                   unsigned char* __restrict__ scratch, uint32_t* __restrict__ tokbuf) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WarpSmem&         sm = reinterpret_cast<WarpSmem*>(smem_raw)[warp];
  const WarpScratch sc = scratch_of(scratch + ((size_t)blockIdx.x * WARPS + warp) * kWarpScratchBytes);
  for (;;) {
    uint32_t k = 0;
    if (lane == 0) k = atomicAdd(ticket, 1u);
    k = simt::shfl(kFull, k, 0);
    if (k >= njobs) break;
    const SplitJob job = jobs[k];
    if (!COUNT && tokbuf == nullptr && job.ntok != kNoTokens) continue;     // replayed by split_replay_kernel
    const SplitStreamDesc& sd = streams[job.stream];
    SplitChunkOut  o;
    TokSink        sink;
    sink.buf = (COUNT && tokbuf) ? tokbuf + job.tok_off : nullptr;
    sink.cap = job.tok_cap; sink.n = 0; sink.ok = sink.buf != nullptr;
    // the chunk that starts the stream has no unknown window before it
    inflate_chunk_warp<uint16_t, COUNT>(sm, sc, lane, sd.src, (uint32_t)sd.srclen, cand + sd.cand_base, sd.ncand, job.cand,
                                        COUNT ? nullptr : sym + job.out_off, COUNT ? 0xFFFFFFFFu : job.cap,
                                        job.cand == 0u ? 0u : 32768u, o, COUNT ? &sink : nullptr,
                                        (!COUNT && tokbuf) ? tokbuf + job.tok_off : nullptr, job.ntok,
                                        sd.mode, job.start_final);
    if (lane == 0) outs[k] = o;
    simt::syncwarp(kFull);
  }
}

// C for the chunks whose tokens the COUNT pass kept (nearly all): only the resolver, in a kernel of its own.  A warp's
// replay is a chain of dependent memory round trips (tokens -> sources -> stores), so what it needs is company: without
// the decode tables (1 KiB of stage per warp instead of 5) and the decoder's registers, 12 CTAs x 4 warps fit an SM --
// 7104 warps, so the ~4700 chunks of a 1 GiB stream run in ONE wave where split_chunk_kernel<false> (4144 warps) needed two.
#ifndef DEFL_REPLAY_CTAS_PER_SM
#define DEFL_REPLAY_CTAS_PER_SM 12
#endif
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32, DEFL_REPLAY_CTAS_PER_SM)
split_replay_kernel(const SplitJob* __restrict__ jobs, uint32_t njobs, uint16_t* __restrict__ sym,
                    SplitChunkOut* __restrict__ outs, unsigned int* __restrict__ ticket,
                    const uint32_t* __restrict__ tokbuf) {
  __shared__ __align__(16) uint16_t stage[WARPS][kStageV3];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (;;) {
    uint32_t k = 0;
    if (lane == 0) k = atomicAdd(ticket, 1u);
    k = simt::shfl(kFull, k, 0);
    if (k >= njobs) break;
    const SplitJob job = jobs[k];
    if (job.ntok == kNoTokens) continue;                 // decoded again by split_chunk_kernel<false>
    uint32_t p = 0;
    simt::syncwarp(kFull);
    const uint32_t r = resolve_tokens<uint16_t>(stage[warp], lane, tokbuf + job.tok_off, job.ntok, sym + job.out_off,
                                                job.cap, p, job.tok_cap);      // DECODE jobs: tok_cap = symbols of unknown history
    if (lane == 0) {
      SplitChunkOut o;
      o.npiece = 0;
      o.link_final = FIN_NO; o.has_eob = 0; o.eob_bytes = 0; o.eob_ntok = kNoTokens; o.eob_bit = 0;
      o.end_bit = 0; o.link = 0; o.ntok = job.ntok; o.out_bytes = p;
      o.status = r ? (CH_FAIL | (r << 8)) : (uint32_t)CH_REPLAYED;
      outs[k] = o;
    }
    simt::syncwarp(kFull);
  }
}

// P for every granule of the streams in MODE_FIXED_MID: one lane per granule, the warp's fixed-code tables shared
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
fixed_probe_kernel(const SplitStreamDesc* __restrict__ streams, uint32_t nstreams, uint32_t ngran_total,
                   uint32_t granule_bytes, uint32_t probe_bits, uint64_t* __restrict__ found) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WarpSmem& sm = reinterpret_cast<WarpSmem*>(smem_raw)[warp];
  Lane<32> L;
  L.gmask = kFull; L.lane = lane; L.wlane = lane;
  BitReader none;
  none.w = nullptr; none.wend = 0;
  bool fixed_ready = false;
  huffman_block_setup<32>(L, sm.lit, sm.dist, sm.s, none, 0, 1u, fixed_ready);
  simt::syncwarp(kFull);
  const uint32_t gg = (blockIdx.x * WARPS + warp) * 32u + (uint32_t)lane;
  if (gg >= ngran_total) return;
  const SplitStreamDesc& sd = streams[stream_of(nstreams, gg, [&](uint32_t i) { return (uint64_t)streams[i].gran_base; })];
  if (sd.mode != MODE_FIXED_MID) return;
  const uint32_t g = gg - sd.gran_base;
  uint64_t result = kNoCandidate;
  if (g != 0u) {                                            // granule 0 starts the stream: a known block start
    const uint32_t mis      = (uint32_t)((uintptr_t)sd.src & 3u);
    const uint64_t base_bit = (uint64_t)mis * 8u;
    const uint64_t end_bit  = base_bit + sd.srclen * 8u;
    const uint32_t* words   = reinterpret_cast<const uint32_t*>((uintptr_t)sd.src - mis);
    const uint32_t  wend    = (uint32_t)(((uint64_t)mis + sd.srclen + 3u) >> 2);
    const uint64_t  p       = base_bit + ((uint64_t)sd.body + (uint64_t)g * granule_bytes) * 8u;
    const uint64_t  r       = fixed_probe_lane(sm.lit, sm.dist, words, wend, p, end_bit, probe_bits);
    if (r != kNoCandidate) result = r - base_bit;
  }
  found[gg] = result;
}

// first bytes of every stream (zlib header check, first block type, on the host) / the four bytes at a given offset (Adler trailer)
__global__ void split_peek_kernel(const SplitStreamDesc* __restrict__ streams, uint32_t nstreams,
                                  const uint64_t* __restrict__ at, uint32_t nbytes, uint8_t* __restrict__ out) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= nstreams) return;
  const uint64_t a = at ? at[s] : 0u;
  for (uint32_t i = 0; i < nbytes; i++)
    out[s * nbytes + i] = a + i < streams[s].srclen ? streams[s].src[a + i] : (uint8_t)0;
}

#endif  // __CUDACC__

// ---- F: block-start search -------------------------------------------------------------------------------

// The search is a funnel, every stage lane-parallel over bit positions (nothing here is warp-serial):
//   1. cheap   : BTYPE = dynamic, HLIT <= 286, HDIST <= 30                       (22 % of positions pass)
//   2. precode : the code-length code is Kraft-complete (what every encoder emits) (0.09 % pass), on a
//                compacted list so that all 32 lanes work; 3 lengths per lookup in a 512-entry table
//   3. header  : the lane decodes the whole code-length sequence in registers (canonical decode by
//                comparisons, no table) and keeps Kraft sums: literal/length code complete, an end-of-block
//                code, a non-empty distance code that is not over-subscribed.  32 candidates per round.
// The first position of the granule that passes all three goes to found[g].
#ifdef __CUDACC__
DEFL_DEV bool header_cheap(uint32_t x0) {
  return ((x0 >> 1) & 3u) == 2u && ((x0 >> 3) & 31u) <= 29u && ((x0 >> 8) & 31u) <= 29u;
}

// x0/x1/x2 = the 96 bits at the candidate; kraft9[v] = sum over the three 3-bit lengths in v of (l ? 128 >> l : 0)
DEFL_DEV bool precode_complete(uint32_t x0, uint32_t x1, uint32_t x2, const uint8_t* kraft9) {
  const uint32_t nb = (((x0 >> 13) & 15u) + 4u) * 3u;              // 12..57 bits of lengths from bit 17 on
  uint32_t f0 = simt::funnel_r(x0, x1, 17u) & 0x7FFFFFFu;          // lengths 0..8
  uint32_t f1 = simt::funnel_r(x1, x2, 12u) & 0x7FFFFFFu;          // lengths 9..17
  uint32_t f2 = (x2 >> 7) & 7u;                                    // length 18
  if (nb < 27u) f0 &= (1u << nb) - 1u;
  f1 = nb <= 27u ? 0u : (nb < 54u ? f1 & ((1u << (nb - 27u)) - 1u) : f1);
  f2 = nb > 54u ? f2 : 0u;
  const uint32_t k = kraft9[f0 & 511u] + kraft9[(f0 >> 9) & 511u] + kraft9[f0 >> 18] +
                     kraft9[f1 & 511u] + kraft9[(f1 >> 9) & 511u] + kraft9[f1 >> 18] + kraft9[f2];
  return k == 128u;
}

struct LaneBits {        // one lane's private bit reader over the input words
  const uint32_t* w;
  uint32_t        wend, widx, nb;
  uint64_t        bb;
  DEFL_DEV uint32_t ld(uint32_t i) const { return i < wend ? simt::ld_in32(w + i) : 0u; }
  DEFL_DEV void init(const uint32_t* words, uint32_t we, uint64_t bit) {
    w = words; wend = we;
    widx = (uint32_t)(bit >> 5);
    const uint32_t sh = (uint32_t)(bit & 31u);
    bb = (((uint64_t)ld(widx + 1) << 32) | ld(widx)) >> sh;
    nb = 64u - sh;
    widx += 2;
  }
  DEFL_DEV uint32_t peek() const { return (uint32_t)bb; }
  DEFL_DEV void skip(uint32_t n) {       // n <= 17
    bb >>= n; nb -= n;
    if (nb <= 32u) { bb |= (uint64_t)ld(widx++) << nb; nb += 32u; }
  }
};

// Stage 3 for one lane: pc = absolute bit of the 3-bit block header.  The precode is known to be complete.
DEFL_DEV bool header_full_check(const uint32_t* words, uint32_t wend, uint64_t pc, uint64_t end_bit) {
  LaneBits br;
  br.init(words, wend, pc);
  uint32_t v = br.peek();
  const uint32_t hlit = ((v >> 3) & 31u) + 257u, hdist = ((v >> 8) & 31u) + 1u, hclen = ((v >> 13) & 15u) + 4u;
  const uint32_t n = hlit + hdist;
  br.skip(17);
  uint64_t used = 17u + 3u * hclen;
  // the 19 code-length-code lengths, packed 3 bits each in symbol order (RFC 1951 3.2.7 permutation)
  uint64_t clpack = 0;
#pragma unroll 1
  for (uint32_t i = 0; i < hclen; i++) {
    clpack |= (uint64_t)(br.peek() & 7u) << (3u * cl_order(i));
    br.skip(3);
  }
  // canonical code in a 7-bit MSB-aligned code space: lim(l) = upper end of the codes of length <= l
  uint64_t cnt = 0;                        // byte l = symbols of length l
#pragma unroll 1
  for (uint32_t s = 0; s < 19u; s++) cnt += 1ull << (8u * ((uint32_t)(clpack >> (3u * s)) & 7u));
  uint64_t lims = 0, offs = 0;             // byte l = lim(l) / number of coded symbols shorter than l
  {
    uint32_t acc = 0, o = 0;
#pragma unroll
    for (uint32_t l = 1; l <= 7u; l++) {
      const uint32_t c = (uint32_t)(cnt >> (8u * l)) & 255u;
      offs |= (uint64_t)o << (8u * l);
      acc += c << (7u - l);
      o += c;
      lims |= (uint64_t)(acc > 255u ? 255u : acc) << (8u * l);
    }
  }
  uint64_t slo = 0, shi = 0;               // symbols sorted by (length, symbol), 5 bits each, 12 per word
  {
    uint64_t cursor = offs;
#pragma unroll 1
    for (uint32_t s = 0; s < 19u; s++) {
      const uint32_t l = (uint32_t)(clpack >> (3u * s)) & 7u;
      if (l) {
        const uint32_t pos = (uint32_t)(cursor >> (8u * l)) & 255u;
        cursor += 1ull << (8u * l);
        if (pos < 12u) slo |= (uint64_t)s << (5u * pos);
        else shi |= (uint64_t)s << (5u * (pos - 12u));
      }
    }
  }
  const uint32_t limA = (uint32_t)(lims >> 8);                       // lim(1..4)
  const uint32_t limB = ((uint32_t)(lims >> 40) & 0xFFFFu) | 0xFFFF0000u;   // lim(5..6), then never
  uint32_t i = 0, prev = 0, kl = 0, kd = 0;
  bool eob = false;
#pragma unroll 1
  while (i < n) {
    v = br.peek();
    const uint32_t x  = simt::brev(v) >> 25;                         // 7 code bits, first bit on top
    const uint32_t xx = x * 0x01010101u;
    const uint32_t l  = 1u + (uint32_t)((simt::popc(__vcmpgeu4(xx, limA)) + simt::popc(__vcmpgeu4(xx, limB))) >> 3);
    const uint32_t base = (uint32_t)(lims >> (8u * (l - 1u))) & 255u;
    const uint32_t idx  = ((uint32_t)(offs >> (8u * l)) & 255u) + ((x - base) >> (7u - l));
    const uint32_t sym  = (uint32_t)(idx < 12u ? slo >> (5u * idx) : shi >> (5u * (idx - 12u))) & 31u;
    v >>= l;
    uint32_t rep = 1, val = sym, xb = 0;
    if (sym >= 16u) {
      if (sym == 16u)      { if (i == 0u) return false; xb = 2; rep = 3u + (v & 3u); val = prev; }
      else if (sym == 17u) { xb = 3; rep = 3u + (v & 7u); val = 0; }
      else                 { xb = 7; rep = 11u + (v & 127u); val = 0; }
      if (i + rep > n) return false;
      if (sym != 16u) prev = 0;
    } else {
      prev = sym;
    }
    if (val) {
      const uint32_t c  = 32768u >> val;
      const uint32_t nl = i >= hlit ? 0u : (hlit - i < rep ? hlit - i : rep);
      kl += nl * c;
      kd += (rep - nl) * c;
      if (i <= 256u && 256u < i + rep) eob = true;
    }
    i += rep;
    br.skip(l + xb);
    used += l + xb;
    if (kl > 32768u) return false;
  }
  return kl == 32768u && eob && kd != 0u && kd <= 32768u && pc + used <= end_bit;
}

constexpr uint32_t kFindSpan = 1024;   // bit positions per scan step: one 32-bit word per lane

struct FindSmem {
  uint16_t list[kFindSpan];   // positions of the current span that pass stage 1, in increasing order
  uint32_t pend[64];          // granule-relative positions that passed stage 2, in increasing order
};

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
find_block_starts_kernel(const SplitStreamDesc* __restrict__ streams, uint32_t nstreams, uint32_t ngran_total,
                         uint32_t granule_bytes, uint64_t* __restrict__ found) {
  __shared__ FindSmem fsm[WARPS];
  __shared__ uint8_t  kraft9[512];
  for (uint32_t i = threadIdx.x; i < 512u; i += WARPS * 32) {
    const uint32_t a = i & 7u, b = (i >> 3) & 7u, c = i >> 6;
    kraft9[i] = (uint8_t)((a ? 128u >> a : 0u) + (b ? 128u >> b : 0u) + (c ? 128u >> c : 0u));
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  FindSmem& fs = fsm[warp];
  const uint32_t gg = blockIdx.x * WARPS + warp;          // granule number over all streams
  if (gg >= ngran_total) return;
  const SplitStreamDesc& sd = streams[stream_of(nstreams, gg, [&](uint32_t i) { return (uint64_t)streams[i].gran_base; })];
  if (sd.mode != MODE_BLOCK_STARTS) return;               // P fills this stream's entries
  const uint8_t* const src = sd.src;
  const uint32_t srclen = (uint32_t)sd.srclen, body = sd.body;
  const uint32_t g = gg - sd.gran_base;
  const unsigned lt = simt::lanemask_lt();
  const uint32_t mis      = (uint32_t)((uintptr_t)src & 3u);
  const uint64_t base_bit = (uint64_t)mis * 8u;
  const uint64_t end_bit  = base_bit + (uint64_t)srclen * 8u;
  const uint32_t* words   = reinterpret_cast<const uint32_t*>((uintptr_t)src - mis);
  const uint32_t  wend    = (uint32_t)(((uint64_t)mis + srclen + 3u) >> 2);
  const uint64_t p0 = base_bit + ((uint64_t)body + (uint64_t)g * granule_bytes) * 8u;
  uint64_t       p1 = p0 + (uint64_t)granule_bytes * 8u;
  if (p1 + 96u > end_bit) p1 = end_bit > 96u ? end_bit - 96u : 0u;     // a header needs room
  uint64_t result = kNoCandidate;
  uint32_t npend = 0;

  // stage 3 over the first (up to) 32 pending candidates
  auto parse_round = [&]() {
    const uint32_t take = npend < 32u ? npend : 32u;
    bool pass = false;
    if ((uint32_t)lane < take) pass = header_full_check(words, wend, p0 + fs.pend[lane], end_bit);
    const unsigned m = simt::ballot(kFull, pass);
    if (m) {
      const int l = simt::ffs(m) - 1;
      result = p0 + fs.pend[l] - base_bit;
    } else {
      const uint32_t keep = npend - take;
      const uint32_t v = (uint32_t)lane < keep ? fs.pend[32 + lane] : 0u;
      simt::syncwarp(kFull);
      if ((uint32_t)lane < keep) fs.pend[lane] = v;
      npend = keep;
    }
    simt::syncwarp(kFull);
  };

#pragma unroll 1
  for (uint64_t pb = p0 & ~31ull; pb < p1 && result == kNoCandidate; pb += kFindSpan) {
    // stage 1, 32 positions per lane with plain bit logic: bit k of sK = input bit (position + K)
    const uint64_t q  = pb + 32u * (uint32_t)lane;             // this lane's first position (word aligned)
    const uint32_t wi = (uint32_t)(q >> 5);
    const uint32_t w0 = wi < wend ? simt::ld_in32(words + wi) : 0u;
    const uint32_t w1 = wi + 1 < wend ? simt::ld_in32(words + wi + 1) : 0u;
#define DEFL_S(K) simt::funnel_r(w0, w1, K)
    uint32_t m = ~DEFL_S(1) & DEFL_S(2)                                   // BTYPE = 2
                 & ~(DEFL_S(4) & DEFL_S(5) & DEFL_S(6) & DEFL_S(7))        // HLIT  < 30
                 & ~(DEFL_S(9) & DEFL_S(10) & DEFL_S(11) & DEFL_S(12));    // HDIST < 30
#undef DEFL_S
    if (q < p0) m = p0 - q >= 32u ? 0u : m & (0xFFFFFFFFu << (uint32_t)(p0 - q));
    if (q + 32u > p1) m = q >= p1 ? 0u : m & ((1u << (uint32_t)(p1 - q)) - 1u);
    // compact: every lane appends its positions at its prefix offset
    const uint32_t c = (uint32_t)simt::popc(m);
    uint32_t inc = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = simt::shfl_up(kFull, inc, d);
      if (lane >= d) inc += t;
    }
    const uint32_t cnt = simt::shfl(kFull, inc, 31);
    uint32_t o = inc - c;
    while (m) {
      const uint32_t b = (uint32_t)simt::ffs(m) - 1u;
      m &= m - 1u;
      fs.list[o++] = (uint16_t)(32u * (uint32_t)lane + b);
    }
    simt::syncwarp(kFull);
#pragma unroll 1
    for (uint32_t k0 = 0; k0 < cnt && result == kNoCandidate; k0 += 32u) {
      const bool has = k0 + (uint32_t)lane < cnt;
      bool good = false;
      uint64_t p = 0;
      if (has) {
        p = pb + fs.list[k0 + (uint32_t)lane];
        const uint32_t xi = (uint32_t)(p >> 5), sh = (uint32_t)(p & 31u);
        const uint32_t x0 = xi < wend ? simt::ld_in32(words + xi) : 0u;
        const uint32_t x1 = xi + 1 < wend ? simt::ld_in32(words + xi + 1) : 0u;
        const uint32_t x2 = xi + 2 < wend ? simt::ld_in32(words + xi + 2) : 0u;
        const uint32_t x3 = xi + 3 < wend ? simt::ld_in32(words + xi + 3) : 0u;
        good = precode_complete(simt::funnel_r(x0, x1, sh), simt::funnel_r(x1, x2, sh), simt::funnel_r(x2, x3, sh), kraft9);
      }
      const unsigned gm = simt::ballot(kFull, good);
      if (good) fs.pend[npend + (uint32_t)simt::popc(gm & lt)] = (uint32_t)(p - p0);
      npend += (uint32_t)simt::popc(gm);
      simt::syncwarp(kFull);
      if (npend >= 32u) parse_round();
    }
    simt::syncwarp(kFull);
  }
#pragma unroll 1
  while (result == kNoCandidate && npend > 0u) parse_round();
  if (lane == 0) found[gg] = result;
}

#endif  // __CUDACC__

// ---- D: markers -> bytes, Adler-32 partial sums --------------------------------------------------------------
// sym[i] for output position i: a byte, or 0x8000 | k = "byte k of the 32 KiB before my chunk".  off[c] =
// first output position of chunk c (off[nchunk] = total).  Every chunk but the last produces >= 32 KiB, so
// the window before chunk c normally lies inside chunk c-1 and a marker found there refers to the tail before that chunk, ...
// Each block reduces its bytes to (sum x, sum i*x mod 65521) for the host to combine.
// the byte at output position i (inside chunk cc): follow markers into the tails before the chunks
DEFL_DEV uint32_t resolve_marker(const uint16_t* sym, const uint64_t* off, uint32_t cc, uint64_t i) {
  uint32_t s = sym[i];
  while (s & 0x8000u) {
    const uint64_t at = off[cc] - 32768u + (s & 0x7FFFu);
    do { cc--; } while (off[cc] > at);          // normally one step: chunks produce >= 32 KiB
    s = sym[at];
  }
  return s;
}

#ifdef __CUDACC__
// D runs as two kernels so that no marker needs more than one look-up in fast memory:
//   D1 window_resolve_kernel : for every chunk c >= 1 the 32 KiB of output before it, as resolved bytes
//                              (wbuf[c-1]); only these tail symbols follow marker chains through HBM.
//   D2 marker_resolve_kernel : a block owns 64 KiB of output; per chunk it touches it stages wbuf[c-1] in
//                              shared memory, streams the symbols (16-byte loads), replaces markers from the
//                              staged window, stores 8 bytes at a time and keeps Adler-32 partial sums.
constexpr uint32_t kWinBytes        = 32768;
constexpr uint32_t kResolvePerBlock = 65536;      // output bytes per D2 block

__global__ void __launch_bounds__(256)
window_resolve_kernel(const SplitStreamDesc* __restrict__ streams, uint32_t nstreams, const uint16_t* __restrict__ sym_all,
                      const uint64_t* __restrict__ off_all, uint8_t* __restrict__ wbuf_all) {
  const uint32_t wi = blockIdx.x / 16u;                           // window number over all streams; 16 blocks x 2048 symbols each
  const SplitStreamDesc& sd = streams[stream_of(nstreams, wi, [&](uint32_t i) { return (uint64_t)streams[i].win_base; })];
  const uint16_t* const sym  = sym_all + sd.sym_base;
  const uint64_t* const off  = off_all + sd.off_base;
  uint8_t* const        wbuf = wbuf_all + (size_t)sd.win_base * kWinBytes;
  const uint32_t nchunk = sd.nchunk;
  const uint32_t c  = 1u + (wi - sd.win_base);
  const uint32_t k0 = (blockIdx.x % 16u) * 2048u + threadIdx.x * 8u;
  if (c >= nchunk) return;
  const uint64_t base = off[c] - kWinBytes;                       // the host made sure off[c] >= 32 KiB
  uint32_t s[8];
#pragma unroll
  for (int k = 0; k < 8; k++) s[k] = sym[base + k0 + (uint32_t)k];
  if ((s[0] | s[1] | s[2] | s[3] | s[4] | s[5] | s[6] | s[7]) & 0x8000u) {
#pragma unroll 1
    for (int k = 0; k < 8; k++) {
      if (!(s[k] & 0x8000u)) continue;
      const uint64_t i = base + k0 + (uint32_t)k;
      uint32_t cc = c - 1u;
      while (off[cc] > i) cc--;
      s[k] = resolve_marker(sym, off, cc, i);
    }
  }
  const uint32_t lo = s[0] | (s[1] << 8) | (s[2] << 16) | (s[3] << 24);
  const uint32_t hi = s[4] | (s[5] << 8) | (s[6] << 16) | (s[7] << 24);
  *reinterpret_cast<uint2*>(wbuf + (size_t)(c - 1u) * kWinBytes + k0) = make_uint2(lo, hi);
}

// D1 for streams with long chains of chunks (one deflate history over hundreds of MB): following every marker
// back chunk by chunk, as window_resolve_kernel does, costs (markers) x (chain length) serial look-ups -- a byte
// of a common word can travel through thousands of windows.  Pointer doubling instead: T[w][k] is the 32 KiB
// before chunk w + 1 as 16-bit symbols whose markers index the window `hop` windows earlier (hop = 1 at first:
// a chunk's markers index the window right before it; needs every inner chunk to produce >= 32 KiB, which the
// host checks).  One round replaces every marker by the symbol it points at in T[w - hop], which indexes the
// window 2 * hop back; after ceil(log2(windows)) rounds nothing is left to follow.
__global__ void __launch_bounds__(256)
window_double_init_kernel(const SplitStreamDesc* __restrict__ streams, uint32_t nstreams, const uint16_t* __restrict__ sym_all,
                          const uint64_t* __restrict__ off_all, uint16_t* __restrict__ T) {
  const uint32_t wi = blockIdx.x / 16u;
  const SplitStreamDesc& sd = streams[stream_of(nstreams, wi, [&](uint32_t i) { return (uint64_t)streams[i].win_base; })];
  const uint32_t c  = 1u + (wi - sd.win_base);
  if (c >= sd.nchunk) return;
  const uint32_t k0 = (blockIdx.x % 16u) * 2048u + threadIdx.x * 8u;
  const uint16_t* src = sym_all + sd.sym_base + (off_all + sd.off_base)[c] - kWinBytes + k0;
  uint16_t*       dst = T + (size_t)wi * kWinBytes + k0;
#pragma unroll
  for (int k = 0; k < 8; k++) dst[k] = src[k];           // the tail need not be 16-byte aligned
}

__global__ void __launch_bounds__(256)
window_double_round_kernel(const SplitStreamDesc* __restrict__ streams, uint32_t nstreams, const uint16_t* __restrict__ Tin,
                           uint16_t* __restrict__ Tout, uint32_t hop, uint8_t* __restrict__ wbuf_all) {
  const uint32_t wi = blockIdx.x / 16u;
  const SplitStreamDesc& sd = streams[stream_of(nstreams, wi, [&](uint32_t i) { return (uint64_t)streams[i].win_base; })];
  const uint32_t w  = wi - sd.win_base;
  if (w + 1u >= sd.nchunk) return;
  const uint32_t k0 = (blockIdx.x % 16u) * 2048u + threadIdx.x * 8u;
  const uint4 v = *reinterpret_cast<const uint4*>(Tin + (size_t)wi * kWinBytes + k0);
  uint32_t s[8] = {v.x & 0xFFFFu, v.x >> 16, v.y & 0xFFFFu, v.y >> 16, v.z & 0xFFFFu, v.z >> 16, v.w & 0xFFFFu, v.w >> 16};
  if (w >= hop && ((v.x | v.y | v.z | v.w) & 0x80008000u)) {
    const uint16_t* back = Tin + (size_t)(wi - hop) * kWinBytes;
#pragma unroll
    for (int k = 0; k < 8; k++)
      if (s[k] & 0x8000u) s[k] = back[s[k] & 0x7FFFu];
  }
  if (wbuf_all) {                                           // the last round leaves the windows as bytes for D2
    const uint32_t lo = (s[0] & 255u) | ((s[1] & 255u) << 8) | ((s[2] & 255u) << 16) | (s[3] << 24);
    const uint32_t hi = (s[4] & 255u) | ((s[5] & 255u) << 8) | ((s[6] & 255u) << 16) | (s[7] << 24);
    *reinterpret_cast<uint2*>(wbuf_all + (size_t)wi * kWinBytes + k0) = make_uint2(lo, hi);
  } else {
    uint4 o;
    o.x = s[0] | (s[1] << 16); o.y = s[2] | (s[3] << 16); o.z = s[4] | (s[5] << 16); o.w = s[6] | (s[7] << 16);
    *reinterpret_cast<uint4*>(Tout + (size_t)wi * kWinBytes + k0) = o;
  }
}

__global__ void __launch_bounds__(256)
marker_resolve_kernel(const SplitStreamDesc* __restrict__ streams, uint32_t nstreams, const uint16_t* __restrict__ sym_all,
                      const uint64_t* __restrict__ off_all, const uint8_t* __restrict__ wbuf_all,
                      uint32_t* __restrict__ partial) {
  __shared__ __align__(16) uint8_t win[kWinBytes];
  __shared__ uint32_t sa[8], sw[8], sc0;
  const SplitStreamDesc& sd = streams[stream_of(nstreams, blockIdx.x, [&](uint32_t i) { return streams[i].blk_base; })];
  const uint16_t* const sym  = sym_all + sd.sym_base;
  const uint64_t* const off  = off_all + sd.off_base;
  const uint8_t* const  wbuf = wbuf_all + (size_t)sd.win_base * kWinBytes;
  uint8_t* const        dst  = sd.dst;
  const uint32_t nchunk = sd.nchunk;
  const uint64_t total  = sd.total;
  const uint64_t b0 = (uint64_t)(blockIdx.x - sd.blk_base) * kResolvePerBlock;
  const uint64_t b1 = b0 + kResolvePerBlock < total ? b0 + kResolvePerBlock : total;
  if (threadIdx.x == 0) {          // the chunk that holds this block's first position
    uint32_t lo = 0, hi = nchunk;
    while (hi - lo > 1u) { const uint32_t mid = (lo + hi) >> 1; if (off[mid] <= b0) lo = mid; else hi = mid; }
    sc0 = lo;
  }
  __syncthreads();
  uint32_t   c = sc0;
  uint32_t   A = 0;
  uint64_t   W = 0;
  const bool dst_aligned = ((uintptr_t)dst & 7u) == 0u;
  uint64_t   pos = b0;
#pragma unroll 1
  while (pos < b1) {               // one round per chunk the block touches (block-uniform)
    while (c + 1u < nchunk && off[c + 1u] <= pos) c++;
    const uint64_t cend = c + 1u < nchunk ? off[c + 1u] : total;
    const uint64_t send = cend < b1 ? cend : b1;
    __syncthreads();
    if (c >= 1u) {
      const uint4* wsrc = reinterpret_cast<const uint4*>(wbuf + (size_t)(c - 1u) * kWinBytes);
      for (uint32_t i = threadIdx.x; i < kWinBytes / 16u; i += 256u) reinterpret_cast<uint4*>(win)[i] = __ldg(wsrc + i);
    }
    __syncthreads();
#pragma unroll 1
    for (uint64_t i0 = (pos & ~7ull) + threadIdx.x * 8u; i0 < send; i0 += 2048u) {
      const uint32_t j0 = (uint32_t)(i0 - (b0 & ~7ull));     // relative to the block's aligned start
      uint32_t s[8];
      if (i0 >= pos && i0 + 8u <= send) {
        const uint4 v = __ldcs(reinterpret_cast<const uint4*>(sym + i0));
        s[0] = v.x & 0xFFFFu; s[1] = v.x >> 16; s[2] = v.y & 0xFFFFu; s[3] = v.y >> 16;
        s[4] = v.z & 0xFFFFu; s[5] = v.z >> 16; s[6] = v.w & 0xFFFFu; s[7] = v.w >> 16;
        if ((v.x | v.y | v.z | v.w) & 0x80008000u) {
#pragma unroll
          for (int k = 0; k < 8; k++) if (s[k] & 0x8000u) s[k] = win[s[k] & 0x7FFFu];
        }
        const uint32_t lo = s[0] | (s[1] << 8) | (s[2] << 16) | (s[3] << 24);
        const uint32_t hi = s[4] | (s[5] << 8) | (s[6] << 16) | (s[7] << 24);
        if (dst_aligned) {
          __stcs(reinterpret_cast<uint2*>(dst + i0), make_uint2(lo, hi));
        } else {
#pragma unroll
          for (int k = 0; k < 8; k++) dst[i0 + k] = (uint8_t)s[k];
        }
        const uint32_t sum = __dp4a(lo, 0x01010101u, __dp4a(hi, 0x01010101u, 0u));
        const uint32_t q   = __dp4a(lo, 0x03020100u, __dp4a(hi, 0x07060504u, 0u));
        A += sum;
        W += (uint64_t)j0 * sum + q;
      } else {                     // a group cut by the chunk's or the block's ends
#pragma unroll 1
        for (uint32_t k = 0; k < 8u; k++) {
          const uint64_t i = i0 + k;
          if (i < pos || i >= send) continue;
          uint32_t x = sym[i];
          if (x & 0x8000u) x = win[x & 0x7FFFu];
          dst[i] = (uint8_t)x;
          A += x;
          W += (uint64_t)(j0 + k) * x;
        }
      }
    }
    pos = send;
  }
  for (int o = 16; o > 0; o >>= 1) {
    A += __shfl_xor_sync(0xFFFFFFFFu, A, o);
    W += __shfl_xor_sync(0xFFFFFFFFu, W, o);
  }
  if ((threadIdx.x & 31) == 0) { sa[threadIdx.x >> 5] = A; sw[threadIdx.x >> 5] = (uint32_t)(W % kAdlerMod); }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t ta = 0, tw = 0;
    for (int k = 0; k < 8; k++) { ta += sa[k]; tw += sw[k]; }
    ta %= kAdlerMod;
    // sum of i * x over the block = (origin mod M) * sum x + sum j * x, j counted from the aligned origin
    partial[2 * blockIdx.x]     = ta;
    partial[2 * blockIdx.x + 1] = (uint32_t)((((b0 & ~7ull) % kAdlerMod) * ta + tw) % kAdlerMod);
  }
}

#endif  // __CUDACC__

}  // namespace defl
