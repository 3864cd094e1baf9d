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

enum : uint32_t { CH_LINKED = 0, CH_FINAL = 1, CH_FAIL = 2 };   // CH_FAIL | reason << 8

struct SplitChunkOut {
  uint64_t end_bit;     // stream-relative bit where the chunk stopped (a block boundary unless it failed)
  uint32_t status;      // CH_*
  uint32_t link;        // CH_LINKED: index of the candidate at end_bit
  uint32_t out_bytes;   // bytes the chunk produces
  uint32_t pad;
};

struct SplitJob {
  uint32_t cand;        // index into cand[]: where the chunk starts
  uint32_t cap;         // DECODE: exactly the bytes the COUNT pass reported
  uint64_t out_off;     // DECODE: offset of the chunk's first symbol
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
                                 uint32_t hist, SplitChunkOut& out) {
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

  uint32_t pos = 0, reason = R_NONE, next = me + 1u, status = CH_FAIL;
  bool     fixed_ready = false;
#pragma unroll 1
  for (;;) {
    const uint64_t q = br.bitpos() - base_bit;
    if (q != start) {
      while (next < ncand && cand[next] < q) next++;
      if (next < ncand && cand[next] == q) { status = CH_LINKED; break; }
    }
    if ((int64_t)(end_bit - br.bitpos()) < 3) { reason = R_TRUNCATED; break; }
    const uint32_t hdr    = br.peek();
    const uint32_t bfinal = hdr & 1u, btype = (hdr >> 1) & 3u;
    br.skip(3); br.norm();
    if (btype == 0) {
      reason = stored_block_sym<SYM, COUNT>(lane, br, src, srclen, base_bit, dst, pos, cap);
      if (reason) break;
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
        st = decode_window<SYM, COUNT>(sm, sc, lane, br.w, br.wend, end_bit, cur, dst, cap, pos, hist);
      } while (st == LS_OK);
      if (st != LS_EOB) { reason = st >> 8; break; }
      br.seek(cur);
    }
    if (bfinal) { status = CH_FINAL; break; }
  }
  out.end_bit   = br.bitpos() - base_bit;
  out.status    = reason ? (CH_FAIL | (reason << 8)) : status;
  out.link      = next;
  out.out_bytes = pos;
  out.pad       = 0;
}

#ifdef __CUDACC__
template <bool COUNT, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 32 / WARPS)
split_chunk_kernel(const uint8_t* __restrict__ src, uint32_t srclen, const uint64_t* __restrict__ cand,
                   uint32_t ncand, const SplitJob* __restrict__ jobs, uint32_t njobs, uint16_t* __restrict__ sym,
                   SplitChunkOut* __restrict__ outs, unsigned int* __restrict__ ticket,
                   unsigned char* __restrict__ scratch) {
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
    SplitChunkOut  o;
    // the chunk that starts the stream has no unknown window before it
    inflate_chunk_warp<uint16_t, COUNT>(sm, sc, lane, src, srclen, cand, ncand, job.cand,
                                        COUNT ? nullptr : sym + job.out_off, COUNT ? 0xFFFFFFFFu : job.cap,
                                        job.cand == 0u ? 0u : 32768u, o);
    if (lane == 0) outs[k] = o;
    simt::syncwarp(kFull);
  }
}

#endif  // __CUDACC__

// ---- F: block-start search -------------------------------------------------------------------------------
constexpr uint32_t kGranuleBytes = 32768;             // one search result per this much input
constexpr uint64_t kNoCandidate  = ~0ull;

// Cheap per-position filters, one bit position per lane: a non-reserved dynamic header whose
// code-length code is Kraft-complete (what every deflate encoder emits).  x0/x1/x2 = the 96 bits at p.
DEFL_DEV bool header_plausible(uint32_t x0, uint32_t x1, uint32_t x2) {
  if (((x0 >> 1) & 3u) != 2u) return false;                       // BTYPE = dynamic
  if (((x0 >> 3) & 31u) > 29u || ((x0 >> 8) & 31u) > 29u) return false;   // HLIT <= 286, HDIST <= 30
  const uint32_t n = ((x0 >> 13) & 15u) + 4u;
  uint64_t y = (uint64_t)(x0 >> 17) | ((uint64_t)x1 << 15) | ((uint64_t)x2 << 47);
  uint32_t kraft = 0;
#pragma unroll 1
  for (uint32_t i = 0; i < n; i++) {
    const uint32_t l = (uint32_t)y & 7u;
    y >>= 3;
    kraft += l ? (128u >> l) : 0u;
  }
  return kraft == 128u;
}

// One warp per granule [g * 32 KiB, (g + 1) * 32 KiB) of the stream body: the first position that passes
// the filters AND parses as a whole dynamic header with complete literal/length code, an end-of-block
// code and a non-empty distance code goes to found[g].
#ifdef __CUDACC__
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
find_block_starts_kernel(const uint8_t* __restrict__ src, uint32_t srclen, uint32_t body, uint32_t ngran,
                         uint64_t* __restrict__ found) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WarpSmem& sm = reinterpret_cast<WarpSmem*>(smem_raw)[warp];
  const uint32_t g = blockIdx.x * WARPS + warp;
  if (g >= ngran) return;
  Lane<32> L;
  L.gmask = kFull; L.lane = lane; L.wlane = lane;
  const uint32_t mis      = (uint32_t)((uintptr_t)src & 3u);
  const uint64_t base_bit = (uint64_t)mis * 8u;
  const uint64_t end_bit  = base_bit + (uint64_t)srclen * 8u;
  const uint32_t* words   = reinterpret_cast<const uint32_t*>((uintptr_t)src - mis);
  const uint32_t  wend    = (uint32_t)(((uint64_t)mis + srclen + 3u) >> 2);
  const uint64_t p0 = base_bit + ((uint64_t)body + (uint64_t)g * kGranuleBytes) * 8u;
  uint64_t       p1 = p0 + (uint64_t)kGranuleBytes * 8u;
  if (p1 + 96u > end_bit) p1 = end_bit > 96u ? end_bit - 96u : 0u;     // a header needs room
  uint64_t result = kNoCandidate;
#pragma unroll 1
  for (uint64_t pb = p0; pb < p1 && result == kNoCandidate; pb += 32u) {
    const uint64_t p  = pb + (uint64_t)lane;
    const uint32_t wi = (uint32_t)(p >> 5), sh = (uint32_t)(p & 31u);
    const uint32_t w0 = wi < wend ? simt::ld_in32(words + wi) : 0u;
    const uint32_t w1 = wi + 1 < wend ? simt::ld_in32(words + wi + 1) : 0u;
    const uint32_t w2 = wi + 2 < wend ? simt::ld_in32(words + wi + 2) : 0u;
    const uint32_t w3 = wi + 3 < wend ? simt::ld_in32(words + wi + 3) : 0u;
    const bool ok = p < p1 && header_plausible(simt::funnel_r(w0, w1, sh), simt::funnel_r(w1, w2, sh),
                                               simt::funnel_r(w2, w3, sh));
    unsigned m = simt::ballot(kFull, ok);
#pragma unroll 1
    while (m) {
      const int l = simt::ffs(m) - 1;
      m &= m - 1u;
      const uint64_t pc = pb + (uint64_t)l;
      BitReader br;
      br.w = words; br.wend = wend;
      br.seek(pc + 3u);
      bool fixed_ready = false;
      simt::syncwarp(kFull);
      const uint32_t reason = huffman_block_setup<32>(L, sm.lit, sm.dist, sm.s, br, end_bit, 2u, fixed_ready);
      simt::syncwarp(kFull);
      if (reason) continue;
      BitReader hr;
      hr.w = words; hr.wend = wend;
      hr.seek(pc);
      const uint32_t hv = hr.peek();
      const uint32_t hlit = ((hv >> 3) & 31u) + 257u, hdist = ((hv >> 8) & 31u) + 1u;
      uint32_t kl = 0, kd = 0;
      for (uint32_t i = (uint32_t)lane; i < hlit; i += 32u) { const uint32_t len = sm.s.lens[i]; kl += len ? (32768u >> len) : 0u; }
      for (uint32_t i = (uint32_t)lane; i < hdist; i += 32u) { const uint32_t len = sm.s.lens[hlit + i]; kd += len ? (32768u >> len) : 0u; }
      kl = simt::reduce_add(kFull, kl);
      kd = simt::reduce_add(kFull, kd);
      const bool good = kl == 32768u && sm.s.lens[256] != 0u && kd != 0u;
      simt::syncwarp(kFull);
      if (good) { result = pc - base_bit; break; }
    }
  }
  if (lane == 0) found[g] = result;
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
__global__ void __launch_bounds__(256)
marker_resolve_kernel(const uint16_t* __restrict__ sym, const uint64_t* __restrict__ off, uint32_t nchunk,
                      uint8_t* __restrict__ dst, uint64_t total, uint32_t* __restrict__ partial) {
  __shared__ uint32_t sa[8], sw[8];
  const uint64_t per_block = 256u * 16u;
  const uint64_t b0 = (uint64_t)blockIdx.x * per_block;
  uint32_t A = 0;
  uint64_t W = 0;
  // the chunk that holds this block's first position (blocks are small next to chunks: walk forward from it)
  uint32_t c = 0;
  {
    uint32_t lo = 0, hi = nchunk;
    while (hi - lo > 1u) { const uint32_t mid = (lo + hi) >> 1; if (off[mid] <= b0) lo = mid; else hi = mid; }
    c = lo;
  }
#pragma unroll 1
  for (int it = 0; it < 16; it++) {
    const uint64_t i = b0 + (uint64_t)it * 256u + threadIdx.x;
    if (i < total) {
      uint32_t cc = c;
      while (cc + 1u < nchunk && off[cc + 1u] <= i) cc++;
      const uint32_t s = resolve_marker(sym, off, cc, i);
      dst[i] = (uint8_t)s;
      A += s;
      W += (uint64_t)(i % kAdlerMod) * s;
    }
  }
  uint32_t w32 = (uint32_t)(W % kAdlerMod);
  for (int o = 16; o > 0; o >>= 1) {
    A += __shfl_xor_sync(0xFFFFFFFFu, A, o);
    w32 += __shfl_xor_sync(0xFFFFFFFFu, w32, o);
  }
  if ((threadIdx.x & 31) == 0) { sa[threadIdx.x >> 5] = A; sw[threadIdx.x >> 5] = w32 % kAdlerMod; }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t ta = 0, tw = 0;
    for (int k = 0; k < 8; k++) { ta += sa[k]; tw = (tw + sw[k]) % kAdlerMod; }
    partial[2 * blockIdx.x]     = ta % kAdlerMod;
    partial[2 * blockIdx.x + 1] = tw;
  }
}

#endif  // __CUDACC__

}  // namespace defl
