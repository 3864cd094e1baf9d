// kernels.cuh -- __global__ entry points (sm_100a) around the decode core.
#pragma once
#include "simt_cuda.cuh"
#include "inflate_core.cuh"
#ifdef DEFL_LEGACY_ENGINES
#include "pipeline.cuh"
#endif
#include "warp_decode.cuh"
#include "warp_decode4.cuh"
#include "split_decode.cuh"
#include "../../include/defl/infl_batch.h"

namespace defl {

#ifdef DEFL_LEGACY_ENGINES
// K3: batch inflate.  Persistent CTAs; every group of G lanes pulls the next stream from a
// global ticket counter (streams are pre-ordered largest-first by the host so the tail is
// short), decodes it start to finish and writes one 32-byte result record.
template <int G, int GROUPS>
__global__ void __launch_bounds__(G * GROUPS)
inflate_batch_kernel(const infl_batch_item_t* __restrict__ items, const uint32_t* __restrict__ order,
                     uint32_t n, int flags, infl_batch_result_t* __restrict__ results,
                     unsigned int* __restrict__ ticket) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  GroupSmem& sm = reinterpret_cast<GroupSmem*>(smem_raw)[threadIdx.x / G];

  Lane<G> L;
  L.wlane = threadIdx.x & 31;
  L.lane  = threadIdx.x % G;
  L.gmask = (G == 32) ? 0xFFFFFFFFu : (((1u << G) - 1u) << (L.wlane & ~(G - 1)));

  for (;;) {
    uint32_t k = 0;
    if (L.lane == 0) k = atomicAdd(ticket, 1u);
    if (G > 1) k = simt::shfl(L.gmask, k, L.wlane & ~(G - 1));
    if (k >= n) break;
    const uint32_t idx = order ? order[k] : k;
    const infl_batch_item_t it = items[idx];

    StreamOut o;
    if (it.srclen > 0xFFFFFFFFull || it.dstlen > 0xFFFFFFFFull) {
      o.code = -1; o.adler = 1; o.outlen = 0; o.inpos = 0; o.adler_stored = 0; o.info = R_TRUNCATED << 8;
    } else {
      inflate_stream<G>(L, sm, static_cast<const uint8_t*>(it.src), (uint32_t)it.srclen,
                        static_cast<uint8_t*>(it.dst), (uint32_t)it.dstlen, flags, o);
    }
    if (L.lane == 0) {
      infl_batch_result_t r;
      r.code = o.code; r.adler = o.adler; r.outlen = o.outlen; r.inpos = o.inpos;
      r.adler_stored = o.adler_stored; r.info = o.info;
      results[idx] = r;
    }
    simt::syncwarp(L.gmask);
  }
}

// K3 (v2): the lane-decoder pipeline of pipeline.cuh.  Persistent CTAs of D decoder warps (SL stream
// slots each) and K worker warps; decoder lanes pull streams from the global ticket counter.
template <int D, int SL, int K>
__global__ void __launch_bounds__((D + K) * 32)
inflate_pipeline_kernel(const infl_batch_item_t* __restrict__ items, const uint32_t* __restrict__ order,
                        uint32_t n, int flags, infl_batch_result_t* __restrict__ results,
                        unsigned int* __restrict__ ticket) {
  static_assert(D * SL <= 32 * K && D + K <= 32, "every worker polls at most 32 slots; a CTA has at most 32 warps");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Slot*       slots = reinterpret_cast<Slot*>(smem_raw);
  WorkerSmem* wsm   = reinterpret_cast<WorkerSmem*>(smem_raw + sizeof(Slot) * D * SL);
  uint32_t*   live  = reinterpret_cast<uint32_t*>(wsm + K);

  for (int s = threadIdx.x; s < D * SL; s += blockDim.x) {
    Slot& sl = slots[s];
    sl.tail = 0; sl.req = REQ_NONE; sl.req_bit_lo = 0; sl.req_bit_hi = 0; sl.req_reason = 0; sl.stream = 0;
    sl.head = 0; sl.ack = ACK_NONE; sl.ack_bit_lo = 0; sl.ack_bit_hi = 0; sl.ack_pos = 0; sl.ack_bfinal = 0;
    sl.wpos = 0; sl.fixed_ready = 0; sl.dst_lo = 0; sl.dst_hi = 0;
  }
  if (threadIdx.x == 0) *live = D;
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < D) {
    decoder_warp(slots + warp * SL, SL, lane, items, order, n, ticket);
    __syncwarp();
    if (lane == 0) {
      __threadfence_block();
      atomicSub(live, 1u);
    }
  } else {
    worker_warp(slots, D * SL, warp - D, K, wsm[warp - D], lane, live, items, flags, results);
  }
}

template <int D, int SL, int K>
constexpr size_t pipeline_smem_bytes() { return sizeof(Slot) * D * SL + sizeof(WorkerSmem) * K + 16; }

#endif  // DEFL_LEGACY_ENGINES

// K3 (v3, default): one warp per stream, the 32 lanes decode 32 sub-sequences of the current block at
// once (warp_decode.cuh).  Persistent CTAs of WARPS warps; each warp owns a slice of the global token
// scratch and pulls streams from the ticket counter.
template <int WARPS>
#ifndef DEFL_WARP_CTAS_PER_SM
#define DEFL_WARP_CTAS_PER_SM 7      // x 4 warps: 28 resident warps per SM at 72 registers (8 -> 64 registers: the speculative loop recomputes what it cannot hold, 149 GB/s; 6 -> 80 registers: 191; profiles/README.md)
#endif
__global__ void __launch_bounds__(WARPS * 32, DEFL_WARP_CTAS_PER_SM)
inflate_warp_kernel(const infl_batch_item_t* __restrict__ items, const uint32_t* __restrict__ order,
                    uint32_t n, int flags, infl_batch_result_t* __restrict__ results,
                    unsigned int* __restrict__ ticket, unsigned char* __restrict__ scratch) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WarpSmem&         sm = reinterpret_cast<WarpSmem*>(smem_raw)[warp];
  const WarpScratch sc = scratch_of(scratch + ((size_t)blockIdx.x * WARPS + warp) * kWarpScratchBytes);

  for (;;) {
    uint32_t k = 0;
    if (lane == 0) k = atomicAdd(ticket, 1u);
    k = simt::shfl(kFull, k, 0);
    if (k >= n) break;
    const uint32_t idx = order ? order[k] : k;
    const infl_batch_item_t it = items[idx];
    StreamOut o;
    if (it.srclen > 0xFFFFFFFFull || it.dstlen > 0xFFFFFFFFull) {
      o.code = -1; o.adler = 1; o.outlen = 0; o.inpos = 0; o.adler_stored = 0; o.info = R_TRUNCATED << 8;
    } else {
      inflate_stream_warp(sm, sc, lane, static_cast<const uint8_t*>(it.src), (uint32_t)it.srclen,
                          static_cast<uint8_t*>(it.dst), (uint32_t)it.dstlen, flags, o);
    }
    if (lane == 0) {
      infl_batch_result_t r;
      r.code = o.code; r.adler = o.adler; r.outlen = o.outlen; r.inpos = o.inpos;
      r.adler_stored = o.adler_stored; r.info = o.info;
      results[idx] = r;
    }
    simt::syncwarp(kFull);
  }
}

// K3 (v4, default): one warp per stream, 64-bit "literals + match" tokens, shared-memory input rings and
// output stage (warp_decode4.cuh).  Persistent CTAs of ONE warp: its shared memory is then a static object at
// constant addresses (every table lookup is "LDS [R + imm]", nothing is rematerialised from threadIdx), and
// up to 32 such CTAs are resident per SM.  Each warp owns a slice of the global token scratch and pulls
// streams from the ticket counter.
#ifndef DEFL_WARP4_CTAS_PER_SM
#define DEFL_WARP4_CTAS_PER_SM 24
#endif
__global__ void __launch_bounds__(32, DEFL_WARP4_CTAS_PER_SM)
inflate_warp4_kernel(const infl_batch_item_t* __restrict__ items, const uint32_t* __restrict__ order,
                     uint32_t n, int flags, infl_batch_result_t* __restrict__ results,
                     unsigned int* __restrict__ ticket, unsigned char* __restrict__ scratch) {
  __shared__ __align__(16) v4::WarpSmem4 sm;
  const int lane = threadIdx.x;
  const v4::WarpScratch4 sc = v4::scratch4_of(scratch + (size_t)blockIdx.x * v4::kWarpScratchBytes);

  for (;;) {
    uint32_t k = 0;
    if (lane == 0) k = atomicAdd(ticket, 1u);
    k = simt::shfl(kFull, k, 0);
    if (k >= n) break;
    const uint32_t idx = order ? order[k] : k;
    const infl_batch_item_t it = items[idx];
    StreamOut o;
    if (it.srclen > 0xFFFFFFFFull || it.dstlen > 0xFFFFFFFFull) {
      o.code = -1; o.adler = 1; o.outlen = 0; o.inpos = 0; o.adler_stored = 0; o.info = R_TRUNCATED << 8;
    } else {
      v4::inflate_stream_warp4(sm, sc, lane, static_cast<const uint8_t*>(it.src), (uint32_t)it.srclen,
                               static_cast<uint8_t*>(it.dst), (uint32_t)it.dstlen, flags, o);
    }
    if (lane == 0) {
      infl_batch_result_t r;
      r.code = o.code; r.adler = o.adler; r.outlen = o.outlen; r.inpos = o.inpos;
      r.adler_stored = o.adler_stored; r.info = o.info;
      results[idx] = r;
    }
    simt::syncwarp(kFull);
  }
}

// infl_stream(): one stream whose input arrives in pieces, taken up at the last block boundary the previous call
// passed (reference: the resumable state machine of src/infl/stream.c:468-794).  One warp; input and output stay
// on the device between calls.
__global__ void __launch_bounds__(32)
inflate_resume_kernel(const uint8_t* __restrict__ src, uint32_t srclen, uint8_t* __restrict__ dst, uint32_t cap,
                      int flags, StreamResume* __restrict__ rs, infl_batch_result_t* __restrict__ result,
                      unsigned char* __restrict__ scratch) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpSmem&         sm = *reinterpret_cast<WarpSmem*>(smem_raw);
  const WarpScratch sc = scratch_of(scratch);
  const int lane = threadIdx.x;
  StreamOut o;
  inflate_stream_warp(sm, sc, lane, src, srclen, dst, cap, flags, o, rs);
  if (lane == 0) {
    infl_batch_result_t r;
    r.code = o.code; r.adler = o.adler; r.outlen = o.outlen; r.inpos = o.inpos;
    r.adler_stored = o.adler_stored; r.info = o.info;
    *result = r;
  }
}

// K2: descriptor-driven gather of input pieces into contiguous device memory (replaces the
// host-side join of infl_include, reference src/infl/mem.c:121-155, and the chunk walk of
// REFILL, infl.c:1121-1128).  One warp per piece per trip; 16-byte stores when the
// destination allows, byte-exact for any alignment and length (including 0 and 1).
__global__ void __launch_bounds__(256)
gather_pieces_kernel(const void* const* __restrict__ seg_src, const uint64_t* __restrict__ seg_len,
                     const uint64_t* __restrict__ seg_dst_off, uint64_t nseg,
                     uint8_t* __restrict__ out) {
  const uint32_t lane   = threadIdx.x & 31;
  const uint64_t warp   = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  // pieces are taken 32 at a time.  When all 32 are tiny (a PNG cut into 1-byte IDATs, test_files.c:300-307) every
  // lane copies one piece by itself; otherwise the warp copies them one after another, 16 bytes per lane and step.
  for (uint64_t s0 = warp * 32u; s0 < nseg; s0 += nwarps * 32u) {
    const uint64_t mine = s0 + lane;
    const uint64_t mylen = mine < nseg ? seg_len[mine] : 0u;
    if (__all_sync(0xFFFFFFFFu, mylen <= 16u)) {
      if (mylen) {
        const uint8_t* src = static_cast<const uint8_t*>(seg_src[mine]);
        uint8_t*       dst = out + seg_dst_off[mine];
        for (uint32_t j = 0; j < (uint32_t)mylen; j++) dst[j] = __ldg(src + j);
      }
      continue;
    }
   const uint64_t s1 = s0 + 32u < nseg ? s0 + 32u : nseg;
   for (uint64_t s = s0; s < s1; s++) {
    const uint8_t* src = static_cast<const uint8_t*>(seg_src[s]);
    uint8_t*       dst = out + seg_dst_off[s];
    uint64_t       n   = seg_len[s];
    uint64_t head = (16u - ((uintptr_t)dst & 15u)) & 15u;
    if (head > n) head = n;
    for (uint64_t j = lane; j < head; j += 32) dst[j] = __ldg(src + j);
    src += head; dst += head; n -= head;
    const uint64_t nvec = n >> 4;
    if (((uintptr_t)src & 15u) == 0) {
      const uint4* s4 = reinterpret_cast<const uint4*>(src);
      uint4*       d4 = reinterpret_cast<uint4*>(dst);
      for (uint64_t k = lane; k < nvec; k += 32) d4[k] = __ldg(s4 + k);
    } else if (((uintptr_t)src & 3u) == 0) {
      const uint32_t* s1 = reinterpret_cast<const uint32_t*>(src);
      uint4*          d4 = reinterpret_cast<uint4*>(dst);
      for (uint64_t k = lane; k < nvec; k += 32) {
        uint4 v;
        v.x = __ldg(s1 + 4 * k); v.y = __ldg(s1 + 4 * k + 1);
        v.z = __ldg(s1 + 4 * k + 2); v.w = __ldg(s1 + 4 * k + 3);
        d4[k] = v;
      }
    } else {
      const uint32_t  sh = (uint32_t)((uintptr_t)src & 3u) * 8u;
      const uint32_t* s1 = reinterpret_cast<const uint32_t*>((uintptr_t)src & ~(uintptr_t)3);
      uint4*          d4 = reinterpret_cast<uint4*>(dst);
      for (uint64_t k = lane; k < nvec; k += 32) {
        const uint32_t a = __ldg(s1 + 4 * k), b = __ldg(s1 + 4 * k + 1), c = __ldg(s1 + 4 * k + 2),
                       d = __ldg(s1 + 4 * k + 3), e = __ldg(s1 + 4 * k + 4);
        uint4 v;
        v.x = __funnelshift_r(a, b, sh); v.y = __funnelshift_r(b, c, sh);
        v.z = __funnelshift_r(c, d, sh); v.w = __funnelshift_r(d, e, sh);
        d4[k] = v;
      }
    }
    for (uint64_t j = (nvec << 4) + lane; j < n; j += 32) dst[j] = __ldg(src + j);
   }
  }
}

}  // namespace defl
