// simt_cuda.cuh -- the warp-level primitives the inflate core uses, as sm_100a intrinsics.
// (tests/emu/simt_emu.h provides the same names on the CPU for a lock-step emulator that
//  exists only to debug the decode logic without a GPU; it is never part of the product.)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define DEFL_DEV __device__ __forceinline__
#define DEFL_DEV_NOINLINE __device__ __noinline__

namespace simt {

DEFL_DEV void     syncwarp(unsigned mask) { __syncwarp(mask); }
DEFL_DEV uint32_t shfl(unsigned mask, uint32_t v, int src_lane) { return __shfl_sync(mask, v, src_lane); }
DEFL_DEV unsigned match_any(unsigned mask, uint32_t v) { return __match_any_sync(mask, v); }
DEFL_DEV unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
DEFL_DEV uint32_t atomic_add_shared(uint32_t* p, uint32_t v) { return atomicAdd(p, v); }
DEFL_DEV uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) { return __funnelshift_r(lo, hi, sh); }
DEFL_DEV uint32_t brev(uint32_t v) { return __brev(v); }
DEFL_DEV int      popc(uint32_t v) { return __popc(v); }
DEFL_DEV uint32_t bfe(uint32_t v, uint32_t pos, uint32_t len) {
  uint32_t r;
  asm("bfe.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(pos), "r"(len));
  return r;
}
DEFL_DEV uint32_t dp4a(uint32_t a, uint32_t b, uint32_t c) { return __dp4a(a, b, c); }
DEFL_DEV uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t s) { return __byte_perm(a, b, s); }

// compressed input: read-only for the whole kernel -> non-coherent path
DEFL_DEV uint32_t ld_in32(const uint32_t* p) { return __ldg(p); }
DEFL_DEV uint8_t  ld_in8(const uint8_t* p) { return __ldg(p); }
DEFL_DEV uint4    ld_in128(const uint4* p) { return __ldg(p); }
DEFL_DEV void     prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }


// ---- used by the lane-decoder pipeline (pipeline.cuh) -----------------------------------------------
DEFL_DEV unsigned ballot(unsigned mask, bool p) { return __ballot_sync(mask, p); }
DEFL_DEV bool     all(unsigned mask, bool p) { return __all_sync(mask, p); }
DEFL_DEV uint32_t shfl_up(unsigned mask, uint32_t v, unsigned d) { return __shfl_up_sync(mask, v, d); }
DEFL_DEV int      ffs(uint32_t v) { return __ffs(v); }
DEFL_DEV uint32_t reduce_or(unsigned mask, uint32_t v) { return __reduce_or_sync(mask, v); }
DEFL_DEV unsigned lanemask_le() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_le;" : "=r"(m));
  return m;
}
// words shared between warps of a CTA: volatile accesses + block-scope fences
DEFL_DEV uint32_t vload(const uint32_t* p) { return *reinterpret_cast<const volatile uint32_t*>(p); }
DEFL_DEV void     vstore(uint32_t* p, uint32_t v) { *reinterpret_cast<volatile uint32_t*>(p) = v; }
DEFL_DEV void     fence_block() { __threadfence_block(); }
DEFL_DEV void     backoff(unsigned ns) { __nanosleep(ns); }
DEFL_DEV uint32_t atomic_add_global(unsigned int* p, unsigned int v) { return atomicAdd(p, v); }
DEFL_DEV long long ticks() { return clock64(); }
// decoded output read back as a match source: written earlier by this SM, cached at L2 only so the
// L1 stays with the compressed input lines the decoder lanes stream through
DEFL_DEV uint32_t ld_out8(const uint8_t* p) { return __ldcg(p); }
DEFL_DEV uint32_t ld_out(const uint8_t* p) { return __ldcg(p); }
DEFL_DEV uint32_t ld_out(const uint16_t* p) { return __ldcg(p); }
DEFL_DEV uint32_t ld_outw(const uint32_t* p) { return __ldcg(p); }      // the aligned word around a match source

// ---- used by the warp-per-stream decoder (warp_decode.cuh) ------------------------------------------
DEFL_DEV bool     any(unsigned mask, bool p) { return __any_sync(mask, p); }
DEFL_DEV int      clz(uint32_t v) { return __clz((int)v); }
DEFL_DEV uint32_t reduce_add(unsigned mask, uint32_t v) { return __reduce_add_sync(mask, v); }
DEFL_DEV uint32_t reduce_min(unsigned mask, uint32_t v) { return __reduce_min_sync(mask, v); }
DEFL_DEV uint32_t reduce_max(unsigned mask, uint32_t v) { return __reduce_max_sync(mask, v); }
// tokens written to the global scratch by one lane and read back by the whole warp: L2 only
DEFL_DEV uint32_t ld_tok(const uint32_t* p) { return __ldcg(p); }

// global -> shared without a register in between (LDGSTS): the copy is in flight while the lane decodes on, and
// nothing waits for it until cp_async_wait<N>() ("all but my N newest groups have landed").  valid = false zero-fills.
DEFL_DEV void cp_async4(void* smem_dst, const void* gsrc, bool valid) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  const size_t   g = __cvta_generic_to_global(gsrc);
  const int      n = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(s), "l"(g), "r"(n) : "memory");
}
DEFL_DEV void cp_async16(void* smem_dst, const void* gsrc, bool valid) {      // both 16-byte aligned
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  const size_t   g = __cvta_generic_to_global(gsrc);
  const int      n = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(g), "r"(n) : "memory");
}
DEFL_DEV void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
DEFL_DEV void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- used by the v4 warp decoder (warp_decode4.cuh) ---------------------------------------------------
// 64-bit tokens written by one lane, read back by the whole warp: L2 only
DEFL_DEV uint2 ld_tok2(const uint2* p) { return __ldcg(p); }
// global -> shared, 16 bytes, of which only the first `nbytes` (0..16) are read; the rest is zero-filled
DEFL_DEV void cp_async16_n(void* smem_dst, const void* gsrc, uint32_t nbytes) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  const size_t   g = __cvta_generic_to_global(gsrc);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(g), "r"(nbytes) : "memory");
}
DEFL_DEV uint4 ld_out128(const uint4* p) { return __ldcg(p); }
DEFL_DEV void  atomic_min_shared(uint32_t* p, uint32_t v) { atomicMin(p, v); }

}  // namespace simt
