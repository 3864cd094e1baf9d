// tests/emu/simt_emu.h -- DEBUG AID, TEST TREE ONLY.
// Lock-step CPU emulation of the few warp primitives defl_b200/csrc/inflate_core.cuh uses, so the
// decode *logic* (bit reader, table construction, copies, Adler) can be exercised in the CPU
// container, where there is no GPU.  G lanes of one group run as G host threads that meet at a
// barrier wherever the kernel calls __syncwarp / __shfl_sync / __match_any_sync.
// It is not a decode path of the product: nothing under defl_b200/ or include/ references it,
// it is built only by tests/test_emu_logic.py, and the GPU tests never load it.
#pragma once
#include <pthread.h>
#include <sched.h>
#include <stdint.h>
#include <string.h>
#include <algorithm>
using std::max;

#define DEFL_DEV inline
#define DEFL_DEV_NOINLINE inline

namespace simt {

struct Group {
  int               G;
  pthread_barrier_t bar;
  uint32_t          xchg[32];
};
static thread_local Group* t_group = nullptr;
static thread_local int    t_lane  = 0;

inline void barrier() { if (t_group->G > 1) pthread_barrier_wait(&t_group->bar); }

inline void syncwarp(unsigned) { barrier(); }
inline uint32_t shfl(unsigned, uint32_t v, int src) {
  if (t_group->G == 1) return v;
  t_group->xchg[t_lane] = v;
  barrier();
  uint32_t r = t_group->xchg[src];
  barrier();
  return r;
}
inline unsigned match_any(unsigned, uint32_t v) {
  if (t_group->G == 1) return 1u;
  t_group->xchg[t_lane] = v;
  barrier();
  unsigned m = 0;
  for (int i = 0; i < t_group->G; i++)
    if (t_group->xchg[i] == v) m |= 1u << i;
  barrier();
  return m;
}
inline unsigned lanemask_lt() { return (1u << t_lane) - 1u; }
inline uint32_t atomic_add_shared(uint32_t* p, uint32_t v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
inline uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t sh) {
  sh &= 31u;
  return sh ? (lo >> sh) | (hi << (32u - sh)) : lo;
}
inline uint32_t brev(uint32_t v) {
  uint32_t r = 0;
  for (int i = 0; i < 32; i++) { r = (r << 1) | (v & 1u); v >>= 1; }
  return r;
}
inline int popc(uint32_t v) { return __builtin_popcount(v); }
inline uint32_t bfe(uint32_t v, uint32_t pos, uint32_t len) {
  if (len == 0) return 0;
  if (pos > 31) return 0;
  uint64_t x = (uint64_t)v >> pos;
  return len >= 32 ? (uint32_t)x : (uint32_t)(x & ((1ull << len) - 1ull));
}
inline uint32_t dp4a(uint32_t a, uint32_t b, uint32_t c) {
  for (int i = 0; i < 4; i++) c += ((a >> (8 * i)) & 0xFFu) * ((b >> (8 * i)) & 0xFFu);
  return c;
}
inline uint32_t ld_in32(const uint32_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
inline uint8_t  ld_in8(const uint8_t* p) { return *p; }
inline void     prefetch_l2(const void*) {}


// ---- pipeline primitives: one Group per warp, several warps (thread sets) per emulated CTA -----------
inline unsigned ballot(unsigned, bool p) {
  t_group->xchg[t_lane] = p ? 1u : 0u;
  barrier();
  unsigned m = 0;
  for (int i = 0; i < t_group->G; i++) m |= t_group->xchg[i] << i;
  barrier();
  return m;
}
inline bool all(unsigned mk, bool p) { return ballot(mk, p) == (t_group->G == 32 ? 0xFFFFFFFFu : ((1u << t_group->G) - 1u)); }
inline uint32_t shfl_up(unsigned, uint32_t v, unsigned d) {
  t_group->xchg[t_lane] = v;
  barrier();
  uint32_t r = (unsigned)t_lane >= d ? t_group->xchg[t_lane - d] : v;
  barrier();
  return r;
}
inline int ffs(uint32_t v) { return __builtin_ffs((int)v); }
inline uint32_t reduce_or(unsigned, uint32_t v) {
  t_group->xchg[t_lane] = v;
  barrier();
  uint32_t r = 0;
  for (int i = 0; i < t_group->G; i++) r |= t_group->xchg[i];
  barrier();
  return r;
}
inline unsigned lanemask_le() { return t_lane == 31 ? 0xFFFFFFFFu : ((2u << t_lane) - 1u); }
inline uint32_t vload(const uint32_t* p) { return __atomic_load_n(p, __ATOMIC_ACQUIRE); }
inline void vstore(uint32_t* p, uint32_t v) { __atomic_store_n(p, v, __ATOMIC_RELEASE); }
inline void fence_block() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
inline void backoff(unsigned) { sched_yield(); }
inline uint32_t atomic_add_global(unsigned int* p, unsigned int v) { return __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
inline long long ticks() { return 0; }
inline uint32_t ld_out8(const uint8_t* p) { return *(const volatile uint8_t*)p; }
inline uint32_t ld_out(const uint8_t* p) { return *(const volatile uint8_t*)p; }
inline uint32_t ld_out(const uint16_t* p) { return *(const volatile uint16_t*)p; }
inline uint32_t ld_outw(const uint32_t* p) { return *(const volatile uint32_t*)p; }
inline bool any(unsigned mk, bool p) { return ballot(mk, p) != 0u; }
inline int clz(uint32_t v) { return v ? __builtin_clz(v) : 32; }
inline uint32_t reduce_add(unsigned, uint32_t v) {
  t_group->xchg[t_lane] = v;
  barrier();
  uint32_t r = 0;
  for (int i = 0; i < t_group->G; i++) r += t_group->xchg[i];
  barrier();
  return r;
}
inline uint32_t reduce_min(unsigned, uint32_t v) {
  t_group->xchg[t_lane] = v;
  barrier();
  uint32_t r = 0xFFFFFFFFu;
  for (int i = 0; i < t_group->G; i++) r = t_group->xchg[i] < r ? t_group->xchg[i] : r;
  barrier();
  return r;
}
inline uint32_t reduce_max(unsigned, uint32_t v) {
  t_group->xchg[t_lane] = v;
  barrier();
  uint32_t r = 0;
  for (int i = 0; i < t_group->G; i++) r = t_group->xchg[i] > r ? t_group->xchg[i] : r;
  barrier();
  return r;
}
inline uint32_t ld_tok(const uint32_t* p) { return *(const volatile uint32_t*)p; }

inline void cp_async4(void* smem_dst, const void* gsrc, bool valid) {
  if (valid) memcpy(smem_dst, gsrc, 4); else memset(smem_dst, 0, 4);
}
inline void cp_async16(void* smem_dst, const void* gsrc, bool valid) {
  if (valid) memcpy(smem_dst, gsrc, 16); else memset(smem_dst, 0, 16);
}
inline void cp_async_commit() {}
template <int N> inline void cp_async_wait() {}

inline uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t s) {
  const uint64_t v = ((uint64_t)b << 32) | a;
  uint32_t r = 0;
  for (int i = 0; i < 4; i++) r |= (uint32_t)((v >> (8 * ((s >> (4 * i)) & 7u))) & 0xFFu) << (8 * i);
  return r;
}
}  // namespace simt
struct uint2 { uint32_t x, y; };
struct uint4 { uint32_t x, y, z, w; };
inline uint2 make_uint2(uint32_t x, uint32_t y) { uint2 r; r.x = x; r.y = y; return r; }
inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { uint4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
namespace simt {
inline uint2 ld_tok2(const uint2* p) { uint2 r; r.x = *(const volatile uint32_t*)&p->x; r.y = *(const volatile uint32_t*)&p->y; return r; }
inline void cp_async16_n(void* smem_dst, const void* gsrc, uint32_t nbytes) {
  memset(smem_dst, 0, 16);
  if (nbytes) memcpy(smem_dst, gsrc, nbytes > 16 ? 16 : nbytes);
}
inline uint4 ld_out128(const uint4* p) { uint4 r; memcpy(&r, p, 16); return r; }
inline void atomic_min_shared(uint32_t* p, uint32_t v) {
  uint32_t o = __atomic_load_n(p, __ATOMIC_RELAXED);
  while (v < o && !__atomic_compare_exchange_n(p, &o, v, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) {}
}

}  // namespace simt
