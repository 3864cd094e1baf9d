// tests/emu/emu_main.cpp -- DEBUG AID, TEST TREE ONLY (see simt_emu.h).
#include "simt_emu.h"
#include "../../defl_b200/csrc/inflate_core.cuh"

#include <stdlib.h>
#include <vector>

namespace {

struct Job {
  simt::Group*      grp;
  int               lane;
  defl::GroupSmem*  sm;
  const uint8_t*    src;
  uint32_t          srclen;
  uint8_t*          dst;
  uint32_t          cap;
  int               flags;
  defl::StreamOut*  out;
};

template <int G>
void* lane_main(void* arg) {
  Job* j = static_cast<Job*>(arg);
  simt::t_group = j->grp;
  simt::t_lane  = j->lane;
  defl::Lane<G> L;
  L.gmask = G == 32 ? 0xFFFFFFFFu : ((1u << G) - 1u);
  L.lane = j->lane;
  L.wlane = j->lane;
  defl::StreamOut o;
  defl::inflate_stream<G>(L, *j->sm, j->src, j->srclen, j->dst, j->cap, j->flags, o);
  if (j->lane == 0) *j->out = o;
  return nullptr;
}

template <int G>
int run(const uint8_t* src, uint32_t srclen, uint8_t* dst, uint32_t cap, int flags, defl::StreamOut* out) {
  simt::Group grp;
  grp.G = G;
  if (G > 1) pthread_barrier_init(&grp.bar, nullptr, G);
  defl::GroupSmem* sm = static_cast<defl::GroupSmem*>(aligned_alloc(64, (sizeof(defl::GroupSmem) + 63) & ~63ul));
  memset(sm, 0xA5, sizeof(*sm));  // shared memory starts as garbage on the GPU too
  std::vector<Job>       jobs(G);
  std::vector<pthread_t> th(G);
  for (int i = 0; i < G; i++) jobs[i] = Job{&grp, i, sm, src, srclen, dst, cap, flags, out};
  for (int i = 1; i < G; i++) pthread_create(&th[i], nullptr, lane_main<G>, &jobs[i]);
  lane_main<G>(&jobs[0]);
  for (int i = 1; i < G; i++) pthread_join(th[i], nullptr);
  if (G > 1) pthread_barrier_destroy(&grp.bar);
  free(sm);
  return out->code;
}

}  // namespace

extern "C" int emu_inflate(int G, const uint8_t* src, uint32_t srclen, uint8_t* dst, uint32_t cap, int flags,
                           defl::StreamOut* out) {
  switch (G) {
    case 1:  return run<1>(src, srclen, dst, cap, flags, out);
    case 2:  return run<2>(src, srclen, dst, cap, flags, out);
    case 4:  return run<4>(src, srclen, dst, cap, flags, out);
    case 8:  return run<8>(src, srclen, dst, cap, flags, out);
    case 16: return run<16>(src, srclen, dst, cap, flags, out);
    case 32: return run<32>(src, srclen, dst, cap, flags, out);
  }
  return -99;
}
