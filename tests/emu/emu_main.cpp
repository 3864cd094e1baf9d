// tests/emu/emu_main.cpp -- DEBUG AID, TEST TREE ONLY (see simt_emu.h).
#define DEFL_WATCHDOG_SPINS (1u << 21)
#include "simt_emu.h"
#include "../../defl_b200/csrc/inflate_core.cuh"
#include "../../defl_b200/csrc/pipeline.cuh"
#include "../../defl_b200/csrc/warp_decode.cuh"
#include "../../defl_b200/csrc/warp_decode4.cuh"

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

// ---- one emulated CTA of the lane-decoder pipeline: (D + K) warps of 32 host threads -----------------
namespace {

struct PipeCta {
  int                        D, SL, K;
  defl::Slot*                slots;
  defl::WorkerSmem*          wsm;
  uint32_t                   live;
  const infl_batch_item_t*   items;
  uint32_t                   n;
  int                        flags;
  infl_batch_result_t*       results;
  unsigned int               ticket;
  std::vector<simt::Group>   warps;
};

struct PipeThread { PipeCta* cta; int warp, lane; };

void* pipe_main(void* arg) {
  PipeThread* t = static_cast<PipeThread*>(arg);
  PipeCta*    c = t->cta;
  simt::t_group = &c->warps[t->warp];
  simt::t_lane  = t->lane;
  if (t->warp < c->D) {
    defl::decoder_warp(c->slots + t->warp * c->SL, c->SL, t->lane, c->items, nullptr, c->n, &c->ticket);
    simt::syncwarp(0xFFFFFFFFu);
    if (t->lane == 0) __atomic_fetch_sub(&c->live, 1u, __ATOMIC_SEQ_CST);
  } else {
    defl::worker_warp(c->slots, c->D * c->SL, t->warp - c->D, c->K, c->wsm[t->warp - c->D], t->lane, &c->live,
                      c->items, c->flags, c->results);
  }
  return nullptr;
}

}  // namespace

extern "C" int emu_pipeline(int D, int SL, int K, const infl_batch_item_t* items, uint32_t n, int flags,
                            infl_batch_result_t* results) {
  PipeCta c;
  c.D = D; c.SL = SL; c.K = K; c.items = items; c.n = n; c.flags = flags; c.results = results;
  c.ticket = 0; c.live = (uint32_t)D;
  const size_t nslots = (size_t)D * SL;
  c.slots = static_cast<defl::Slot*>(aligned_alloc(64, ((sizeof(defl::Slot) * nslots + 63) & ~63ul)));
  c.wsm   = static_cast<defl::WorkerSmem*>(aligned_alloc(64, ((sizeof(defl::WorkerSmem) * K + 63) & ~63ul)));
  memset(c.slots, 0xA5, sizeof(defl::Slot) * nslots);     // shared memory starts as garbage
  memset(c.wsm, 0x5A, sizeof(defl::WorkerSmem) * K);
  for (size_t s = 0; s < nslots; s++) {
    defl::Slot& sl = c.slots[s];
    sl.tail = 0; sl.req = 0; sl.req_bit_lo = 0; sl.req_bit_hi = 0; sl.req_reason = 0; sl.stream = 0;
    sl.head = 0; sl.ack = 0; sl.ack_bit_lo = 0; sl.ack_bit_hi = 0; sl.ack_pos = 0; sl.ack_bfinal = 0;
    sl.wpos = 0; sl.fixed_ready = 0; sl.dst_lo = 0; sl.dst_hi = 0;
  }
  const int nw = D + K;
  c.warps.resize(nw);
  for (int w = 0; w < nw; w++) { c.warps[w].G = 32; pthread_barrier_init(&c.warps[w].bar, nullptr, 32); }
  std::vector<PipeThread> th(nw * 32);
  std::vector<pthread_t>  tid(nw * 32);
  for (int i = 0; i < nw * 32; i++) {
    th[i] = PipeThread{&c, i / 32, i % 32};
    pthread_create(&tid[i], nullptr, pipe_main, &th[i]);
  }
  for (int i = 0; i < nw * 32; i++) pthread_join(tid[i], nullptr);
  for (int w = 0; w < nw; w++) pthread_barrier_destroy(&c.warps[w].bar);
  free(c.slots); free(c.wsm);
  return 0;
}

// ---- one emulated warp of the warp-per-stream decoder (warp_decode.cuh) --------------------------------
namespace {

struct WarpJob {
  simt::Group*     grp;
  int              lane;
  defl::WarpSmem*  sm;
  void*            scratch;
  const uint8_t*   src;
  uint32_t         srclen;
  uint8_t*         dst;
  uint32_t         cap;
  int              flags;
  defl::StreamOut* out;
};

void* warp_main(void* arg) {
  WarpJob* j = static_cast<WarpJob*>(arg);
  simt::t_group = j->grp;
  simt::t_lane  = j->lane;
  defl::StreamOut o;
  defl::inflate_stream_warp(*j->sm, defl::scratch_of(j->scratch), j->lane, j->src, j->srclen, j->dst, j->cap,
                            j->flags, o);
  if (j->lane == 0) *j->out = o;
  return nullptr;
}

}  // namespace

extern "C" int emu_warp_inflate(const uint8_t* src, uint32_t srclen, uint8_t* dst, uint32_t cap, int flags,
                                defl::StreamOut* out) {
  simt::Group grp;
  grp.G = 32;
  pthread_barrier_init(&grp.bar, nullptr, 32);
  defl::WarpSmem* sm = static_cast<defl::WarpSmem*>(aligned_alloc(64, (sizeof(defl::WarpSmem) + 63) & ~63ul));
  memset(sm, 0xA5, sizeof(*sm));
  void* scratch = aligned_alloc(64, defl::kWarpScratchBytes);
  memset(scratch, 0x5A, defl::kWarpScratchBytes);
  std::vector<WarpJob>   jobs(32);
  std::vector<pthread_t> th(32);
  for (int i = 0; i < 32; i++) jobs[i] = WarpJob{&grp, i, sm, scratch, src, srclen, dst, cap, flags, out};
  for (int i = 1; i < 32; i++) pthread_create(&th[i], nullptr, warp_main, &jobs[i]);
  warp_main(&jobs[0]);
  for (int i = 1; i < 32; i++) pthread_join(th[i], nullptr);
  pthread_barrier_destroy(&grp.bar);
  free(sm); free(scratch);
  return out->code;
}

// ---- one emulated warp of the v4 warp-per-stream decoder (warp_decode4.cuh) ------------------------------
namespace {
struct Warp4Job {
  simt::Group*           grp;
  int                    lane;
  defl::v4::WarpSmem4*   sm;
  void*                  scratch;
  const uint8_t*         src;
  uint32_t               srclen;
  uint8_t*               dst;
  uint32_t               cap;
  int                    flags;
  defl::StreamOut*       out;
};
void* warp4_main(void* arg) {
  Warp4Job* j = static_cast<Warp4Job*>(arg);
  simt::t_group = j->grp;
  simt::t_lane  = j->lane;
  defl::StreamOut o;
  defl::v4::inflate_stream_warp4(*j->sm, defl::v4::scratch4_of(j->scratch), j->lane, j->src, j->srclen, j->dst,
                                 j->cap, j->flags, o);
  if (j->lane == 0) *j->out = o;
  return nullptr;
}
}  // namespace

extern "C" int emu_warp4_inflate(const uint8_t* src, uint32_t srclen, uint8_t* dst, uint32_t cap, int flags,
                                 defl::StreamOut* out) {
  simt::Group grp;
  grp.G = 32;
  pthread_barrier_init(&grp.bar, nullptr, 32);
  using defl::v4::WarpSmem4;
  WarpSmem4* sm = static_cast<WarpSmem4*>(aligned_alloc(64, (sizeof(WarpSmem4) + 63) & ~63ul));
  memset(sm, 0xA5, sizeof(*sm));
  void* scratch = aligned_alloc(64, defl::v4::kWarpScratchBytes);
  memset(scratch, 0x5A, defl::v4::kWarpScratchBytes);
  std::vector<Warp4Job>  jobs(32);
  std::vector<pthread_t> th(32);
  for (int i = 0; i < 32; i++) jobs[i] = Warp4Job{&grp, i, sm, scratch, src, srclen, dst, cap, flags, out};
  for (int i = 1; i < 32; i++) pthread_create(&th[i], nullptr, warp4_main, &jobs[i]);
  warp4_main(&jobs[0]);
  for (int i = 1; i < 32; i++) pthread_join(th[i], nullptr);
  pthread_barrier_destroy(&grp.bar);
  free(sm); free(scratch);
  return out->code;
}

// ---- split decode of one stream (split_decode.cuh): chunks run one after another on one emulated warp ----
#include "../../defl_b200/csrc/split_decode.cuh"
namespace {

struct ChunkJob {
  simt::Group*        grp;
  int                 lane;
  defl::WarpSmem*     sm;
  void*               scratch;
  const uint8_t*      src;
  uint32_t            srclen;
  const uint64_t*     cand;
  uint32_t            ncand, me;
  uint16_t*           dst;
  uint32_t            cap;
  bool                count;
  defl::SplitChunkOut* out;
  uint32_t*           tok;      // count: where tokens are kept; decode: tokens to replay (with ntok)
  uint32_t            tokcap, ntok;
  uint32_t            mode, start_final;
};

void* chunk_main(void* arg) {
  ChunkJob* j = static_cast<ChunkJob*>(arg);
  simt::t_group = j->grp;
  simt::t_lane  = j->lane;
  defl::SplitChunkOut o;
  const uint32_t hist = j->me == 0 ? 0u : 32768u;
  if (j->count) {
    defl::TokSink sink;
    sink.buf = j->tok; sink.cap = j->tokcap; sink.n = 0; sink.ok = j->tok != nullptr;
    defl::inflate_chunk_warp<uint16_t, true>(*j->sm, defl::scratch_of(j->scratch), j->lane, j->src, j->srclen, j->cand,
                                             j->ncand, j->me, nullptr, 0xFFFFFFFFu, hist, o, &sink, nullptr, defl::kNoTokens,
                                             j->mode, j->start_final);
  } else {
    defl::inflate_chunk_warp<uint16_t, false>(*j->sm, defl::scratch_of(j->scratch), j->lane, j->src, j->srclen, j->cand,
                                              j->ncand, j->me, j->dst, j->cap, hist, o, nullptr, j->tok, j->ntok, j->mode,
                                              j->start_final);
  }
  if (j->lane == 0) *j->out = o;
  return nullptr;
}

void run_chunk(defl::WarpSmem* sm, void* scratch, const uint8_t* src, uint32_t srclen, const uint64_t* cand,
               uint32_t ncand, uint32_t me, uint16_t* dst, uint32_t cap, bool count, defl::SplitChunkOut* out,
               uint32_t* tok = nullptr, uint32_t tokcap = 0, uint32_t ntok = defl::kNoTokens,
               uint32_t mode = defl::MODE_BLOCK_STARTS, uint32_t start_final = defl::FIN_UNKNOWN) {
  simt::Group grp;
  grp.G = 32;
  pthread_barrier_init(&grp.bar, nullptr, 32);
  std::vector<ChunkJob>  jobs(32);
  std::vector<pthread_t> th(32);
  for (int i = 0; i < 32; i++) jobs[i] = ChunkJob{&grp, i, sm, scratch, src, srclen, cand, ncand, me, dst, cap, count, out, tok, tokcap, ntok, mode, start_final};
  for (int i = 1; i < 32; i++) pthread_create(&th[i], nullptr, chunk_main, &jobs[i]);
  chunk_main(&jobs[0]);
  for (int i = 1; i < 32; i++) pthread_join(th[i], nullptr);
  pthread_barrier_destroy(&grp.bar);
}

}  // namespace

static long g_replayed = 0;
extern "C" long emu_split_replayed() { return g_replayed; }

// cand[0] must be the first block of the stream.  Returns produced bytes, or -1 - (failing step).
extern "C" long emu_split_inflate(const uint8_t* src, uint32_t srclen, const uint64_t* cand, uint32_t ncand,
                                  uint8_t* dst, uint32_t cap, uint64_t* end_bit) {
  defl::WarpSmem* sm = static_cast<defl::WarpSmem*>(aligned_alloc(64, (sizeof(defl::WarpSmem) + 63) & ~63ul));
  void* scratch = aligned_alloc(64, defl::kWarpScratchBytes);
  std::vector<defl::SplitChunkOut> outs(ncand);
  // every chunk keeps its tokens in the count pass (2 per compressed byte of budget, as the host does);
  // odd chunks replay them in the decode pass, even chunks decode again: both ways are exercised
  std::vector<std::vector<uint32_t>> toks(ncand);
  for (uint32_t i = 0; i < ncand; i++) {
    const uint64_t b0 = cand[i] >> 3, b1 = i + 1 < ncand ? cand[i + 1] >> 3 : srclen;
    toks[i].resize(2 * (b1 - b0) + 1);
    run_chunk(sm, scratch, src, srclen, cand, ncand, i, nullptr, 0, true, &outs[i], toks[i].data(), (uint32_t)toks[i].size() - 1);
  }
  std::vector<uint32_t> chain;
  std::vector<uint64_t> off;
  uint64_t total = 0;
  long rc = 0;
  for (uint32_t i = 0;;) {
    if ((outs[i].status & 0xFF) == defl::CH_FAIL) { rc = -1; break; }
    chain.push_back(i); off.push_back(total); total += outs[i].out_bytes;
    if (total > cap) { rc = -2; break; }
    if (outs[i].status == defl::CH_FINAL) { *end_bit = outs[i].end_bit; break; }
    if (outs[i].link <= i || outs[i].link >= ncand) { rc = -3; break; }
    i = outs[i].link;
  }
  if (rc == 0) {
    off.push_back(total);
    std::vector<uint16_t> sym(total + 8, 0xFFFF);
    for (size_t k = 0; k < chain.size() && rc == 0; k++) {
      defl::SplitChunkOut o2;
      const bool replay = (k & 1) && outs[chain[k]].ntok != defl::kNoTokens;
      run_chunk(sm, scratch, src, srclen, cand, ncand, chain[k], sym.data() + off[k], outs[chain[k]].out_bytes, false, &o2,
                toks[chain[k]].data(), 0, replay ? outs[chain[k]].ntok : defl::kNoTokens);
      if (replay ? (o2.status != defl::CH_REPLAYED || o2.out_bytes != outs[chain[k]].out_bytes)
                 : (o2.status != outs[chain[k]].status || o2.out_bytes != outs[chain[k]].out_bytes)) rc = -4;
      if (replay) g_replayed++;
    }
    if (rc == 0) {
      uint32_t cc = 0;
      for (uint64_t i = 0; i < total; i++) {
        while (cc + 1 < chain.size() && off[cc + 1] <= i) cc++;
        dst[i] = (uint8_t)defl::resolve_marker(sym.data(), off.data(), cc, i);
      }
      rc = (long)total;
    }
  }
  free(sm); free(scratch);
  return rc;
}

// ---- split decode of a fixed-Huffman stream from token boundaries inside its blocks (MODE_FIXED_MID) ----
namespace {
struct ProbeJob {
  simt::Group*    grp;
  int             lane;
  defl::WarpSmem* sm;
  const uint8_t*  src;
  uint32_t        srclen, body, granule, probe_bits, ngran;
  uint64_t*       found;
};

void* probe_main(void* arg) {
  ProbeJob* j = static_cast<ProbeJob*>(arg);
  simt::t_group = j->grp;
  simt::t_lane  = j->lane;
  defl::Lane<32> L;
  L.gmask = defl::kFull; L.lane = j->lane; L.wlane = j->lane;
  defl::BitReader none;
  none.w = nullptr; none.wend = 0;
  bool fixed_ready = false;
  defl::huffman_block_setup<32>(L, j->sm->lit, j->sm->dist, j->sm->s, none, 0, 1u, fixed_ready);
  simt::syncwarp(defl::kFull);
  const uint32_t  mis      = (uint32_t)((uintptr_t)j->src & 3u);
  const uint64_t  base_bit = (uint64_t)mis * 8u;
  const uint64_t  end_bit  = base_bit + (uint64_t)j->srclen * 8u;
  const uint32_t* words    = reinterpret_cast<const uint32_t*>((uintptr_t)j->src - mis);
  const uint32_t  wend     = (uint32_t)(((uint64_t)mis + j->srclen + 3u) >> 2);
  for (uint32_t g = 1 + (uint32_t)j->lane; g < j->ngran; g += 32) {
    const uint64_t p = base_bit + ((uint64_t)j->body + (uint64_t)g * j->granule) * 8u;
    const uint64_t r = defl::fixed_probe_lane(j->sm->lit, j->sm->dist, words, wend, p, end_bit, j->probe_bits);
    j->found[g] = r == defl::kNoCandidate ? r : r - base_bit;
  }
  return nullptr;
}
}  // namespace

// src: a stream whose blocks are all fixed-Huffman, body at byte `body`.  The host side of split_decode_many's
// mid-block mode restated for one stream: probe, count, chain walk carrying block finality, decode, resolve.
// ncand_out: candidates used; nchain_out: chunks on the true chain.
extern "C" long emu_split_inflate_mid(const uint8_t* src, uint32_t srclen, uint32_t body, uint32_t granule,
                                      uint32_t probe_bits, uint8_t* dst, uint32_t cap, uint64_t* end_bit,
                                      uint32_t* ncand_out, uint32_t* nchain_out) {
  using namespace defl;
  WarpSmem* sm = static_cast<WarpSmem*>(aligned_alloc(64, (sizeof(WarpSmem) + 63) & ~63ul));
  void* scratch = aligned_alloc(64, kWarpScratchBytes);
  const uint32_t ngran = (srclen - body + granule - 1) / granule;
  std::vector<uint64_t> found(ngran, kNoCandidate);
  {
    simt::Group grp;
    grp.G = 32;
    pthread_barrier_init(&grp.bar, nullptr, 32);
    std::vector<ProbeJob>  jobs(32);
    std::vector<pthread_t> th(32);
    for (int i = 0; i < 32; i++) jobs[i] = ProbeJob{&grp, i, sm, src, srclen, body, granule, probe_bits, ngran, found.data()};
    for (int i = 1; i < 32; i++) pthread_create(&th[i], nullptr, probe_main, &jobs[i]);
    probe_main(&jobs[0]);
    for (int i = 1; i < 32; i++) pthread_join(th[i], nullptr);
    pthread_barrier_destroy(&grp.bar);
  }
  std::vector<uint64_t> cand{(uint64_t)body * 8u};
  for (uint32_t g = 1; g < ngran; g++)
    if (found[g] != kNoCandidate && found[g] > cand.back()) cand.push_back(found[g]);
  const uint32_t ncand = (uint32_t)cand.size();
  *ncand_out = ncand;

  std::vector<SplitChunkOut> outs(ncand);
  std::vector<std::vector<uint32_t>> toks(ncand);
  for (uint32_t i = 0; i < ncand; i++) {
    const uint64_t b0 = cand[i] >> 3, b1 = i + 1 < ncand ? cand[i + 1] >> 3 : srclen;
    toks[i].resize(3 * (b1 - b0) + 1);
    run_chunk(sm, scratch, src, srclen, cand.data(), ncand, i, nullptr, 0, true, &outs[i], toks[i].data(),
              (uint32_t)toks[i].size() - 1, kNoTokens, MODE_FIXED_MID, FIN_UNKNOWN);
  }
  struct Item { uint32_t cand, bytes, ntok, start_final; };
  std::vector<Item>     chain;
  std::vector<uint64_t> off;
  uint64_t total = 0;
  long     rc = 0;
  uint32_t fin = FIN_NO;
  for (uint32_t k = 0;;) {
    const SplitChunkOut& o = outs[k];
    if ((o.status & 0xFF) == CH_FAIL && !(fin == FIN_YES && o.has_eob)) { rc = -1; break; }
    const bool ends_here = fin == FIN_YES && k != 0 && o.has_eob;
    if (fin == FIN_YES && k != 0 && !o.has_eob && o.status != CH_LINKED) { rc = -5; break; }
    const uint32_t bytes = ends_here ? o.eob_bytes : o.out_bytes;
    chain.push_back({k, bytes, ends_here ? o.eob_ntok : o.ntok, k == 0 ? (uint32_t)FIN_NO : fin});
    off.push_back(total);
    total += bytes;
    if (total > cap) { rc = -2; break; }
    if (ends_here) { *end_bit = o.eob_bit; break; }
    if (o.status == CH_FINAL) { *end_bit = o.end_bit; break; }
    if (o.status != CH_LINKED || o.link <= k || o.link >= ncand) { rc = -3; break; }
    if (o.link_final != FIN_UNKNOWN) fin = o.link_final;
    k = o.link;
  }
  *nchain_out = (uint32_t)chain.size();
  if (rc == 0) {
    off.push_back(total);
    std::vector<uint16_t> sym(total + 8, 0xFFFF);
    for (size_t k = 0; k < chain.size() && rc == 0; k++) {
      SplitChunkOut o2;
      const bool replay = (k & 1) && chain[k].ntok != kNoTokens;
      run_chunk(sm, scratch, src, srclen, cand.data(), ncand, chain[k].cand, sym.data() + off[k], chain[k].bytes, false, &o2,
                toks[chain[k].cand].data(), 0, replay ? chain[k].ntok : kNoTokens, MODE_FIXED_MID, chain[k].start_final);
      if (replay ? (o2.status != CH_REPLAYED || o2.out_bytes != chain[k].bytes)
                 : ((o2.status & 0xFF) == CH_FAIL || o2.out_bytes != chain[k].bytes)) rc = -4 - 10 * (long)k;
      if (replay) g_replayed++;
    }
    if (rc == 0) {
      uint32_t cc = 0;
      for (uint64_t i = 0; i < total; i++) {
        while (cc + 1 < chain.size() && off[cc + 1] <= i) cc++;
        dst[i] = (uint8_t)resolve_marker(sym.data(), off.data(), cc, i);
      }
      rc = (long)total;
    }
  }
  free(sm); free(scratch);
  return rc;
}
