// defl_b200.cu -- C-ABI of the B200 inflate engine (libdefl_b200.so).
//
// Host side of the drop-in boundary: the reference's public functions
// (/root/reference/include/defl/infl.h:31-189, implemented there by src/infl/infl.c:1507-1695,
// src/infl/mem.c:114-361 and src/infl/stream.c:468-794) re-implemented as a thin layer that
// describes the work to the sm_100a kernels in kernels.cuh.  There is deliberately no CPU
// decode path in this file: without a usable CUDA device every call fails.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <numeric>
#include <string>
#include <thread>
#include <functional>
#include <vector>

#include "../../include/defl/infl.h"
#include "../../include/defl/infl_batch.h"
#include "../../include/defl/version.h"
#include "kernels.cuh"
#include "wrap_kernels.cuh"
#include "../../include/defl/infl_wrap.h"

namespace {

using defl::GroupSmem;

thread_local std::string      g_last_error;
std::atomic<uint64_t>         g_launches{0};
std::atomic<int>              g_lanes_per_stream{0};
std::atomic<int>              g_pipeline_shape{0};


int fail(const char* what, cudaError_t e = cudaSuccess) {
  g_last_error = what;
  if (e != cudaSuccess) {
    g_last_error += ": ";
    g_last_error += cudaGetErrorString(e);
  }
  return UNZ_ERR;
}

#define CU_TRY(expr, what)                        \
  do {                                            \
    cudaError_t _e = (expr);                      \
    if (_e != cudaSuccess) return fail(what, _e); \
  } while (0)

// ---- per-device engine state: cached staging buffers, one worker stream ------------------------
struct DevBuf {
  void*  p   = nullptr;
  size_t cap = 0;
  int reserve(size_t n) {
    if (n <= cap) return UNZ_OK;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = std::max<size_t>(n + n / 4, 1 << 20);
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { e = cudaMalloc(&p, n); want = n; }
    if (e != cudaSuccess) return fail("cudaMalloc (staging)", e);
    cap = want;
    return UNZ_OK;
  }
  int reserve_exact(size_t n) {      // no slack: the token scratch is large and never grows
    if (n <= cap) return UNZ_OK;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc(&p, n);
    if (e != cudaSuccess) return fail("cudaMalloc (workspace)", e);
    cap = n;
    return UNZ_OK;
  }
};

struct PinBuf {
  void*  p   = nullptr;
  size_t cap = 0;
  int reserve(size_t n) {
    if (n <= cap) return UNZ_OK;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    size_t want = std::max<size_t>(n + n / 4, 1 << 16);
    cudaError_t e = cudaMallocHost(&p, want);
    if (e != cudaSuccess) return fail("cudaMallocHost", e);
    cap = want;
    return UNZ_OK;
  }
};

struct Engine {
  int          dev      = -1;
  int          sms      = 0;
  cudaStream_t stream   = nullptr;
  cudaStream_t stream2  = nullptr;   // the one-warp kernel of a batch that also has split-decoded members
  cudaStream_t stream3  = nullptr;   // chunk-list batches: the decoded bytes of one group go home while the next group decodes
  std::mutex   mu;
  DevBuf       d_in, d_out, d_items, d_results, d_order, d_ws, d_seg, d_stage;
  DevBuf       d_items2, d_results2, d_order2, d_ws2;
  PinBuf       h_items, h_results, h_order, h_pack, h_seg;
  std::atomic<bool> ready{false};   // set with release once stream / sms are in place (double-checked under g_engine_mu)
};

constexpr int kMaxDevices = 16;
Engine        g_engines[kMaxDevices];
std::mutex    g_engine_mu;

int engine_for_current_device(Engine** out) {
  int dev = 0;
  CU_TRY(cudaGetDevice(&dev), "cudaGetDevice (is a CUDA device visible? this library has no CPU path)");
  if (dev < 0 || dev >= kMaxDevices) return fail("device index out of range");
  Engine& e = g_engines[dev];
  if (!e.ready.load(std::memory_order_acquire)) {
    std::lock_guard<std::mutex> g(g_engine_mu);
    if (!e.ready.load(std::memory_order_relaxed)) {
      cudaDeviceProp prop;
      CU_TRY(cudaGetDeviceProperties(&prop, dev), "cudaGetDeviceProperties");
      if (prop.major < 10) return fail("this build contains sm_100a code only (needs a B200-class GPU)");
      e.dev = dev;
      e.sms = prop.multiProcessorCount;
      CU_TRY(cudaStreamCreateWithFlags(&e.stream, cudaStreamNonBlocking), "cudaStreamCreate");
      CU_TRY(cudaStreamCreateWithFlags(&e.stream2, cudaStreamNonBlocking), "cudaStreamCreate");
      CU_TRY(cudaStreamCreateWithFlags(&e.stream3, cudaStreamNonBlocking), "cudaStreamCreate");
      e.ready.store(true, std::memory_order_release);
    }
  }
  *out = &e;
  return UNZ_OK;
}

// ---- kernel launch ------------------------------------------------------------------------------------
#ifdef DEFL_LEGACY_ENGINES   // v1 (group per stream) and v2 (lane-decoder pipeline): cross-check builds only (python -m defl_b200.build --legacy)
template <int G, int GROUPS>
int launch_batch(int sms, const infl_batch_item_t* d_items, const uint32_t* d_order, uint32_t n, int flags,
                 infl_batch_result_t* d_results, unsigned int* d_ticket, cudaStream_t st) {
  auto         kern = defl::inflate_batch_kernel<G, GROUPS>;
  const size_t smem = sizeof(GroupSmem) * GROUPS;
  static std::atomic<int> ctas_per_sm[kMaxDevices];   // zero-initialised; racing first callers compute the same value
  int          dev = 0;
  cudaGetDevice(&dev);
  if (!ctas_per_sm[dev]) {
    CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
           "cudaFuncSetAttribute(smem)");
    int occ = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, G * GROUPS, smem), "occupancy query");
    if (occ < 1) return fail("inflate kernel does not fit on an SM");
    ctas_per_sm[dev] = occ;
  }
  // persistent grid: a whole number of CTAs per SM, never more groups than streams
  uint64_t grid = (uint64_t)sms * ctas_per_sm[dev];
  grid = std::min<uint64_t>(grid, (n + GROUPS - 1) / GROUPS);
  grid = std::max<uint64_t>(grid, 1);
  CU_TRY(cudaMemsetAsync(d_ticket, 0, sizeof(unsigned int), st), "cudaMemsetAsync(ticket)");
  kern<<<(unsigned)grid, G * GROUPS, smem, st>>>(d_items, d_order, n, flags, d_results, d_ticket);
  CU_TRY(cudaGetLastError(), "inflate_batch_kernel launch");
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return UNZ_OK;
}


template <int D, int SL, int K>
int launch_pipeline(int sms, const infl_batch_item_t* d_items, const uint32_t* d_order, uint32_t n, int flags,
                    infl_batch_result_t* d_results, unsigned int* d_ticket, cudaStream_t st) {
  auto             kern = defl::inflate_pipeline_kernel<D, SL, K>;
  constexpr size_t smem = defl::pipeline_smem_bytes<D, SL, K>();
  static std::atomic<int> ctas_per_sm[kMaxDevices];
  int              dev = 0;
  cudaGetDevice(&dev);
  if (!ctas_per_sm[dev]) {
    CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
           "cudaFuncSetAttribute(smem)");
    int occ = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, (D + K) * 32, smem), "occupancy query");
    if (occ < 1) return fail("inflate pipeline kernel does not fit on an SM");
    ctas_per_sm[dev] = occ;
  }
  uint64_t grid = (uint64_t)sms * ctas_per_sm[dev];
  grid = std::min<uint64_t>(grid, (n + D * SL - 1) / (D * SL));
  grid = std::max<uint64_t>(grid, 1);
  CU_TRY(cudaMemsetAsync(d_ticket, 0, sizeof(unsigned int), st), "cudaMemsetAsync(ticket)");
  kern<<<(unsigned)grid, (D + K) * 32, smem, st>>>(d_items, d_order, n, flags, d_results, d_ticket);
  CU_TRY(cudaGetLastError(), "inflate_pipeline_kernel launch");
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return UNZ_OK;
}

#endif  // DEFL_LEGACY_ENGINES

// ---- v3 (default engine): one warp per stream --------------------------------------------------------------
constexpr int    kWarpsPerCta = 4;
constexpr size_t kWsHeader    = 256;     // ticket counter and friends live in front of the token scratch

int warp_kernel_grid(int sms, int* grid_out) {
  auto         kern = defl::inflate_warp_kernel<kWarpsPerCta>;
  const size_t smem = sizeof(defl::WarpSmem) * kWarpsPerCta;
  static std::atomic<int> ctas_per_sm[kMaxDevices];   // zero-initialised; racing first callers compute the same value
  int          dev = 0;
  cudaGetDevice(&dev);
  if (!ctas_per_sm[dev]) {
    CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
           "cudaFuncSetAttribute(smem)");
    int occ = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kWarpsPerCta * 32, smem), "occupancy query");
    if (occ < 1) return fail("inflate warp kernel does not fit on an SM");
    ctas_per_sm[dev] = occ;
  }
  *grid_out = sms * ctas_per_sm[dev];
  return UNZ_OK;
}

// ---- v4 (default engine): warp per stream, 64-bit tokens, shared-memory rings and stage ----------------------
int warp4_kernel_grid(int sms, int* grid_out) {        // in warps = CTAs
  auto       kern = defl::inflate_warp4_kernel;
  static std::atomic<int> ctas_per_sm[kMaxDevices];
  int        dev = 0;
  cudaGetDevice(&dev);
  if (!ctas_per_sm[dev]) {
    int occ = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32, 0), "occupancy query");
    if (occ < 1) return fail("inflate warp4 kernel does not fit on an SM");
    ctas_per_sm[dev] = occ;
  }
  *grid_out = sms * ctas_per_sm[dev];
  return UNZ_OK;
}

// one workspace serves whichever engine runs: the larger of the two
size_t warp_kernel_workspace(int sms) {
  int grid = 0, grid4 = 0;
  if (warp_kernel_grid(sms, &grid) != UNZ_OK || warp4_kernel_grid(sms, &grid4) != UNZ_OK) return 0;
  return kWsHeader + std::max((size_t)grid * kWarpsPerCta * defl::kWarpScratchBytes,
                              (size_t)grid4 * defl::v4::kWarpScratchBytes);
}

int launch_warp4(int sms, const infl_batch_item_t* d_items, const uint32_t* d_order, uint32_t n, int flags,
                 infl_batch_result_t* d_results, void* d_ws, cudaStream_t st) {
  int grid = 0;
  if (warp4_kernel_grid(sms, &grid) != UNZ_OK) return UNZ_ERR;
  grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)grid, n));
  auto* ticket  = static_cast<unsigned int*>(d_ws);
  auto* scratch = static_cast<unsigned char*>(d_ws) + kWsHeader;
  CU_TRY(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st), "cudaMemsetAsync(ticket)");
  defl::inflate_warp4_kernel<<<(unsigned)grid, 32, 0, st>>>(d_items, d_order, n, flags, d_results, ticket, scratch);
  CU_TRY(cudaGetLastError(), "inflate_warp4_kernel launch");
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return UNZ_OK;
}

int launch_warp(int sms, const infl_batch_item_t* d_items, const uint32_t* d_order, uint32_t n, int flags,
                infl_batch_result_t* d_results, void* d_ws, cudaStream_t st) {
  int grid = 0;
  if (warp_kernel_grid(sms, &grid) != UNZ_OK) return UNZ_ERR;
  grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)grid, (n + kWarpsPerCta - 1) / kWarpsPerCta));
  auto* ticket  = static_cast<unsigned int*>(d_ws);
  auto* scratch = static_cast<unsigned char*>(d_ws) + kWsHeader;
  CU_TRY(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st), "cudaMemsetAsync(ticket)");
  defl::inflate_warp_kernel<kWarpsPerCta><<<(unsigned)grid, kWarpsPerCta * 32, sizeof(defl::WarpSmem) * kWarpsPerCta, st>>>(
      d_items, d_order, n, flags, d_results, ticket, scratch);
  CU_TRY(cudaGetLastError(), "inflate_warp_kernel launch");
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return UNZ_OK;
}

std::atomic<int> g_engine{3};   // 3 = warp per stream (default), 4 = v4 (selectable; not faster yet), 2 = lane-decoder pipeline, 1 = group per stream

int launch_batch_any(int sms, const infl_batch_item_t* d_items, const uint32_t* d_order, size_t n, int flags,
                     infl_batch_result_t* d_results, void* d_ws, cudaStream_t st) {
  if (n == 0) return UNZ_OK;
  if (n > 0xFFFFFFFFull) return fail("batch too large (n must fit 32 bits)");
  int  lanes = g_lanes_per_stream.load();
  auto t = static_cast<unsigned int*>(d_ws);
  if (!lanes && g_engine.load() == 4)
    return launch_warp4(sms, d_items, d_order, (uint32_t)n, flags, d_results, d_ws, st);
  if (!lanes && g_engine.load() == 3)
    return launch_warp(sms, d_items, d_order, (uint32_t)n, flags, d_results, d_ws, st);
#ifdef DEFL_LEGACY_ENGINES
  if (!lanes) {   // the lane-decoder pipeline (v2)
    switch (g_pipeline_shape.load()) {
      case 0:  return launch_pipeline<2, 24, 14>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
      case 1:  return launch_pipeline<2, 24, 8>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
      case 2:  return launch_pipeline<3, 16, 10>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
      case 3:  return launch_pipeline<4, 12, 10>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
      case 4:  return launch_pipeline<6, 8, 10>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
      case 5:  return launch_pipeline<12, 4, 8>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
      case 6:  return launch_pipeline<2, 24, 19>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
      case 7:  return launch_pipeline<1, 32, 16>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
      default: return fail("unknown pipeline shape");
    }
  }
  switch (lanes) {
    case 1:  return launch_batch<1, 32>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
    case 2:  return launch_batch<2, 16>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
    case 4:  return launch_batch<4, 16>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
    case 8:  return launch_batch<8, 16>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
    case 16: return launch_batch<16, 8>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
    case 32: return launch_batch<32, 4>(sms, d_items, d_order, (uint32_t)n, flags, d_results, t, st);
    default: return fail("lanes per stream must be 1, 2, 4, 8, 16 or 32");
  }
#else
  (void)t;
  return fail("this build has the warp-per-stream engines only (engines 3 and 4); build with -DDEFL_LEGACY_ENGINES for 1 and 2");
#endif
}

// longest serial decode first.  The work of a stream is its tokens (about its compressed size) plus its
// output words: a few KiB that expand to megabytes of one repeated byte are not a short job.
void size_order(const infl_batch_item_t* items, size_t n, uint32_t* order) {
  std::iota(order, order + n, 0u);
  auto work = [&](uint32_t i) { return items[i].srclen + (items[i].dstlen >> 4); };
  std::stable_sort(order, order + n, [&](uint32_t a, uint32_t b) { return work(a) > work(b); });
}

constexpr size_t kCopyThreshold = 8192;   // pieces up to this size are copied by the reference (src/common.h:46)
constexpr size_t kGatherPiece   = 65536;  // bytes one gather descriptor covers, at most
constexpr size_t kAlign = 16;
inline size_t align_up(size_t v) { return (v + kAlign - 1) & ~(kAlign - 1); }

// ---- one large stream, decoded by the whole GPU (split_decode.cuh) ----------------------------------------------
constexpr uint64_t kSplitMinBytes = 4ull << 20;      // compressed size from which a stream is worth splitting
uint32_t env_u32(const char* name, uint32_t dflt, uint32_t lo, uint32_t hi) {
  const char* v = std::getenv(name);
  if (!v || !*v) return dflt;
  const unsigned long long x = std::strtoull(v, nullptr, 0);
  return (uint32_t)std::min<unsigned long long>(hi, std::max<unsigned long long>(lo, x));
}
std::atomic<uint64_t> g_split_decodes{0};

struct SplitBufs {
  DevBuf d_found, d_cand, d_jobs, d_outs, d_outs2, d_sym, d_off, d_partial, d_wbuf, d_wdbl, d_tok, d_streams, d_peek, d_at;
};
SplitBufs g_split[kMaxDevices];

struct SplitInput {          // one stream handed to the split decoder: device pointers
  const uint8_t* d_src;
  uint64_t       srclen;
  uint8_t*       d_dst;
  uint64_t       dstlen;
};

// Split decode of MANY streams at once (every kernel works on all of them: see split_decode.cuh).
// status[i] = UNZ_OK with res[i] filled when stream i was decoded here, UNZ_NOOP when it has to go through
// the one-warp decoder instead (too few block starts, any failure: that path defines error behaviour).
// Returns UNZ_ERR only when the engine itself failed.  d_ws is the token scratch of the batch engine.
// after_peek (optional) is called once, with one flag per stream, as soon as the cheap stages (a look at each
// stream's first bytes, the search for split points) have ruled some out: the caller can start those elsewhere
// at once.
int split_decode_many(Engine* E, const SplitInput* in, size_t nin, int flags, void* d_ws, cudaStream_t st,
                      infl_batch_result_t* res, int* status,
                      const std::function<int(const std::vector<char>&)>* after_peek = nullptr) {
  using namespace defl;
  SplitBufs& SB = g_split[E->dev];
  const bool zlib = (flags & 0xFF) == INFL_ZLIB;
  const uint32_t ns = (uint32_t)nin;
  for (uint32_t i = 0; i < ns; i++) status[i] = UNZ_NOOP;
  if (!ns) return UNZ_OK;
  // DEFL_SPLIT_TRACE=1: per-phase wall times on stderr (adds a stream synchronise per phase)
  static const bool trace = std::getenv("DEFL_SPLIT_TRACE") != nullptr;
  auto tnow = [] { return std::chrono::steady_clock::now(); };
  auto tprev = tnow();
  auto mark = [&](const char* what) {
    if (!trace) return;
    cudaStreamSynchronize(st);
    const auto t = tnow();
    std::fprintf(stderr, "[split] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t - tprev).count());
    tprev = t;
  };
  static const uint32_t granule = env_u32("DEFL_SPLIT_GRANULE", 65536u, 4096u, 1u << 24);
  static const uint64_t min_gap = (uint64_t)env_u32("DEFL_SPLIT_GAP", 0u, 0u, 1u << 28) * 8u;   // bits
  static const uint32_t tok_per_byte = env_u32("DEFL_SPLIT_TOKENS_PER_BYTE", 3u, 0u, 32u);
  static const bool     fixed_mid = env_u32("DEFL_SPLIT_FIXED", 1u, 0u, 1u) != 0u;
  static const uint32_t probe_bits = env_u32("DEFL_SPLIT_PROBE_BITS", 32768u, 256u, 1u << 20);

  int grid_warps = 0;
  if (warp_kernel_grid(E->sms, &grid_warps) != UNZ_OK) return UNZ_ERR;      // CTAs of kWarpsPerCta warps
  auto* ticket  = static_cast<unsigned int*>(d_ws);
  auto* scratch = static_cast<unsigned char*>(d_ws) + kWsHeader;
  const size_t smem = sizeof(WarpSmem) * kWarpsPerCta;

  std::vector<SplitStreamDesc> sd(ns);
  std::vector<char> alive(ns, 1);
  for (uint32_t i = 0; i < ns; i++) {
    std::memset(&sd[i], 0, sizeof(sd[i]));
    sd[i].src = in[i].d_src; sd[i].dst = in[i].d_dst; sd[i].srclen = in[i].srclen;
    if (in[i].srclen > 0xFFFFFFFFull || in[i].dstlen > 0xFFFFFFFFull || in[i].srclen < 64) alive[i] = 0;
  }
  if (SB.d_streams.reserve((size_t)ns * sizeof(SplitStreamDesc)) || SB.d_peek.reserve((size_t)ns * 4 + 16) || SB.d_at.reserve((size_t)ns * 8))
    return UNZ_ERR;
  auto* d_sd = static_cast<SplitStreamDesc*>(SB.d_streams.p);
  auto upload_streams = [&]() -> int {
    CU_TRY(cudaMemcpyAsync(d_sd, sd.data(), (size_t)ns * sizeof(SplitStreamDesc), cudaMemcpyHostToDevice, st), "H2D stream table");
    return UNZ_OK;
  };

  // ---- zlib headers, and the type of each stream's first block
  {
    if (upload_streams() != UNZ_OK) return UNZ_ERR;
    std::vector<uint8_t> h3((size_t)ns * 3);
    split_peek_kernel<<<(ns + 127) / 128, 128, 0, st>>>(d_sd, ns, nullptr, 3u, static_cast<uint8_t*>(SB.d_peek.p));
    CU_TRY(cudaGetLastError(), "split_peek_kernel launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU_TRY(cudaMemcpyAsync(h3.data(), SB.d_peek.p, h3.size(), cudaMemcpyDeviceToHost, st), "D2H stream heads");
    CU_TRY(cudaStreamSynchronize(st), "D2H stream heads");
    for (uint32_t i = 0; i < ns; i++) {
      const uint8_t c0 = h3[3 * i], c1 = h3[3 * i + 1];
      if (zlib) {
        if ((c0 & 15) != 8 || (c0 >> 4) > 7 || ((c0 << 8) + c1) % 31 != 0 || (c1 & 0x20)) alive[i] = 0;
        sd[i].body = 2;
      }
      const uint32_t btype = (h3[3 * i + (zlib ? 2 : 0)] >> 1) & 3u;
      // a stream that opens with a fixed-Huffman block has no block headers worth searching for (Z_FIXED output is
      // fixed blocks throughout): its chunks start on token boundaries inside the blocks instead
      sd[i].mode = (btype == 1u && fixed_mid) ? MODE_FIXED_MID : MODE_BLOCK_STARTS;
      // a stream that opens with a stored block is almost always stored throughout (incompressible data): nothing
      // to find, and one warp copies it at memory speed anyway
      if (btype == 0u || btype == 3u) alive[i] = 0;
    }
  }

  // ---- F: block starts
  uint32_t ngran_total = 0;
  for (uint32_t i = 0; i < ns; i++) {
    sd[i].gran_base = ngran_total;
    sd[i].ngran = alive[i] ? (uint32_t)((in[i].srclen - sd[i].body + granule - 1) / granule) : 0u;
    ngran_total += sd[i].ngran;
  }
  if (!ngran_total) return UNZ_OK;
  // F's answers in the first ngran_total entries of d_found, P's (streams in mid-block mode) in the second
  if (SB.d_found.reserve((size_t)ngran_total * 16) || upload_streams() != UNZ_OK) return UNZ_ERR;
  find_block_starts_kernel<kWarpsPerCta><<<(ngran_total + kWarpsPerCta - 1) / kWarpsPerCta, kWarpsPerCta * 32, 0, st>>>(
      d_sd, ns, ngran_total, granule, static_cast<uint64_t*>(SB.d_found.p));
  CU_TRY(cudaGetLastError(), "find_block_starts_kernel launch");
  g_launches.fetch_add(1, std::memory_order_relaxed);
  bool any_fixed = false;
  for (uint32_t i = 0; i < ns; i++) any_fixed = any_fixed || (alive[i] && sd[i].mode == MODE_FIXED_MID);
  if (any_fixed) {
    const unsigned per_cta = kWarpsPerCta * 32;
    fixed_probe_kernel<kWarpsPerCta><<<(ngran_total + per_cta - 1) / per_cta, per_cta, sizeof(WarpSmem) * kWarpsPerCta, st>>>(
        d_sd, ns, ngran_total, granule, probe_bits, static_cast<uint64_t*>(SB.d_found.p) + ngran_total);
    CU_TRY(cudaGetLastError(), "fixed_probe_kernel launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
  }
  std::vector<uint64_t> found((size_t)ngran_total * 2, kNoCandidate);
  CU_TRY(cudaMemcpyAsync(found.data(), SB.d_found.p, (size_t)ngran_total * (any_fixed ? 16 : 8), cudaMemcpyDeviceToHost, st), "D2H block starts");
  CU_TRY(cudaStreamSynchronize(st), "find_block_starts_kernel");
  mark("F find_block_starts");
  std::vector<uint64_t> cand;
  for (uint32_t i = 0; i < ns; i++) {
    sd[i].cand_base = (uint32_t)cand.size();
    if (!alive[i]) continue;
    const uint64_t* fnd = found.data() + (sd[i].mode == MODE_FIXED_MID ? ngran_total : 0u) + sd[i].gran_base;
    cand.push_back((uint64_t)sd[i].body * 8);
    for (uint32_t g = 0; g < sd[i].ngran; g++) {
      const uint64_t f = fnd[g];
      if (f != kNoCandidate && f >= cand.back() + min_gap && f > cand.back()) cand.push_back(f);
    }
    sd[i].ncand = (uint32_t)cand.size() - sd[i].cand_base;
    if (sd[i].ncand < 4) {
      if (trace) std::fprintf(stderr, "[split] given back: stream %u (%llu -> %llu bytes, mode %u) has %u candidates in %u granules\n", i,
                              (unsigned long long)in[i].srclen, (unsigned long long)in[i].dstlen, sd[i].mode, sd[i].ncand, sd[i].ngran);
      alive[i] = 0; cand.resize(sd[i].cand_base); sd[i].ncand = 0;
    }
  }
  // from here on only a broken chain (damaged input, a false block start that decodes) sends a stream back
  if (after_peek && (*after_peek)(alive) != UNZ_OK) return UNZ_ERR;
  const uint32_t ncand = (uint32_t)cand.size();
  if (!ncand) return UNZ_OK;

  // ---- A: every candidate counts its chunk.  The count pass keeps each chunk's tokens (tok_per_byte tokens of
  // budget per compressed byte, regions laid out by compressed position) so that the decode pass only has to
  // turn them into bytes.
  uint64_t tok_total = 0;
  for (uint32_t i = 0; i < ns; i++) {
    sd[i].tok_base = tok_total;
    if (alive[i]) tok_total += (uint64_t)tok_per_byte * in[i].srclen + 16;
  }
  uint32_t* d_tok = nullptr;
  if (tok_per_byte) {
    if (SB.d_tok.reserve_exact((size_t)tok_total * 4 + 64) == UNZ_OK) d_tok = static_cast<uint32_t*>(SB.d_tok.p);
    else cudaGetLastError();       // no room for it: the decode pass decodes again
  }
  std::vector<SplitJob> jobs(ncand);
  {
    uint32_t j = 0;
    for (uint32_t i = 0; i < ns; i++)
      for (uint32_t k = 0; k < sd[i].ncand; k++, j++) {
        const uint64_t* c = cand.data() + sd[i].cand_base;
        const uint64_t b0 = c[k] >> 3, b1 = k + 1 < sd[i].ncand ? c[k + 1] >> 3 : in[i].srclen;
        jobs[j].stream = i; jobs[j].start_final = FIN_UNKNOWN; jobs[j].cand = k; jobs[j].cap = 0; jobs[j].out_off = 0;
        jobs[j].tok_off = sd[i].tok_base + (uint64_t)tok_per_byte * b0;
        jobs[j].tok_cap = (uint32_t)std::min<uint64_t>((uint64_t)tok_per_byte * (b1 - b0), 0xFFFFFFF0ull);
        jobs[j].ntok = kNoTokens;
      }
  }
  // longest chunk first (fixed-Huffman chunks count double: they decode slowly): the pass ends when its slowest
  // warp does.  outs[] stays in candidate order: a_slot[j] says which record job j fills.
  std::vector<uint32_t> a_order(ncand);
  std::iota(a_order.begin(), a_order.end(), 0u);
  {
    std::vector<uint64_t> w(ncand);
    for (uint32_t j = 0; j < ncand; j++) {
      const SplitStreamDesc& d = sd[jobs[j].stream];
      const uint64_t* c = cand.data() + d.cand_base;
      const uint64_t e = jobs[j].cand + 1 < d.ncand ? c[jobs[j].cand + 1] : in[jobs[j].stream].srclen * 8;
      w[j] = (e - c[jobs[j].cand]) * (d.mode == MODE_FIXED_MID ? 2u : 1u);
    }
    std::stable_sort(a_order.begin(), a_order.end(), [&](uint32_t a, uint32_t b) { return w[a] > w[b]; });
    std::vector<SplitJob> sorted(ncand);
    for (uint32_t j = 0; j < ncand; j++) sorted[j] = jobs[a_order[j]];
    jobs.swap(sorted);
  }
  if (SB.d_cand.reserve((size_t)ncand * 8) || SB.d_jobs.reserve((size_t)ncand * sizeof(SplitJob)) ||
      SB.d_outs.reserve((size_t)ncand * sizeof(SplitChunkOut)) || SB.d_outs2.reserve((size_t)ncand * sizeof(SplitChunkOut)) ||
      upload_streams() != UNZ_OK)
    return UNZ_ERR;
  CU_TRY(cudaMemcpyAsync(SB.d_cand.p, cand.data(), (size_t)ncand * 8, cudaMemcpyHostToDevice, st), "H2D candidates");
  CU_TRY(cudaMemcpyAsync(SB.d_jobs.p, jobs.data(), (size_t)ncand * sizeof(SplitJob), cudaMemcpyHostToDevice, st), "H2D jobs");
  auto launch_chunks = [&](bool count, uint32_t njobs, uint16_t* sym) -> int {
    static bool attr_done[kMaxDevices][2] = {};
    if (!attr_done[E->dev][count]) {
      if (count) CU_TRY(cudaFuncSetAttribute(split_chunk_kernel<true, kWarpsPerCta>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "cudaFuncSetAttribute");
      else CU_TRY(cudaFuncSetAttribute(split_chunk_kernel<false, kWarpsPerCta>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "cudaFuncSetAttribute");
      attr_done[E->dev][count] = true;
    }
    const unsigned grid = (unsigned)std::max<uint32_t>(1, std::min<uint32_t>((uint32_t)grid_warps, (njobs + kWarpsPerCta - 1) / kWarpsPerCta));
    CU_TRY(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st), "cudaMemsetAsync(ticket)");
    if (count)
      split_chunk_kernel<true, kWarpsPerCta><<<grid, kWarpsPerCta * 32, smem, st>>>(
          d_sd, static_cast<uint64_t*>(SB.d_cand.p), static_cast<SplitJob*>(SB.d_jobs.p), njobs,
          sym, static_cast<SplitChunkOut*>(SB.d_outs.p), ticket, scratch, d_tok);
    else {
      // chunks whose tokens were kept: the resolver alone, in its own kernel (more warps per SM, see split_replay_kernel);
      // the others (over their token budget): decoded again by the full kernel, which skips the replayed ones
      uint32_t nreplay = 0;
      for (uint32_t j = 0; j < njobs; j++) nreplay += jobs[j].ntok != kNoTokens;
      if (nreplay) {
        static std::atomic<int> replay_ctas[kMaxDevices];
        if (!replay_ctas[E->dev]) {
          int occ = 0;
          CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, split_replay_kernel<kWarpsPerCta>, kWarpsPerCta * 32, 0), "occupancy query");
          replay_ctas[E->dev] = occ > 0 ? occ : 1;
        }
        const unsigned rgrid = (unsigned)std::max<uint32_t>(1, std::min<uint32_t>((uint32_t)(E->sms * replay_ctas[E->dev]), (njobs + kWarpsPerCta - 1) / kWarpsPerCta));
        split_replay_kernel<kWarpsPerCta><<<rgrid, kWarpsPerCta * 32, 0, st>>>(
            static_cast<SplitJob*>(SB.d_jobs.p), njobs, sym, static_cast<SplitChunkOut*>(SB.d_outs2.p), ticket, d_tok);
        CU_TRY(cudaGetLastError(), "split_replay_kernel launch");
        g_launches.fetch_add(1, std::memory_order_relaxed);
      }
      if (nreplay == njobs) return UNZ_OK;
      if (nreplay) CU_TRY(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st), "cudaMemsetAsync(ticket)");
      split_chunk_kernel<false, kWarpsPerCta><<<grid, kWarpsPerCta * 32, smem, st>>>(
          d_sd, static_cast<uint64_t*>(SB.d_cand.p), static_cast<SplitJob*>(SB.d_jobs.p), njobs,
          sym, static_cast<SplitChunkOut*>(SB.d_outs2.p), ticket, scratch, nullptr);
    }
    CU_TRY(cudaGetLastError(), "split_chunk_kernel launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    return UNZ_OK;
  };
  if (launch_chunks(true, ncand, nullptr) != UNZ_OK) return UNZ_ERR;
  std::vector<SplitChunkOut> outs(ncand);
  {
    std::vector<SplitChunkOut> by_job(ncand);
    CU_TRY(cudaMemcpyAsync(by_job.data(), SB.d_outs.p, (size_t)ncand * sizeof(SplitChunkOut), cudaMemcpyDeviceToHost, st), "D2H chunk sizes");
    CU_TRY(cudaStreamSynchronize(st), "split_chunk_kernel (count)");
    for (uint32_t j = 0; j < ncand; j++) outs[a_order[j]] = by_job[j];      // back to candidate order
  }
  mark("A count pass");

  // ---- per stream, the chain of true chunks from its first block to its final block
  // A chunk that starts inside a block (MODE_FIXED_MID) learns from its predecessor whether that block is the
  // stream's last one; if it is, the chunk ends at that block's end-of-block code (eob_* of its record).
  // off = offset inside the stream; tok0 / sub: a chunk whose tokens were kept is replayed in pieces (cut where the COUNT
  // pass marked the end of a decode window about every kPieceBytes of output): tok0 = first token of the piece, sub = it
  // is not the chunk's first piece.  Every piece is a chunk of its own for the C and D passes -- the 32 KiB before it are
  // unknown to its replay and become markers -- so that one long chunk no longer sets the time of the C pass.
  struct ChainItem { uint32_t stream, cand; uint64_t off; uint32_t bytes, ntok, start_final; bool ends; uint32_t tok0; bool sub; };
  // Pieces pay when the chunks alone do not fill the replay kernel (one big stream: 4700 chunks of uneven length for
  // ~7100 warps, C pass 10.5 -> 5.2 ms per GiB); a batch of many split members has chunks enough, and there the extra
  // windows only cost (mixed workload 83.5 -> 75.9 GB/s with pieces; a 1 GiB single-history stream with 6738 candidates
  // 61.2 -> 57.9: its D1 by doubling pays per window).  DEFL_SPLIT_PIECES: 0 never, 1 always, 2 = when the candidates
  // fill less than 0.8 of the replay kernel's 48 warps per SM.
  const uint32_t pieces_mode = env_u32("DEFL_SPLIT_PIECES", 2, 0, 2);
  const bool pieces = pieces_mode == 1u || (pieces_mode == 2u && (uint64_t)ncand * 5u <= (uint64_t)E->sms * 48u * 4u);
  std::vector<ChainItem> chain;
  std::vector<uint64_t>  off;                                      // per stream nchunk + 1 entries
  std::vector<uint64_t>  end_bit(ns, 0);
  uint64_t sym_total = 0, blk_total = 0;
  uint32_t win_total = 0;
  for (uint32_t i = 0; i < ns; i++) {
    sd[i].off_base = (uint32_t)off.size();
    sd[i].win_base = win_total;
    sd[i].blk_base = blk_total;
    sd[i].sym_base = sym_total;
    sd[i].nchunk = 0; sd[i].nblk = 0; sd[i].total = 0;
    if (!alive[i]) continue;
    const SplitChunkOut* o = outs.data() + sd[i].cand_base;
    const size_t chain0 = chain.size(), off0 = off.size();
    uint64_t total = 0;
    bool ok = true;
    uint32_t fin = FIN_NO;                 // finality of the block chunk k starts in (meaningful for mid-block starts)
    for (uint32_t k = 0;;) {
      if ((o[k].status & 0xFF) == CH_FAIL && !(fin == FIN_YES && o[k].has_eob)) { ok = false; break; }
      const bool ends_here = fin == FIN_YES && k != 0 && o[k].has_eob;       // its first block was the last one
      if (fin == FIN_YES && k != 0 && !o[k].has_eob && o[k].status != CH_LINKED) { ok = false; break; }
      const uint32_t bytes = ends_here ? o[k].eob_bytes : o[k].out_bytes;
      if (bytes >= 0x80000000u) { ok = false; break; }   // resolve_tokens<uint16_t> compares chunk-relative positions as int32
      const uint32_t ntok_k = ends_here ? o[k].eob_ntok : o[k].ntok;
      const uint32_t sf = k == 0 ? (uint32_t)FIN_NO : fin;
      uint32_t p_pos = 0, p_tok = 0;
      bool     sub = false;
      if (pieces && ntok_k != defl::kNoTokens) {
        for (uint32_t m = 0; m < o[k].npiece && m < defl::kMaxPieces; m++) {
          const uint32_t mp = o[k].piece_pos[m], mt = o[k].piece_tok[m];
          if (mp >= bytes || mt >= ntok_k || bytes - mp < kWinBytes) break;       // every piece but a stream's last holds a whole window
          if (mp - p_pos < kWinBytes || mt <= p_tok) continue;
          chain.push_back({i, k, total + p_pos, mp - p_pos, mt - p_tok, sf, false, p_tok, sub});
          off.push_back(total + p_pos);
          p_pos = mp; p_tok = mt; sub = true;
        }
      }
      chain.push_back({i, k, total + p_pos, bytes - p_pos, ntok_k == defl::kNoTokens ? ntok_k : ntok_k - p_tok, sf, ends_here, p_tok, sub});
      off.push_back(total + p_pos);
      total += bytes;
      if (total > in[i].dstlen || total > 0xFFFFFFFFull) { ok = false; break; }
      if (ends_here) { end_bit[i] = o[k].eob_bit; break; }
      if (o[k].status == CH_FINAL) { end_bit[i] = o[k].end_bit; break; }
      if (o[k].status != CH_LINKED || o[k].link <= k || o[k].link >= sd[i].ncand) { ok = false; break; }
      if (o[k].link_final != FIN_UNKNOWN) fin = o[k].link_final;            // else: still inside the same block
      k = o[k].link;
    }
    // markers index the 32 KiB before a chunk: they must exist
    if (ok && chain.size() - chain0 > 1 && off[off0 + 1] < kWinBytes) ok = false;
    if (!ok) {
      if (trace) std::fprintf(stderr, "[split] given back: stream %u (%llu -> %llu bytes, mode %u): chain of %zu chunks broke after %llu bytes\n", i,
                              (unsigned long long)in[i].srclen, (unsigned long long)in[i].dstlen, sd[i].mode, chain.size() - chain0, (unsigned long long)total);
      alive[i] = 0; chain.resize(chain0); off.resize(off0); continue;
    }
    off.push_back(total);
    sd[i].nchunk = (uint32_t)(chain.size() - chain0);
    sd[i].total  = total;
    sd[i].nblk   = (uint32_t)((total + kResolvePerBlock - 1) / kResolvePerBlock);
    win_total += sd[i].nchunk - 1;
    blk_total += sd[i].nblk;
    sym_total += (total + 7) & ~7ull;
  }
  const uint32_t nchunk = (uint32_t)chain.size();
  if (!nchunk) return UNZ_OK;

  // ---- C: the chains decode into 16-bit symbols at final offsets; largest chunk first (the pass ends when
  // its slowest warp does)
  if (jobs.size() < nchunk) jobs.resize(nchunk);
  if (SB.d_jobs.reserve((size_t)nchunk * sizeof(SplitJob)) || SB.d_outs2.reserve((size_t)nchunk * sizeof(SplitChunkOut))) return UNZ_ERR;
  std::vector<uint32_t> perm(nchunk);
  std::iota(perm.begin(), perm.end(), 0u);
  auto out_of = [&](uint32_t c) -> const SplitChunkOut& { return outs[sd[chain[c].stream].cand_base + chain[c].cand]; };
  // a replayed chunk costs about its tokens, a re-decoded one (no kept tokens) about its bytes
  auto c_weight = [&](uint32_t c) -> uint64_t {
    return chain[c].ntok != kNoTokens ? (uint64_t)chain[c].ntok * 4u + chain[c].bytes / 4u : (uint64_t)chain[c].bytes * 2u;
  };
  std::stable_sort(perm.begin(), perm.end(), [&](uint32_t a, uint32_t b) { return c_weight(a) > c_weight(b); });
  for (uint32_t j = 0; j < nchunk; j++) {
    const ChainItem& ci = chain[perm[j]];
    const SplitStreamDesc& d = sd[ci.stream];
    jobs[j].stream = ci.stream; jobs[j].start_final = ci.start_final; jobs[j].cand = ci.cand;
    jobs[j].cap = ci.bytes; jobs[j].out_off = d.sym_base + ci.off;
    jobs[j].tok_off = d.tok_base + (uint64_t)tok_per_byte * (cand[d.cand_base + ci.cand] >> 3) + ci.tok0;
    jobs[j].tok_cap = (ci.sub || ci.cand != 0u) ? 32768u : 0u;                 // replay: symbols of unknown history
    jobs[j].ntok = d_tok ? ci.ntok : kNoTokens;
  }
  if (SB.d_sym.reserve((size_t)sym_total * 2 + 64) || SB.d_off.reserve(off.size() * 8) ||
      SB.d_wbuf.reserve((size_t)win_total * kWinBytes + 64) || SB.d_partial.reserve((size_t)blk_total * 8 + 8) ||
      upload_streams() != UNZ_OK)
    return UNZ_ERR;
  CU_TRY(cudaMemcpyAsync(SB.d_jobs.p, jobs.data(), (size_t)nchunk * sizeof(SplitJob), cudaMemcpyHostToDevice, st), "H2D jobs");
  CU_TRY(cudaMemcpyAsync(SB.d_off.p, off.data(), off.size() * 8, cudaMemcpyHostToDevice, st), "H2D offsets");
  if (launch_chunks(false, nchunk, static_cast<uint16_t*>(SB.d_sym.p)) != UNZ_OK) return UNZ_ERR;
  mark("C decode pass");
  if (trace) {
    uint64_t nre = 0, bre = 0, nfx = 0;
    for (uint32_t c = 0; c < nchunk; c++) {
      if (chain[c].ntok == kNoTokens) { nre++; bre += chain[c].bytes; if (sd[chain[c].stream].mode == MODE_FIXED_MID) nfx++; }
    }
    std::fprintf(stderr, "[split] streams %u ncand %u nchunk %u symbols %llu; decoded again (tokens not kept): %llu chunks "
                 "(%llu in fixed-Huffman streams), %llu bytes\n", ns, ncand, nchunk, (unsigned long long)sym_total,
                 (unsigned long long)nre, (unsigned long long)nfx, (unsigned long long)bre);
  }
  std::vector<SplitChunkOut> outs2(nchunk);
  CU_TRY(cudaMemcpyAsync(outs2.data(), SB.d_outs2.p, (size_t)nchunk * sizeof(SplitChunkOut), cudaMemcpyDeviceToHost, st), "D2H chunk results");

  // ---- D: windows, then markers -> bytes + Adler-32 partial sums
  std::vector<uint32_t> partial((size_t)blk_total * 2);
  if (win_total) {
    // long chains of chunks (one history over hundreds of MB): pointer doubling over the windows; else every
    // marker of a window is followed back directly (few hops)
    uint32_t max_win = 0;
    bool     frames_ok = true;
    for (size_t p = 0, run = 0; p < chain.size(); p++) {
      const bool first = p == 0 || chain[p - 1].stream != chain[p].stream;
      const bool lastc = p + 1 == chain.size() || chain[p + 1].stream != chain[p].stream;
      run = first ? 0 : run + 1;                              // index of the window before this chunk, + 1
      max_win = std::max<uint32_t>(max_win, (uint32_t)run);
      if (!first && !lastc && chain[p].bytes < kWinBytes) frames_ok = false;
    }
    const uint32_t dbl_from = env_u32("DEFL_SPLIT_DOUBLING_FROM", 96, 1, 1u << 30);
    if (frames_ok && max_win >= dbl_from) {
      const size_t tbytes = (size_t)win_total * kWinBytes * 2;
      if (SB.d_wdbl.reserve(2 * tbytes + 64)) return UNZ_ERR;
      uint16_t* T[2] = {static_cast<uint16_t*>(SB.d_wdbl.p), reinterpret_cast<uint16_t*>(static_cast<uint8_t*>(SB.d_wdbl.p) + tbytes)};
      window_double_init_kernel<<<win_total * 16u, 256, 0, st>>>(d_sd, ns, static_cast<uint16_t*>(SB.d_sym.p),
                                                                 static_cast<uint64_t*>(SB.d_off.p), T[0]);
      CU_TRY(cudaGetLastError(), "window_double_init_kernel launch");
      g_launches.fetch_add(1, std::memory_order_relaxed);
      int cur = 0;
      for (uint32_t hop = 1;; hop <<= 1) {
        const bool last = (uint64_t)hop * 2 > max_win;      // after this round hop' = 2 * hop > every window index
        window_double_round_kernel<<<win_total * 16u, 256, 0, st>>>(d_sd, ns, T[cur], T[cur ^ 1], hop,
                                                                    last ? static_cast<uint8_t*>(SB.d_wbuf.p) : nullptr);
        CU_TRY(cudaGetLastError(), "window_double_round_kernel launch");
        g_launches.fetch_add(1, std::memory_order_relaxed);
        cur ^= 1;
        if (last) break;
      }
    } else {
      window_resolve_kernel<<<win_total * 16u, 256, 0, st>>>(d_sd, ns, static_cast<uint16_t*>(SB.d_sym.p),
                                                             static_cast<uint64_t*>(SB.d_off.p), static_cast<uint8_t*>(SB.d_wbuf.p));
      CU_TRY(cudaGetLastError(), "window_resolve_kernel launch");
      g_launches.fetch_add(1, std::memory_order_relaxed);
    }
  }
  if (blk_total) {
    marker_resolve_kernel<<<(unsigned)blk_total, 256, 0, st>>>(d_sd, ns, static_cast<uint16_t*>(SB.d_sym.p),
                                                                static_cast<uint64_t*>(SB.d_off.p), static_cast<uint8_t*>(SB.d_wbuf.p),
                                                                static_cast<uint32_t*>(SB.d_partial.p));
    CU_TRY(cudaGetLastError(), "marker_resolve_kernel launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU_TRY(cudaMemcpyAsync(partial.data(), SB.d_partial.p, (size_t)blk_total * 8, cudaMemcpyDeviceToHost, st), "D2H adler partials");
  }
  std::vector<uint8_t>  trailer((size_t)ns * 4, 0);
  std::vector<uint64_t> inpos(ns, 0);
  if (zlib) {
    for (uint32_t i = 0; i < ns; i++) inpos[i] = (end_bit[i] + 7) >> 3;
    CU_TRY(cudaMemcpyAsync(SB.d_at.p, inpos.data(), (size_t)ns * 8, cudaMemcpyHostToDevice, st), "H2D trailer offsets");
    split_peek_kernel<<<(ns + 127) / 128, 128, 0, st>>>(d_sd, ns, static_cast<uint64_t*>(SB.d_at.p), 4u, static_cast<uint8_t*>(SB.d_peek.p));
    CU_TRY(cudaGetLastError(), "split_peek_kernel launch");
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU_TRY(cudaMemcpyAsync(trailer.data(), SB.d_peek.p, trailer.size(), cudaMemcpyDeviceToHost, st), "D2H trailers");
  }
  CU_TRY(cudaStreamSynchronize(st), "split decode");
  mark("D marker resolve + adler");

  // the decode pass must have reproduced the count pass (it cannot differ for a deterministic decode; the
  // one-warp decoder has the last word if it ever does)
  for (uint32_t j = 0; j < nchunk; j++) {
    const ChainItem& ci = chain[perm[j]];
    const SplitChunkOut& a = out_of(perm[j]);
    bool same;
    if (outs2[j].status == CH_REPLAYED) same = outs2[j].out_bytes == ci.bytes;
    else if (ci.ends) same = outs2[j].status == CH_FINAL && outs2[j].out_bytes == ci.bytes && outs2[j].end_bit == a.eob_bit;
    else same = outs2[j].status == a.status && outs2[j].out_bytes == a.out_bytes && outs2[j].end_bit == a.end_bit;
    if (!same) alive[ci.stream] = 0;
  }
  for (uint32_t i = 0; i < ns; i++) {
    if (!alive[i]) continue;
    uint64_t sa = 0, sw = 0;
    for (uint64_t b = sd[i].blk_base; b < sd[i].blk_base + sd[i].nblk; b++) { sa += partial[2 * b]; sw += partial[2 * b + 1]; }
    sa %= defl::kAdlerMod; sw %= defl::kAdlerMod;
    const uint64_t total = sd[i].total;
    const uint64_t nm = total % defl::kAdlerMod;
    const uint32_t a = (uint32_t)((1 + sa) % defl::kAdlerMod);
    const uint32_t b = (uint32_t)((nm + nm * sa + defl::kAdlerMod - sw) % defl::kAdlerMod);
    infl_batch_result_t* r = &res[i];
    std::memset(r, 0, sizeof(*r));
    r->code = UNZ_OK;
    r->adler = (b << 16) | a;
    r->outlen = total;
    r->inpos = (end_bit[i] + 7) >> 3;
    r->info = INFL_INFO_SPLIT_DECODE;
    const bool has_trailer = zlib && in[i].srclen - r->inpos >= 4;
    if (has_trailer) {
      const uint8_t* t = &trailer[(size_t)i * 4];
      r->adler_stored = ((uint32_t)t[0] << 24) | ((uint32_t)t[1] << 16) | ((uint32_t)t[2] << 8) | t[3];
      r->info |= INFL_INFO_HAS_TRAILER;
      if (r->adler_stored == r->adler) r->info |= INFL_INFO_ADLER_MATCH;
    }
    if (zlib && (flags & INFL_BATCH_VERIFY_ADLER) && (r->info & 3u) != 3u) {
      r->code = UNZ_ERR;
      r->info |= defl::R_ADLER << INFL_INFO_REASON_SHIFT;
    }
    status[i] = UNZ_OK;
    g_split_decodes.fetch_add(1, std::memory_order_relaxed);
  }
  return UNZ_OK;
}

// one stream: UNZ_OK (decoded, *res filled) / UNZ_NOOP (take the one-warp decoder) / UNZ_ERR (engine failure)
int split_decode_device(Engine* E, const uint8_t* d_src, uint64_t srclen, uint8_t* d_dst, uint64_t dstlen, int flags,
                        void* d_ws, cudaStream_t st, infl_batch_result_t* res) {
  SplitInput in{d_src, srclen, d_dst, dstlen};
  int status = UNZ_NOOP;
  if (split_decode_many(E, &in, 1, flags, d_ws, st, res, &status) != UNZ_OK) return UNZ_ERR;
  return status;
}

// Which streams of a batch go through the split decoder: the very large ones always; in a batch, those whose
// one-warp decode would outlast the rest of the batch (bytes x resident warps > half the batch's bytes).
static const uint64_t kSplitMinInBatch = 192ull << 10;    // compressed bytes below which splitting never pays
bool wants_split(const infl_batch_item_t& it, uint64_t batch_bytes, uint64_t resident_warps) {
  if (it.srclen > 0xFFFFFFFFull || it.dstlen > 0xFFFFFFFFull) return false;
  if (it.srclen >= kSplitMinBytes) return true;
  if (it.srclen < kSplitMinInBatch) return false;
  // DEFL_SPLIT_TAIL_PCT: a stream is split when its share of the batch exceeds this percentage of 1 / resident warps
  static const uint32_t pct = env_u32("DEFL_SPLIT_TAIL_PCT", 55u, 1u, 100000u);
  return (it.srclen + it.dstlen) * resident_warps * 100u > batch_bytes * pct;
}

// ---- host-pointer batch on ONE device: stage in, decode, stage out ----------------------------------------
// The batch is cut into chunks that run through a three-slot pipeline: while chunk c decodes, chunk c+1's
// input is on its way to the device and chunk c-1's output on its way back (separate copy engines).
// Inputs that are (nearly) adjacent in host memory travel as one copy; adjacent outputs do too.
struct HostRun { const uint8_t* h; size_t len; size_t dev_off; size_t first; size_t count; };

struct PipeSlot {
  cudaStream_t st = nullptr;
  DevBuf       d_in, d_out, d_items, d_results, d_order, d_ws;
  PinBuf       h_items, h_results, h_order;
  std::vector<HostRun> out_runs;
  std::vector<size_t>  out_off;
  size_t first = 0, count = 0;
};

constexpr int kPipeSlots = 6;    // at most; DEFL_PIPE_SLOTS (default 3) are used: chunk c+1 is on its way in while chunk c
                                 // decodes and chunk c-1 is on its way out
struct HostPipe {
  PipeSlot slot[kPipeSlots];
  bool     ready = false;
};
HostPipe g_pipes[kMaxDevices];

constexpr size_t kGapMerge   = 512;          // input runs this close in host memory are copied as one
constexpr size_t kChunkMin   = 64u << 20;    // bytes (in + out) per chunk, at least
constexpr size_t kChunkCount = 32;           // aim for this many chunks in a large batch (the first chunk's way in and the
                                             // last one's way out overlap with nothing: keep them small)

int chunk_submit(Engine* E, PipeSlot& S, const infl_batch_item_t* items, size_t first, size_t count, int flags) {
  cudaStream_t st = S.st;
  std::vector<HostRun> in_runs;
  std::vector<size_t>  in_off(count);
  S.out_runs.clear();
  S.out_off.assign(count, 0);
  S.first = first; S.count = count;
  size_t in_total = 0, out_total = 0;
  for (size_t k = 0; k < count; k++) {
    const infl_batch_item_t& it = items[first + k];
    const uint8_t* s = static_cast<const uint8_t*>(it.src);
    if (it.srclen == 0) {
      // nothing to copy, and its pointer may be anything: it neither starts nor extends a run
      in_off[k] = 0;
    } else if (!in_runs.empty() && s >= in_runs.back().h + in_runs.back().len &&
        (size_t)(s - (in_runs.back().h + in_runs.back().len)) <= kGapMerge) {
      // the gap (<= 512 bytes) lies between two non-empty buffers of the caller, so inside the page of the last byte
      // of the one or of the first byte of the other: mapped memory.  Copy it along.
      HostRun& r = in_runs.back();
      in_off[k] = r.dev_off + (size_t)(s - r.h);
      r.len = (size_t)(s - r.h) + it.srclen;
      r.count++;
    } else {
      // keep the device copy congruent to the host pointer mod 16 (same alignment behaviour)
      in_total = align_up(in_total) + ((uintptr_t)s & (kAlign - 1));
      in_runs.push_back({s, (size_t)it.srclen, in_total, k, 1});
      in_off[k] = in_total;
    }
    if (!in_runs.empty()) in_total = in_runs.back().dev_off + in_runs.back().len;

    const uint8_t* d = static_cast<const uint8_t*>(it.dst);
    if (!S.out_runs.empty() && d == S.out_runs.back().h + S.out_runs.back().len) {
      S.out_off[k] = S.out_runs.back().dev_off + S.out_runs.back().len;
      S.out_runs.back().len += it.dstlen;
      S.out_runs.back().count++;
    } else {
      out_total = align_up(out_total) + ((uintptr_t)d & (kAlign - 1));
      S.out_runs.push_back({d, (size_t)it.dstlen, out_total, k, 1});
      S.out_off[k] = out_total;
    }
    out_total = S.out_runs.back().dev_off + S.out_runs.back().len;
  }

  if (S.d_in.reserve(in_total + 64) || S.d_out.reserve(out_total + 64) ||
      S.d_items.reserve(count * sizeof(infl_batch_item_t)) || S.d_results.reserve(count * sizeof(infl_batch_result_t)) ||
      S.d_order.reserve(count * sizeof(uint32_t)) || S.d_ws.reserve_exact(warp_kernel_workspace(E->sms)) ||
      S.h_items.reserve(count * sizeof(infl_batch_item_t)) || S.h_results.reserve(count * sizeof(infl_batch_result_t)) ||
      S.h_order.reserve(count * sizeof(uint32_t)))
    return UNZ_ERR;

  auto* hi = static_cast<infl_batch_item_t*>(S.h_items.p);
  auto* ho = static_cast<uint32_t*>(S.h_order.p);
  uint8_t* din  = static_cast<uint8_t*>(S.d_in.p);
  uint8_t* dout = static_cast<uint8_t*>(S.d_out.p);
  for (size_t k = 0; k < count; k++) {
    hi[k].src = din + in_off[k];     hi[k].srclen = items[first + k].srclen;
    hi[k].dst = dout + S.out_off[k]; hi[k].dstlen = items[first + k].dstlen;
  }
  size_order(items + first, count, ho);

  for (const HostRun& r : in_runs)
    if (r.len) CU_TRY(cudaMemcpyAsync(din + r.dev_off, r.h, r.len, cudaMemcpyHostToDevice, st), "H2D input");
  CU_TRY(cudaMemcpyAsync(S.d_items.p, hi, count * sizeof(*hi), cudaMemcpyHostToDevice, st), "H2D descriptors");
  CU_TRY(cudaMemcpyAsync(S.d_order.p, ho, count * sizeof(*ho), cudaMemcpyHostToDevice, st), "H2D order");
  if (launch_batch_any(E->sms, static_cast<infl_batch_item_t*>(S.d_items.p), static_cast<uint32_t*>(S.d_order.p),
                       count, flags, static_cast<infl_batch_result_t*>(S.d_results.p), S.d_ws.p, st) != UNZ_OK)
    return UNZ_ERR;
  CU_TRY(cudaMemcpyAsync(S.h_results.p, S.d_results.p, count * sizeof(infl_batch_result_t), cudaMemcpyDeviceToHost, st),
         "D2H results");
  // the decoded bytes follow in the same stream, whole runs of adjacent destinations at once: no host round trip
  // between the kernel and the copy.  A stream that failed or produced less than its buffer holds leaves
  // unspecified bytes after its outlen, as the reference does after its produced length (infl.c:269-376 write
  // past the end of a match inside dst).
  for (const HostRun& r : S.out_runs)
    if (r.len)
      CU_TRY(cudaMemcpyAsync(const_cast<uint8_t*>(r.h), dout + r.dev_off, r.len, cudaMemcpyDeviceToHost, st), "D2H output");
  return UNZ_OK;
}

// wait for the chunk's results, then send its outputs home (asynchronously on the slot's stream)
int chunk_finish(PipeSlot& S, const infl_batch_item_t* items, infl_batch_result_t* results) {
  cudaStream_t st = S.st;
  CU_TRY(cudaStreamSynchronize(st), "inflate kernel");
  auto* hr = static_cast<infl_batch_result_t*>(S.h_results.p);
  uint8_t* dout = static_cast<uint8_t*>(S.d_out.p);
  std::memcpy(results + S.first, hr, S.count * sizeof(*hr));
  return UNZ_OK;
}

int run_device_batch_locked(Engine* E, const infl_batch_item_t* items, size_t n, int flags, infl_batch_result_t* results);

int run_host_batch_on_device(const infl_batch_item_t* items, size_t n, int flags,
                             infl_batch_result_t* results) {
  Engine* E = nullptr;
  if (engine_for_current_device(&E) != UNZ_OK) return UNZ_ERR;
  std::unique_lock<std::mutex> guard(E->mu);
  HostPipe& HP = g_pipes[E->dev];
  if (!HP.ready) {
    for (PipeSlot& S : HP.slot) CU_TRY(cudaStreamCreateWithFlags(&S.st, cudaStreamNonBlocking), "cudaStreamCreate");
    HP.ready = true;
  }

  size_t total = 0;
  for (size_t i = 0; i < n; i++) {
    if (items[i].srclen > 0xFFFFFFFFull || items[i].dstlen > 0xFFFFFFFFull)
      return fail("stream larger than 4 GiB - 1 (the reference API limit)");
    total += items[i].srclen + items[i].dstlen;
  }
  // large streams first: groups of them (about 512 MiB of output each) go through the device batch path TOGETHER -- one
  // split decode for the whole group, which fills the GPU and pays the decoder's host round trips once -- three slots
  // deep: while group g decodes, group g + 1 is already on its way in and group g - 1 on its way home (each slot has its
  // own stream and buffers; a slot is reused once the stream of its last user has drained)
  if (!(flags & INFL_BATCH_NO_SPLIT)) {
    bool any = false;
    for (size_t i = 0; i < n; i++) any = any || items[i].srclen >= kSplitMinBytes;
    if (any) {
      std::vector<infl_batch_item_t> rest;
      std::vector<size_t>            rest_idx, big;
      for (size_t i = 0; i < n; i++) {
        if (items[i].srclen >= kSplitMinBytes) big.push_back(i);
        else { rest.push_back(items[i]); rest_idx.push_back(i); }
      }
      constexpr size_t kBigSlots = 3;
      static_assert(kBigSlots <= (size_t)kPipeSlots, "slots of the host pipeline");
      const size_t kGroupOut = (size_t)env_u32("DEFL_BIG_GROUP_MIB", 512u, 16u, 1u << 20) << 20;     // experiments: profiles/README.md
      std::vector<size_t> gfirst;                                    // group g = big[gfirst[g] .. gfirst[g + 1])
      for (size_t k = 0, out = 0; k < big.size(); k++) {
        if (k == 0 || out + items[big[k]].dstlen > kGroupOut) { gfirst.push_back(k); out = 0; }
        out += items[big[k]].dstlen;
      }
      gfirst.push_back(big.size());
      const size_t ngroups = gfirst.size() - 1;
      std::vector<std::vector<infl_batch_item_t>> dev(ngroups);
      static cudaEvent_t in_done[kMaxDevices][kBigSlots] = {};
      auto upload = [&](size_t g) -> int {                          // H2D of group g into its slot
        PipeSlot& S = HP.slot[g % kBigSlots];
        CU_TRY(cudaStreamSynchronize(S.st), "large members: slot drain");         // group g - 3 is home
        size_t in_bytes = 0, out_bytes = 0;
        for (size_t k = gfirst[g]; k < gfirst[g + 1]; k++) {
          in_bytes  = align_up(in_bytes) + kAlign + items[big[k]].srclen;
          out_bytes = align_up(out_bytes) + items[big[k]].dstlen;
        }
        if (S.d_in.reserve(in_bytes + 64) || S.d_out.reserve(align_up(out_bytes) + 64)) return UNZ_ERR;
        size_t ia = 0, oa = 0;
        for (size_t k = gfirst[g]; k < gfirst[g + 1]; k++) {
          const infl_batch_item_t& it = items[big[k]];
          // same alignment of the device copy as the host pointer's
          ia = align_up(ia);
          uint8_t* din = static_cast<uint8_t*>(S.d_in.p) + ia + ((uintptr_t)it.src & (kAlign - 1));
          oa = align_up(oa);
          CU_TRY(cudaMemcpyAsync(din, it.src, it.srclen, cudaMemcpyHostToDevice, S.st), "H2D input");
          infl_batch_item_t dv;
          dv.src = din; dv.srclen = it.srclen; dv.dst = static_cast<uint8_t*>(S.d_out.p) + oa; dv.dstlen = it.dstlen;
          dev[g].push_back(dv);
          ia += kAlign + it.srclen;
          oa += it.dstlen;
        }
        cudaEvent_t& ev = in_done[E->dev][g % kBigSlots];
        if (!ev) CU_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming), "cudaEventCreate");
        CU_TRY(cudaEventRecord(ev, S.st), "cudaEventRecord");
        return UNZ_OK;
      };
      if (ngroups && upload(0) != UNZ_OK) return UNZ_ERR;
      for (size_t g = 0; g < ngroups; g++) {
        if (g + 1 < ngroups && upload(g + 1) != UNZ_OK) return UNZ_ERR;
        PipeSlot& S = HP.slot[g % kBigSlots];
        CU_TRY(cudaStreamWaitEvent(E->stream, in_done[E->dev][g % kBigSlots], 0), "cudaStreamWaitEvent");
        std::vector<infl_batch_result_t> gr(dev[g].size());
        if (run_device_batch_locked(E, dev[g].data(), dev[g].size(), flags, gr.data()) != UNZ_OK) return UNZ_ERR;
        for (size_t k = gfirst[g], j = 0; k < gfirst[g + 1]; k++, j++) {
          results[big[k]] = gr[j];
          if (gr[j].outlen)                                          // what was produced goes home, failed or not
            CU_TRY(cudaMemcpyAsync(items[big[k]].dst, dev[g][j].dst, gr[j].outlen, cudaMemcpyDeviceToHost, S.st), "D2H output");
        }
      }
      for (size_t s2 = 0; s2 < kBigSlots; s2++) CU_TRY(cudaStreamSynchronize(HP.slot[s2].st), "D2H output");
      if (rest.empty()) return UNZ_OK;
      std::vector<infl_batch_result_t> rr(rest.size());
      guard.unlock();
      const int rc2 = run_host_batch_on_device(rest.data(), rest.size(), flags | INFL_BATCH_NO_SPLIT, rr.data());
      for (size_t k = 0; k < rr.size(); k++) results[rest_idx[k]] = rr[k];
      return rc2;
    }
  }
  static const size_t chunk_count = env_u32("DEFL_CHUNK_COUNT", (uint32_t)kChunkCount, 1, 4096);
  static const int    slots = (int)env_u32("DEFL_PIPE_SLOTS", 3, 1, kPipeSlots);
  const size_t chunk_bytes = std::max(kChunkMin, total / chunk_count + 1);

  size_t first = 0;
  int    c = 0, rc = UNZ_OK;
  while (first < n && rc == UNZ_OK) {
    size_t count = 0, bytes = 0;
    while (first + count < n && (count == 0 || bytes < chunk_bytes)) {
      bytes += items[first + count].srclen + items[first + count].dstlen;
      count++;
    }
    PipeSlot& S = HP.slot[c % slots];
    if (c >= slots) rc = chunk_finish(S, items, results);           // the chunk that used this slot is home
    if (rc == UNZ_OK) rc = chunk_submit(E, S, items, first, count, flags);
    first += count;
    c++;
  }
  for (int k = std::max(0, c - slots); k < c && rc == UNZ_OK; k++) rc = chunk_finish(HP.slot[k % slots], items, results);
  for (PipeSlot& S : HP.slot) {
    const cudaError_t e = cudaStreamSynchronize(S.st);
    if (e != cudaSuccess && rc == UNZ_OK) rc = fail("D2H output", e);
  }
  return rc;
}

// ---- device-pointer batch: only descriptors and results cross the bus ---------------------------------------
// device-resident streams -> results (host array).  The engine's mutex is held by the caller.
int run_device_batch_locked(Engine* E, const infl_batch_item_t* items, size_t n, int flags, infl_batch_result_t* results) {
  cudaStream_t st = E->stream;
  if (E->d_items.reserve(n * sizeof(infl_batch_item_t)) || E->d_results.reserve(n * sizeof(infl_batch_result_t)) ||
      E->d_order.reserve(n * sizeof(uint32_t)) || E->d_ws.reserve_exact(warp_kernel_workspace(E->sms)) ||
      E->h_order.reserve(n * sizeof(uint32_t)))
    return UNZ_ERR;
  // streams that would outlast the rest of the batch on one warp go through the split decoder, all of them
  // together; the rest is one batch launch
  std::vector<infl_batch_item_t> rest;
  std::vector<size_t>            rest_idx;
  bool any_split = false;
  uint64_t batch_bytes = 0;
  int grid_ctas = 0;
  if (warp_kernel_grid(E->sms, &grid_ctas) != UNZ_OK) return UNZ_ERR;
  const uint64_t resident = (uint64_t)grid_ctas * kWarpsPerCta;
  for (size_t i = 0; i < n; i++) batch_bytes += items[i].srclen + items[i].dstlen;
  if (!(flags & INFL_BATCH_NO_SPLIT))
    for (size_t i = 0; i < n && !any_split; i++) any_split = wants_split(items[i], batch_bytes, resident);
  if (any_split) {
    std::vector<SplitInput> sin;
    std::vector<size_t>     sin_idx;
    for (size_t i = 0; i < n; i++) {
      if (wants_split(items[i], batch_bytes, resident)) {
        sin.push_back({static_cast<const uint8_t*>(items[i].src), items[i].srclen, static_cast<uint8_t*>(items[i].dst), items[i].dstlen});
        sin_idx.push_back(i);
      } else {
        rest.push_back(items[i]); rest_idx.push_back(i);
      }
    }
    // The members that stay on one warp start first, on their own stream and scratch: the batch kernel is
    // persistent, so the split kernels (queued behind it) take over the SMs as its CTAs run out of streams, and
    // the long tail of its largest members is hidden behind them.
    const bool trace = std::getenv("DEFL_SPLIT_TRACE") != nullptr;
    const auto t_begin = std::chrono::steady_clock::now();
    std::vector<uint32_t> ord;
    std::vector<char>     early(sin.size(), 0);      // given back right after the peek: they join the one-warp launch
    bool launched = false;
    auto launch_rest = [&](const std::vector<char>& alive) -> int {
      launched = true;
      for (size_t k = 0; k < alive.size() && k < sin.size(); k++)
        if (!alive[k]) { early[k] = 1; rest.push_back(items[sin_idx[k]]); rest_idx.push_back(sin_idx[k]); }
      if (rest.empty()) return UNZ_OK;
      ord.resize(rest.size());
      if (E->d_items2.reserve(rest.size() * sizeof(infl_batch_item_t)) || E->d_results2.reserve(rest.size() * sizeof(infl_batch_result_t)) ||
          E->d_order2.reserve(rest.size() * sizeof(uint32_t)) || E->d_ws2.reserve_exact(warp_kernel_workspace(E->sms)))
        return UNZ_ERR;
      size_order(rest.data(), rest.size(), ord.data());
      CU_TRY(cudaMemcpyAsync(E->d_items2.p, rest.data(), rest.size() * sizeof(rest[0]), cudaMemcpyHostToDevice, E->stream2), "H2D descriptors");
      CU_TRY(cudaMemcpyAsync(E->d_order2.p, ord.data(), rest.size() * sizeof(ord[0]), cudaMemcpyHostToDevice, E->stream2), "H2D order");
      return launch_batch_any(E->sms, static_cast<infl_batch_item_t*>(E->d_items2.p), static_cast<uint32_t*>(E->d_order2.p),
                              rest.size(), flags, static_cast<infl_batch_result_t*>(E->d_results2.p), E->d_ws2.p, E->stream2);
    };
    const std::function<int(const std::vector<char>&)> hook = launch_rest;
    std::vector<infl_batch_result_t> sres(sin.size());
    std::vector<int>                 sstat(sin.size(), UNZ_NOOP);
    int src = split_decode_many(E, sin.data(), sin.size(), flags, E->d_ws.p, st, sres.data(), sstat.data(), &hook);
    if (!launched && src == UNZ_OK) src = launch_rest(std::vector<char>(sin.size(), 1));   // the split path left before its peek
    const auto t_split = std::chrono::steady_clock::now();
    if (!rest.empty()) {
      std::vector<infl_batch_result_t> rr(rest.size());
      CU_TRY(cudaMemcpyAsync(rr.data(), E->d_results2.p, rr.size() * sizeof(rr[0]), cudaMemcpyDeviceToHost, E->stream2), "D2H results");
      CU_TRY(cudaStreamSynchronize(E->stream2), "inflate kernel");
      for (size_t k = 0; k < rr.size(); k++) results[rest_idx[k]] = rr[k];
    }
    if (src != UNZ_OK) return UNZ_ERR;
    const auto t_rest = std::chrono::steady_clock::now();
    // members the split path gave back (no usable split points, or damaged): one warp each, as if never tried
    std::vector<infl_batch_item_t> back;
    std::vector<size_t>            back_idx;
    for (size_t k = 0; k < sin.size(); k++) {
      if (sstat[k] == UNZ_OK) results[sin_idx[k]] = sres[k];
      else if (!early[k]) { back.push_back(items[sin_idx[k]]); back_idx.push_back(sin_idx[k]); }
    }
    if (!back.empty()) {
      std::vector<infl_batch_result_t> rr(back.size());
      auto* ho2 = static_cast<uint32_t*>(E->h_order.p);
      size_order(back.data(), back.size(), ho2);
      CU_TRY(cudaMemcpyAsync(E->d_items.p, back.data(), back.size() * sizeof(back[0]), cudaMemcpyHostToDevice, st), "H2D descriptors");
      CU_TRY(cudaMemcpyAsync(E->d_order.p, ho2, back.size() * sizeof(*ho2), cudaMemcpyHostToDevice, st), "H2D order");
      if (launch_batch_any(E->sms, static_cast<infl_batch_item_t*>(E->d_items.p), static_cast<uint32_t*>(E->d_order.p),
                           back.size(), flags, static_cast<infl_batch_result_t*>(E->d_results.p), E->d_ws.p, st) != UNZ_OK)
        return UNZ_ERR;
      CU_TRY(cudaMemcpyAsync(rr.data(), E->d_results.p, rr.size() * sizeof(rr[0]), cudaMemcpyDeviceToHost, st), "D2H results");
      CU_TRY(cudaStreamSynchronize(st), "inflate kernel");
      for (size_t k = 0; k < rr.size(); k++) results[back_idx[k]] = rr[k];
    }
    if (trace) {
      auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double, std::milli>(b - a).count();
      };
      std::fprintf(stderr, "[split] batch: %zu members on one warp each beside %zu split-decoded: split path done at %.3f ms, "
                           "one-warp kernel at %.3f ms, %zu given back done at %.3f ms\n", rest.size(), sin.size() - back.size(),
                   ms(t_begin, t_split), ms(t_begin, t_rest), back.size(), ms(t_begin, std::chrono::steady_clock::now()));
    }
    return UNZ_OK;
  }
  auto* ho = static_cast<uint32_t*>(E->h_order.p);
  size_order(items, n, ho);
  CU_TRY(cudaMemcpyAsync(E->d_items.p, items, n * sizeof(*items), cudaMemcpyHostToDevice, st), "H2D descriptors");
  CU_TRY(cudaMemcpyAsync(E->d_order.p, ho, n * sizeof(*ho), cudaMemcpyHostToDevice, st), "H2D order");
  if (launch_batch_any(E->sms, static_cast<infl_batch_item_t*>(E->d_items.p), static_cast<uint32_t*>(E->d_order.p),
                       n, flags, static_cast<infl_batch_result_t*>(E->d_results.p), E->d_ws.p, st) != UNZ_OK)
    return UNZ_ERR;
  CU_TRY(cudaMemcpyAsync(results, E->d_results.p, n * sizeof(*results), cudaMemcpyDeviceToHost, st), "D2H results");
  CU_TRY(cudaStreamSynchronize(st), "inflate kernel");
  return UNZ_OK;
}

int run_device_batch(const infl_batch_item_t* items, size_t n, int flags, infl_batch_result_t* results) {
  Engine* E = nullptr;
  if (engine_for_current_device(&E) != UNZ_OK) return UNZ_ERR;
  std::lock_guard<std::mutex> guard(E->mu);
  return run_device_batch_locked(E, items, n, flags, results);
}

// ---- multi-GPU sharding of a host batch: byte-balanced, no collective ----------------------------
// Greedy longest-processing-time assignment on B_alg = compressed + decompressed bytes.
void partition_lpt(const infl_batch_item_t* items, size_t n, int ngpu, std::vector<std::vector<uint32_t>>& parts) {
  std::vector<uint32_t> idx(n);
  std::iota(idx.begin(), idx.end(), 0u);
  auto weight = [&](uint32_t i) { return items[i].srclen + items[i].dstlen; };
  std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return weight(a) > weight(b); });
  std::vector<uint64_t> load(ngpu, 0);
  parts.assign(ngpu, {});
  for (uint32_t i : idx) {
    int g = (int)(std::min_element(load.begin(), load.end()) - load.begin());
    parts[g].push_back(i);
    load[g] += weight(i) + 1;
  }
  for (auto& p : parts) std::sort(p.begin(), p.end());  // keep host adjacency => few large copies
}

int run_host_batch_all_gpus(const infl_batch_item_t* items, size_t n, int flags, infl_batch_result_t* results) {
  int ngpu = 0;
  CU_TRY(cudaGetDeviceCount(&ngpu), "cudaGetDeviceCount");
  if (ngpu <= 1 || n < 2) return run_host_batch_on_device(items, n, flags, results);
  ngpu = std::min(ngpu, kMaxDevices);
  std::vector<std::vector<uint32_t>> parts;
  partition_lpt(items, n, ngpu, parts);
  std::vector<int>         rc(ngpu, UNZ_OK);
  std::vector<std::string> err(ngpu);
  std::vector<std::thread> th;
  for (int g = 0; g < ngpu; g++) {
    th.emplace_back([&, g]() {
      if (parts[g].empty()) return;
      if (cudaSetDevice(g) != cudaSuccess) { rc[g] = UNZ_ERR; err[g] = "cudaSetDevice"; return; }
      std::vector<infl_batch_item_t>   sub(parts[g].size());
      std::vector<infl_batch_result_t> res(parts[g].size());
      for (size_t k = 0; k < sub.size(); k++) sub[k] = items[parts[g][k]];
      rc[g] = run_host_batch_on_device(sub.data(), sub.size(), flags, res.data());
      if (rc[g] != UNZ_OK) { err[g] = g_last_error; return; }
      for (size_t k = 0; k < sub.size(); k++) results[parts[g][k]] = res[k];
    });
  }
  for (auto& t : th) t.join();
  for (int g = 0; g < ngpu; g++)
    if (rc[g] != UNZ_OK) return fail(err[g].c_str());
  return UNZ_OK;
}

}  // namespace

// =====================================================================================================
//                                         exported C ABI
// =====================================================================================================
extern "C" {

UNZ_EXPORT int infl_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

UNZ_EXPORT const char* infl_last_error(void) { return g_last_error.c_str(); }

UNZ_EXPORT const char* infl_engine_info(void) {
  static char buf[256];
  int lanes = g_lanes_per_stream.load();
  const int engine = lanes ? 1 : g_engine.load();
  std::snprintf(buf, sizeof(buf),
                "defl_b200 %d.%d sm_100a; engine=%d (%s); lanes/stream=%d; smem/warp=%zu B; window<=32x%u bits; "
                "lit root %d bits, dist root %d bits",
                DEFL_B200_VERSION_MAJOR, DEFL_B200_VERSION_MINOR, engine,
                engine == 4 ? "warp per stream v4: literals+match tokens, smem rings and stage"
                : engine == 3 ? "warp per stream, speculative sub-sequences" : engine == 2 ? "lane-decoder pipeline"
                                                                                            : "group per stream",
                lanes, engine == 4 ? sizeof(defl::v4::WarpSmem4) : sizeof(defl::WarpSmem), defl::kSubMax, defl::kLitRootBits, defl::kDistRootBits);
  return buf;
}

UNZ_EXPORT uint64_t infl_kernel_launch_count(void) { return g_launches.load(); }

UNZ_EXPORT int infl_set_lanes_per_stream(int lanes) {
#ifndef DEFL_LEGACY_ENGINES
  if (lanes != 0) return fail("the group-per-stream engine is not in this build (-DDEFL_LEGACY_ENGINES)");
#endif
  if (lanes != 0 && lanes != 1 && lanes != 2 && lanes != 4 && lanes != 8 && lanes != 16 && lanes != 32)
    return fail("lanes per stream must be 0 (default), 1, 2, 4, 8, 16 or 32");
  g_lanes_per_stream.store(lanes);
  return UNZ_OK;
}

UNZ_EXPORT int infl_set_pipeline_shape(int shape) {
  if (shape < 0 || shape > 7) return fail("pipeline shape must be 0..7");
  g_pipeline_shape.store(shape);
  return UNZ_OK;
}

UNZ_EXPORT size_t infl_batch_workspace_size(size_t n) {
  (void)n;   // the scratch is per resident warp, not per stream
  Engine* E = nullptr;
  if (engine_for_current_device(&E) != UNZ_OK) return 0;
  return warp_kernel_workspace(E->sms);
}

UNZ_EXPORT int infl_set_engine(int engine) {
#ifndef DEFL_LEGACY_ENGINES
  if (engine == 1 || engine == 2) return fail("engines 1 and 2 are not in this build (-DDEFL_LEGACY_ENGINES)");
#endif
  if (engine < 1 || engine > 4) return fail("engine must be 1 (group per stream), 2 (lane-decoder pipeline), 3 (warp per stream, v3) or 4 (v4)");
  g_engine.store(engine);
  return UNZ_OK;
}

UNZ_EXPORT int infl_batch_decode_device(const infl_batch_item_t* d_items, const uint32_t* d_order, size_t n,
                                        int flags, infl_batch_result_t* d_results, void* d_workspace,
                                        void* cuda_stream) {
  if (n == 0) return UNZ_OK;
  if (!d_items || !d_results || !d_workspace) return fail("null device pointer");
  Engine* E = nullptr;
  if (engine_for_current_device(&E) != UNZ_OK) return UNZ_ERR;
  return launch_batch_any(E->sms, d_items, d_order, n, flags, d_results, d_workspace,
                          static_cast<cudaStream_t>(cuda_stream));
}

// ---- infl_batch_decode_device_static: the batch kernel with NO allocation anywhere on the call path ----------
// The token scratch of the one-warp-per-stream kernel lives in a file-scope device array instead of a caller's
// workspace: kStaticWarps warps' worth (37 CTAs x 4 warps = one warp per SM of a B200), about 38.2 MiB that every
// GPU which loads this library holds until the process exits.  In HOST code `g_static_ws` is only a stand-in for
// the symbol, not a device address: the address comes from cudaGetSymbolAddress, once per device, with that
// device current, and that pointer is what the kernel gets.
constexpr int      kStaticCtas       = 37;
constexpr int      kStaticWarps      = kStaticCtas * kWarpsPerCta;                    // 148
constexpr size_t   kStaticWsBytes    = kWsHeader + (size_t)kStaticWarps * defl::kWarpScratchBytes;
constexpr size_t   kStaticMaxStreams = INFL_BATCH_STATIC_MAX_STREAMS;
static __device__ __align__(256) unsigned char g_static_ws[kStaticWsBytes];

struct StaticWs {
  std::atomic<unsigned char*> ptr{nullptr};      // device address of g_static_ws on this device
};
static StaticWs   g_static[kMaxDevices];
static std::mutex g_static_mu;

UNZ_EXPORT int infl_batch_decode_device_static(const infl_batch_item_t* d_items, const uint32_t* d_order, size_t n,
                                               int flags, infl_batch_result_t* d_results, void* cuda_stream) {
  if (n == 0) return UNZ_OK;
  if (!d_items || !d_results) return fail("infl_batch_decode_device_static: null device pointer");
  if (n > kStaticMaxStreams)
    return fail("infl_batch_decode_device_static: more than INFL_BATCH_STATIC_MAX_STREAMS (65536) streams in one call; "
                "use infl_batch_decode_device with a workspace");
  int dev = 0;
  CU_TRY(cudaGetDevice(&dev), "cudaGetDevice (is a CUDA device visible? this library has no CPU path)");
  if (dev < 0 || dev >= kMaxDevices) return fail("device index out of range");
  unsigned char* ws = g_static[dev].ptr.load(std::memory_order_acquire);
  if (!ws) {                                       // set-up, once per device, with that device current
    std::lock_guard<std::mutex> g(g_static_mu);
    ws = g_static[dev].ptr.load(std::memory_order_relaxed);
    if (!ws) {
      int major = 0;
      CU_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev), "cudaDeviceGetAttribute");
      if (major < 10) return fail("this build contains sm_100a code only (needs a B200-class GPU)");
      void* p = nullptr;
      CU_TRY(cudaGetSymbolAddress(&p, g_static_ws), "cudaGetSymbolAddress(g_static_ws)");
      if (!p) return fail("cudaGetSymbolAddress(g_static_ws) gave a null address");
      CU_TRY(cudaFuncSetAttribute(defl::inflate_warp_kernel<kWarpsPerCta>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)(sizeof(defl::WarpSmem) * kWarpsPerCta)), "cudaFuncSetAttribute(smem)");
      ws = static_cast<unsigned char*>(p);
      g_static[dev].ptr.store(ws, std::memory_order_release);
    }
  }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const unsigned grid = (unsigned)std::min<uint64_t>((uint64_t)kStaticCtas, (n + kWarpsPerCta - 1) / kWarpsPerCta);
  auto* ticket = reinterpret_cast<unsigned int*>(ws);
  // the array keeps the last call's contents: the ticket counter must start at zero (token scratch need not)
  CU_TRY(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), st), "cudaMemsetAsync(ticket)");
  defl::inflate_warp_kernel<kWarpsPerCta><<<grid, kWarpsPerCta * 32, sizeof(defl::WarpSmem) * kWarpsPerCta, st>>>(
      d_items, d_order, (uint32_t)n, flags, d_results, ticket, ws + kWsHeader);
  CU_TRY(cudaGetLastError(), "inflate_warp_kernel launch (static scratch)");
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return UNZ_OK;
}

UNZ_EXPORT int infl_gather_device(const void* const* d_seg_src, const uint64_t* d_seg_len,
                                  const uint64_t* d_seg_dst_off, size_t nseg, void* d_out, void* cuda_stream) {
  if (nseg == 0) return UNZ_OK;
  Engine* E = nullptr;
  if (engine_for_current_device(&E) != UNZ_OK) return UNZ_ERR;
  const uint64_t warps  = std::min<uint64_t>((nseg + 31) / 32, (uint64_t)E->sms * 64);   // a warp takes 32 pieces at a time
  const unsigned blocks = (unsigned)((warps * 32 + 255) / 256);
  defl::gather_pieces_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(cuda_stream)>>>(
      d_seg_src, d_seg_len, d_seg_dst_off, nseg, static_cast<uint8_t*>(d_out));
  CU_TRY(cudaGetLastError(), "gather_pieces_kernel launch");
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return UNZ_OK;
}

UNZ_EXPORT int infl_batch_partition(const infl_batch_item_t* items, size_t n, int nparts, uint32_t* part_of,
                                    uint64_t* shard_bytes) {
  if ((!items && n) || (!part_of && n) || nparts < 1 || n > 0xFFFFFFFFull) return fail("infl_batch_partition: bad arguments");
  std::vector<std::vector<uint32_t>> parts;
  partition_lpt(items, n, nparts, parts);
  for (int g = 0; g < nparts; g++) {
    uint64_t b = 0;
    for (uint32_t i : parts[g]) { part_of[i] = (uint32_t)g; b += items[i].srclen + items[i].dstlen; }
    if (shard_bytes) shard_bytes[g] = b;
  }
  return UNZ_OK;
}

UNZ_EXPORT int infl_buf_batch(const infl_batch_item_t* items, size_t n, int flags, infl_batch_result_t* results) {
  if (n == 0) return UNZ_OK;
  if (!items || !results) return fail("null argument");
  if (flags & INFL_BATCH_DEVICE_PTRS) return run_device_batch(items, n, flags, results);
  if (flags & INFL_BATCH_ALL_GPUS) return run_host_batch_all_gpus(items, n, flags, results);
  return run_host_batch_on_device(items, n, flags, results);
}

// Chunk-list batch.  Pieces are never joined by the host CPU:
//  * device pieces: a (src, len, dst offset) descriptor table drives gather_pieces_kernel;
//  * host pieces that lie next to each other in memory (gaps <= 512 bytes: IDAT chunks of one file) travel as one
//    DMA into a staging area, pieces >= 8 KiB elsewhere as one DMA each, and scattered pieces < 8 KiB are packed
//    into a pinned bounce buffer that travels as one DMA (the reference copies exactly these itself,
//    mem.c:121-155, threshold common.h:46); gather_pieces_kernel then puts every piece at its logical offset.
// One group of a chunk-list batch.  out_base: where this group's outputs start in E->d_out (the buffer holds the
// whole batch, so the copy back of a group can run on E->stream3 while the next group decodes).
static int iov_group(Engine* E, const infl_batch_iov_item_t* items, size_t n, int flags, infl_batch_result_t* results,
                     size_t out_base) {

  const bool dev_ptrs = (flags & INFL_BATCH_DEVICE_PTRS) != 0;
  std::vector<size_t> in_off(n + 1, 0), out_off(n + 1, 0);
  size_t nseg = 0;
  for (size_t i = 0; i < n; i++) {
    size_t tot = 0;
    for (uint64_t k = 0; k < items[i].niov; k++) tot += items[i].iov[k].len;
    if (tot > 0xFFFFFFFFull || items[i].dstlen > 0xFFFFFFFFull)
      return fail("stream larger than 4 GiB - 1 (the reference API limit)");
    in_off[i + 1]  = align_up(in_off[i] + tot);
    out_off[i + 1] = dev_ptrs ? 0 : align_up(out_off[i] + items[i].dstlen);
    nseg += items[i].niov;
  }

  std::vector<infl_batch_item_t>   flat(n);
  std::vector<infl_batch_result_t> res(n);
  int rc = UNZ_OK;
  {
    cudaStream_t st = E->stream;
    if (E->d_in.reserve(in_off[n] + 64)) return UNZ_ERR;
    uint8_t* din  = static_cast<uint8_t*>(E->d_in.p);
    uint8_t* dout = static_cast<uint8_t*>(E->d_out.p) + out_base;

    if (dev_ptrs) {
      // descriptor table -> K2
      // a descriptor covers at most 64 KiB: a warp of the gather kernel takes 32 descriptors at a time, so a
      // few huge pieces still spread over the whole GPU
      nseg = 0;
      for (size_t i = 0; i < n; i++)
        for (uint64_t k = 0; k < items[i].niov; k++) nseg += (items[i].iov[k].len + kGatherPiece - 1) / kGatherPiece;
      if (E->h_seg.reserve(nseg * 24 + 64) || E->d_seg.reserve(nseg * 24 + 64)) return UNZ_ERR;
      auto* hs_src = static_cast<const void**>(E->h_seg.p);
      auto* hs_len = reinterpret_cast<uint64_t*>(hs_src + nseg);
      auto* hs_off = hs_len + nseg;
      size_t s = 0;
      for (size_t i = 0; i < n; i++) {
        size_t at = in_off[i];
        for (uint64_t k = 0; k < items[i].niov; k++) {
          const uint8_t* p = static_cast<const uint8_t*>(items[i].iov[k].ptr);
          for (uint64_t o = 0; o < items[i].iov[k].len; o += kGatherPiece, s++) {
            hs_src[s] = p + o; hs_len[s] = std::min<uint64_t>(kGatherPiece, items[i].iov[k].len - o); hs_off[s] = at + o;
          }
          at += items[i].iov[k].len;
        }
      }
      if (nseg) {
        CU_TRY(cudaMemcpyAsync(E->d_seg.p, E->h_seg.p, nseg * 24, cudaMemcpyHostToDevice, st), "H2D gather table");
        auto* ds_src = static_cast<const void* const*>(E->d_seg.p);
        auto* ds_len = reinterpret_cast<const uint64_t*>(ds_src + nseg);
        if (infl_gather_device(ds_src, ds_len, ds_len + nseg, nseg, din, st) != UNZ_OK) return UNZ_ERR;
      }
    } else {
      // Host pieces.  Pieces that lie (nearly) next to each other in host memory -- the IDAT chunks of a PNG file are
      // 12 bytes apart -- travel as ONE copy, gaps included, into a staging area that keeps the host layout; the
      // gather kernel then joins them on the device.  Pieces that are exactly adjacent share one descriptor.
      // Pieces scattered over the heap each take a copy of their own (one cudaMemcpyAsync per span).
      // A small piece (< 8 KiB, the reference's own copy threshold, common.h:46 / mem.c:121-155) that does not
      // continue the previous span is packed into a pinned bounce buffer, and all of those travel as one copy.
      struct Span { const uint8_t* h; size_t len; size_t dev_off; };
      struct Seg  { uint32_t span; size_t span_off; size_t len; size_t dst_off; };   // span == kBounce: offset into the bounce area
      constexpr uint32_t kBounce = 0xFFFFFFFFu;
      std::vector<Span> spans;
      std::vector<Seg>  segs;
      size_t stage_total = 0, small_total = 0;
      for (size_t i = 0; i < n; i++)
        for (uint64_t k = 0; k < items[i].niov; k++)
          if (items[i].iov[k].len < kCopyThreshold) small_total += items[i].iov[k].len;
      if (E->h_pack.reserve(small_total + 64)) return UNZ_ERR;
      uint8_t* const pack = static_cast<uint8_t*>(E->h_pack.p);
      size_t bounce_used = 0;
      for (size_t i = 0; i < n; i++) {
        size_t at = in_off[i];
        bool   first = true;
        for (uint64_t k = 0; k < items[i].niov; k++) {
          const uint8_t* p  = static_cast<const uint8_t*>(items[i].iov[k].ptr);
          const size_t  len = items[i].iov[k].len;
          if (!len) continue;
          bool extend = false;
          if (!spans.empty()) {                         // continues the last span (whatever went to the bounce area since)
            const uint8_t* end = spans.back().h + spans.back().len;
            extend = p >= end && (size_t)(p - end) <= kGapMerge;
          }
          if (extend) {
            Span& sp = spans.back();
            const bool glued = !first && !segs.empty() && segs.back().span == spans.size() - 1 &&
                               sp.h + segs.back().span_off + segs.back().len == p;
            const size_t off = (size_t)(p - sp.h);
            sp.len = off + len;
            if (glued) segs.back().len += len;
            else segs.push_back({(uint32_t)(spans.size() - 1), off, len, at});
            stage_total = sp.dev_off + sp.len;
          } else if (len < kCopyThreshold) {
            std::memcpy(pack + bounce_used, p, len);
            const bool glued = !first && !segs.empty() && segs.back().span == kBounce &&
                               segs.back().span_off + segs.back().len == bounce_used;
            if (glued) segs.back().len += len;
            else segs.push_back({kBounce, bounce_used, len, at});
            bounce_used += len;
          } else {
            stage_total = align_up(stage_total);
            spans.push_back({p, len, stage_total});
            segs.push_back({(uint32_t)(spans.size() - 1), 0, len, at});
            stage_total = spans.back().dev_off + spans.back().len;
          }
          at += len;
          first = false;
        }
      }
      const size_t bounce_base = align_up(stage_total);
      {   // descriptors of at most 64 KiB (see above)
        std::vector<Seg> cut;
        cut.reserve(segs.size());
        for (const Seg& g : segs)
          for (size_t o = 0; o < g.len; o += kGatherPiece)
            cut.push_back({g.span, g.span_off + o, std::min<size_t>(kGatherPiece, g.len - o), g.dst_off + o});
        segs.swap(cut);
      }
      const size_t ns2 = segs.size();
      if (ns2) {
        if (E->d_stage.reserve(bounce_base + bounce_used + 64) || E->h_seg.reserve(ns2 * 24 + 64) || E->d_seg.reserve(ns2 * 24 + 64)) return UNZ_ERR;
        uint8_t* dstage = static_cast<uint8_t*>(E->d_stage.p);
        for (const Span& sp : spans)
          CU_TRY(cudaMemcpyAsync(dstage + sp.dev_off, sp.h, sp.len, cudaMemcpyHostToDevice, st), "H2D pieces");
        if (bounce_used)
          CU_TRY(cudaMemcpyAsync(dstage + bounce_base, pack, bounce_used, cudaMemcpyHostToDevice, st), "H2D small pieces");
        auto* hs_src = static_cast<const void**>(E->h_seg.p);
        auto* hs_len = reinterpret_cast<uint64_t*>(hs_src + ns2);
        auto* hs_off = hs_len + ns2;
        for (size_t k = 0; k < ns2; k++) {
          hs_src[k] = segs[k].span == kBounce ? dstage + bounce_base + segs[k].span_off
                                              : dstage + spans[segs[k].span].dev_off + segs[k].span_off;
          hs_len[k] = segs[k].len;
          hs_off[k] = segs[k].dst_off;
        }
        CU_TRY(cudaMemcpyAsync(E->d_seg.p, E->h_seg.p, ns2 * 24, cudaMemcpyHostToDevice, st), "H2D gather table");
        auto* ds_src = static_cast<const void* const*>(E->d_seg.p);
        auto* ds_len = reinterpret_cast<const uint64_t*>(ds_src + ns2);
        if (infl_gather_device(ds_src, ds_len, ds_len + ns2, ns2, din, st) != UNZ_OK) return UNZ_ERR;
      }
    }

    for (size_t i = 0; i < n; i++) {
      size_t tot = 0;
      for (uint64_t k = 0; k < items[i].niov; k++) tot += items[i].iov[k].len;
      flat[i].src = din + in_off[i];
      flat[i].srclen = tot;
      flat[i].dst = dev_ptrs ? items[i].dst : static_cast<void*>(dout + out_off[i]);
      flat[i].dstlen = items[i].dstlen;
    }
    // the gathered streams are an ordinary device-resident batch from here on (large members take the split decoder)
    rc = run_device_batch_locked(E, flat.data(), n, flags, res.data());
    if (rc != UNZ_OK) return rc;
    if (!dev_ptrs) {
      // the group is decoded (run_device_batch_locked returned its results): its bytes go home on the copy stream,
      // neighbours in host memory as one copy
      size_t i = 0;
      while (i < n) {
        size_t j = i;
        uint8_t* h0 = static_cast<uint8_t*>(items[i].dst);
        size_t   len = res[i].outlen;
        while (j + 1 < n && res[j].outlen == items[j].dstlen && out_off[j + 1] == out_off[j] + items[j].dstlen &&
               static_cast<uint8_t*>(items[j + 1].dst) == static_cast<uint8_t*>(items[j].dst) + items[j].dstlen) {
          j++;
          len = (size_t)(static_cast<uint8_t*>(items[j].dst) - h0) + res[j].outlen;
        }
        if (len) CU_TRY(cudaMemcpyAsync(h0, dout + out_off[i], len, cudaMemcpyDeviceToHost, E->stream3), "D2H output");
        i = j + 1;
      }
    }
  }
  for (size_t i = 0; i < n; i++) {
    results[i] = res[i];
    if (items[i].niov == 0) {   // nothing was ever included: UNZ_NOOP (infl.c:1544-1545,1671-1672)
      std::memset(&results[i], 0, sizeof(results[i]));
      results[i].code = UNZ_NOOP;
      results[i].adler = 1;
    }
  }
  return UNZ_OK;
}

UNZ_EXPORT int infl_buf_batch_iov(const infl_batch_iov_item_t* items, size_t n, int flags,
                                  infl_batch_result_t* results) {
  if (n == 0) return UNZ_OK;
  if (!items || !results) return fail("null argument");
  Engine* E = nullptr;
  if (engine_for_current_device(&E) != UNZ_OK) return UNZ_ERR;
  std::lock_guard<std::mutex> guard(E->mu);
  const bool dev_ptrs = (flags & INFL_BATCH_DEVICE_PTRS) != 0;
  if (dev_ptrs) return iov_group(E, items, n, flags, results, 0);
  // host pieces: groups of about equal output; while group g decodes, group g - 1 is on its way home
  std::vector<size_t> obase(n + 1, 0);
  for (size_t i = 0; i < n; i++) {
    if (items[i].dstlen > 0xFFFFFFFFull) return fail("stream larger than 4 GiB - 1 (the reference API limit)");
    obase[i + 1] = align_up(obase[i] + items[i].dstlen);
  }
  if (E->d_out.reserve(obase[n] + 64)) return UNZ_ERR;
  // (a group wants about a GiB of output: the split-point decoder needs thousands of chunks to fill the GPU --
  //  four groups of 256 MiB took 105 ms where one took 63)
  const size_t groups = std::min<size_t>(8, std::max<size_t>(1, obase[n] >> 30));
  int    rc = UNZ_OK;
  size_t a = 0;
  for (size_t g = 0; g < groups && rc == UNZ_OK && a < n; g++) {
    size_t b = a;
    const size_t upto = g + 1 == groups ? obase[n] : obase[n] / groups * (g + 1);
    while (b < n && (b == a || obase[b + 1] <= upto)) b++;
    if (g + 1 == groups) b = n;
    rc = iov_group(E, items + a, b - a, flags, results + a, obase[a]);
    a = b;
  }
  const cudaError_t e = cudaStreamSynchronize(E->stream3);
  if (e != cudaSuccess && rc == UNZ_OK) rc = fail("D2H output", e);
  return rc;
}

// ---------------------------------------------------------------------------------------------------
// The reference's handle API.  The handle is opaque to callers (reference include/defl/common.h:94-96),
// so its layout is ours: destination, flags and the list of included pieces.
// ---------------------------------------------------------------------------------------------------
}  // extern "C"

struct unz__stream_t {
  uint8_t* dst    = nullptr;
  uint32_t dstlen = 0;
  int      flags  = 0;
  uint32_t dstpos = 0;      // infl_output_pos
  uint32_t inpos  = 0;      // infl_input_pos
  uint32_t adler = 1, adler_stored = 0;
  bool     has_trailer = false;
  bool     done = false;    // infl_stream(): final block decoded

  // pieces in call order.  Small ones (<= 8 KiB, reference threshold src/common.h:46) are copied
  // into pages owned by the handle, because the reference lets callers release them right after
  // infl_include (mem.c:135-151); larger ones are borrowed (README.md:47).
  std::vector<infl_iovec_t>               pieces;
  std::vector<std::unique_ptr<uint8_t[]>> pages;
  size_t page_used = 0, page_cap = 0;
  bool   last_is_page_tail = false;
  uint64_t total_in = 0;
  bool   any_piece = false;

  // infl_stream() state: the input so far (host copy + device copy), the output on the device, and -- on the
  // device -- the last block boundary a call passed, from which the next call goes on (stream.c:468-794)
  std::vector<uint8_t> acc;       // accumulated compressed input
  bool                 strm_started = false;
  DevBuf   sd_in, sd_out, sd_meta;   // meta: StreamResume + infl_batch_result_t
  int      sd_dev = -1;              // device the three live on
  size_t   uploaded = 0;             // bytes of acc that are in sd_in
  size_t   tried_at = 0;             // acc.size() at the last decode attempt
  uint32_t out_on_host = 0;          // dst[0, out_on_host) has been copied back
  uint32_t last_feed = 0;            // size of the previous piece
  bool     out_moved = false;        // infl_resize_output() gave a new dst since the last attempt
  int      last_rc = UNZ_UNFINISHED; // what the last look said
  ~unz__stream_t() {
    if (sd_in.p || sd_out.p || sd_meta.p) {
      int cur = 0;
      cudaGetDevice(&cur);
      if (sd_dev >= 0 && sd_dev != cur) cudaSetDevice(sd_dev);
      if (sd_in.p) cudaFree(sd_in.p);
      if (sd_out.p) cudaFree(sd_out.p);
      if (sd_meta.p) cudaFree(sd_meta.p);
      if (sd_dev >= 0 && sd_dev != cur) cudaSetDevice(cur);
    }
  }
};

namespace {

constexpr size_t kPiecePage      = 1 << 20;

void stream_clear_pieces(unz__stream_t* s) {
  s->pieces.clear();
  s->pages.clear();
  s->page_used = s->page_cap = 0;
  s->last_is_page_tail = false;
  s->total_in = 0;
  s->any_piece = false;
}

void stream_reset_state(unz__stream_t* s) {
  s->dstpos = 0; s->inpos = 0; s->adler = 1; s->adler_stored = 0; s->has_trailer = false;
  s->done = false; s->acc.clear(); s->strm_started = false;
  s->uploaded = 0; s->tried_at = 0; s->out_on_host = 0; s->last_feed = 0; s->out_moved = false;
  s->last_rc = UNZ_UNFINISHED;
  if (s->sd_meta.p) {                       // forget the resume point (the device buffers are kept)
    int cur = 0;
    cudaGetDevice(&cur);
    if (s->sd_dev >= 0 && s->sd_dev != cur) cudaSetDevice(s->sd_dev);
    cudaMemset(s->sd_meta.p, 0, sizeof(defl::StreamResume));
    if (s->sd_dev >= 0 && s->sd_dev != cur) cudaSetDevice(cur);
  }
}

int decode_pieces(unz__stream_t* s, const infl_iovec_t* iov, size_t niov, infl_batch_result_t* r) {
  infl_batch_iov_item_t it;
  it.iov = iov; it.niov = niov; it.dst = s->dst; it.dstlen = s->dstlen;
  return infl_buf_batch_iov(&it, 1, s->flags == INFL_ZLIB ? INFL_ZLIB : 0, r);
}

// One look at a stream that arrives through infl_stream(): bring the new input to the device, decode from the
// last block boundary an earlier look passed, bring the new output back.  Work per look: the new bytes plus one
// block (the one the previous look ended in).
int stream_attempt(unz__stream_t* s, infl_batch_result_t* r) {
  Engine* E = nullptr;
  if (engine_for_current_device(&E) != UNZ_OK) return UNZ_ERR;
  std::lock_guard<std::mutex> g(E->mu);
  cudaStream_t st = E->stream;
  if (s->sd_dev >= 0 && s->sd_dev != E->dev) return fail("infl_stream: the handle was started on another device");
  s->sd_dev = E->dev;
  const size_t meta_bytes = sizeof(defl::StreamResume) + sizeof(infl_batch_result_t);
  if (!s->sd_meta.p) {
    if (s->sd_meta.reserve(256)) return UNZ_ERR;
    CU_TRY(cudaMemsetAsync(s->sd_meta.p, 0, meta_bytes, st), "cudaMemsetAsync(stream state)");
  }
  auto* d_rs  = static_cast<defl::StreamResume*>(s->sd_meta.p);
  auto* d_res = reinterpret_cast<infl_batch_result_t*>(static_cast<uint8_t*>(s->sd_meta.p) + sizeof(defl::StreamResume));
  // input: only what is new, unless the buffer had to grow
  const size_t n = s->acc.size();
  if (n > 0xFFFFFFFFull) return fail("infl_stream: more than 4 GiB of input");
  if (n > s->sd_in.cap) {
    if (s->sd_in.p) { cudaStreamSynchronize(st); }
    if (s->sd_in.reserve(std::max<size_t>(2 * n, 1 << 16))) return UNZ_ERR;
    s->uploaded = 0;
  }
  if (n > s->uploaded)
    CU_TRY(cudaMemcpyAsync(static_cast<uint8_t*>(s->sd_in.p) + s->uploaded, s->acc.data() + s->uploaded, n - s->uploaded,
                           cudaMemcpyHostToDevice, st), "cudaMemcpyAsync H2D (infl_stream)");
  s->uploaded = n;
  // output: lives on the device; a larger or different caller buffer brings what it holds so far
  const size_t want_out = std::max<size_t>(s->dstlen, 16);
  if (want_out > s->sd_out.cap || s->out_moved) {
    if (want_out > s->sd_out.cap) {
      if (s->sd_out.p) cudaStreamSynchronize(st);
      if (s->sd_out.reserve(want_out)) return UNZ_ERR;
    }
    if (s->out_on_host)
      CU_TRY(cudaMemcpyAsync(s->sd_out.p, s->dst, s->out_on_host, cudaMemcpyHostToDevice, st), "cudaMemcpyAsync H2D (output so far)");
    s->out_moved = false;
  }
  if (E->d_ws.reserve_exact(warp_kernel_workspace(E->sms))) return UNZ_ERR;
  static bool attr_done[kMaxDevices] = {false};
  if (!attr_done[E->dev]) {
    CU_TRY(cudaFuncSetAttribute(defl::inflate_resume_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)sizeof(defl::WarpSmem)), "cudaFuncSetAttribute(smem)");
    attr_done[E->dev] = true;
  }
  defl::inflate_resume_kernel<<<1, 32, sizeof(defl::WarpSmem), st>>>(
      static_cast<const uint8_t*>(s->sd_in.p), (uint32_t)n, static_cast<uint8_t*>(s->sd_out.p), s->dstlen,
      s->flags == INFL_ZLIB ? INFL_ZLIB : 0, d_rs, d_res, static_cast<unsigned char*>(E->d_ws.p) + kWsHeader);
  CU_TRY(cudaGetLastError(), "inflate_resume_kernel launch");
  g_launches.fetch_add(1, std::memory_order_relaxed);
  CU_TRY(cudaMemcpyAsync(r, d_res, sizeof(*r), cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync D2H (result)");
  CU_TRY(cudaStreamSynchronize(st), "cudaStreamSynchronize (infl_stream)");
  s->tried_at = n;
  const uint32_t outlen = (uint32_t)std::min<uint64_t>(r->outlen, s->dstlen);
  if (outlen > s->out_on_host) {
    CU_TRY(cudaMemcpyAsync(s->dst + s->out_on_host, static_cast<uint8_t*>(s->sd_out.p) + s->out_on_host,
                           outlen - s->out_on_host, cudaMemcpyDeviceToHost, st), "cudaMemcpyAsync D2H (output)");
    CU_TRY(cudaStreamSynchronize(st), "cudaStreamSynchronize (infl_stream output)");
    s->out_on_host = outlen;
  }
  return UNZ_OK;
}

}  // namespace

extern "C" {

UNZ_EXPORT infl_stream_t* infl_init(void* dst, uint32_t dstlen, int flags) {
  // the reference only allocates here (infl.c:1677-1695); so do we -- the device is touched in infl()
  unz__stream_t* s = new (std::nothrow) unz__stream_t();
  if (!s) return nullptr;
  s->dst = static_cast<uint8_t*>(dst);
  s->dstlen = dstlen;
  s->flags = flags;
  return s;
}

UNZ_EXPORT void infl_reset_pool(infl_stream_t* s) {     // mem.c:220-232
  if (!s) return;
  stream_clear_pieces(s);
}

UNZ_EXPORT void infl_reset(infl_stream_t* s, void* dst, uint32_t dstlen, int flags) {  // mem.c:257-275
  if (!s) return;
  stream_clear_pieces(s);
  stream_reset_state(s);
  s->dst = static_cast<uint8_t*>(dst);
  s->dstlen = dstlen;
  s->flags = flags;
}

UNZ_EXPORT int infl_resize_output(infl_stream_t* s, void* dst, uint32_t dstlen) {      // mem.c:277-288
  if (!s || !dst || dstlen < s->dstpos) return UNZ_ERR;
  s->out_moved = s->out_moved || static_cast<uint8_t*>(dst) != s->dst;
  s->dst = static_cast<uint8_t*>(dst);
  s->dstlen = dstlen;
  return UNZ_OK;
}

UNZ_EXPORT void infl_include(infl_stream_t* s, const void* ptr, uint32_t len) {        // mem.c:114-216
  if (!s) return;
  s->any_piece = true;
  s->total_in += len;
  if (len <= kCopyThreshold) {
    if (s->pages.empty() || len > s->page_cap - s->page_used) {
      std::unique_ptr<uint8_t[]> pg(new (std::nothrow) uint8_t[kPiecePage]);
      if (!pg) return;   // allocation failure silently drops the piece, as mem.c:191 does
      s->pages.push_back(std::move(pg));
      s->page_used = 0; s->page_cap = kPiecePage;
      s->last_is_page_tail = false;
    }
    uint8_t* at = s->pages.back().get() + s->page_used;
    if (len) std::memcpy(at, ptr, len);
    s->page_used += len;
    if (s->last_is_page_tail) s->pieces.back().len += len;   // extend the run inside the page
    else { s->pieces.push_back({at, len}); s->last_is_page_tail = true; }
  } else {
    s->pieces.push_back({ptr, len});
    s->last_is_page_tail = false;
  }
}

UNZ_EXPORT int infl(infl_stream_t* s) {                                                 // infl.c:1507-1675
  if (!s) return UNZ_ERR;
  if (!s->any_piece) return UNZ_NOOP;                      // infl.c:1544-1545
  if (s->total_in == 0) return UNZ_OK;                     // infl.c:1547-1550
  infl_batch_result_t r;
  if (decode_pieces(s, s->pieces.data(), s->pieces.size(), &r) != UNZ_OK) return UNZ_ERR;
  if (r.code != UNZ_OK) return UNZ_ERR;                    // every failure collapses to UNZ_ERR (infl.c:1673-1674)
  s->dstpos = (uint32_t)r.outlen;
  s->inpos  = (uint32_t)r.inpos;
  s->adler = r.adler; s->adler_stored = r.adler_stored;
  s->has_trailer = (r.info & INFL_INFO_HAS_TRAILER) != 0;
  s->done = true;
  return UNZ_OK;
}

UNZ_EXPORT void infl_destroy(infl_stream_t* s) { delete s; }                             // mem.c:331-361

UNZ_EXPORT int infl_buf(const void* src, uint32_t srclen, void* dst, uint32_t dstlen, int flags) {  // infl.h:125-145
  infl_batch_item_t   it;
  infl_batch_result_t r;
  if (srclen == 0) return UNZ_OK;                          // one empty piece: infl.c:1547-1550
  it.src = src; it.srclen = srclen; it.dst = dst; it.dstlen = dstlen;
  if (infl_buf_batch(&it, 1, flags == INFL_ZLIB ? INFL_ZLIB : 0, &r) != UNZ_OK) return UNZ_ERR;
  return r.code == UNZ_OK ? UNZ_OK : UNZ_ERR;
}

UNZ_EXPORT uint32_t infl_output_pos(const infl_stream_t* s) { return s ? s->dstpos : 0u; }  // mem.c:290-294
UNZ_EXPORT uint32_t infl_input_pos(const infl_stream_t* s) { return s ? s->inpos : 0u; }    // mem.c:296-329

UNZ_EXPORT uint32_t infl_adler32(const infl_stream_t* s, uint32_t* stored, int* present) {
  if (stored) *stored = s ? s->adler_stored : 0u;
  if (present) *present = (s && s->has_trailer) ? 1 : 0;
  return s ? s->adler : 1u;
}

// Resumable decode (reference src/infl/stream.c:468-794) as a host shim: input is accumulated
// and the whole prefix is offered to the device decoder; a stream that is not complete yet
// reports UNZ_UNFINISHED, one whose output does not fit reports UNZ_EFULL, and the call after
// infl_resize_output() decodes again into the larger buffer.  See INTEGRATION.md.
UNZ_EXPORT int infl_stream(infl_stream_t* s, const void* src, uint32_t srclen) {
  if (!s) return UNZ_ERR;
  const bool poll = !(src && srclen);
  if (!poll) {
    const uint8_t* p = static_cast<const uint8_t*>(src);
    s->acc.insert(s->acc.end(), p, p + srclen);
    s->strm_started = true;
  } else if (!s->strm_started) {
    return UNZ_OK;                                          // stream.c:489-493: nothing yet => UNZ_OK
  }
  if (s->done) return UNZ_OK;                               // stream.c:496-499
  if (s->acc.empty()) return UNZ_NOOP;
  if (s->flags == INFL_ZLIB && s->acc.size() < 2) return UNZ_UNFINISHED;   // stream.c:576-579

  // When to look: a decode attempt is a kernel launch and two small copies, so a caller that feeds single bytes
  // must not pay for one per call.  Look when the caller polls (NULL, 0), while the stream is small, when it has
  // grown by a sixteenth since the last look (by 1/128 when the pieces are not tiny), or when this piece is shorter
  // than the one before (the last piece of a file usually is).  Otherwise the answer is "go on": UNZ_UNFINISHED, as the reference gives for input
  // that ends inside a block (stream.c:614-616).
  const bool shorter = !poll && srclen < s->last_feed;
  if (!poll) s->last_feed = srclen;
  const size_t fresh = s->acc.size() - s->tried_at;
  if (!(poll || s->acc.size() <= 512 || shorter || fresh >= (s->tried_at >> 4) + 1 ||
        (srclen >= 64 && fresh >= (s->tried_at >> 7) + 1)))
    return UNZ_UNFINISHED;
  if (poll && s->tried_at == s->acc.size() && !s->out_moved) {
    // nothing new since the last look: the same verdict without a launch
    return s->last_rc;
  }
  infl_batch_result_t r;
  if (stream_attempt(s, &r) != UNZ_OK) return UNZ_ERR;
  const uint32_t reason = (r.info >> INFL_INFO_REASON_SHIFT) & 0xFF;
  s->dstpos = (uint32_t)r.outlen;
  if (r.code == UNZ_OK) {
    s->inpos = (uint32_t)r.inpos;
    s->adler = r.adler; s->adler_stored = r.adler_stored;
    s->has_trailer = (r.info & INFL_INFO_HAS_TRAILER) != 0;
    s->done = true;
    return UNZ_OK;
  }
  s->last_rc = reason == defl::R_TRUNCATED ? UNZ_UNFINISHED      // stream.c:614-616 and friends
             : reason == defl::R_OUTPUT_FULL ? UNZ_EFULL       // stream.c:625-627, :769-770
                                             : UNZ_ERR;
  return s->last_rc;
}

}  // extern "C"

#include "wrap_host.inc"
