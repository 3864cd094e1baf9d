/*
 * defl/infl_batch.h -- NEW entry points (not in the reference): batched inflate on B200.
 *
 * The reference decodes one stream per infl_buf() call on one CPU thread
 * (include/defl/infl.h:125-145).  These calls hand many independent streams (or one very
 * large stream) to the sm_100a kernels in one go.  Per item the result `code` is what the
 * reference's infl_buf() returns for the same bytes (UNZ_OK / UNZ_ERR), unless Adler-32
 * verification is explicitly requested with INFL_BATCH_VERIFY_ADLER.
 *
 * Plain C ABI: pointers and sizes only.  "device pointer" = an address valid on the CUDA
 * device that is current when the call is made (e.g. cudaMalloc / a torch CUDA tensor's
 * data_ptr()).
 */
#ifndef infl_batch_h
#define infl_batch_h

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- flag bits (low byte keeps the reference's meaning: == INFL_ZLIB -> zlib wrapper) ---- */
#define INFL_BATCH_VERIFY_ADLER  0x0100  /* zlib only: missing or wrong trailer => UNZ_ERR        */
#define INFL_BATCH_DEVICE_PTRS   0x0200  /* src/dst (and iov ptr) are device pointers             */
#define INFL_BATCH_ALL_GPUS      0x0400  /* host-pointer batches: shard over every visible GPU    */
#define INFL_BATCH_NO_SPLIT      0x0800  /* never use the split-point decoder for large streams   */

/* one stream: src[0..srclen) -> dst[0..dstlen).  srclen, dstlen < 4 GiB (the reference's
   uint32_t limits, infl.h:31-33,78-82).  32 bytes; this is also the device descriptor. */
typedef struct infl_batch_item_t {
  const void *src;
  uint64_t    srclen;
  void       *dst;
  uint64_t    dstlen;
} infl_batch_item_t;

/* one piece of a non-contiguous input (what one infl_include() call passes, infl.h:78-82) */
typedef struct infl_iovec_t {
  const void *ptr;
  uint64_t    len;
} infl_iovec_t;

/* one stream given as a chunk list: logical input = concatenation of iov[0..niov) */
typedef struct infl_batch_iov_item_t {
  const infl_iovec_t *iov;     /* host array, even with INFL_BATCH_DEVICE_PTRS */
  uint64_t            niov;    /* 0 => UNZ_NOOP, as infl() with nothing included (infl.c:1544-1545) */
  void               *dst;
  uint64_t            dstlen;
} infl_batch_iov_item_t;

/* info bits */
#define INFL_INFO_HAS_TRAILER   0x0001u  /* 4 trailer bytes were present after the body  */
#define INFL_INFO_ADLER_MATCH   0x0002u  /* computed == stored                           */
#define INFL_INFO_SPLIT_DECODE  0x0004u  /* decoded by the split-point (intra-stream) path */
#define INFL_INFO_REASON_SHIFT  8        /* bits 8..15: engine-specific failure reason   */

typedef struct infl_batch_result_t {
  int32_t  code;          /* UNZ_OK / UNZ_ERR / UNZ_NOOP                                  */
  uint32_t adler;         /* Adler-32 of dst[0..outlen)                                   */
  uint64_t outlen;        /* bytes produced  (what infl_output_pos() reports, infl.h:175) */
  uint64_t inpos;         /* bytes consumed  (what infl_input_pos() reports, infl.h:187)  */
  uint32_t adler_stored;  /* big-endian trailer after the body, when present              */
  uint32_t info;          /* INFL_INFO_* bits                                             */
} infl_batch_result_t;

/* ---- host-facing batch calls -------------------------------------------------------- */

/* Decode n independent streams.  Host pointers by default: inputs are staged to the GPU,
   decoded, and outputs copied back before the call returns.  With INFL_BATCH_DEVICE_PTRS the
   data never leaves HBM (items/results themselves stay host arrays).  Returns UNZ_OK when the
   batch ran (look at results[i].code per stream), UNZ_ERR when the engine could not run. */
UNZ_EXPORT int infl_buf_batch(const infl_batch_item_t *items, size_t n, int flags,
                              infl_batch_result_t *results);

/* Same, each stream given as a chunk list (PNG IDAT style).  Pieces are gathered on the
   device by a descriptor-driven kernel; they are never joined on the host. */
UNZ_EXPORT int infl_buf_batch_iov(const infl_batch_iov_item_t *items, size_t n, int flags,
                                  infl_batch_result_t *results);

/* ---- device-resident, stream-ordered calls (everything already in HBM) --------------- */

/* bytes of device scratch infl_batch_decode_device needs for n streams */
UNZ_EXPORT size_t infl_batch_workspace_size(size_t n);

/* d_items, d_results, d_workspace: device arrays.  d_order: optional device permutation
   (largest stream first keeps the tail short); NULL = index order.  Asynchronous on
   `cuda_stream` (a cudaStream_t, NULL = default stream).  Returns UNZ_OK if launched. */
UNZ_EXPORT int infl_batch_decode_device(const infl_batch_item_t *d_items, const uint32_t *d_order,
                                        size_t n, int flags, infl_batch_result_t *d_results,
                                        void *d_workspace, void *cuda_stream);

/* The same decode with NO allocation on the call path.  This entry point has no counterpart in the
   reference; it was added for callers that cannot call
   cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The
   kernel's token scratch is a file-scope `static __device__` array inside the library (about
   38.2 MiB: 148 warps' worth, held by every GPU that loads the library until the process exits), so
   no workspace is passed and neither cudaMalloc / cudaFree nor any allocating library routine is
   called.  Limits: n <= INFL_BATCH_STATIC_MAX_STREAMS per call (more => UNZ_ERR and a message in
   infl_last_error(); use infl_batch_decode_device), 148 warps work at a time however large the GPU,
   always the default (v3) engine.  Every call on a device shares the one array: calls that overlap
   on the GPU (two streams, two threads) race on it -- order them on one stream or with events.
   d_items, d_results, d_order as for infl_batch_decode_device; asynchronous on `cuda_stream`. */
#define INFL_BATCH_STATIC_MAX_STREAMS 65536
UNZ_EXPORT int infl_batch_decode_device_static(const infl_batch_item_t *d_items, const uint32_t *d_order,
                                               size_t n, int flags, infl_batch_result_t *d_results,
                                               void *cuda_stream);

/* Gather pieces into contiguous device memory: for i < nseg copies
   seg_src[i][0..seg_len[i]) to d_out + seg_dst_off[i].  All arrays are device arrays. */
UNZ_EXPORT int infl_gather_device(const void *const *d_seg_src, const uint64_t *d_seg_len,
                                  const uint64_t *d_seg_dst_off, size_t nseg, void *d_out,
                                  void *cuda_stream);

/* ---- sharding (host only, no device needed) ------------------------------------------- */

/* Byte-balanced partition of n independent streams into nparts shards: greedy longest-processing-time
 * on srclen + dstlen, largest first.  part_of[i] = shard of stream i.  It is what INFL_BATCH_ALL_GPUS
 * uses inside one process and what one-process-per-GPU callers use to pick their share (rank r decodes
 * the streams with part_of[i] == r); there is no exchange between shards.  shard_bytes (optional,
 * nparts entries) receives each shard's srclen + dstlen total.  Returns UNZ_OK / UNZ_ERR. */
UNZ_EXPORT int infl_batch_partition(const infl_batch_item_t *items, size_t n, int nparts,
                                    uint32_t *part_of, uint64_t *shard_bytes);

/* ---- engine control ------------------------------------------------------------------ */
UNZ_EXPORT int         infl_device_count(void);
UNZ_EXPORT const char *infl_last_error(void);       /* thread-local, "" when none */
UNZ_EXPORT const char *infl_engine_info(void);      /* build / kernel configuration string */
/* number of kernels launched by this library in this process (bench.py's gpu_launches) */
UNZ_EXPORT uint64_t    infl_kernel_launch_count(void);
/* tuning knobs, mainly for benchmarks.  lanes per stream: 0 = default engine (the lane-decoder
   pipeline: one thread decodes, warps copy); 1,2,4,8,16,32 = the group-per-stream kernel.
   pipeline shape: 0..3 = compiled (decoder warps, slots, worker warps) variants */
UNZ_EXPORT int         infl_set_lanes_per_stream(int lanes);
UNZ_EXPORT int         infl_set_pipeline_shape(int shape);
/* decode engine for batches: 3 = one warp per stream, lanes decode sub-sequences speculatively (default);
   2 = lane-decoder pipeline; 1 = group per stream (uses infl_set_lanes_per_stream) */
UNZ_EXPORT int         infl_set_engine(int engine);

#ifdef __cplusplus
}
#endif
#endif /* infl_batch_h */
