/*
 * oracle/ref_driver.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Thin driver linked into oracle/_ref/libdefl_ref.so next to the reference's
 * unmodified sources.  It only *calls* the reference's public API
 * (/root/reference/include/defl/infl.h:31-189); no reference code is restated
 * here.  Gives Python/ctypes real symbols for the header-inline infl_buf
 * (infl.h:125-145), a chunk-list call (infl_init + N x infl_include + infl,
 * README.md:120-155) and a pthread work queue used by bench.py's
 * `--impl reference` / `cpu_baseline` leg ("defl infl_buf on all host cores").
 */
#define _GNU_SOURCE
#include <defl/infl.h>
#include <pthread.h>
#include <stdatomic.h>
#include <time.h>

int
ref_infl_buf(const void *src, uint32_t srclen, void *dst, uint32_t dstlen, int flags) {
  return infl_buf(src, srclen, dst, dstlen, flags);
}

/* one-shot decode that also reports produced/consumed byte counts */
int
ref_infl_buf_ex(const void *src, uint32_t srclen, void *dst, uint32_t dstlen, int flags,
                uint32_t *outlen, uint32_t *inpos) {
  infl_stream_t *st;
  int            rc;

  if (!(st = infl_init(dst, dstlen, flags))) return -1;
  infl_include(st, src, srclen);
  rc = infl(st);
  if (outlen) *outlen = infl_output_pos(st);
  if (inpos)  *inpos  = infl_input_pos(st);
  infl_destroy(st);
  return rc;
}

/* chunk-list decode: chunk i is src[offs[i] .. offs[i+1]) */
int
ref_infl_chunks(const uint8_t *src, const uint64_t *offs, uint32_t nchunks,
                void *dst, uint32_t dstlen, int flags, uint32_t *outlen) {
  infl_stream_t *st;
  uint32_t       i;
  int            rc;

  if (!(st = infl_init(dst, dstlen, flags))) return -1;
  for (i = 0; i < nchunks; i++)
    infl_include(st, src + offs[i], (uint32_t)(offs[i + 1] - offs[i]));
  rc = infl(st);
  if (outlen) *outlen = infl_output_pos(st);
  infl_destroy(st);
  return rc;
}

/* feed [src, src+srclen) through infl_stream() in pieces of `piece` bytes */
int
ref_infl_stream_pieces(const uint8_t *src, uint32_t srclen, uint32_t piece,
                       void *dst, uint32_t dstlen, int flags, uint32_t *outlen) {
  infl_stream_t *st;
  uint32_t       off = 0;
  int            rc  = UNZ_UNFINISHED;

  if (!(st = infl_init(dst, dstlen, flags))) return -1;
  if (!piece) piece = 1;
  while (off < srclen) {
    uint32_t n = srclen - off < piece ? srclen - off : piece;
    rc = infl_stream(st, src + off, n);
    off += n;
    if (rc != UNZ_UNFINISHED) break;
  }
  if (outlen) *outlen = infl_output_pos(st);
  infl_destroy(st);
  return rc;
}

typedef struct ref_job_t {
  const uint8_t  *src;
  const uint64_t *src_off;   /* n+1 offsets into src */
  uint8_t        *dst;
  const uint64_t *dst_off;   /* n+1 offsets into dst */
  int32_t        *codes;
  const uint32_t *order;     /* optional processing order */
  uint64_t        n;
  int             flags;
  atomic_ullong   next;
} ref_job_t;

static void *
ref_worker(void *arg) {
  ref_job_t *job = (ref_job_t *)arg;

  for (;;) {
    uint64_t k = atomic_fetch_add(&job->next, 1), i;
    if (k >= job->n) break;
    i = job->order ? job->order[k] : k;
    job->codes[i] = infl_buf(job->src + job->src_off[i],
                             (uint32_t)(job->src_off[i + 1] - job->src_off[i]),
                             job->dst + job->dst_off[i],
                             (uint32_t)(job->dst_off[i + 1] - job->dst_off[i]),
                             job->flags);
  }
  return NULL;
}

/* Decode n independent streams with `nthreads` host threads; returns seconds.
   One single-threaded warm-up call is made first because the reference lazily
   initialises static tables without synchronisation (infl.c:1004-1006,1044-1052). */
double
ref_infl_buf_batch_mt(const uint8_t *src, const uint64_t *src_off,
                      uint8_t *dst, const uint64_t *dst_off,
                      uint64_t n, int flags, int nthreads,
                      const uint32_t *order, int32_t *codes) {
  static const uint8_t warm[] = {0x4B, 0x04, 0x00}; /* fixed block "a" */
  uint8_t         wout[8];
  pthread_t       th[256];
  ref_job_t       job;
  struct timespec t0, t1;
  int             i;

  (void)infl_buf(warm, sizeof(warm), wout, sizeof(wout), 0);

  if (nthreads < 1)   nthreads = 1;
  if (nthreads > 256) nthreads = 256;

  job.src = src; job.src_off = src_off; job.dst = dst; job.dst_off = dst_off;
  job.codes = codes; job.order = order; job.n = n; job.flags = flags;
  atomic_init(&job.next, 0);

  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (i = 1; i < nthreads; i++) pthread_create(&th[i], NULL, ref_worker, &job);
  ref_worker(&job);
  for (i = 1; i < nthreads; i++) pthread_join(th[i], NULL);
  clock_gettime(CLOCK_MONOTONIC, &t1);

  return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
