/*
 * oracle/inflate_oracle.h -- TEST INFRASTRUCTURE ONLY.
 * CPU restatement of the reference's one-shot inflate path (see inflate_oracle.c).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this; the product (defl_b200/) never does.
 */
#ifndef inflate_oracle_h
#define inflate_oracle_h

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORACLE_OK    0
#define ORACLE_ERR  (-1)
#define ORACLE_NOOP  1

typedef struct oracle_result_t {
  int32_t  code;          /* 0 / -1 / 1, what the reference's infl() returns            */
  uint32_t adler;         /* Adler-32 of dst[0..outlen) (new capability, RFC 1950)      */
  uint64_t outlen;        /* infl_output_pos()                                          */
  uint64_t inpos;         /* infl_input_pos(): compressed bytes consumed                */
  uint32_t adler_stored;  /* big-endian trailer after the deflate body, if present      */
  uint32_t has_trailer;   /* 1 when 4 trailer bytes were available                      */
  uint32_t nblocks;       /* deflate blocks decoded                                     */
  uint32_t reserved;
} oracle_result_t;

uint32_t oracle_adler32(uint32_t adler, const uint8_t *buf, size_t len);

int oracle_inflate(const uint8_t *src, size_t srclen, uint8_t *dst, size_t dstcap,
                   int flags, oracle_result_t *res);

/* logical input = concatenation of chunks (src + offs[i] .. src + offs[i+1]) */
int oracle_inflate_chunks(const uint8_t *src, const uint64_t *offs, uint32_t nchunks,
                          uint8_t *dst, size_t dstcap, int flags, oracle_result_t *res);

/* accept/reject of a code-length set, as infl_ft_build decides it */
int oracle_lens_ok(const uint8_t *lens, int n);
int oracle_table_fits(const uint8_t *lens, int n, unsigned rootbits, unsigned cap);

/* bit offsets (from the start of the deflate body) and output offsets of every block
   header in a valid stream; returns the number of blocks or -1 */
long oracle_block_index(const uint8_t *src, size_t srclen, int flags,
                        uint64_t *bitoffs, uint64_t *outoffs, size_t maxblocks);

double oracle_inflate_batch_mt(const uint8_t *src, const uint64_t *src_off,
                               uint8_t *dst, const uint64_t *dst_off,
                               uint64_t n, int flags, int nthreads, int32_t *codes);

#ifdef __cplusplus
}
#endif
#endif
