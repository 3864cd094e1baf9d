/*
 * oracle/inflate_oracle.c -- TEST INFRASTRUCTURE ONLY (the checker, never the product).
 *
 * Plain-C restatement of what the reference's one-shot decoder computes for a
 * contiguous input, i.e. the path infl_buf() takes:
 *   infl()               /root/reference/src/infl/infl.c:1507-1675
 *   infl_stored_direct   infl.c:206-267   (stored-only prefix)
 *   infl_ft_full         infl.c:1002-1104 (block loop, zlib header)
 *   infl_ft_stored       infl.c:659-715
 *   infl_ft_dynamic      infl.c:903-1000  (HLIT/HDIST/HCLEN + code-length decode)
 *   infl_ft_build        infl.c:452-578   (accept/reject of a code-length set)
 *   infl_ft_block        infl.c:717-901   (symbol decode + LZ77 copy)
 * It states the *semantics* (which bytes, which return code), not the reference's
 * table layout or copy tricks; decode is the RFC 1951 canonical walk behind a
 * direct-lookup table.  Adler-32 (RFC 1950) is added on top: the reference never
 * computes it (no "adler" anywhere under /root/reference).
 *
 * Pinning: tests/test_oracle.py checks this file against (a) the reference's whole
 * test corpus and inline KATs via committed fixtures under tests/golden/, (b) the
 * compiled reference oracle/_ref/libdefl_ref.so when present, (c) Python zlib.
 *
 * Known, deliberate scope limit: for inputs whose first block is stored and a later
 * block is not, the reference leaves its fast path and decodes through the absent
 * third-party huff library (infl.c:256-258 -> :1576-1664).  Valid streams give the
 * same bytes (RFC 1951); on *malformed* input that path's return code depends on huff
 * and on unsigned wrap-around in REFILL (infl.c:1106-1130) -- parity unpinned there.
 * This file applies the fast-path rules uniformly.
 */
#define _GNU_SOURCE
#include "inflate_oracle.h"
#include <string.h>
#include <stdlib.h>
#include <pthread.h>
#include <stdatomic.h>
#include <time.h>

#define MAXBITS   15
#define FASTBITS  11

/* RFC 1951 3.2.5 tables == reference lvals/dvals (src/infl/apicommon.h:36-49) */
static const uint16_t k_len_base[29] = {3,4,5,6,7,8,9,10,11,13,15,17,19,23,27,31,35,43,51,59,
                                        67,83,99,115,131,163,195,227,258};
static const uint8_t  k_len_xbits[29] = {0,0,0,0,0,0,0,0,1,1,1,1,2,2,2,2,3,3,3,3,4,4,4,4,5,5,5,5,0};
static const uint16_t k_dist_base[30] = {1,2,3,4,5,7,9,13,17,25,33,49,65,97,129,193,257,385,513,769,
                                         1025,1537,2049,3073,4097,6145,8193,12289,16385,24577};
static const uint8_t  k_dist_xbits[30] = {0,0,0,0,1,1,2,2,3,3,4,4,5,5,6,6,7,7,8,8,9,9,10,10,11,11,
                                          12,12,13,13};
/* src/infl/apicommon.h:51-52 */
static const uint8_t  k_clorder[19] = {16,17,18,0,8,7,9,6,10,5,11,4,12,3,13,2,14,1,15};

typedef struct bitreader_t {
  const uint8_t *p, *end;
  uint64_t       bits;   /* zero above nbits, like infl_ft_bits_t after its last refill */
  unsigned       nbits;
} bitreader_t;

static inline void
br_refill(bitreader_t *br) {
  while (br->nbits <= 56 && br->p < br->end) {
    br->bits  |= (uint64_t)*br->p++ << br->nbits;
    br->nbits += 8;
  }
}

static inline void
br_drop(bitreader_t *br, unsigned n) {
  br->bits >>= n;
  br->nbits -= n;
}

typedef struct hufftab_t {
  uint16_t count[MAXBITS + 1];
  uint16_t sorted[288];
  uint16_t fast[1 << FASTBITS]; /* sym << 4 | len, 0 = take the slow walk */
} hufftab_t;

static unsigned
bitrev(unsigned v, unsigned n) {
  unsigned r = 0;
  while (n--) { r = (r << 1) | (v & 1u); v >>= 1; }
  return r;
}

/* infl_ft_build's acceptance rule, infl.c:478-490: every length <= 15, Kraft sum not
   exceeded ("left < 0" => fail); incomplete and empty sets pass. */
int
oracle_lens_ok(const uint8_t *lens, int n) {
  int count[MAXBITS + 1] = {0}, left = 1, l, i;

  for (i = 0; i < n; i++) {
    if (lens[i] > MAXBITS) return 0;
    count[lens[i]]++;
  }
  for (l = 1; l <= MAXBITS; l++) {
    left = (left << 1) - count[l];
    if (left < 0) return 0;
  }
  return 1;
}

/* infl_ft_build's table budget, infl.c:500-529: a root table of 2^rootbits entries plus, for every root
   prefix that holds codes longer than the root, a sub-table of 2^(longest such code - rootbits) entries;
   more than `cap` entries in all => fail (infl.c:524-525; cap 2048 for literal/length at rootbits 10, 512 for
   distance at rootbits 8, infl.c:382-383).  Reachable: tests/test_table_caps.py has a 32-symbol distance set
   that needs 534.  The set is known to pass oracle_lens_ok. */
int
oracle_table_fits(const uint8_t *lens, int n, unsigned rootbits, unsigned cap) {
  unsigned count[MAXBITS + 1] = {0}, next[MAXBITS + 1], code = 0, used, l, p;
  uint8_t        needl[1u << 10];
  int            i;

  for (i = 0; i < n; i++) count[lens[i]]++;
  count[0] = 0;
  for (l = 1; l <= MAXBITS; l++) { code = (code + count[l - 1]) << 1; next[l] = code; }
  memset(needl, 0, sizeof(needl));
  for (i = 0; i < n; i++) {
    l = lens[i];
    if (!l) continue;
    code = next[l]++;
    if (l <= rootbits) continue;
    p = bitrev(code, l) & ((1u << rootbits) - 1u);
    if (needl[p] < l - rootbits) needl[p] = (uint8_t)(l - rootbits);
  }
  used = 1u << rootbits;
  for (p = 0; p < (1u << rootbits); p++)
    if (needl[p]) {
      if (used + (1u << needl[p]) > cap) return 0;
      used += 1u << needl[p];
    }
  return 1;
}

static int
huff_build(hufftab_t *t, const uint8_t *lens, int n) {
  uint16_t offs[MAXBITS + 2];
  unsigned code, idx, l, k, j;
  int      i;

  if (!oracle_lens_ok(lens, n)) return 0;

  memset(t->count, 0, sizeof(t->count));
  for (i = 0; i < n; i++) t->count[lens[i]]++;
  t->count[0] = 0;

  offs[1] = 0;
  for (l = 1; l <= MAXBITS; l++) offs[l + 1] = (uint16_t)(offs[l] + t->count[l]);
  for (i = 0; i < n; i++)
    if (lens[i]) t->sorted[offs[lens[i]]++] = (uint16_t)i;

  memset(t->fast, 0, sizeof(t->fast));
  code = idx = 0;
  for (l = 1; l <= FASTBITS; l++) {
    for (k = 0; k < t->count[l]; k++, code++, idx++) {
      unsigned rev = bitrev(code, l);
      for (j = rev; j < (1u << FASTBITS); j += 1u << l)
        t->fast[j] = (uint16_t)((t->sorted[idx] << 4) | l);
    }
    code <<= 1;
  }
  return 1;
}

/* decode one symbol from zero-padded bits; returns code length, 0 when no code of the
   set is a prefix of the bits (the reference's "entry == 0", infl.c:736-737,798-799) */
static inline unsigned
huff_decode(const hufftab_t *t, uint64_t bits, unsigned *sym) {
  unsigned e = t->fast[bits & ((1u << FASTBITS) - 1u)];
  unsigned code, first, index, l, cnt;

  if (e) { *sym = e >> 4; return e & 15u; }

  code = first = index = 0;
  for (l = 1; l <= MAXBITS; l++) {
    code |= (unsigned)(bits & 1u);
    bits >>= 1;
    cnt = t->count[l];
    if (code < first + cnt) { *sym = t->sorted[index + (code - first)]; return l; }
    index += cnt;
    first  = (first + cnt) << 1;
    code <<= 1;
  }
  return 0;
}

/* RFC 1950 checksum; not in the reference. */
uint32_t
oracle_adler32(uint32_t adler, const uint8_t *buf, size_t len) {
  uint32_t a = adler & 0xffffu, b = adler >> 16;

  while (len) {
    size_t n = len < 5552 ? len : 5552, i;
    for (i = 0; i < n; i++) { a += buf[i]; b += a; }
    a %= 65521u; b %= 65521u;
    buf += n; len -= n;
  }
  return (b << 16) | a;
}

/* infl_ft_stored, infl.c:659-715 */
static int
stored_block(bitreader_t *br, uint8_t *dst, size_t *pos, size_t cap) {
  unsigned len, nlen;
  size_t   take;

  br_drop(br, br->nbits & 7u);                 /* :669-671 */
  br_refill(br);
  if (br->nbits < 32) return ORACLE_ERR;       /* :673-675 */
  len  = (unsigned)(br->bits & 0xffffu);
  nlen = (unsigned)((br->bits >> 16) & 0xffffu);
  br_drop(br, 32);
  if ((len ^ (~nlen & 0xffffu)) || len > cap - *pos) return ORACLE_ERR;   /* :682-683 */

  /* whole bytes still sitting in the bit buffer first (:685-704), then the rest */
  while (len && br->nbits >= 8) {
    dst[(*pos)++] = (uint8_t)br->bits;
    br_drop(br, 8);
    len--;
  }
  take = len;
  if ((size_t)(br->end - br->p) < take) return ORACLE_ERR;  /* :706-707 */
  memcpy(dst + *pos, br->p, take);
  br->p += take;
  *pos  += take;
  return ORACLE_OK;
}

/* infl_ft_dynamic, infl.c:903-1000 */
static int
dynamic_header(bitreader_t *br, hufftab_t *tlit, hufftab_t *tdist) {
  uint8_t   cl[19] = {0}, lens[320];
  hufftab_t tcl;
  int       hlit, hdist, hclen, n, i;

  memset(lens, 0, sizeof(lens));
  br_refill(br);
  if (br->nbits < 14) return ORACLE_ERR;                 /* :917-919 */
  hlit  = (int)(br->bits & 0x1f) + 257;
  hdist = (int)((br->bits >> 5) & 0x1f) + 1;
  hclen = (int)((br->bits >> 10) & 0xf) + 4;
  n     = hlit + hdist;
  br_drop(br, 14);
  if (n > 320) return ORACLE_ERR;                        /* :927-928 (unreachable) */

  for (i = 0; i < hclen; i++) {                          /* :930-936 */
    br_refill(br);
    if (br->nbits < 3) return ORACLE_ERR;
    cl[k_clorder[i]] = (uint8_t)(br->bits & 7u);
    br_drop(br, 3);
  }
  /* huff_init_fast_lsb (:938): third-party; over-subscribed sets are rejected
     (test_files.c:581), incomplete ones decode until an unassigned pattern shows up */
  if (!huff_build(&tcl, cl, 19)) return ORACLE_ERR;

  for (i = 0; i < n;) {                                  /* :945-989 */
    unsigned sym, clen, rep;

    br_refill(br);
    if (br->nbits < 7) return ORACLE_ERR;                /* :947-948: 7 bits must be left even
                                                            for a shorter code */
    clen = huff_decode(&tcl, br->bits, &sym);
    if (!clen) return ORACLE_ERR;                        /* :951-952 */
    br_drop(br, clen);

    if (sym < 16) {
      lens[i++] = (uint8_t)sym;
    } else if (sym == 16) {
      if (br->nbits < 2 || i == 0) return ORACLE_ERR;    /* :960-961 */
      rep = 3 + (unsigned)(br->bits & 3u);
      br_drop(br, 2);
      if (i + (int)rep > n) return ORACLE_ERR;           /* :964-965 */
      while (rep--) { lens[i] = lens[i - 1]; i++; }
    } else if (sym == 17) {
      if (br->nbits < 3) return ORACLE_ERR;
      rep = 3 + (unsigned)(br->bits & 7u);
      br_drop(br, 3);
      if (i + (int)rep > n) return ORACLE_ERR;
      i += (int)rep;
    } else {
      if (br->nbits < 7) return ORACLE_ERR;
      rep = 11 + (unsigned)(br->bits & 0x7fu);
      br_drop(br, 7);
      if (i + (int)rep > n) return ORACLE_ERR;
      i += (int)rep;
    }
  }

  /* :991-997 -- no HLIT<=286 / HDIST<=30 / "has end-of-block" checks, unlike zlib */
  if (!huff_build(tlit, lens, hlit) || !huff_build(tdist, lens + hlit, hdist) ||
      !oracle_table_fits(lens, (int)hlit, 10, 2048) || !oracle_table_fits(lens + hlit, (int)hdist, 8, 512))
    return ORACLE_ERR;
  return ORACLE_OK;
}

/* infl_ft_block, infl.c:717-901 */
static int
huffman_block(bitreader_t *br, uint8_t *dst, size_t *dpos, size_t cap,
              const hufftab_t *tlit, const hufftab_t *tdist) {
  size_t pos = *dpos;

  for (;;) {
    unsigned sym, clen, xb, len, dist, i;

    br_refill(br);
    clen = huff_decode(tlit, br->bits, &sym);
    if (!clen || sym > 285) return ORACLE_ERR;           /* :736-737; 286/287 have no entry :433-438 */

    if (sym < 256) {
      if (br->nbits < clen) return ORACLE_ERR;           /* :741-745 */
      if (pos >= cap) return ORACLE_ERR;                 /* :751-752 (EFULL -> ERR :1078-1087) */
      br_drop(br, clen);
      dst[pos++] = (uint8_t)sym;
      continue;
    }
    if (sym == 256) {
      if (br->nbits < clen) return ORACLE_ERR;
      br_drop(br, clen);                                 /* :782-785 */
      break;
    }

    xb = k_len_xbits[sym - 257];
    if (br->nbits < clen + xb) return ORACLE_ERR;        /* :741-745, total = code + extra */
    len = k_len_base[sym - 257] + (unsigned)((br->bits >> clen) & ((1u << xb) - 1u));
    br_drop(br, clen + xb);

    br_refill(br);
    clen = huff_decode(tdist, br->bits, &sym);
    if (!clen || sym > 29) return ORACLE_ERR;            /* :798-799; 30/31 have no entry :445-446 */
    xb = k_dist_xbits[sym];
    if (br->nbits < clen + xb) return ORACLE_ERR;        /* :806-810 */
    dist = k_dist_base[sym] + (unsigned)((br->bits >> clen) & ((1u << xb) - 1u));
    br_drop(br, clen + xb);

    if ((size_t)dist > pos) return ORACLE_ERR;           /* :816-817 */
    if (len > cap - pos) return ORACLE_ERR;              /* :819-821 */
    for (i = 0; i < len; i++, pos++) dst[pos] = dst[pos - dist];
  }

  *dpos = pos;
  return ORACLE_OK;
}

static void
fixed_tables(hufftab_t *tlit, hufftab_t *tdist) {
  uint8_t lens[320];
  int     i;

  /* src/infl/apicommon.h:53-64 (fxd): 144x8, 112x9, 24x7, 8x8, then 32x5 */
  for (i = 0; i < 144; i++) lens[i] = 8;
  for (; i < 256; i++)      lens[i] = 9;
  for (; i < 280; i++)      lens[i] = 7;
  for (; i < 288; i++)      lens[i] = 8;
  for (; i < 320; i++)      lens[i] = 5;
  huff_build(tlit, lens, 288);
  huff_build(tdist, lens + 288, 32);
}

static int
inflate_core(const uint8_t *src, size_t srclen, uint8_t *dst, size_t cap, int flags,
             oracle_result_t *res, uint64_t *bitoffs, uint64_t *outoffs, size_t maxblocks) {
  hufftab_t   tlit, tdist, flit, fdist;
  bitreader_t br;
  size_t      pos = 0, body = 0;
  unsigned    bfinal = 0, btype, nblocks = 0;
  int         fixed_ready = 0, rc;

  memset(res, 0, sizeof(*res));
  res->adler = 1;

  /* all chunks empty => UNZ_OK with nothing produced (infl.c:1547-1550), flags 0 or 1 */
  if (srclen == 0) { res->code = ORACLE_OK; return ORACLE_OK; }

  if (flags == 1) {                                     /* INFL_ZLIB by equality, infl.c:1026 */
    if (srclen < 2) goto err;                           /* :1030-1031 */
    if ((src[0] & 0x0f) != 8 || (src[0] >> 4) > 7 ||
        ((((unsigned)src[0] << 8) + src[1]) % 31) != 0 || (src[1] & 0x20))
      goto err;                                         /* :1033-1038, zlib.h:45-96 */
    body = 2;
  }

  br.p = src + body; br.end = src + srclen; br.bits = 0; br.nbits = 0;

  while (!bfinal) {                                     /* :1063-1092 */
    br_refill(&br);
    if (br.nbits < 3) goto err;
    if (bitoffs && nblocks < maxblocks) {
      bitoffs[nblocks] = (uint64_t)(br.p - (src + body)) * 8u - br.nbits;
      outoffs[nblocks] = pos;
    }
    bfinal = (unsigned)(br.bits & 1u);
    btype  = (unsigned)((br.bits >> 1) & 3u);
    br_drop(&br, 3);
    nblocks++;

    if (btype == 0) {
      rc = stored_block(&br, dst, &pos, cap);
    } else if (btype == 1) {
      if (!fixed_ready) { fixed_tables(&flit, &fdist); fixed_ready = 1; }
      rc = huffman_block(&br, dst, &pos, cap, &flit, &fdist);
    } else if (btype == 2) {
      rc = dynamic_header(&br, &tlit, &tdist);
      if (rc == ORACLE_OK) rc = huffman_block(&br, dst, &pos, cap, &tlit, &tdist);
    } else {
      rc = ORACLE_ERR;
    }
    if (rc != ORACLE_OK) goto err;
  }

  res->code    = ORACLE_OK;
  res->outlen  = pos;
  res->nblocks = nblocks;
  /* infl_input_pos, mem.c:296-329: loaded bytes minus whole unread bytes in the buffer */
  res->inpos   = (uint64_t)(br.p - src) - (br.nbits >> 3);
  res->adler   = oracle_adler32(1, dst, pos);
  if (flags == 1 && srclen - res->inpos >= 4) {
    const uint8_t *t = src + res->inpos;
    res->adler_stored = ((uint32_t)t[0] << 24) | ((uint32_t)t[1] << 16) | ((uint32_t)t[2] << 8) | t[3];
    res->has_trailer  = 1;
  }
  return ORACLE_OK;

err:
  res->code    = ORACLE_ERR;
  res->outlen  = pos;
  res->nblocks = nblocks;
  return ORACLE_ERR;
}

int
oracle_inflate(const uint8_t *src, size_t srclen, uint8_t *dst, size_t dstcap, int flags,
               oracle_result_t *res) {
  oracle_result_t tmp;
  return inflate_core(src, srclen, dst, dstcap, flags, res ? res : &tmp, NULL, NULL, 0);
}

long
oracle_block_index(const uint8_t *src, size_t srclen, int flags,
                   uint64_t *bitoffs, uint64_t *outoffs, size_t maxblocks) {
  oracle_result_t res;
  uint8_t        *tmp;
  size_t          cap = srclen * 1100 + 65536; /* deflate's max expansion is ~1032x */
  long            n;

  if (!(tmp = malloc(cap))) return -1;
  n = inflate_core(src, srclen, tmp, cap, flags, &res, bitoffs, outoffs, maxblocks) == ORACLE_OK
        ? (long)res.nblocks : -1;
  free(tmp);
  return n;
}

/* infl_include semantics, mem.c:114-216: the logical input is the concatenation of the
   chunks in call order; zero-length chunks are legal; no chunk at all => UNZ_NOOP
   (infl.c:1544-1545). */
int
oracle_inflate_chunks(const uint8_t *src, const uint64_t *offs, uint32_t nchunks,
                      uint8_t *dst, size_t dstcap, int flags, oracle_result_t *res) {
  oracle_result_t tmp;
  uint8_t        *joined;
  size_t          total = 0, at = 0;
  uint32_t        i;
  int             rc;

  if (!res) res = &tmp;
  if (nchunks == 0) {
    memset(res, 0, sizeof(*res));
    res->code = ORACLE_NOOP;
    return ORACLE_NOOP;
  }
  for (i = 0; i < nchunks; i++) total += (size_t)(offs[i + 1] - offs[i]);
  if (!(joined = malloc(total ? total : 1))) return ORACLE_ERR;
  for (i = 0; i < nchunks; i++) {
    size_t n = (size_t)(offs[i + 1] - offs[i]);
    memcpy(joined + at, src + offs[i], n);
    at += n;
  }
  rc = oracle_inflate(joined, total, dst, dstcap, flags, res);
  free(joined);
  return rc;
}

typedef struct port_job_t {
  const uint8_t  *src;
  const uint64_t *src_off;
  uint8_t        *dst;
  const uint64_t *dst_off;
  int32_t        *codes;
  uint64_t        n;
  int             flags;
  atomic_ullong   next;
} port_job_t;

static void *
port_worker(void *arg) {
  port_job_t *job = (port_job_t *)arg;

  for (;;) {
    uint64_t        i = atomic_fetch_add(&job->next, 1);
    oracle_result_t r;
    if (i >= job->n) break;
    job->codes[i] = inflate_core(job->src + job->src_off[i],
                                 (size_t)(job->src_off[i + 1] - job->src_off[i]),
                                 job->dst + job->dst_off[i],
                                 (size_t)(job->dst_off[i + 1] - job->dst_off[i]),
                                 job->flags, &r, NULL, NULL, 0);
  }
  return NULL;
}

double
oracle_inflate_batch_mt(const uint8_t *src, const uint64_t *src_off,
                        uint8_t *dst, const uint64_t *dst_off,
                        uint64_t n, int flags, int nthreads, int32_t *codes) {
  pthread_t       th[256];
  port_job_t      job;
  struct timespec t0, t1;
  int             i;

  if (nthreads < 1)   nthreads = 1;
  if (nthreads > 256) nthreads = 256;
  job.src = src; job.src_off = src_off; job.dst = dst; job.dst_off = dst_off;
  job.codes = codes; job.n = n; job.flags = flags;
  atomic_init(&job.next, 0);

  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (i = 1; i < nthreads; i++) pthread_create(&th[i], NULL, port_worker, &job);
  port_worker(&job);
  for (i = 1; i < nthreads; i++) pthread_join(th[i], NULL);
  clock_gettime(CLOCK_MONOTONIC, &t1);
  return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
