/*
 * oracle/huff/huff.h -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Clean-room stand-in for the third-party Huffman library `recp/huff`
 * (pinned by the reference at commit ee4138944a6718da84b0d6aa043fedaab3d6efbf,
 * /root/reference/.gitmodules:1-3), whose sources are NOT present under
 * /root/reference/deps/huff.  It exists so that the reference's own
 * src/infl/{infl,mem,stream}.c can be compiled *unmodified* into
 * oracle/_ref/libdefl_ref.so (see oracle/Makefile).
 *
 * It was written only from the way the reference calls the library
 * (src/infl/infl.c:460-469,911-957,1127,1251-1283,1555-1556,1620,1654-1655;
 *  src/infl/stream.c:49,271-330,511-512,642,690,759-760; src/common.h:80,133-134;
 *  src/infl/apicommon.h:36-49; test/test_files.c:550-601) and from RFC 1951
 *  section 3.2.2 (canonical Huffman codes, packed LSB-first).  The decode is the
 * textbook count/first-code walk, one bit at a time, preceded by an 8-bit
 * direct-lookup table.
 *
 * What this means for parity: the reference's contiguous fast path
 * (infl_stored_direct / infl_ft_full) uses this header only for the 19-symbol
 * code-length table (huff_init_fast_lsb), so its decode loops, speed and error
 * behaviour are the reference's own.  The chunk-list path (infl_block / REFILL)
 * decodes through this stand-in: bytes for valid streams are fixed by RFC 1951,
 * but its speed and its accept/reject decisions on malformed Huffman code sets
 * are this file's, not upstream huff's ("parity unpinned" for those).
 */
#ifndef oracle_huff_shim_h
#define oracle_huff_shim_h

#include <stdint.h>
#include <stdbool.h>
#include <stddef.h>
#include <string.h>

typedef uint64_t bitstream_t;

#define HUFF_MAX_CODE_LENGTH 15
#define HUFF_FAST_TABLE_BITS 8
#define HUFF_FAST_TABLE_SIZE (1 << HUFF_FAST_TABLE_BITS)
#define HUFF_SHIM_MAX_SYMS   320

typedef struct huff_ext_t {
  uint16_t base;
  uint8_t  bits;
  uint16_t mask;
} huff_ext_t;

typedef struct huff_fast_entry_t {
  uint8_t  len;
  uint16_t sym;
} huff_fast_entry_t;

typedef huff_fast_entry_t huff_fast_entry_ext_t;

typedef struct huff_table_ext_t {
  huff_fast_entry_ext_t fast[HUFF_FAST_TABLE_SIZE];
  uint16_t              count[HUFF_MAX_CODE_LENGTH + 1];
  uint16_t              sorted[HUFF_SHIM_MAX_SYMS];
  const huff_ext_t     *ext;
  int                   extof;
  int                   nsorted;
} huff_table_ext_t;

typedef huff_table_ext_t huff_table_t;

static inline unsigned
huff__bitrev(unsigned v, unsigned n) {
  unsigned r = 0;
  while (n--) { r = (r << 1) | (v & 1u); v >>= 1; }
  return r;
}

/* lengths -> per-length counts + symbols sorted by (length, index); false if a
   length exceeds HUFF_MAX_CODE_LENGTH or the set is over-subscribed.
   Incomplete and all-zero sets are accepted. */
static inline bool
huff__prepare(huff_table_ext_t *t, const uint8_t *lens, const uint16_t *syms, int n) {
  uint16_t start[HUFF_MAX_CODE_LENGTH + 2];
  int      i, l, left;

  memset(t->count, 0, sizeof(t->count));
  t->nsorted = 0;
  if (n < 0 || n > HUFF_SHIM_MAX_SYMS) return false;

  for (i = 0; i < n; i++) {
    if (lens[i] > HUFF_MAX_CODE_LENGTH) return false;
    t->count[lens[i]]++;
  }
  t->count[0] = 0;

  for (left = 1, l = 1; l <= HUFF_MAX_CODE_LENGTH; l++) {
    left = (left << 1) - (int)t->count[l];
    if (left < 0) return false;
  }

  start[1] = 0;
  for (l = 1; l <= HUFF_MAX_CODE_LENGTH; l++)
    start[l + 1] = (uint16_t)(start[l] + t->count[l]);
  t->nsorted = start[HUFF_MAX_CODE_LENGTH + 1];

  for (i = 0; i < n; i++)
    if (lens[i])
      t->sorted[start[lens[i]]++] = syms ? syms[i] : (uint16_t)i;

  return true;
}

static inline void
huff__fill_fast(huff_fast_entry_t *fast, const huff_table_ext_t *t) {
  unsigned code = 0, idx = 0, l, k, j;

  memset(fast, 0, sizeof(huff_fast_entry_t) * HUFF_FAST_TABLE_SIZE);
  for (l = 1; l <= HUFF_FAST_TABLE_BITS; l++) {
    for (k = 0; k < t->count[l]; k++, code++, idx++) {
      unsigned rev = huff__bitrev(code, l);
      for (j = rev; j < HUFF_FAST_TABLE_SIZE; j += 1u << l) {
        fast[j].len = (uint8_t)l;
        fast[j].sym = t->sorted[idx];
      }
    }
    code <<= 1;
  }
}

/* canonical walk: returns symbol, *used = code length, 0 if no code matches
   within `avail` bits. */
static inline unsigned
huff__walk(const huff_table_ext_t *t, bitstream_t bits, unsigned avail, uint8_t *used) {
  const huff_fast_entry_ext_t *fe = &t->fast[(uint8_t)bits];
  unsigned code, first, index, l, cnt;

  if (fe->len) {
    if (fe->len > avail) { *used = 0; return 0; }
    *used = fe->len;
    return fe->sym;
  }

  code = first = index = 0;
  for (l = 1; l <= HUFF_MAX_CODE_LENGTH && l <= avail; l++) {
    code |= (unsigned)(bits & 1u);
    bits >>= 1;
    cnt = t->count[l];
    if (code < first + cnt) {
      *used = (uint8_t)l;
      return t->sorted[index + (code - first)];
    }
    index += cnt;
    first  = (first + cnt) << 1;
    code <<= 1;
  }
  *used = 0;
  return 0;
}

static inline bool
huff_init_lsb_extof(huff_table_ext_t *t, const uint8_t *lens, const uint16_t *syms,
                    const huff_ext_t *ext, int extof, int n) {
  t->ext   = ext;
  t->extof = extof;
  if (!huff__prepare(t, lens, syms, n)) return false;
  huff__fill_fast(t->fast, t);
  return true;
}

static inline bool
huff_init_lsb_ext(huff_table_ext_t *t, const uint8_t *lens, const uint16_t *syms,
                  const huff_ext_t *ext, int n) {
  return huff_init_lsb_extof(t, lens, syms, ext, 0, n);
}

static inline bool
huff_init_lsb(huff_table_t *t, const uint8_t *lens, const uint16_t *syms, int n) {
  return huff_init_lsb_extof(t, lens, syms, NULL, 0x7fffffff, n);
}

static inline bool
huff_init_fast_lsb(huff_fast_entry_t *fast, const uint8_t *lens, const uint16_t *syms, int n) {
  huff_table_ext_t t;
  if (!huff__prepare(&t, lens, syms, n)) return false;
  huff__fill_fast(fast, &t);
  return true;
}

static inline unsigned
huff_decode_lsb(const huff_table_t *t, bitstream_t bits, unsigned nbits, uint8_t *used) {
  return huff__walk(t, bits, nbits, used);
}

static inline unsigned
huff_decode_lsb_extof_safe(const huff_table_ext_t *t, bitstream_t bits, uint8_t nbits,
                           uint8_t *used, unsigned *val, int extof) {
  uint8_t  cl;
  unsigned sym = huff__walk(t, bits, nbits, &cl);

  if (!cl) { *used = 0; return 0; }
  if ((int)sym >= extof && t->ext) {
    const huff_ext_t *e = &t->ext[(int)sym - extof];
    if ((unsigned)cl + e->bits > nbits) { *used = 0; return 0; }
    *val  = e->base + (unsigned)((bits >> cl) & e->mask);
    *used = (uint8_t)(cl + e->bits);
    return sym;
  }
  *used = cl;
  return sym;
}

static inline unsigned
huff_decode_lsb_extof(const huff_table_ext_t *t, bitstream_t bits,
                      uint8_t *used, unsigned *val, int extof) {
  return huff_decode_lsb_extof_safe(t, bits, 64, used, val, extof);
}

static inline unsigned
huff_decode_lsb_ext_safe(const huff_table_ext_t *t, bitstream_t bits, uint8_t nbits,
                         uint8_t *used) {
  unsigned val = 0;
  (void)huff_decode_lsb_extof_safe(t, bits, nbits, used, &val, 0);
  return *used ? val : 0;
}

static inline unsigned
huff_decode_lsb_ext(const huff_table_ext_t *t, bitstream_t bits, uint8_t *used) {
  return huff_decode_lsb_ext_safe(t, bits, 64, used);
}

/* load up to 8 little-endian bytes, advance *p, return the number of bits. */
static inline unsigned
huff_read(const uint8_t **p, bitstream_t *out, const uint8_t *end) {
  size_t      n = (size_t)(end - *p), i;
  bitstream_t v = 0;

  if (n > 8) n = 8;
  for (i = 0; i < n; i++) v |= (bitstream_t)(*p)[i] << (8 * i);
  *p  += n;
  *out = v;
  return (unsigned)(n * 8);
}

#endif /* oracle_huff_shim_h */
