/* oracle/wrap_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into or loaded by the product).
 *
 * CPU restatement of what sits around the deflate body for the reference's callers: the reference
 * itself has no such code (its README.md:22 names recp/unz for gzip / ZIP and recp/im for PNG, README.md:120-155
 * shows the PNG flow "collect IDAT chunks, infl(), undo filters"), so the published algorithms are restated:
 *   CRC-32      RFC 1952 section 8 (the bit-at-a-time form of its sample code)
 *   gzip member RFC 1952 section 2.3
 *   PNG         ISO/IEC 15948 (PNG 1.2) sections 5.2-5.3 (signature, chunks, IHDR) and 9.2 (filter types 0-4)
 * Pinned by tests/test_wrap_oracle.py against Python's zlib.crc32 / gzip module and against forward filtering.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

uint32_t
wrap_oracle_crc32(const uint8_t *p, uint64_t n) {
  uint32_t c = 0xFFFFFFFFu;
  uint64_t i;
  int      k;

  for (i = 0; i < n; i++) {
    c ^= p[i];
    for (k = 0; k < 8; k++) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
  }
  return c ^ 0xFFFFFFFFu;
}

/* -> offset of the deflate body, 0 when this is not a gzip member with room for the 8-byte trailer */
uint32_t
wrap_oracle_gzip_body(const uint8_t *s, uint64_t n, uint32_t *crc, uint32_t *isize) {
  uint64_t at = 10;
  unsigned flg;

  if (n < 18 || s[0] != 0x1F || s[1] != 0x8B || s[2] != 8 || (s[3] & 0xE0)) return 0;
  flg = s[3];
  if (flg & 4) { if (at + 2 > n) return 0; at += 2u + (s[at] | ((unsigned)s[at + 1] << 8)); }
  if (flg & 8)  { while (at < n && s[at]) at++; at++; }
  if (flg & 16) { while (at < n && s[at]) at++; at++; }
  if (flg & 2) at += 2;
  if (at + 8 > n) return 0;
  *crc   = s[n - 8] | ((uint32_t)s[n - 7] << 8) | ((uint32_t)s[n - 6] << 16) | ((uint32_t)s[n - 5] << 24);
  *isize = s[n - 4] | ((uint32_t)s[n - 3] << 8) | ((uint32_t)s[n - 2] << 16) | ((uint32_t)s[n - 1] << 24);
  return (uint32_t)at;
}

static unsigned
paeth(unsigned a, unsigned b, unsigned c) {
  int p = (int)a + (int)b - (int)c;
  int pa = abs(p - (int)a), pb = abs(p - (int)b), pc = abs(p - (int)c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

/* raw: height x (1 + stride) -> pix: height x stride.  Returns 0, or 1 + row of the first bad filter byte. */
int
wrap_oracle_png_unfilter(const uint8_t *raw, uint8_t *pix, uint32_t height, uint32_t stride, uint32_t bpp) {
  uint32_t y, x;

  for (y = 0; y < height; y++) {
    const uint8_t *src = raw + (uint64_t)y * (stride + 1u);
    uint8_t       *dst = pix + (uint64_t)y * stride;
    const uint8_t *up  = y ? dst - stride : NULL;
    unsigned       ft  = src[0];
    if (ft > 4) return 1 + (int)y;
    for (x = 0; x < stride; x++) {
      unsigned r = src[1 + x], a = x >= bpp ? dst[x - bpp] : 0, b = up ? up[x] : 0, c = (up && x >= bpp) ? up[x - bpp] : 0, v;
      switch (ft) {
        case 1:  v = r + a; break;
        case 2:  v = r + b; break;
        case 3:  v = r + ((a + b) >> 1); break;
        case 4:  v = r + paeth(a, b, c); break;
        default: v = r; break;
      }
      dst[x] = (uint8_t)v;
    }
  }
  return 0;
}

static uint32_t be32(const uint8_t *p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

/* Walk a PNG file: IHDR fields into hdr[0..4] = width, height, bit depth, colour type, interlace; the IDAT payloads
   joined into z (capacity zcap).  Returns the zlib stream's length, or -1 when the file is not well formed. */
int64_t
wrap_oracle_png_idat(const uint8_t *f, uint64_t n, uint32_t *hdr, uint8_t *z, uint64_t zcap, uint32_t *nidat) {
  uint64_t at = 8, zl = 0;
  int      end = 0;

  *nidat = 0;
  if (n < 33 || be32(f) != 0x89504E47u || be32(f + 4) != 0x0D0A1A0Au || be32(f + 8) != 13 || be32(f + 12) != 0x49484452u) return -1;
  hdr[0] = be32(f + 16); hdr[1] = be32(f + 20); hdr[2] = f[24]; hdr[3] = f[25]; hdr[4] = f[28];
  if (f[26] || f[27]) return -1;
  while (!end && at + 12 <= n) {
    uint32_t len = be32(f + at), type = be32(f + at + 4);
    if (at + 12 + (uint64_t)len > n) return -1;
    if (type == 0x49444154u) {
      if (zl + len > zcap) return -1;
      memcpy(z + zl, f + at + 8, len);
      zl += len;
      (*nidat)++;
    } else if (type == 0x49454E44u) {
      end = 1;
    }
    at += 12 + (uint64_t)len;
  }
  return (end && *nidat) ? (int64_t)zl : -1;
}

/* ZIP (PKWARE APPNOTE 4.3): the central directory of an archive.  Fills up to cap records of
   {local_off, data_off, csize, usize, crc32, method | gpflags << 16, name_off, name_len | ok << 16} (8 x uint64)
   and returns the number of members, or -1 without a (non-ZIP64) end-of-central-directory record. */
static uint32_t le16(const uint8_t *p) { return p[0] | ((uint32_t)p[1] << 8); }
static uint32_t le32(const uint8_t *p) { return le16(p) | (le16(p + 2) << 16); }

int64_t
wrap_oracle_zip_list(const uint8_t *z, uint64_t n, uint64_t *out, uint32_t cap) {
  uint64_t e, stop, at;
  uint32_t total, cdsize, cdoff, k;

  if (n < 22) return -1;
  e = n - 22;
  stop = n > 22 + 65535 ? n - 22 - 65535 : 0;
  for (;; e--) {
    if (le32(z + e) == 0x06054B50u && e + 22 + le16(z + e + 20) <= n) break;
    if (e == stop) return -1;
  }
  total = le16(z + e + 10); cdsize = le32(z + e + 12); cdoff = le32(z + e + 16);
  if (total == 0xFFFFu || cdoff == 0xFFFFFFFFu || (uint64_t)cdoff + cdsize > e) return -1;
  at = cdoff;
  for (k = 0; k < total; k++) {
    uint32_t nl, xl, cl;
    if (at + 46 > e || le32(z + at) != 0x02014B50u) return -1;
    nl = le16(z + at + 28); xl = le16(z + at + 30); cl = le16(z + at + 32);
    if (k < cap) {
      uint64_t *r = out + 8u * k;
      uint32_t gp = le16(z + at + 8), method = le16(z + at + 10);
      int      ok;
      r[0] = le32(z + at + 42); r[1] = 0; r[2] = le32(z + at + 20); r[3] = le32(z + at + 24); r[4] = le32(z + at + 16);
      r[5] = method | ((uint64_t)gp << 16); r[6] = at + 46;
      ok = (method == 0 || method == 8) && !(gp & 1u) && r[2] != 0xFFFFFFFFu && r[3] != 0xFFFFFFFFu && r[0] != 0xFFFFFFFFu;
      if (ok) {
        ok = r[0] + 30 <= n && le32(z + r[0]) == 0x04034B50u;
        if (ok) { r[1] = r[0] + 30 + le16(z + r[0] + 26) + le16(z + r[0] + 28); ok = r[1] + r[2] <= n; }
      }
      r[7] = nl | ((uint64_t)(ok ? 1 : 0) << 16);
    }
    at += 46 + (uint64_t)nl + xl + cl;
  }
  return total;
}
