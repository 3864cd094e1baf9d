/*
 * defl/infl_wrap.h -- NEW entry points (not in the reference): what sits either side of the inflate hot path
 * for the reference's callers (README.md:22: `unz` for gzip / ZIP members, `im` for PNG; README.md:120-155 is
 * the PNG example: "collect IDAT chunks, infl(), undo filters").  SURVEY.md 8(f)-1, f-2, f-4.
 *
 *   infl_gzip_batch   RFC 1952 members: header skipped on the device, body through the batch inflate engine
 *                     (large members take the split-point decoder), CRC-32 and ISIZE of the trailer checked
 *                     against a CRC-32 computed on the device.
 *   infl_crc32_batch  CRC-32 (IEEE, as zlib's crc32()) of n buffers on the device.
 *   infl_png_batch    whole PNG files: the container is walked on the device (signature, IHDR, IDAT chunk list ->
 *                     gather descriptors: the IDATs are never joined by the host), the zlib stream is inflated,
 *                     the scanlines are un-filtered (None / Sub / Up / Average / Paeth) into packed pixels rows.
 *
 * Plain C ABI: pointers and sizes only; host pointers by default, device pointers with INFL_BATCH_DEVICE_PTRS.
 */
#ifndef infl_wrap_h
#define infl_wrap_h

#include "infl_batch.h"

#ifdef __cplusplus
extern "C" {
#endif

#define INFL_WRAP_VERIFY        0x1000  /* gzip: wrong CRC-32 / ISIZE => UNZ_ERR; PNG: wrong Adler-32 => UNZ_ERR */

#define INFL_WRAP_INFO_CRC_MATCH   0x0010u  /* gzip: computed CRC-32 == stored and ISIZE == outlen mod 2^32 */
#define INFL_WRAP_INFO_BAD_HEADER  0x0020u  /* not a gzip member / not a PNG file this path takes            */
#define INFL_WRAP_INFO_INTERLACED  0x0040u  /* PNG: Adam7 files are not un-filtered here (code = UNZ_ERR)     */
#define INFL_WRAP_INFO_BAD_FILTER  0x0080u  /* PNG: a scanline's filter byte was > 4                          */

typedef struct infl_gzip_result_t {
  int32_t  code;           /* UNZ_OK / UNZ_ERR                                              */
  uint32_t crc32;          /* CRC-32 of dst[0..outlen)                                      */
  uint64_t outlen;         /* bytes produced                                                */
  uint64_t inpos;          /* bytes of the member consumed, header and trailer included     */
  uint32_t crc32_stored;   /* the trailer's CRC-32                                          */
  uint32_t isize_stored;   /* the trailer's ISIZE                                           */
  uint32_t info;           /* INFL_INFO_SPLIT_DECODE, INFL_WRAP_INFO_*, reason << 8         */
  uint32_t pad_;
} infl_gzip_result_t;

/* items[i].src[0..srclen) = one gzip member -> items[i].dst[0..dstlen).  Returns UNZ_OK when the batch ran. */
UNZ_EXPORT int infl_gzip_batch(const infl_batch_item_t *items, size_t n, int flags, infl_gzip_result_t *results);

/* crc[i] = CRC-32 of items[i].src[0..srclen) (dst, dstlen ignored). */
UNZ_EXPORT int infl_crc32_batch(const infl_batch_item_t *items, size_t n, int flags, uint32_t *crc);

/* ---- ZIP archives (PKWARE APPNOTE 4.3; the `unz` caller) ------------------------------------------------------ */
typedef struct infl_zip_entry_t {
  uint64_t local_off;      /* offset of the local file header                                  */
  uint64_t data_off;       /* offset of the member's data                                      */
  uint64_t csize, usize;   /* compressed / uncompressed size (central directory)               */
  uint32_t crc32;
  uint16_t method;         /* 0 stored, 8 deflate                                              */
  uint16_t gpflags;
  uint32_t name_off;       /* of the file name inside the archive                              */
  uint16_t name_len;
  uint16_t ok;             /* 1: extractable here (method 0 / 8, not encrypted, no ZIP64 sizes) */
} infl_zip_entry_t;

/* Walk the central directory on the device.  entries[0..min(*count, cap)) are filled; *count = members in the
   archive.  UNZ_ERR when there is no (non-ZIP64) end-of-central-directory record. */
UNZ_EXPORT int infl_zip_list(const void *zip, uint64_t ziplen, int flags, infl_zip_entry_t *entries, size_t cap,
                             size_t *count);

/* Extract n members: entries[i] -> dst[i][0..entries[i].usize).  Deflated members go through the batch engine,
   stored ones are copied; CRC-32 is computed on the device and compared (results[i].crc32 / crc32_stored;
   INFL_WRAP_VERIFY makes a mismatch UNZ_ERR).  isize_stored receives the directory's uncompressed size. */
UNZ_EXPORT int infl_zip_extract(const void *zip, uint64_t ziplen, const infl_zip_entry_t *entries, size_t n,
                                void *const *dst, int flags, infl_gzip_result_t *results);

typedef struct infl_png_item_t {
  const void *png;         /* the whole file                                                 */
  uint64_t    pnglen;
  void       *pixels;      /* height rows of `stride` bytes, no filter bytes                 */
  uint64_t    pixcap;
} infl_png_item_t;

typedef struct infl_png_result_t {
  int32_t  code;           /* UNZ_OK / UNZ_ERR                                               */
  uint32_t width, height;
  uint8_t  bit_depth, color_type, interlace, bpp;   /* bpp = bytes per complete pixel, >= 1   */
  uint32_t stride;         /* bytes per row of pixels                                        */
  uint32_t nidat;          /* IDAT chunks the stream was spread over                         */
  uint64_t outlen;         /* height * stride when code == UNZ_OK                            */
  uint32_t info;
  uint32_t pad_;
} infl_png_result_t;

UNZ_EXPORT int infl_png_batch(const infl_png_item_t *items, size_t n, int flags, infl_png_result_t *results);

/* The un-filter step alone, device resident: raw[i] = height rows of (1 filter byte + stride bytes) ->
   pix[i] = height rows of stride bytes.  bad[i] != 0 when a filter byte was > 4.  Arrays are host arrays of
   device pointers / sizes. */
UNZ_EXPORT int infl_png_unfilter_device(const void *const *d_raw, void *const *d_pix, const uint32_t *height,
                                        const uint32_t *stride, const uint32_t *bpp, size_t n, uint32_t *bad,
                                        void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* infl_wrap_h */
