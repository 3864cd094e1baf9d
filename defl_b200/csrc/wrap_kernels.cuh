// wrap_kernels.cuh -- what sits around the deflate body for the callers either side of the hot path
// (SURVEY.md 8f): CRC-32 (gzip / ZIP / PNG chunks), the gzip member header, the PNG container walk and the PNG
// scanline un-filter.  The reference has none of this (its callers recp/unz and recp/im do, README.md:22,120-155);
// results are defined by RFC 1952, the PNG specification and oracle/wrap_oracle.c.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace defl {

// ---- CRC-32 (IEEE 802.3, reflected, polynomial 0xEDB88320) ---------------------------------------------------
// "Raw" CRC (register starts at 0, no final xor) is linear over GF(2): raw(A || B) = shift(raw(A), len B) ^ raw(B),
// where shift multiplies by x^(8 len) modulo the polynomial.  A block of 256 threads takes one job of up to
// 64 KiB: every thread runs the byte-at-a-time table over its own 256-byte slice, the slices are combined in a
// tree with the precomputed shift matrices, the block writes raw(job).  The host combines a stream's jobs and
// applies the init / final xor terms (crc = raw ^ shift(0xFFFFFFFF, total) ^ 0xFFFFFFFF).
constexpr uint32_t kCrcJobBytes  = 65536;
constexpr uint32_t kCrcSlice     = 256;

struct CrcJob {
  const uint8_t* p;
  uint32_t       len;      // <= kCrcJobBytes
  uint32_t       pad;
};

// shift_mat[k][i]: image of bit i under "append 2^k zero bytes" (k = 0..31); filled by the host once per device
__constant__ uint32_t c_crc_shift[32][32];
__constant__ uint32_t c_crc_table[256];

__device__ __forceinline__ uint32_t crc_mat_apply(const uint32_t* mat, uint32_t v) {
  uint32_t r = 0;
#pragma unroll 8
  for (int i = 0; i < 32; i++) r ^= (v >> i) & 1u ? mat[i] : 0u;
  return r;
}
// append n zero bytes to a raw CRC
__device__ __forceinline__ uint32_t crc_shift(uint32_t crc, uint32_t n) {
  for (int k = 0; n; k++, n >>= 1)
    if (n & 1u) crc = crc_mat_apply(c_crc_shift[k], crc);
  return crc;
}

__global__ void __launch_bounds__(256)
crc32_jobs_kernel(const CrcJob* __restrict__ jobs, uint32_t njobs, uint32_t* __restrict__ raw_out) {
  __shared__ uint32_t tab[256];
  __shared__ uint32_t part[256];
  tab[threadIdx.x] = c_crc_table[threadIdx.x];
  __syncthreads();
  for (uint32_t j = blockIdx.x; j < njobs; j += gridDim.x) {
    const CrcJob job = jobs[j];
    // The job is taken as if it had leading zero bytes up to 64 KiB (they do not change a raw CRC): slice t then
    // always ENDS 256 * (255 - t) bytes before the job's end, and the slices combine with fixed shifts.
    const int64_t  s0 = (int64_t)job.len - (int64_t)kCrcJobBytes + (int64_t)threadIdx.x * kCrcSlice;   // may be < 0
    const uint32_t lo = s0 < 0 ? 0u : (uint32_t)s0;
    const int64_t  hi = s0 + kCrcSlice;
    const uint32_t n  = hi > 0 ? (uint32_t)hi - lo : 0u;
    const uint8_t* p  = job.p + lo;
    uint32_t crc = 0;
    uint32_t i = 0;
    for (; i < n && (((uintptr_t)(p + i)) & 15u); i++) crc = tab[(crc ^ __ldg(p + i)) & 255u] ^ (crc >> 8);
    for (; i + 16u <= n; i += 16u) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(p + i));
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        uint32_t x = w[k];
#pragma unroll
        for (int b = 0; b < 4; b++) { crc = tab[(crc ^ x) & 255u] ^ (crc >> 8); x >>= 8; }
      }
    }
    for (; i < n; i++) crc = tab[(crc ^ __ldg(p + i)) & 255u] ^ (crc >> 8);
    part[threadIdx.x] = crc;
    __syncthreads();
    // tree: raw(A || B) = shift(raw(A), |B|) ^ raw(B) with |B| = 256 * 2^k at level k (matrix 8 + k, one broadcast
    // constant-memory read per row)
#pragma unroll 1
    for (uint32_t k = 0; k < 8; k++) {
      const uint32_t step = 1u << k;
      if ((threadIdx.x & (2u * step - 1u)) == 0u)
        part[threadIdx.x] = crc_mat_apply(c_crc_shift[8 + k], part[threadIdx.x]) ^ part[threadIdx.x + step];
      __syncthreads();
    }
    if (threadIdx.x == 0) raw_out[j] = part[0];
    __syncthreads();
  }
}

// ---- gzip member header (RFC 1952 2.3) ------------------------------------------------------------------------
// One thread per stream: where the deflate body starts, and the trailer (CRC-32, ISIZE) from the last 8 bytes.
struct GzipInfo {
  uint32_t body;       // offset of the deflate body; 0 = not a gzip member / truncated header
  uint32_t crc, isize; // the trailer's fields (little endian), valid when trailer != 0
  uint32_t trailer;
};

__device__ __forceinline__ GzipInfo gzip_parse(const uint8_t* s, uint64_t n) {
  GzipInfo g;
  g.body = 0; g.crc = 0; g.isize = 0; g.trailer = 0;
  if (n < 18 || s[0] != 0x1F || s[1] != 0x8B || s[2] != 8 || (s[3] & 0xE0)) return g;
  const uint32_t flg = s[3];
  uint64_t at = 10;
  if (flg & 4) {                                            // FEXTRA
    if (at + 2 > n) return g;
    at += 2u + (s[at] | ((uint32_t)s[at + 1] << 8));
  }
  for (int f = 0; f < 2; f++)                               // FNAME, FCOMMENT: zero-terminated
    if (flg & (f ? 16 : 8)) {
      while (at < n && s[at]) at++;
      at++;
    }
  if (flg & 2) at += 2;                                     // FHCRC
  if (at + 8 > n) return g;
  g.body = (uint32_t)at;
  const uint8_t* t = s + n - 8;
  g.crc   = t[0] | ((uint32_t)t[1] << 8) | ((uint32_t)t[2] << 16) | ((uint32_t)t[3] << 24);
  g.isize = t[4] | ((uint32_t)t[5] << 8) | ((uint32_t)t[6] << 16) | ((uint32_t)t[7] << 24);
  g.trailer = 1;
  return g;
}

__global__ void gzip_header_kernel(const uint8_t* const* __restrict__ src, const uint64_t* __restrict__ len, uint32_t n,
                                   GzipInfo* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = gzip_parse(src[i], len[i]);
}

// ---- ZIP archives (APPNOTE 4.3: end of central directory, central file headers, local file headers) ------------
// One thread walks the central directory of an archive that lives on the device and fills one record per member.
struct ZipEntry {
  uint64_t local_off;     // offset of the local file header
  uint64_t data_off;      // offset of the member's (compressed) data
  uint64_t csize, usize;
  uint32_t crc32;
  uint16_t method;        // 0 stored, 8 deflate
  uint16_t gpflags;
  uint32_t name_off;      // of the name inside the archive (central directory copy)
  uint16_t name_len;
  uint16_t ok;            // 1: a member this path can extract (method 0 / 8, not encrypted, no ZIP64 sizes)
};

__device__ __forceinline__ uint32_t le16(const uint8_t* p) { return p[0] | ((uint32_t)p[1] << 8); }
__device__ __forceinline__ uint32_t le32(const uint8_t* p) { return le16(p) | (le16(p + 2) << 16); }

// out[0].. up to cap entries; *count = members in the directory (may exceed cap), or 0xFFFFFFFF when no directory
__global__ void zip_walk_kernel(const uint8_t* __restrict__ z, uint64_t n, ZipEntry* __restrict__ out, uint32_t cap,
                                uint32_t* __restrict__ count) {
  if (blockIdx.x || threadIdx.x) return;
  *count = 0xFFFFFFFFu;
  if (n < 22) return;
  uint64_t e = n - 22;                                      // end of central directory record: scan back over the comment
  const uint64_t stop = n > 22 + 65535 ? n - 22 - 65535 : 0;
  for (;; e--) {
    if (le32(z + e) == 0x06054B50u && e + 22 + le16(z + e + 20) <= n) break;
    if (e == stop) return;
  }
  const uint32_t total = le16(z + e + 10), cdsize = le32(z + e + 12), cdoff = le32(z + e + 16);
  if (total == 0xFFFFu || cdoff == 0xFFFFFFFFu || (uint64_t)cdoff + cdsize > e) return;     // ZIP64: not here
  uint64_t at = cdoff;
  uint32_t k = 0;
  for (; k < total; k++) {
    if (at + 46 > e || le32(z + at) != 0x02014B50u) return;
    const uint32_t nl = le16(z + at + 28), xl = le16(z + at + 30), cl = le16(z + at + 32);
    if (k < cap) {
      ZipEntry r;
      r.gpflags = (uint16_t)le16(z + at + 8); r.method = (uint16_t)le16(z + at + 10);
      r.crc32 = le32(z + at + 16); r.csize = le32(z + at + 20); r.usize = le32(z + at + 24);
      r.local_off = le32(z + at + 42);
      r.name_off = (uint32_t)(at + 46); r.name_len = (uint16_t)nl;
      r.data_off = 0;
      bool ok = (r.method == 0 || r.method == 8) && !(r.gpflags & 1u) && r.csize != 0xFFFFFFFFu && r.usize != 0xFFFFFFFFu &&
                r.local_off != 0xFFFFFFFFu;
      if (ok) {
        const uint64_t lo = r.local_off;
        ok = lo + 30 <= n && le32(z + lo) == 0x04034B50u;
        if (ok) {
          r.data_off = lo + 30 + le16(z + lo + 26) + le16(z + lo + 28);
          ok = r.data_off + r.csize <= n;
        }
      }
      r.ok = ok ? 1 : 0;
      out[k] = r;
    }
    at += 46 + (uint64_t)nl + xl + cl;
  }
  *count = total;
}

// ---- PNG container walk (PNG spec 5.3; the reference's motivating caller, README.md:120-155) ---------------------
// One thread per file: check the signature, read IHDR, and list the IDAT chunks' payloads as gather descriptors
// (pointer, length, offset in the joined zlib stream) -- the host never touches the file's bytes.
struct PngInfo {
  uint32_t width, height;
  uint8_t  bit_depth, color_type, interlace, ok;    // ok = 0: not a PNG this path takes
  uint32_t nidat;
  uint64_t zlen;                                      // bytes of the zlib stream (all IDAT payloads)
};

__device__ __forceinline__ uint32_t be32(const uint8_t* p) {
  return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}

// seg_base == nullptr: count only (IHDR fields, number of IDATs, bytes of the zlib stream).  Otherwise file i
// writes its descriptors from seg_base[i] on, with offsets into the joined buffer starting at z_base[i].
__global__ void png_walk_kernel(const uint8_t* const* __restrict__ files, const uint64_t* __restrict__ flen, uint32_t n,
                                PngInfo* __restrict__ info, const uint64_t* __restrict__ seg_base,
                                const uint64_t* __restrict__ z_base, const void** __restrict__ seg_src,
                                uint64_t* __restrict__ seg_len, uint64_t* __restrict__ seg_off) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint8_t* f = files[i];
  const uint64_t L = flen[i];
  PngInfo g;
  g.width = g.height = 0; g.bit_depth = g.color_type = g.interlace = g.ok = 0; g.nidat = 0; g.zlen = 0;
  const uint64_t s0 = seg_base ? seg_base[i] : 0u, z0 = seg_base ? z_base[i] : 0u;
  bool good = L >= 8 + 25 && be32(f) == 0x89504E47u && be32(f + 4) == 0x0D0A1A0Au && be32(f + 8) == 13u && be32(f + 12) == 0x49484452u;
  if (good) {
    g.width = be32(f + 16); g.height = be32(f + 20);
    g.bit_depth = f[24]; g.color_type = f[25]; g.interlace = f[28];
    good = f[26] == 0 && f[27] == 0;
    uint64_t at = 8;
    bool end = false;
    while (good && !end && at + 12 <= L) {
      const uint32_t clen = be32(f + at), type = be32(f + at + 4);
      if (at + 12 + (uint64_t)clen > L) { good = false; break; }
      if (type == 0x49444154u) {                       // IDAT
        if (seg_base) {
          seg_src[s0 + g.nidat] = f + at + 8;
          seg_len[s0 + g.nidat] = clen;
          seg_off[s0 + g.nidat] = z0 + g.zlen;
        }
        g.nidat++;
        g.zlen += clen;
      } else if (type == 0x49454E44u) {                // IEND
        end = true;
      }
      at += 12 + (uint64_t)clen;
    }
    good = good && end && g.nidat > 0;
  }
  g.ok = good ? 1 : 0;
  if (!seg_base) info[i] = g;
}

// ---- PNG scanline un-filter (PNG spec 9.2: None / Sub / Up / Average / Paeth) -------------------------------------
// raw = height x (1 filter byte + stride bytes), as the zlib stream decodes; pix = height x stride.
// One WARP per image: lane l takes row (32 k + l) of a group of 32 rows and the lanes walk along their rows as a
// wavefront, lane l one pixel behind lane l - 1, so the row above (the bytes at x and x - bpp of it) arrives by
// shuffle from the neighbouring lane when it is needed; only lane 0 reads the previous group's last row from
// memory.  bpp = bytes per complete pixel, at least 1 (PNG spec 9.2); one step = one pixel = bpp bytes.
struct PngJob {
  const uint8_t* raw;
  uint8_t*       pix;
  uint32_t       height, stride, bpp, bad;    // bad: set by the kernel when a filter byte is > 4
};

__device__ __forceinline__ uint32_t paeth(uint32_t a, uint32_t b, uint32_t c) {
  const int p = (int)a + (int)b - (int)c;
  const int pa = abs(p - (int)a), pb = abs(p - (int)b), pc = abs(p - (int)c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

__global__ void __launch_bounds__(128)
png_unfilter_kernel(PngJob* __restrict__ jobs, uint32_t njobs) {
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const uint32_t nw = (gridDim.x * blockDim.x) >> 5;
  for (uint32_t j = warp; j < njobs; j += nw) {
    const PngJob job = jobs[j];
    const uint32_t bpp = job.bpp, stride = job.stride, npx = (stride + bpp - 1) / bpp;   // steps per row
    const size_t   rawstride = (size_t)stride + 1;
    bool bad = false;
    if (bpp == 4u && (stride & 3u) == 0u && (((uintptr_t)job.pix) & 3u) == 0u) {
      // 4-byte pixels (RGBA8, grey + alpha 16): a pixel is one 32-bit word.  The scanline data is read as aligned
      // words two steps ahead of their use (the row starts 1 byte after the filter byte, so a pixel straddles two
      // words: one funnel shift), the result is stored as one word; the four byte lanes are computed on packed bytes.
      for (uint32_t y0 = 0; y0 < job.height; y0 += 32) {
        const uint32_t y = y0 + lane;
        const bool     live = y < job.height;
        const uint8_t* src = job.raw + (size_t)(live ? y : 0) * rawstride;
        uint32_t*      dst = reinterpret_cast<uint32_t*>(job.pix + (size_t)(live ? y : 0) * stride);
        const uint32_t ft = live ? src[0] : 0u;
        bad = bad || ft > 4u;
        const uint32_t* up_row = (y0 > 0) ? reinterpret_cast<const uint32_t*>(job.pix + (size_t)(y0 - 1) * stride) : nullptr;
        const uint32_t* sw = reinterpret_cast<const uint32_t*>((uintptr_t)(src + 1) & ~(uintptr_t)3);
        const uint32_t  sh = (uint32_t)((uintptr_t)(src + 1) & 3u) * 8u;
        const uint32_t  nwords = npx + 1;                      // aligned words that hold the row's bytes (one more when sh != 0)
        uint32_t w0 = live ? __ldg(sw) : 0u, w1 = (live && 1u < nwords) ? __ldg(sw + 1) : 0u, w2 = 0u;
        uint32_t A = 0, C = 0, CUR = 0;
        for (uint32_t t = 0; t < npx + 31; t++) {
          const uint32_t s = t - lane;
          const bool act = live && t >= lane && s < npx;
          if (act && s + 2u < nwords) w2 = __ldg(sw + s + 2u);  // lands while the next two pixels are computed
          uint32_t UP = __shfl_up_sync(0xFFFFFFFFu, CUR, 1);
          if (lane == 0) UP = (act && up_row) ? up_row[s] : 0u;
          if (act) {
            const uint32_t r = __funnelshift_r(w0, w1, sh);
            uint32_t pred;
            switch (ft) {
              case 1:  pred = A; break;
              case 2:  pred = UP; break;
              case 3:  pred = (A & UP) + (((A ^ UP) & 0xFEFEFEFEu) >> 1); break;      // per-byte floor((a + b) / 2)
              case 4: {
                pred = 0;
#pragma unroll
                for (int k = 0; k < 4; k++)
                  pred |= paeth((A >> (8 * k)) & 255u, (UP >> (8 * k)) & 255u, (C >> (8 * k)) & 255u) << (8 * k);
                break;
              }
              default: pred = 0; break;
            }
            // per-byte add modulo 256
            const uint32_t v = ((r & 0x7F7F7F7Fu) + (pred & 0x7F7F7F7Fu)) ^ ((r ^ pred) & 0x80808080u);
            dst[s] = v;
            A = v; C = UP; CUR = v;
            w0 = w1; w1 = w2;
          }
        }
        __syncwarp();
      }
      if (__any_sync(0xFFFFFFFFu, bad) && lane == 0) jobs[j].bad = 1;
      continue;
    }
    for (uint32_t y0 = 0; y0 < job.height; y0 += 32) {
      const uint32_t y = y0 + lane;
      const bool     live = y < job.height;
      const uint8_t* src = job.raw + (size_t)(live ? y : 0) * rawstride;
      uint8_t*       dst = job.pix + (size_t)(live ? y : 0) * stride;
      const uint32_t ft = live ? src[0] : 0u;
      bad = bad || ft > 4u;
      const uint8_t* up_row = (y0 > 0) ? job.pix + (size_t)(y0 - 1) * stride : nullptr;   // lane 0's row above
      // a pixel's bytes (up to 8) packed in 64 bits: A = the pixel to the left, C = upper left, CUR = what this lane
      // produced last, UP = the pixel above (from lane - 1's CUR of one step ago, or from memory for lane 0)
      uint64_t A = 0, C = 0, CUR = 0;
      // lane l runs its step s at global step t = s + l: steps 0 .. npx - 1 + 31
      for (uint32_t t = 0; t < npx + 31; t++) {
        const uint32_t s = t - lane;                         // my pixel index (wraps when t < lane: not active)
        const bool act = live && t >= lane && s < npx;
        uint64_t UP = __shfl_up_sync(0xFFFFFFFFu, CUR, 1);
        if (lane == 0) {
          UP = 0;
          if (act && up_row)
            for (uint32_t k = 0; k < bpp; k++) {
              const uint32_t x = s * bpp + k;
              if (x < stride) UP |= (uint64_t)up_row[x] << (8u * k);
            }
        }
        if (act) {
          uint64_t NEW = 0;
          for (uint32_t k = 0; k < bpp; k++) {
            const uint32_t x = s * bpp + k;
            if (x < stride) {
              const uint32_t r = src[1 + x];
              const uint32_t a = (uint32_t)(A >> (8u * k)) & 255u, B = (uint32_t)(UP >> (8u * k)) & 255u,
                             c = (uint32_t)(C >> (8u * k)) & 255u;
              uint32_t v;
              switch (ft) {
                case 1:  v = r + a; break;
                case 2:  v = r + B; break;
                case 3:  v = r + ((a + B) >> 1); break;
                case 4:  v = r + paeth(a, B, c); break;
                default: v = r; break;
              }
              v &= 255u;
              dst[x] = (uint8_t)v;
              NEW |= (uint64_t)v << (8u * k);
            }
          }
          A = NEW; C = UP; CUR = NEW;
        }
      }
      __syncwarp();                                          // the group's last row is in memory for the next group
    }
    if (__any_sync(0xFFFFFFFFu, bad) && lane == 0) jobs[j].bad = 1;
  }
}

}  // namespace defl
