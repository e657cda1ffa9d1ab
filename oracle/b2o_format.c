/* oracle/b2o_format.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the blosc2 chunk layer: header,
 * block/stream geometry, split rule, per-stream framing, chunk-level fallbacks.
 * Follows /root/reference/thirdparty/c-blosc2/blosc/blosc2.c (b2c) and blosc/stune.c:
 *   header            b2c:654-766, 893-931          split rule     stune.c:185-210
 *   blocksize         stune.c:46-164                compress       b2c:1080-1332, 1941-2002, 2393-2573
 *   decompress        b2c:1568-1937, 2298-2391, 2922-2951
 * Serial (nthreads = 1) order only: blocks are laid out in index order. */
#include <stdlib.h>
#include <string.h>
#include "b2o.h"

enum { HDR_MIN = 16, HDR_EXT = 32, MIN_BUFFERSIZE = 32, MAX_STREAMS = 16, MAXBLOCKSIZE = 536866816 };
enum { F_SHUFFLE = 1, F_MEMCPYED = 2, F_BITSHUFFLE = 4, F_DELTA = 8, F_DONT_SPLIT = 0x10 };
enum { FMT_BLOSCLZ = 0, FMT_LZ4 = 1, FMT_ZLIB = 3, FMT_ZSTD = 4 };

static int32_t ld32(const uint8_t* p) { return (int32_t)(p[0] | (p[1] << 8) | (p[2] << 16) | ((uint32_t)p[3] << 24)); }
static void st32(uint8_t* p, int32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); }

/* b2c:3696-3712 + read_chunk_header checks b2c:672-711 */
int b2o_chunk_sizes(const void* chunk, int32_t* nbytes, int32_t* cbytes, int32_t* blocksize) {
  const uint8_t* s = (const uint8_t*)chunk;
  int32_t nb = ld32(s + 4), bs = ld32(s + 8), cb = ld32(s + 12);
  int rc = 0;
  if (s[0] > 5) rc = B2O_ERR_VERSION_SUPPORT;
  else if (cb < HDR_MIN) rc = B2O_ERR_INVALID_HEADER;
  else if (bs <= 0 || (nb > 0 && bs > nb)) rc = B2O_ERR_INVALID_HEADER;
  else if (bs > MAXBLOCKSIZE) rc = B2O_ERR_INVALID_HEADER;
  else if (s[3] == 0) rc = B2O_ERR_INVALID_HEADER;
  if (rc < 0) { nb = cb = bs = 0; } /* b2c:3700: header is zeroed on failure */
  if (nbytes) *nbytes = nb;
  if (cbytes) *cbytes = cb;
  if (blocksize) *blocksize = bs;
  return rc;
}

/* stune.c:185-210 with BLOSC_AUTO_SPLIT (compressed-image always passes it, wrapper.h:352) */
static int split_block(int codec, int clevel, int filter, int typesize, int32_t blocksize) {
  return ((codec == B2O_BLOSCLZ || codec == B2O_LZ4 || (codec == B2O_ZSTD && clevel <= 5)) &&
          filter == B2O_SHUFFLE && typesize <= MAX_STREAMS && (blocksize / typesize) >= MIN_BUFFERSIZE);
}

/* stune.c:46-164.  compressed-image always passes a user blocksize; the automatic table is restated too. */
static int32_t next_blocksize(int codec, int clevel, int filter, int typesize, int32_t nbytes, int32_t user) {
  int32_t bs = nbytes;
  if (nbytes < typesize) return 1;
  if (user) {
    bs = user;
  } else {
    int hcr = (codec == B2O_LZ4HC || codec == B2O_ZSTD || codec == 4 /*zlib*/);
    int splitmode = split_block(codec, clevel, filter, typesize, bs);
    if (nbytes >= 32 * 1024) {
      bs = 32 * 1024;
      if (hcr) bs *= 2;
      switch (clevel) {
        case 0: bs /= 4; break;
        case 1: bs /= 2; break;
        case 2: break;
        case 3: bs *= 2; break;
        case 4: case 5: bs *= 4; break;
        case 6: case 7: case 8: bs *= 8; break;
        case 9: bs *= 8; if (hcr) bs *= 2; break;
      }
    }
    if (clevel > 0 && splitmode) {
      switch (clevel) {
        case 1: case 2: case 3: bs = 32 * 1024; break;
        case 4: case 5: case 6: bs = 64 * 1024; break;
        case 7: bs = 128 * 1024; break;
        case 8: bs = 256 * 1024; break;
        default: bs = 512 * 1024; break;
      }
      bs *= typesize;
      if (bs > 4 * 1024 * 1024) bs = 4 * 1024 * 1024;
      if (bs < 32 * 1024) bs = 32 * 1024;
    }
  }
  if (bs > nbytes) bs = nbytes;
  if (bs > typesize) bs = bs / typesize * typesize;
  return bs;
}

/* b2c:1054-1076 */
static int is_run(const uint8_t* p, int32_t n) {
  for (int32_t i = 1; i < n; i++)
    if (p[i] != p[0]) return 0;
  return 1;
}

static void filter_forward(int filter, int ts, int32_t bsize, const uint8_t* src, uint8_t* dst) {
  if (filter == B2O_SHUFFLE) b2o_shuffle(ts, bsize, src, dst);
  else if (filter == B2O_BITSHUFFLE) b2o_bitshuffle(ts, bsize, src, dst);
  else memcpy(dst, src, (size_t)bsize);
}

/* b2c:1080-1332 blosc_c. Returns ctbytes, 0 = does not fit, <0 error. */
static int32_t compress_block(int codec, int clevel, int ts, int filter, int dont_split, int32_t bsize, int leftover,
                              int32_t ntbytes, int32_t destsize, const uint8_t* src, uint8_t* dest, uint8_t* tmp) {
  int32_t ctbytes = 0;
  filter_forward(filter, ts, bsize, src, tmp);
  int nstreams = (!dont_split && !leftover) ? ts : 1;
  int32_t neblock = bsize / nstreams;
  for (int j = 0; j < nstreams; j++) {
    dest += 4; ntbytes += 4; ctbytes += 4;
    const uint8_t* ip = tmp + (size_t)j * neblock;
    if (is_run(ip, neblock)) { /* b2c:1157-1200 */
      int32_t value = ip[0];
      if (ntbytes > destsize) return 0;
      st32(dest - 4, -value);
      if (value > 0) {
        ntbytes += 1; ctbytes += 1;
        if (ntbytes > destsize) return 0;
        dest[0] = 0x1;
        dest += 1;
      }
      continue;
    }
    int32_t maxout = neblock, cbytes;
    if (ntbytes + maxout > destsize) {
      maxout = destsize - ntbytes;
      if (maxout <= 0) return 0;
    }
    if (codec == B2O_BLOSCLZ) cbytes = b2o_blosclz_compress(clevel, ip, neblock, dest, maxout);
    else if (codec == B2O_LZ4) cbytes = b2o_lz4_compress(ip, neblock, dest, maxout);
    else return B2O_ERR_CODEC_SUPPORT; /* zstd / lz4hc encoders are not restated */
    if (cbytes > maxout) return B2O_ERR_WRITE_BUFFER;
    if (cbytes < 0) return B2O_ERR_DATA;
    if (cbytes == 0) cbytes = neblock;
    if (cbytes == neblock) { /* b2c:1314-1324 */
      if (ntbytes + neblock > destsize) return 0;
      memcpy(dest, ip, (size_t)neblock);
    }
    st32(dest - 4, cbytes);
    dest += cbytes; ntbytes += cbytes; ctbytes += cbytes;
  }
  return ctbytes;
}

static uint8_t codec_format(int codec) {
  switch (codec) {
    case B2O_BLOSCLZ: return FMT_BLOSCLZ;
    case B2O_LZ4: case B2O_LZ4HC: return FMT_LZ4;
    case B2O_ZSTD: return FMT_ZSTD;
    default: return 7;
  }
}

/* b2c:2577-2699 blosc2_compress_ctx (no dict, extended header) */
int b2o_compress(int codec, int clevel, int typesize, int filter, int32_t blocksize, const void* src_, int32_t nbytes,
                 void* dst_, int32_t destsize) {
  const uint8_t* src = (const uint8_t*)src_;
  uint8_t* dest = (uint8_t*)dst_;
  if (nbytes < 0 || nbytes > 0x7fffffff - HDR_EXT) return B2O_ERR_MAX_BUFSIZE;
  if (destsize < HDR_EXT) return B2O_ERR_MAX_BUFSIZE;
  if (clevel < 0 || clevel > 9) return B2O_ERR_CODEC_PARAM;
  if (typesize <= 0) return B2O_ERR_INVALID_PARAM;
  if (typesize > 255) typesize = 1; /* b2c:2285-2288 */
  if (codec != B2O_BLOSCLZ && codec != B2O_LZ4) return B2O_ERR_CODEC_SUPPORT;

  blocksize = next_blocksize(codec, clevel, filter, typesize, nbytes, blocksize);
  int32_t nblocks = nbytes / blocksize, leftover = nbytes % blocksize;
  if (leftover > 0) nblocks++;

  /* write_compression_header, b2c:2393-2468 */
  uint8_t flags = 0;
  if (clevel == 0 || nbytes < MIN_BUFFERSIZE) flags |= F_MEMCPYED;
  int memcpyed = flags & F_MEMCPYED;
  flags |= F_SHUFFLE | F_BITSHUFFLE; /* extended-header marker */
  int dont_split = 0;
  if (!memcpyed) {
    /* shuffle bit 0 / bitshuffle bit 2 are already set by the marker; only split + format are added */
    dont_split = !split_block(codec, clevel, filter, typesize, blocksize);
    flags |= (uint8_t)(dont_split << 4);
    flags |= (uint8_t)(codec_format(codec) << 5);
  }
  memset(dest, 0, HDR_EXT);
  dest[0] = 5;                              /* BLOSC2_VERSION_FORMAT_STABLE */
  dest[1] = 1;                              /* versionlz: blosclz / lz4 / zstd all report 1 */
  dest[2] = flags;
  dest[3] = (uint8_t)typesize;
  st32(dest + 4, nbytes);
  st32(dest + 8, blocksize);
  dest[16 + 5] = (uint8_t)filter;           /* filters[5] */
  dest[22] = (uint8_t)codec;                /* udcompcode */

  int32_t ntbytes = 0;
  if (!memcpyed) {
    uint8_t* tmp = (uint8_t*)malloc((size_t)blocksize + 16);
    ntbytes = HDR_EXT + 4 * nblocks;
    for (int32_t j = 0; j < nblocks; j++) { /* serial_blosc b2c:1941-2002 */
      int32_t bsize = blocksize, lo = 0;
      if (j == nblocks - 1 && leftover > 0) { bsize = leftover; lo = 1; }
      /* the reference stores bstarts[j] unchecked (b2c:1958); guard the store, the chunk is doomed anyway */
      if (HDR_EXT + 4 * (j + 1) <= destsize) st32(dest + HDR_EXT + 4 * j, ntbytes);
      int32_t cb = compress_block(codec, clevel, typesize, filter, dont_split, bsize, lo, ntbytes, destsize,
                                  src + (size_t)j * blocksize, dest + ntbytes, tmp);
      if (cb == 0) { ntbytes = 0; break; }
      if (cb < 0) { free(tmp); return cb; }
      ntbytes += cb;
    }
    free(tmp);
    if (ntbytes == 0) { flags |= F_MEMCPYED; memcpyed = 1; } /* b2c:2484-2490 */
  }
  if (memcpyed) {
    if (nbytes + HDR_EXT > destsize) {
      ntbytes = 0;
    } else {
      memcpy(dest + HDR_EXT, src, (size_t)nbytes);
      ntbytes = HDR_EXT + nbytes;
      dest[2] = flags;
    }
  } else {
    int32_t nstreams = nblocks; /* b2c:2492-2502 */
    if (!dont_split) nstreams = leftover ? (nblocks - 1) * typesize + 1 : nblocks * typesize;
    if (ntbytes == HDR_EXT + 4 * nblocks + 4 * nstreams) { /* all zero runs, b2c:2520-2529 */
      dest[31] |= 0x1 << 4;
      ntbytes = HDR_EXT;
    }
  }
  st32(dest + 12, ntbytes);
  return ntbytes;
}

/* one stream, b2c:1783-1923 */
static int32_t decode_stream(int fmt, const uint8_t** psrc, int32_t* psrcsize, uint8_t* out, int32_t neblock) {
  const uint8_t* src = *psrc;
  int32_t srcsize = *psrcsize, cbytes, nbytes;
  if (srcsize < 4) return B2O_ERR_READ_BUFFER;
  srcsize -= 4;
  cbytes = ld32(src);
  if (cbytes > 0) {
    if (srcsize < cbytes) return B2O_ERR_READ_BUFFER;
    srcsize -= cbytes;
  }
  src += 4;
  if (cbytes == 0) {
    memset(out, 0, (size_t)neblock);
    nbytes = neblock;
  } else if (cbytes < 0) {
    if (srcsize < 1) return B2O_ERR_READ_BUFFER;
    srcsize -= 1;
    uint8_t token = *src++;
    if (!(token & 1)) return B2O_ERR_RUN_LENGTH;
    if (cbytes < -255) return B2O_ERR_RUN_LENGTH;
    memset(out, (uint8_t)(-cbytes), (size_t)neblock);
    nbytes = neblock;
    cbytes = 0;
  } else if (cbytes == neblock) {
    memcpy(out, src, (size_t)neblock);
    nbytes = neblock;
  } else {
    if (fmt == FMT_BLOSCLZ) nbytes = b2o_blosclz_decompress(src, cbytes, out, neblock);
    else if (fmt == FMT_LZ4) nbytes = b2o_lz4_decompress(src, cbytes, out, neblock);
    else if (fmt == FMT_ZSTD) {
      long r = b2o_zstd_decompress(src, cbytes, out, neblock);
      nbytes = r < 0 ? 0 : (int32_t)r; /* b2c:545-548: zstd errors become 0 bytes */
    } else return B2O_ERR_CODEC_SUPPORT;
    if (nbytes != neblock) return B2O_ERR_DATA;
  }
  *psrc = src + cbytes;
  *psrcsize = srcsize;
  return nbytes;
}

/* b2c:2922-2951 blosc_run_decompression_with_context + b2c:1568-1937 blosc_d, serial */
int b2o_decompress(const void* src_, int32_t srcsize, void* dst_, int32_t destsize) {
  const uint8_t* src = (const uint8_t*)src_;
  uint8_t* dest = (uint8_t*)dst_;
  if (srcsize < HDR_MIN) return B2O_ERR_READ_BUFFER;
  int version = src[0];
  uint8_t flags = src[2];
  int ts = src[3];
  int32_t nbytes = ld32(src + 4), blocksize = ld32(src + 8), cbytes = ld32(src + 12);
  if (version > 5) return B2O_ERR_VERSION_SUPPORT;
  if (cbytes < HDR_MIN) return B2O_ERR_INVALID_HEADER;
  if (blocksize <= 0 || (nbytes > 0 && blocksize > nbytes)) return B2O_ERR_INVALID_HEADER;
  if (blocksize > MAXBLOCKSIZE) return B2O_ERR_INVALID_HEADER;
  if (ts == 0) return B2O_ERR_INVALID_HEADER;

  int ext = (flags & F_SHUFFLE) && (flags & F_BITSHUFFLE);
  int hdr = ext ? HDR_EXT : HDR_MIN;
  int filter = B2O_NOFILTER, special = 0;
  uint8_t b2flags = 0;
  if (ext) {
    if (cbytes < HDR_EXT) return B2O_ERR_INVALID_HEADER;
    if (srcsize < HDR_EXT) return B2O_ERR_READ_BUFFER;
    b2flags = src[31];
    special = (b2flags >> 4) & 7;
    if (special == 3) { /* repeated value: typesize comes from cbytes, b2c:730-748 */
      int32_t vts = cbytes - HDR_EXT;
      if (vts <= 0 || vts > nbytes || nbytes % vts) return B2O_ERR_INVALID_HEADER;
    } else if (special != 0) {
      if (nbytes % ts) return B2O_ERR_INVALID_HEADER;
    }
    for (int i = 0; i < 6; i++) {
      uint8_t f = src[16 + i];
      if (version == 4 && i == 5) f = 0; /* alpha series, b2c:758-761 */
      if (f == 0) continue;
      if ((f == B2O_SHUFFLE || f == B2O_BITSHUFFLE) && filter == B2O_NOFILTER) filter = f;
      else return B2O_ERR_FILTER_PIPELINE; /* delta / trunc-prec / user filters: outside the hot path */
    }
  } else { /* Blosc1 header: filters come from the flags, b2c flags_to_filters */
    if (flags & F_DELTA) return B2O_ERR_FILTER_PIPELINE;
    if (flags & F_SHUFFLE) filter = B2O_SHUFFLE;
    if (flags & F_BITSHUFFLE) filter = B2O_BITSHUFFLE;
  }
  if (nbytes > destsize) return B2O_ERR_WRITE_BUFFER;
  if (cbytes > srcsize) return B2O_ERR_INVALID_HEADER; /* b2c:850 */
  if (special > 4) return B2O_ERR_DATA;
  int memcpyed = flags & F_MEMCPYED;
  if (memcpyed && cbytes != nbytes + hdr) return B2O_ERR_DATA;
  if (nbytes == 0 && cbytes == hdr && !special) return 0;
  if (ext && (b2flags & 0x08) && !special) return B2O_ERR_INVALID_PARAM; /* lazy chunk without a frame */
  if (ext && (b2flags & 0x01)) return B2O_ERR_CODEC_SUPPORT;             /* dictionaries: not on this path */

  int32_t nblocks = nbytes / blocksize, leftover = nbytes % blocksize;
  if (leftover > 0) nblocks++;

  if (special) { /* b2c:1694-1718 */
    if (special == 1) memset(dest, 0, (size_t)nbytes);
    else if (special == 2) { /* NaN */
      if (ts == 4) { for (int32_t i = 0; i < nbytes / 4; i++) st32(dest + 4 * i, 0x7fc00000); }
      else if (ts == 8) { for (int32_t i = 0; i < nbytes / 8; i++) { st32(dest + 8 * i, 0); st32(dest + 8 * i + 4, 0x7ff80000); } }
      else return B2O_ERR_DATA;
    } else if (special == 3) {
      int32_t vts = cbytes - HDR_EXT;
      for (int32_t i = 0; i < nbytes; i++) dest[i] = src[HDR_EXT + i % vts];
    }
    return nbytes;
  }
  if (memcpyed) {
    memcpy(dest, src + hdr, (size_t)nbytes);
    return nbytes;
  }
  int32_t bstarts_end = hdr + 4 * nblocks;
  if (srcsize < bstarts_end) return B2O_ERR_READ_BUFFER;
  int fmt = flags >> 5;
  int dont_split = (flags & F_DONT_SPLIT) >> 4;
  uint8_t* tmp = (uint8_t*)malloc((size_t)blocksize + 16);
  int32_t ntbytes = 0;
  for (int32_t j = 0; j < nblocks; j++) {
    int32_t bsize = blocksize, lo = 0;
    if (j == nblocks - 1 && leftover > 0) { bsize = leftover; lo = 1; }
    int32_t off = ld32(src + hdr + 4 * j);
    if (off <= 0 || off >= srcsize) { free(tmp); return B2O_ERR_DATA; }
    const uint8_t* p = src + off;
    int32_t avail = srcsize - off;
    int nstreams = (!dont_split && !lo) ? ts : 1;
    int32_t neblock = bsize / nstreams;
    if (neblock == 0) { free(tmp); return B2O_ERR_WRITE_BUFFER; }
    uint8_t* out = (filter == B2O_NOFILTER) ? dest + (size_t)j * blocksize : tmp;
    for (int s = 0; s < nstreams; s++) {
      int32_t r = decode_stream(fmt, &p, &avail, out + (size_t)s * neblock, neblock);
      if (r < 0) { free(tmp); return r; }
    }
    if (filter == B2O_SHUFFLE) b2o_unshuffle(ts, bsize, tmp, dest + (size_t)j * blocksize);
    else if (filter == B2O_BITSHUFFLE) b2o_bitunshuffle(ts, bsize, tmp, dest + (size_t)j * blocksize, version);
    ntbytes += bsize;
  }
  free(tmp);
  return ntbytes;
}
