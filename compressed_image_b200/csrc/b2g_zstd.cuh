// b2g_zstd.cuh -- warp-cooperative zstd frame + block DEcoder (row C5 of SURVEY.md 8a).
//
// Behavioural reference: ZSTD_decompressDCtx of the vendored zstd 1.5.7 as c-blosc2 calls it per stream
// (blosc/blosc2.c:524-549); format per RFC 8878 (frame header zstd_decompress.c:447-560, block header
// zstd_decompress_block.c:63-77, literals :134-340, sequence headers :695-775, FSE table construction :485-640,
// sequence decoding :1229-1346, Huffman weights common/entropy_common.c:234-340).  No dictionaries, checksum not
// verified (the reference never enables either).
//
// One warp owns one stream.  Entropy tables live in a per-warp shared-memory work area; table construction is
// serial on lane 0; the four Huffman literal streams are decoded by four lanes; sequences are decoded redundantly by
// all lanes (uniform control flow, broadcast loads) and executed warp-wide.  Literals are staged right-aligned at the
// end of the output window: output can never overtake unread literals in a valid frame, so no second buffer is
// needed (see zstd_exec bounds).
#pragma once
#include "b2g_lz4.cuh"

namespace b2g {

constexpr int kZMaxLL = 35, kZMaxML = 52, kZMaxOff = 31, kZLLLog = 9, kZMLLog = 9, kZOFLog = 8, kZHufLogMax = 12;

struct ZWork {
  uint32_t ll[512], ml[512], of[256];  // FSE decode tables: sym | nbits << 8 | base << 16
  uint16_t huf[1 << kZHufLogMax];      // Huffman X1 table: sym | nbits << 8
  uint8_t symbols[512];
  uint8_t weights[256];
  int16_t norm[64];
  uint16_t next[64];
  uint32_t rank[16], start[16];
  uint32_t rep[3];
  int32_t ll_log, ml_log, of_log, huf_log;
  int32_t ll_valid, ml_valid, of_valid, huf_valid;
};

__constant__ int16_t kZLLDef[36] = {4, 3, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 3, 2, 1, 1, 1, 1, 1, -1, -1, -1, -1};
__constant__ int16_t kZMLDef[53] = {1, 4, 3, 2, 2, 2, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1};
__constant__ int16_t kZOFDef[29] = {1, 1, 1, 1, 1, 1, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1};
__constant__ uint32_t kZLLBase[36] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 28, 32, 40, 48, 64, 0x80, 0x100, 0x200, 0x400, 0x800, 0x1000, 0x2000, 0x4000, 0x8000, 0x10000};
__constant__ uint8_t kZLLBits[36] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};
__constant__ uint32_t kZMLBase[53] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35, 37, 39, 41, 43, 47, 51, 59, 67, 83, 99, 0x83, 0x103, 0x203, 0x403, 0x803, 0x1003, 0x2003, 0x4003, 0x8003, 0x10003};
__constant__ uint8_t kZMLBits[53] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 4, 5, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};

__device__ __forceinline__ int z_hbit(uint32_t v) { return 31 - __clz((int)v); }  // v != 0

// 64 bits of the little-endian stream starting at byte `byte`; bytes outside [0, n) read as zero.
__device__ __forceinline__ uint64_t z_ld64(const uint8_t* src, long n, long byte) {
  if (byte >= 0 && byte + 8 <= n) {
    return (uint64_t)ld32_any(src + byte) | ((uint64_t)ld32_any(src + byte + 4) << 32);
  }
  uint64_t v = 0;
  for (int i = 0; i < 8; i++) {
    long b = byte + i;
    if (b >= 0 && b < n) v |= (uint64_t)src[b] << (8 * i);
  }
  return v;
}

// nb (<= 32) bits of the stream seen as one little-endian integer, starting at bit `lo` (may be negative: bits
// below 0 are zero).  This is what a backward bitstream read of nb bits ending at bit lo+nb returns.
__device__ __forceinline__ uint32_t z_bits(const uint8_t* src, long n, long lo, int nb) {
  if (nb <= 0) return 0;
  if (lo < 0) {
    int have = nb + (int)lo;
    if (have <= 0) return 0;
    uint64_t v = z_ld64(src, n, 0);
    return (uint32_t)((v & ((1ull << have) - 1)) << (-lo));
  }
  uint64_t v = z_ld64(src, n, lo >> 3);
  return (uint32_t)((v >> (lo & 7)) & ((1ull << nb) - 1));
}

// Backward bit reader (common/bitstream.h:248-400).  `bits` counts the unread bits of the stream (negative = overrun,
// exactly as before); the next `avail` of them sit left-aligned in the 64-bit container `cont`, so that a field costs a
// shift and a mask, and memory is touched once per ~57 bits instead of once per field (the Huffman and FSE loops are
// chains of dependent reads: the four loads + three funnel shifts of an unaligned 64-bit fetch per field were most of
// their time).
struct ZBack {
  const uint8_t* p;
  long n;
  long bits;
  uint64_t cont;
  int avail;
};
__device__ __forceinline__ void zb_refill(ZBack& b) {
  const long end = (b.bits + 7) >> 3;                       // first byte boundary at or above the read position
  const int extra = (int)(end * 8 - b.bits);                // 0..7 bits of `end`'s byte are already consumed
  const uint64_t v = z_ld64(b.p, b.n, end - 8);             // bytes below 0 read as zero
  b.cont = v << extra;
  b.avail = 64 - extra;
}
__device__ __forceinline__ bool zb_init(ZBack& b, const uint8_t* p, long n) {
  if (n < 1) return false;
  uint32_t last = p[n - 1];
  if (last == 0) return false;
  b.p = p; b.n = n;
  b.bits = n * 8 - (8 - z_hbit(last));
  zb_refill(b);
  return true;
}
// the next nb (<= 32) bits without consuming them
__device__ __forceinline__ uint32_t zb_peek(ZBack& b, int nb) {
  if (b.avail < nb) zb_refill(b);
  return nb > 0 ? (uint32_t)(b.cont >> (64 - nb)) : 0u;
}
__device__ __forceinline__ void zb_skip(ZBack& b, int nb) {   // after zb_peek(b, >= nb)
  b.cont <<= nb;
  b.avail -= nb;
  b.bits -= nb;
}
__device__ __forceinline__ uint32_t zb_read(ZBack& b, int nb) {
  const uint32_t v = zb_peek(b, nb);
  zb_skip(b, nb);
  return v;
}

// ---- serial (lane 0) table construction ---------------------------------------------------------------------------
// FSE_readNCount (common/entropy_common.c:42-230).  Returns bytes consumed or -1.
__device__ __noinline__ long z_read_ncount(const uint8_t* src, long n, int16_t* norm, int* maxsym, int* tablelog,
                                           int maxlog) {
  if (n < 1) return -1;
  long bitpos = 0;
  int al = (int)z_bits(src, n, bitpos, 4) + 5;
  bitpos += 4;
  if (al > maxlog) return -1;
  int remaining = 1 << al, sym = 0;
  const int limit = *maxsym + 1;
  while (remaining > 0 && sym < limit) {
    int bits = z_hbit((uint32_t)remaining + 1) + 1;
    uint32_t val = z_bits(src, n, bitpos, bits);
    uint32_t lower = (1u << (bits - 1)) - 1;
    uint32_t thresh = (1u << bits) - 1 - ((uint32_t)remaining + 1);
    if ((val & lower) < thresh) { bitpos += bits - 1; val &= lower; }
    else { bitpos += bits; if (val > lower) val -= thresh; }
    int proba = (int)val - 1;
    remaining -= proba < 0 ? -proba : proba;
    norm[sym++] = (int16_t)proba;
    if (proba == 0) {
      for (;;) {
        uint32_t rep = z_bits(src, n, bitpos, 2);
        bitpos += 2;
        for (uint32_t i = 0; i < rep && sym < limit; i++) norm[sym++] = 0;
        if (rep != 3) break;
        if ((bitpos >> 3) > n) return -1;
      }
    }
    if ((bitpos >> 3) > n) return -1;
  }
  if (remaining != 0) return -1;
  if (sym > limit) return -1;
  *maxsym = sym - 1;
  *tablelog = al;
  long used = (bitpos + 7) >> 3;
  if (used > n) return -1;
  return used;
}

// ZSTD_buildFSETable (zstd_decompress_block.c:485-640) / RFC 8878 4.1.1
__device__ __noinline__ int z_build_fse(ZWork* z, uint32_t* tab, const int16_t* norm, int maxsym, int log) {
  const int size = 1 << log;
  int high = size;
  if (log > 9 || maxsym > 63) return -1;
  for (int s = 0; s <= maxsym; s++) {
    if (norm[s] == -1) { z->symbols[--high] = (uint8_t)s; z->next[s] = 1; }
    else z->next[s] = (uint16_t)norm[s];
  }
  const int step = (size >> 1) + (size >> 3) + 3, mask = size - 1;
  int pos = 0;
  for (int s = 0; s <= maxsym; s++) {
    for (int i = 0; i < norm[s]; i++) {
      z->symbols[pos] = (uint8_t)s;
      do { pos = (pos + step) & mask; } while (pos >= high);
    }
  }
  if (pos != 0) return -1;
  for (int i = 0; i < size; i++) {
    uint32_t s = z->symbols[i];
    uint32_t ns = z->next[s]++;
    int nb = log - z_hbit(ns);
    uint32_t base = ((ns << nb) - (uint32_t)size) & 0xffffu;
    tab[i] = s | ((uint32_t)nb << 8) | (base << 16);
  }
  return 0;
}

// ZSTD_buildSeqTable (zstd_decompress_block.c:647-693).  Lane 0 only.  Returns bytes consumed or -1.
__device__ __noinline__ long z_build_seq_table(ZWork* z, uint32_t* tab, int32_t* plog, int32_t* pvalid, int mode,
                                               int maxsym, int maxlog, const uint8_t* src, long n,
                                               const int16_t* defnorm, int defmax, int deflog) {
  switch (mode) {
    case 0:
      for (int i = 0; i <= defmax; i++) z->norm[i] = defnorm[i];
      if (z_build_fse(z, tab, z->norm, defmax, deflog)) return -1;
      *plog = deflog; *pvalid = 1;
      return 0;
    case 1:
      if (n < 1 || src[0] > maxsym) return -1;
      tab[0] = src[0];
      *plog = 0; *pvalid = 1;
      return 1;
    case 2: {
      int ms = maxsym, log = 0;
      long used = z_read_ncount(src, n, z->norm, &ms, &log, maxlog);
      if (used < 0) return -1;
      if (z_build_fse(z, tab, z->norm, ms, log)) return -1;
      *plog = log; *pvalid = 1;
      return used;
    }
    default:
      return *pvalid ? 0 : -1;
  }
}

// HUF_readStats + X1 table fill (common/entropy_common.c:234-340, decompress/huf_decompress.c:385-500).  Lane 0.
__device__ __noinline__ long z_read_huf_table(ZWork* z, const uint8_t* src, long n) {
  uint8_t* w = z->weights;
  int nw = 0;
  long used;
  if (n < 1) return -1;
  int hb = src[0];
  if (hb >= 128) {
    nw = hb - 127;
    used = 1 + (nw + 1) / 2;
    if (used > n) return -1;
    for (int i = 0; i < nw; i += 2) {
      uint8_t v = src[1 + i / 2];
      w[i] = v >> 4;
      w[i + 1] = v & 15;
    }
  } else {
    used = 1 + hb;
    if (used > n || hb < 1) return -1;
    int ms = 12, log = 0;
    long hs = z_read_ncount(src + 1, hb, z->norm, &ms, &log, 6);
    if (hs < 0) return -1;
    uint32_t* ft = (uint32_t*)z->huf;  // the Huffman table is about to be rebuilt: its storage is free scratch
    if (z_build_fse(z, ft, z->norm, ms, log)) return -1;
    ZBack b;
    if (!zb_init(b, src + 1 + hs, hb - hs)) return -1;
    uint32_t s1 = zb_read(b, log), s2 = zb_read(b, log);
    if (b.bits < 0) return -1;
    for (;;) {
      if (nw > 253) return -1;
      uint32_t e1 = ft[s1];
      w[nw++] = (uint8_t)e1;
      s1 = (e1 >> 16) + zb_read(b, (int)((e1 >> 8) & 0xff));
      if (b.bits < 0) { w[nw++] = (uint8_t)ft[s2]; break; }
      if (nw > 253) return -1;
      uint32_t e2 = ft[s2];
      w[nw++] = (uint8_t)e2;
      s2 = (e2 >> 16) + zb_read(b, (int)((e2 >> 8) & 0xff));
      if (b.bits < 0) { w[nw++] = (uint8_t)ft[s1]; break; }
    }
  }
  uint32_t total = 0;
  for (int i = 0; i < 16; i++) z->rank[i] = 0;
  for (int i = 0; i < nw; i++) {
    if (w[i] > kZHufLogMax) return -1;
    z->rank[w[i]]++;
    total += (1u << w[i]) >> 1;
  }
  if (total == 0) return -1;
  int log = z_hbit(total) + 1;
  if (log > kZHufLogMax) return -1;
  uint32_t rest = (1u << log) - total;
  if (rest & (rest - 1)) return -1;
  int last = z_hbit(rest) + 1;
  w[nw++] = (uint8_t)last;
  z->rank[last]++;
  if (z->rank[1] < 2 || (z->rank[1] & 1)) return -1;
  uint32_t nxt = 0;
  for (int r = 1; r <= log; r++) { z->start[r] = nxt; nxt += z->rank[r] << (r - 1); }
  for (int s = 0; s < nw; s++) {
    int wt = w[s];
    if (!wt) continue;
    uint32_t len = (1u << wt) >> 1;
    uint16_t e = (uint16_t)(s | ((log + 1 - wt) << 8));
    uint32_t st = z->start[wt];
    for (uint32_t k = 0; k < len; k++) z->huf[st + k] = e;
    z->start[wt] = st + len;
  }
  z->huf_log = log;
  z->huf_valid = 1;
  return used;
}

// One Huffman stream, decoded by ONE lane: exactly `count` symbols, the stream must end exactly.
__device__ __forceinline__ bool z_huf_stream(const ZWork* z, const uint8_t* src, long n, uint8_t* out, long count) {
  ZBack b;
  if (!zb_init(b, src, n)) return false;
  const int log = z->huf_log;
  long i = 0;
  // four symbols per refill check: a symbol takes at most kZHufLogMax = 12 bits, a refilled container holds >= 57
  for (; i + 4 <= count; i += 4) {
    if (b.avail < 4 * kZHufLogMax) zb_refill(b);
    uint64_t c = b.cont;
    int used = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t e = z->huf[(uint32_t)(c >> (64 - log))];
      const int nb = (int)(e >> 8);
      out[i + k] = (uint8_t)e;
      c <<= nb;
      used += nb;
    }
    b.cont = c;
    b.avail -= used;
    b.bits -= used;
  }
  for (; i < count; i++) {
    const uint32_t e = z->huf[zb_peek(b, log)];
    out[i] = (uint8_t)e;
    zb_skip(b, (int)(e >> 8));
  }
  return b.bits == 0;
}

// ZSTD_decodeLiteralsBlock (zstd_decompress_block.c:134-340).  Literals are written to win[cap - litsize .. cap).
// Returns bytes consumed or -1; sets *plitsize.
__device__ __forceinline__ long z_decode_literals(ZWork* z, const uint8_t* src, long n, uint8_t* win, long cap,
                                                  long floor_op, long* plitsize, long blockmax, int lane) {
  if (n < 2) return -1;
  const uint32_t b0 = src[0];
  const int type = (int)(b0 & 3), sf = (int)((b0 >> 2) & 3);
  long lh, regen, comp = 0;
  if (type < 2) {
    if (sf == 0 || sf == 2) { lh = 1; regen = b0 >> 3; }
    else if (sf == 1) { lh = 2; regen = (b0 >> 4) + ((long)src[1] << 4); }
    else { if (n < 3) return -1; lh = 3; regen = (b0 >> 4) + ((long)src[1] << 4) + ((long)src[2] << 12); }
    if (regen > blockmax || regen > cap - floor_op) return -1;
    uint8_t* lit = win + (cap - regen);
    if (type == 0) {
      if (lh + regen > n) return -1;
      copy_bytes(lit, src + lh, (int)regen, lane, 32);
      *plitsize = regen;
      return lh + regen;
    }
    if (lh + 1 > n) return -1;
    fill_bytes(lit, src[lh], (int)regen, lane, 32);
    *plitsize = regen;
    return lh + 1;
  }
  if (n < 5) return -1;
  int streams = 4;
  const uint32_t v = b0 | ((uint32_t)src[1] << 8) | ((uint32_t)src[2] << 16) | ((uint32_t)src[3] << 24);
  if (sf < 2) { lh = 3; streams = sf ? 4 : 1; regen = (v >> 4) & 0x3FF; comp = (v >> 14) & 0x3FF; }
  else if (sf == 2) { lh = 4; regen = (v >> 4) & 0x3FFF; comp = v >> 18; }
  else { lh = 5; regen = (v >> 4) & 0x3FFFF; comp = (v >> 22) + ((long)src[4] << 10); }
  if (regen > blockmax || regen > cap - floor_op) return -1;
  if (streams == 4 && regen < 6) return -1;
  if (lh + comp > n) return -1;
  const uint8_t* p = src + lh;
  long left = comp;
  if (type == 2) {
    long hs = 0;
    if (lane == 0) hs = z_read_huf_table(z, p, left);
    hs = __shfl_sync(kFull, hs, 0);
    if (hs < 0) return -1;
    p += hs;
    left -= hs;
  } else {
    __syncwarp();
    if (!z->huf_valid) return -1;
  }
  __syncwarp();
  uint8_t* lit = win + (cap - regen);
  bool ok = true;
  if (streams == 1) {
    if (lane == 0) ok = z_huf_stream(z, p, left, lit, regen);
  } else {
    if (left < 10) return -1;
    long s1 = p[0] | (p[1] << 8), s2 = p[2] | (p[3] << 8), s3 = p[4] | (p[5] << 8);
    long s4 = left - 6 - s1 - s2 - s3;
    if (s4 < 1 || s1 + s2 + s3 + 6 > left) return -1;
    long seg = (regen + 3) / 4, lastseg = regen - 3 * seg;
    if (lastseg < 0) return -1;
    p += 6;
    if (lane == 0) ok = z_huf_stream(z, p, s1, lit, seg);
    else if (lane == 1) ok = z_huf_stream(z, p + s1, s2, lit + seg, seg);
    else if (lane == 2) ok = z_huf_stream(z, p + s1 + s2, s3, lit + 2 * seg, seg);
    else if (lane == 3) ok = z_huf_stream(z, p + s1 + s2 + s3, s4, lit + 3 * seg, lastseg);
  }
  if (!__all_sync(kFull, ok)) return -1;
  *plitsize = regen;
  return lh + comp;
}

// Forward-overlapping literal move inside the window: dst <= src.
__device__ __forceinline__ void z_move_literals(uint8_t* dst, const uint8_t* src, int len, int lane) {
  const long delta = src - dst;
  if (delta == 0 || len <= 0) return;
  if (delta >= len) { copy_bytes(dst, src, len, lane, 32); return; }
  // overlapping: every 32-byte step reads before it writes; a step never overwrites bytes a later step reads
  for (int i = 0; i < len; i += 32) {
    uint8_t v = 0;
    if (i + lane < len) v = src[i + lane];
    __syncwarp();
    if (i + lane < len) dst[i + lane] = v;
  }
}

// One compressed block (ZSTD_decompressBlock_internal, zstd_decompress_block.c:2066-2200).  `win` = frame output
// start, `op` = bytes already produced, cap = capacity.  Returns the new op or -1.
__device__ __forceinline__ long z_decode_block(ZWork* z, const uint8_t* src, long n, uint8_t* win, long op, long cap,
                                               long blockmax, int lane, bool batch) {
  long litsize = 0;
  long used = z_decode_literals(z, src, n, win, cap, op, &litsize, blockmax, lane);
  if (used < 0) return -1;
  const uint8_t* ip = src + used;
  const uint8_t* iend = src + n;
  if (ip >= iend) return -1;
  long nbseq = *ip++;
  if (nbseq > 0x7F) {
    if (nbseq == 0xFF) { if (ip + 2 > iend) return -1; nbseq = (long)(ip[0] | (ip[1] << 8)) + 0x7F00; ip += 2; }
    else { if (ip >= iend) return -1; nbseq = ((nbseq - 0x80) << 8) + *ip++; }
  }
  const uint8_t* lit = win + (cap - litsize);
  long litpos = 0;
  __syncwarp();  // staged literals visible to every lane
  if (nbseq == 0) {
    if (ip != iend) return -1;
  } else {
    if (ip + 1 > iend) return -1;
    const uint32_t modes = *ip++;
    if (modes & 3) return -1;
    long r = 0;
    if (lane == 0) {
      long a = z_build_seq_table(z, z->ll, &z->ll_log, &z->ll_valid, (int)(modes >> 6), kZMaxLL, kZLLLog, ip,
                                 iend - ip, kZLLDef, 35, 6);
      r = a;
      if (a >= 0) {
        long b = z_build_seq_table(z, z->of, &z->of_log, &z->of_valid, (int)((modes >> 4) & 3), kZMaxOff, kZOFLog,
                                   ip + a, iend - ip - a, kZOFDef, 28, 5);
        r = b < 0 ? -1 : a + b;
        if (b >= 0) {
          long c = z_build_seq_table(z, z->ml, &z->ml_log, &z->ml_valid, (int)((modes >> 2) & 3), kZMaxML, kZMLLog,
                                     ip + a + b, iend - ip - a - b, kZMLDef, 52, 6);
          r = c < 0 ? -1 : a + b + c;
        }
      }
    }
    r = __shfl_sync(kFull, r, 0);
    if (r < 0) return -1;
    ip += r;
    __syncwarp();
    ZBack b;
    if (!zb_init(b, ip, iend - ip)) return -1;
    const int lllog = z->ll_log, oflog = z->of_log, mllog = z->ml_log;
    uint32_t sl = zb_read(b, lllog), so = zb_read(b, oflog), sm = zb_read(b, mllog);
    if (b.bits < 0) return -1;
    uint32_t rep0 = z->rep[0], rep1 = z->rep[1], rep2 = z->rep[2];
    // Sequences are DECODED one after the other (the three FSE states and the repeat offsets form one chain, every lane
    // computes it: uniform control flow, broadcast table reads) but EXECUTED up to 32 at a time: lane k keeps sequence k
    // of the pending ones, prefix sums give the literal sources and the output positions, and the copies run lane-parallel
    // (lz_batch_execute, shared with the LZ4 decoder).  `batch` is off for windows without slack behind `cap` (blocks
    // decoded straight into the caller's buffer): the lane copies may read a few bytes past a literal run.
    const MemGeneric W(win);
    const long litbase = cap - litsize;       // literal k of the block sits at win[litbase + k]
    long dec = 0;                             // sequences decoded so far
    int npend = 0;                            // decoded, not yet executed: lanes 0 .. npend - 1
    const int fill = (batch && nbseq >= 64) ? 32 : 1;   // few sequences are few long matches: one by one, warp-wide
    int my_ll = 0, my_ml = 0;
    uint32_t my_off = 0;
    while (dec < nbseq || npend > 0) {
      // ---- decode until 32 are pending ------------------------------------------------------------------------------------
      while (npend < fill && dec < nbseq) {
        const uint32_t el = z->ll[sl], eo = z->of[so], em = z->ml[sm];
        const uint32_t llc = el & 0xff, ofc = eo & 0xff, mlc = em & 0xff;
        if (llc > kZMaxLL || mlc > kZMaxML || ofc > kZMaxOff) return -1;
        const uint32_t ll_base = kZLLBase[llc];
        // The fields of a sequence in stream order: offset bits (ofc of them, also for the repeat codes 0 and 1), match
        // length bits, literal length bits, then the three state updates.  Their widths are all known from the table
        // entries, so the fields are cut out of ONE 64-bit look at the stream side by side instead of one after the
        // other through the reader (seven dependent container updates per sequence were most of this loop).
        const int more = dec + 1 < nbseq;
        const int n_of = (int)ofc, n_ml = kZMLBits[mlc], n_ll = kZLLBits[llc];
        const int n_sl = more ? (int)((el >> 8) & 0xff) : 0, n_sm = more ? (int)((em >> 8) & 0xff) : 0,
                  n_so = more ? (int)((eo >> 8) & 0xff) : 0;
        const int c1 = n_of, c2 = c1 + n_ml, c3 = c2 + n_ll, c4 = c3 + n_sl, c5 = c4 + n_sm, c6 = c5 + n_so;
        uint32_t f_of, f_ml, f_ll, f_sl, f_sm, f_so;
        if (b.avail < 57) zb_refill(b);
        if (c6 <= b.avail) {
          const uint64_t c = b.cont;
          f_of = n_of ? (uint32_t)(c >> (64 - n_of)) : 0u;
          f_ml = n_ml ? (uint32_t)((c << c1) >> (64 - n_ml)) : 0u;
          f_ll = n_ll ? (uint32_t)((c << c2) >> (64 - n_ll)) : 0u;
          f_sl = n_sl ? (uint32_t)((c << c3) >> (64 - n_sl)) : 0u;
          f_sm = n_sm ? (uint32_t)((c << c4) >> (64 - n_sm)) : 0u;
          f_so = n_so ? (uint32_t)((c << c5) >> (64 - n_so)) : 0u;
          b.cont = c6 < 64 ? c << c6 : 0ull;
          b.avail -= c6;
          b.bits -= c6;
        } else {                                  // more than a container's worth (long offsets + long lengths): in turn
          f_of = zb_read(b, n_of); f_ml = zb_read(b, n_ml); f_ll = zb_read(b, n_ll);
          f_sl = zb_read(b, n_sl); f_sm = zb_read(b, n_sm); f_so = zb_read(b, n_so);
        }
        uint32_t offset;
        if (ofc > 1) {
          offset = ((1u << ofc) - 3) + f_of;
          rep2 = rep1; rep1 = rep0; rep0 = offset;
        } else {
          const uint32_t ll0 = (ll_base == 0);
          if (ofc == 0) {
            offset = ll0 ? rep1 : rep0;
            if (ll0) { rep1 = rep0; rep0 = offset; }
          } else {
            const uint32_t idx = 1 + ll0 + f_of;
            uint32_t t = (idx == 3) ? rep0 - 1 : (idx == 1 ? rep1 : rep2);
            if (t == 0) return -1;
            if (idx != 1) rep2 = rep1;
            rep1 = rep0;
            rep0 = offset = t;
          }
        }
        const uint32_t mlen = kZMLBase[mlc] + f_ml;
        const uint32_t llen = ll_base + f_ll;
        if (more) {
          sl = (el >> 16) + f_sl;
          sm = (em >> 16) + f_sm;
          so = (eo >> 16) + f_so;
        }
        if (b.bits < 0) return -1;
        if (lane == npend) { my_ll = (int)llen; my_ml = (int)mlen; my_off = offset; }
        npend++;
        dec++;
      }
      // ---- positions of the pending sequences -------------------------------------------------------------------------------
      const bool have = lane < npend;
      const int ll = have ? my_ll : 0, ml = have ? my_ml : 0;
      const int incl = warp_incl_scan(ll + ml, lane), incl_ll = warp_incl_scan(ll, lane);
      const long myop = op + incl - ll - ml, cpy = myop + ll, mend = cpy + ml;
      const long mylit = litpos + incl_ll - ll;            // where this sequence's literals start in the literal buffer
      // what the one-by-one decoder refuses (zstd_decompress_block.c:896-1010): more literals than the block has, output
      // beyond the first literal that is still unread after the sequence, an offset beyond the start of the output
      const bool broken = have && (mylit + ll > litsize || mend > litbase + mylit + ll || (long)my_off > cpy);
      // lane-parallel copies also need every destination of the batch below the batch's first literal source
      const bool safe = have && !broken && mend <= litbase + litpos;
      const unsigned sm_ = __ballot_sync(kFull, safe), bm_ = __ballot_sync(kFull, broken);
      int cnt = sm_ == kFull ? 32 : __ffs((int)~sm_) - 1;
      if (bm_ && (int)(__ffs((int)bm_) - 1) <= cnt) return -1;   // the first sequence that cannot run is a broken one
      if (cnt > 1) {
        lz_batch_execute(W, W, win, lane, cnt, (int)(litbase + mylit), (int)myop, ll, (int)cpy, (int)my_off, ml);
      } else {
        // one sequence the old way (its output may overlap its own literals: z_move_literals reads before it writes)
        cnt = 1;
        const int l0 = __shfl_sync(kFull, ll, 0), m0 = __shfl_sync(kFull, ml, 0);
        const uint32_t o0 = __shfl_sync(kFull, my_off, 0);
        z_move_literals(win + op, lit + litpos, l0, lane);
        __syncwarp();
        lz_match_copy(win, (int)(op + l0), (int)o0, m0, lane);
        __syncwarp();
      }
      op = __shfl_sync(kFull, (int)mend, cnt - 1);
      litpos = __shfl_sync(kFull, (int)(mylit + ll), cnt - 1);
      my_ll = __shfl_down_sync(kFull, my_ll, (unsigned)cnt);
      my_ml = __shfl_down_sync(kFull, my_ml, (unsigned)cnt);
      my_off = __shfl_down_sync(kFull, my_off, (unsigned)cnt);
      npend -= cnt;
    }
    if (b.bits != 0) return -1;  // must reach the exact start of the bitstream (:1581)
    if (lane == 0) { z->rep[0] = rep0; z->rep[1] = rep1; z->rep[2] = rep2; }
  }
  const long rest = litsize - litpos;
  if (rest > cap - op) return -1;
  __syncwarp();
  z_move_literals(win + op, lit + litpos, (int)rest, lane);
  __syncwarp();
  return op + rest;
}

// One or more frames back to back (zstd_decompress.c:447-560, 953-1060).  Returns the decoded size or -1.
// batch: the window has at least 16 bytes of slack behind dst + cap (see z_decode_block).
__device__ __forceinline__ long zstd_decode_warp(ZWork* z, const uint8_t* src, long n, uint8_t* dst, long cap,
                                                 bool batch = false) {
  const int lane = lane_id();
  long ip = 0, total = 0;
  while (ip < n) {
    if (n - ip < 4) return -1;
    const uint32_t magic = (uint32_t)ld_le32(src + ip);
    if ((magic & 0xFFFFFFF0u) == 0x184D2A50u) {  // skippable frame
      if (n - ip < 8) return -1;
      uint32_t sz = (uint32_t)ld_le32(src + ip + 4);
      if ((long)sz > n - ip - 8) return -1;
      ip += 8 + (long)sz;
      continue;
    }
    if (magic != 0xFD2FB528u) return -1;
    ip += 4;
    if (ip >= n) return -1;
    const uint32_t fhd = src[ip++];
    const int fcs_flag = (int)(fhd >> 6), single = (int)((fhd >> 5) & 1), checksum = (int)((fhd >> 2) & 1);
    const int did = (int)(fhd & 3);
    if (fhd & 0x08) return -1;
    uint64_t window = 0, fcs = 0;
    int have_fcs = 0;
    if (!single) {
      if (ip >= n) return -1;
      uint32_t wd = src[ip++];
      int wlog = (int)(wd >> 3) + 10;
      if (wlog > 31) return -1;
      window = 1ull << wlog;
      window += (window >> 3) * (wd & 7);
    }
    if (did) {
      const int dsz = did == 3 ? 4 : did;
      uint32_t id = 0;
      if (ip + dsz > n) return -1;
      for (int i = 0; i < dsz; i++) id |= (uint32_t)src[ip + i] << (8 * i);
      ip += dsz;
      if (id != 0) return -1;
    }
    const int fcs_size = fcs_flag == 0 ? single : (1 << fcs_flag);
    if (fcs_size) {
      if (ip + fcs_size > n) return -1;
      for (int i = 0; i < fcs_size; i++) fcs |= (uint64_t)src[ip + i] << (8 * i);
      if (fcs_size == 2) fcs += 256;
      ip += fcs_size;
      have_fcs = 1;
    }
    if (single) window = fcs;
    const long blockmax = window < 128 * 1024 ? (long)window : 128 * 1024;
    __syncwarp();
    if (lane == 0) {
      z->rep[0] = 1; z->rep[1] = 4; z->rep[2] = 8;
      z->ll_valid = z->ml_valid = z->of_valid = z->huf_valid = 0;
    }
    __syncwarp();
    uint8_t* fdst = dst + total;
    const long fcap = cap - total;
    long op = 0;
    for (;;) {
      if (ip + 3 > n) return -1;
      const uint32_t bh = (uint32_t)src[ip] | ((uint32_t)src[ip + 1] << 8) | ((uint32_t)src[ip + 2] << 16);
      ip += 3;
      const int last = (int)(bh & 1), type = (int)((bh >> 1) & 3);
      const long bsize = (long)(bh >> 3);
      if (type == 3) return -1;
      if (type == 0) {
        if (bsize > n - ip || bsize > fcap - op) return -1;
        copy_bytes(fdst + op, src + ip, (int)bsize, lane, 32);
        op += bsize;
        ip += bsize;
      } else if (type == 1) {
        if (ip >= n || bsize > fcap - op) return -1;
        fill_bytes(fdst + op, src[ip], (int)bsize, lane, 32);
        op += bsize;
        ip += 1;
      } else {
        if (bsize > n - ip || bsize > blockmax) return -1;
        long r = z_decode_block(z, src + ip, bsize, fdst, op, fcap, blockmax, lane, batch);
        if (r < 0) return -1;
        op = r;
        ip += bsize;
      }
      __syncwarp();
      if (last) break;
    }
    if (have_fcs && (uint64_t)op != fcs) return -1;
    if (checksum) { if (ip + 4 > n) return -1; ip += 4; }
    total += op;
  }
  return total;
}

}  // namespace b2g
