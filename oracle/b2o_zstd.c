/* oracle/b2o_zstd.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the zstd frame + block DEcoder as the
 * reference reaches it (b2c:524-549 zstd_wrap_decompress -> ZSTD_decompressDCtx).  Z/ =
 * /root/reference/thirdparty/c-blosc2/internal-complibs/zstd-1.5.7/.  Follows the published format (RFC 8878)
 * and these reference functions:
 *   frame header   Z/decompress/zstd_decompress.c:447-560     frame loop   :953-1060
 *   block header   Z/decompress/zstd_decompress_block.c:63-77 literals     :134-340
 *   FSE NCount     Z/common/entropy_common.c:42-230           FSE tables   zstd_decompress_block.c:485-640
 *   HUF weights    Z/common/entropy_common.c:234-340          HUF X1 table Z/decompress/huf_decompress.c:385-500
 *   sequences      zstd_decompress_block.c:695-775, 1229-1346, 1500-1590 (decode), 896-1010 (exec)
 * No dictionaries, no legacy frames, content checksum skipped (not verified): the reference never enables it. */
#include <string.h>
#include "b2o.h"

#define ERR (-1L)
enum { MAX_LL = 35, MAX_ML = 52, MAX_OFF = 31, LL_LOG = 9, ML_LOG = 9, OF_LOG = 8, HUF_LOG_MAX = 12 };

typedef struct { uint8_t sym, nbits; uint16_t base; } fse_ent;          /* newState = base + readbits(nbits) */
typedef struct { fse_ent e[512]; int log; int valid; } fse_tab;
typedef struct { uint8_t sym[1 << HUF_LOG_MAX], nb[1 << HUF_LOG_MAX]; int log; int valid; } huf_tab;

typedef struct {
  fse_tab ll, of, ml;
  huf_tab huf;
  uint32_t rep[3];
} frame_state;

static int hbit(uint32_t v) { int r = -1; while (v) { r++; v >>= 1; } return r; }

/* ---- forward (LSB-first) bit reader used by FSE table descriptions ---- */
typedef struct { const uint8_t* p; long n; long bitpos; } fbits;
static uint32_t fpeek(const fbits* b, int nb) {
  uint32_t v = 0;
  for (int i = 0; i < nb; i++) {
    long bp = b->bitpos + i;
    if ((bp >> 3) < b->n) v |= (uint32_t)((b->p[bp >> 3] >> (bp & 7)) & 1) << i;
  }
  return v;
}

/* Z/common/entropy_common.c:42-230 FSE_readNCount.  Returns bytes consumed or ERR. */
static long read_ncount(const uint8_t* src, long n, int16_t* norm, int* maxsym, int* tablelog, int maxlog) {
  fbits b = {src, n, 0};
  if (n < 1) return ERR;
  int al = (int)fpeek(&b, 4) + 5;
  b.bitpos += 4;
  if (al > maxlog) return ERR;
  int remaining = 1 << al, sym = 0, limit = *maxsym + 1;
  while (remaining > 0 && sym < limit) {
    int bits = hbit((uint32_t)remaining + 1) + 1;
    uint32_t val = fpeek(&b, bits);
    uint32_t lower = (1u << (bits - 1)) - 1;
    uint32_t thresh = (1u << bits) - 1 - ((uint32_t)remaining + 1);
    if ((val & lower) < thresh) { b.bitpos += bits - 1; val &= lower; }
    else { b.bitpos += bits; if (val > lower) val -= thresh; }
    int proba = (int)val - 1;
    remaining -= proba < 0 ? -proba : proba;
    norm[sym++] = (int16_t)proba;
    if (proba == 0) {
      for (;;) {
        uint32_t rep = fpeek(&b, 2);
        b.bitpos += 2;
        for (uint32_t i = 0; i < rep && sym < limit; i++) norm[sym++] = 0;
        if (rep != 3) break;
      }
    }
    if ((b.bitpos >> 3) > n) return ERR;
  }
  if (remaining != 0) return ERR;
  if (sym > limit) return ERR;
  *maxsym = sym - 1;
  *tablelog = al;
  long used = (b.bitpos + 7) >> 3;
  if (used > n) return ERR;
  return used;
}

/* zstd_decompress_block.c:485-640 / RFC 8878 4.1.1: spread symbols, then assign nbits + baseline */
static int build_fse(fse_tab* t, const int16_t* norm, int maxsym, int log) {
  int size = 1 << log, high = size;
  uint8_t symbols[512];
  uint16_t next[64];
  if (log > 9 || maxsym > 63) return -1;
  for (int s = 0; s <= maxsym; s++) {
    if (norm[s] == -1) { symbols[--high] = (uint8_t)s; next[s] = 1; }
    else next[s] = (uint16_t)norm[s];
  }
  int step = (size >> 1) + (size >> 3) + 3, mask = size - 1, pos = 0;
  for (int s = 0; s <= maxsym; s++) {
    for (int i = 0; i < norm[s]; i++) {
      symbols[pos] = (uint8_t)s;
      do { pos = (pos + step) & mask; } while (pos >= high);
    }
  }
  if (pos != 0) return -1;
  for (int i = 0; i < size; i++) {
    uint8_t s = symbols[i];
    uint32_t ns = next[s]++;
    int nb = log - hbit(ns);
    t->e[i].sym = s;
    t->e[i].nbits = (uint8_t)nb;
    t->e[i].base = (uint16_t)((ns << nb) - (uint32_t)size);
  }
  t->log = log;
  t->valid = 1;
  return 0;
}

/* ---- backward bit reader (Z/common/bitstream.h:248-400): bits are consumed from the top of the stream ---- */
typedef struct { const uint8_t* p; long bits; } bbits; /* bits = number of unread bits; may go negative */
static int binit(bbits* b, const uint8_t* p, long n) {
  if (n < 1 || p[n - 1] == 0) return -1;
  b->p = p;
  b->bits = n * 8 - (8 - hbit(p[n - 1]));
  return 0;
}
static uint32_t bread(bbits* b, int nb) { /* reads nb bits; bits beyond the start read as 0 */
  uint32_t v = 0;
  for (int i = 0; i < nb; i++) {
    long bp = b->bits - 1 - i;
    v <<= 1;
    if (bp >= 0) v |= (uint32_t)((b->p[bp >> 3] >> (bp & 7)) & 1);
  }
  b->bits -= nb;
  return v;
}

static const int16_t LL_DEF[36] = {4, 3, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2, 1, 1, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 3, 2, 1, 1, 1, 1, 1, -1, -1, -1, -1};
static const int16_t ML_DEF[53] = {1, 4, 3, 2, 2, 2, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1, -1, -1};
static const int16_t OF_DEF[29] = {1, 1, 1, 1, 1, 1, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1};
static const uint32_t LL_BASE[36] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 18, 20, 22, 24, 28, 32, 40, 48, 64, 0x80, 0x100, 0x200, 0x400, 0x800, 0x1000, 0x2000, 0x4000, 0x8000, 0x10000};
static const uint8_t LL_BITS[36] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};
static const uint32_t ML_BASE[53] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16, 17, 18, 19, 20, 21, 22, 23, 24, 25, 26, 27, 28, 29, 30, 31, 32, 33, 34, 35, 37, 39, 41, 43, 47, 51, 59, 67, 83, 99, 0x83, 0x103, 0x203, 0x403, 0x803, 0x1003, 0x2003, 0x4003, 0x8003, 0x10003};
static const uint8_t ML_BITS[53] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 3, 3, 4, 4, 5, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16};

/* zstd_decompress_block.c:647-693 ZSTD_buildSeqTable. Returns bytes consumed or ERR. */
static long build_seq_table(fse_tab* t, int mode, int maxsym, int maxlog, const uint8_t* src, long n,
                            const int16_t* defnorm, int deflog) {
  int16_t norm[64];
  switch (mode) {
    case 0: return build_fse(t, defnorm, maxsym == MAX_OFF ? 28 : maxsym, deflog) ? ERR : 0;
    case 1:
      if (n < 1 || src[0] > maxsym) return ERR;
      t->e[0].sym = src[0]; t->e[0].nbits = 0; t->e[0].base = 0; t->log = 0; t->valid = 1;
      return 1;
    case 2: {
      int ms = maxsym, log;
      long used = read_ncount(src, n, norm, &ms, &log, maxlog);
      if (used < 0) return ERR;
      if (build_fse(t, norm, ms, log)) return ERR;
      return used;
    }
    default: return t->valid ? 0 : ERR;
  }
}

/* Z/common/entropy_common.c:234-340 HUF_readStats + huf_decompress.c:385-500 (X1 table). Returns bytes used. */
static long read_huf_table(huf_tab* h, const uint8_t* src, long n) {
  uint8_t w[256];
  int nw = 0;
  long used;
  if (n < 1) return ERR;
  int hb = src[0];
  if (hb >= 128) { /* direct 4-bit weights */
    nw = hb - 127;
    used = 1 + (nw + 1) / 2;
    if (used > n) return ERR;
    for (int i = 0; i < nw; i += 2) {
      w[i] = src[1 + i / 2] >> 4;
      w[i + 1] = src[1 + i / 2] & 15;
    }
  } else { /* FSE-compressed weights, two interleaved states (Z/common/fse_decompress.c:170-230) */
    used = 1 + hb;
    if (used > n || hb < 1) return ERR;
    int16_t norm[16];
    int ms = 12, log;
    long hs = read_ncount(src + 1, hb, norm, &ms, &log, 6);
    if (hs < 0) return ERR;
    fse_tab ft;
    if (build_fse(&ft, norm, ms, log)) return ERR;
    bbits b;
    if (binit(&b, src + 1 + hs, hb - hs)) return ERR;
    uint32_t s1 = bread(&b, log), s2 = bread(&b, log);
    if (b.bits < 0) return ERR;
    for (;;) {
      if (nw > 253) return ERR;
      w[nw++] = ft.e[s1].sym;
      s1 = ft.e[s1].base + bread(&b, ft.e[s1].nbits);
      if (b.bits < 0) { w[nw++] = ft.e[s2].sym; break; }
      if (nw > 253) return ERR;
      w[nw++] = ft.e[s2].sym;
      s2 = ft.e[s2].base + bread(&b, ft.e[s2].nbits);
      if (b.bits < 0) { w[nw++] = ft.e[s1].sym; break; }
    }
  }
  uint32_t total = 0, rank[HUF_LOG_MAX + 2] = {0};
  for (int i = 0; i < nw; i++) {
    if (w[i] > HUF_LOG_MAX) return ERR;
    rank[w[i]]++;
    total += (1u << w[i]) >> 1;
  }
  if (total == 0) return ERR;
  int log = hbit(total) + 1;
  if (log > HUF_LOG_MAX) return ERR;
  uint32_t rest = (1u << log) - total;
  if (rest & (rest - 1)) return ERR; /* last weight must complete a power of two */
  int last = hbit(rest) + 1;
  w[nw++] = (uint8_t)last;
  rank[last]++;
  if (rank[1] < 2 || (rank[1] & 1)) return ERR;
  uint32_t start[HUF_LOG_MAX + 2], next = 0;
  for (int r = 1; r <= log; r++) { start[r] = next; next += rank[r] << (r - 1); }
  for (int s = 0; s < nw; s++) {
    int wt = w[s];
    if (!wt) continue;
    uint32_t len = (1u << wt) >> 1;
    for (uint32_t k = 0; k < len; k++) {
      h->sym[start[wt] + k] = (uint8_t)s;
      h->nb[start[wt] + k] = (uint8_t)(log + 1 - wt);
    }
    start[wt] += len;
  }
  h->log = log;
  h->valid = 1;
  return used;
}

/* huf_decompress.c:558-606 (1 stream): decode exactly `count` symbols, stream must end exactly */
static int huf_stream(const huf_tab* h, const uint8_t* src, long n, uint8_t* out, long count) {
  bbits b;
  if (binit(&b, src, n)) return -1;
  for (long i = 0; i < count; i++) {
    long save = b.bits;
    uint32_t idx = bread(&b, h->log);
    out[i] = h->sym[idx];
    b.bits = save - h->nb[idx];
  }
  return b.bits == 0 ? 0 : -1;
}

/* zstd_decompress_block.c:134-340 ZSTD_decodeLiteralsBlock. Returns bytes consumed; fills lit/litsize. */
static long decode_literals(frame_state* fs, const uint8_t* src, long n, uint8_t* lit, long* litsize, long blockmax) {
  if (n < 2) return ERR; /* MIN_CBLOCK_SIZE */
  int type = src[0] & 3, sf = (src[0] >> 2) & 3;
  long lh, regen, comp = 0;
  if (type < 2) {
    if (sf == 0 || sf == 2) { lh = 1; regen = src[0] >> 3; }
    else if (sf == 1) { lh = 2; regen = (src[0] >> 4) + ((long)src[1] << 4); }
    else { if (n < 3) return ERR; lh = 3; regen = (src[0] >> 4) + ((long)src[1] << 4) + ((long)src[2] << 12); }
    if (regen > blockmax) return ERR;
    if (type == 0) {
      if (lh + regen > n) return ERR;
      memcpy(lit, src + lh, (size_t)regen);
      *litsize = regen;
      return lh + regen;
    }
    if (lh + 1 > n) return ERR;
    memset(lit, src[lh], (size_t)regen);
    *litsize = regen;
    return lh + 1;
  }
  if (n < 5) return ERR;
  int streams = 4;
  uint32_t v = src[0] | (src[1] << 8) | (src[2] << 16) | ((uint32_t)src[3] << 24);
  if (sf < 2) { lh = 3; streams = sf ? 4 : 1; regen = (v >> 4) & 0x3FF; comp = (v >> 14) & 0x3FF; }
  else if (sf == 2) { lh = 4; regen = (v >> 4) & 0x3FFF; comp = v >> 18; }
  else { lh = 5; regen = (v >> 4) & 0x3FFFF; comp = (v >> 22) + ((long)src[4] << 10); }
  if (regen > blockmax) return ERR;
  if (streams == 4 && regen < 6) return ERR;
  if (lh + comp > n) return ERR;
  const uint8_t* p = src + lh;
  long left = comp;
  if (type == 2) {
    long hs = read_huf_table(&fs->huf, p, left);
    if (hs < 0) return ERR;
    p += hs;
    left -= hs;
  } else if (!fs->huf.valid) {
    return ERR;
  }
  if (streams == 1) {
    if (huf_stream(&fs->huf, p, left, lit, regen)) return ERR;
  } else {
    if (left < 10) return ERR;
    long s1 = p[0] | (p[1] << 8), s2 = p[2] | (p[3] << 8), s3 = p[4] | (p[5] << 8);
    long s4 = left - 6 - s1 - s2 - s3;
    if (s4 < 1 || s1 + s2 + s3 + 6 > left) return ERR;
    long seg = (regen + 3) / 4, lastseg = regen - 3 * seg;
    if (lastseg < 0) return ERR;
    p += 6;
    if (huf_stream(&fs->huf, p, s1, lit, seg)) return ERR;
    if (huf_stream(&fs->huf, p + s1, s2, lit + seg, seg)) return ERR;
    if (huf_stream(&fs->huf, p + s1 + s2, s3, lit + 2 * seg, seg)) return ERR;
    if (huf_stream(&fs->huf, p + s1 + s2 + s3, s4, lit + 3 * seg, lastseg)) return ERR;
  }
  *litsize = regen;
  return lh + comp;
}

/* zstd_decompress_block.c:2066-2200 ZSTD_decompressBlock_internal for one compressed block.
 * `dst` is the frame output start, `op` the current write index, `cap` the capacity. */
static long decode_block(frame_state* fs, const uint8_t* src, long n, uint8_t* dst, long op, long cap, uint8_t* lit,
                         long blockmax) {
  long litsize = 0;
  long used = decode_literals(fs, src, n, lit, &litsize, blockmax);
  if (used < 0) return ERR;
  const uint8_t* ip = src + used;
  const uint8_t* iend = src + n;
  if (ip >= iend) return ERR;
  long nbseq = *ip++;
  if (nbseq > 0x7F) {
    if (nbseq == 0xFF) { if (ip + 2 > iend) return ERR; nbseq = (ip[0] | (ip[1] << 8)) + 0x7F00; ip += 2; }
    else { if (ip >= iend) return ERR; nbseq = ((nbseq - 0x80) << 8) + *ip++; }
  }
  long litpos = 0;
  if (nbseq == 0) {
    if (ip != iend) return ERR;
  } else {
    if (ip + 1 > iend) return ERR;
    if (*ip & 3) return ERR;
    int llm = *ip >> 6, ofm = (*ip >> 4) & 3, mlm = (*ip >> 2) & 3;
    ip++;
    long r;
    if ((r = build_seq_table(&fs->ll, llm, MAX_LL, LL_LOG, ip, iend - ip, LL_DEF, 6)) < 0) return ERR;
    ip += r;
    if ((r = build_seq_table(&fs->of, ofm, MAX_OFF, OF_LOG, ip, iend - ip, OF_DEF, 5)) < 0) return ERR;
    ip += r;
    if ((r = build_seq_table(&fs->ml, mlm, MAX_ML, ML_LOG, ip, iend - ip, ML_DEF, 6)) < 0) return ERR;
    ip += r;
    bbits b;
    if (binit(&b, ip, iend - ip)) return ERR;
    uint32_t sl = bread(&b, fs->ll.log), so = bread(&b, fs->of.log), sm = bread(&b, fs->ml.log);
    if (b.bits < 0) return ERR;
    for (long i = 0; i < nbseq; i++) {
      fse_ent el = fs->ll.e[sl], eo = fs->of.e[so], em = fs->ml.e[sm];
      if (el.sym > MAX_LL || em.sym > MAX_ML || eo.sym > MAX_OFF) return ERR;
      uint32_t ofcode = eo.sym;
      uint32_t offset;
      uint32_t ll_base = LL_BASE[el.sym];
      if (ofcode > 1) {
        offset = ((1u << ofcode) - 3) + bread(&b, (int)ofcode);
        fs->rep[2] = fs->rep[1]; fs->rep[1] = fs->rep[0]; fs->rep[0] = offset;
      } else {
        uint32_t ll0 = (ll_base == 0);
        if (ofcode == 0) {
          offset = fs->rep[ll0];
          fs->rep[1] = fs->rep[!ll0];
          fs->rep[0] = offset;
        } else {
          uint32_t idx = 1 + ll0 + bread(&b, 1);
          uint32_t t = (idx == 3) ? fs->rep[0] - 1 : fs->rep[idx];
          if (t == 0) return ERR;
          if (idx != 1) fs->rep[2] = fs->rep[1];
          fs->rep[1] = fs->rep[0];
          fs->rep[0] = offset = t;
        }
      }
      uint32_t ml = ML_BASE[em.sym] + bread(&b, ML_BITS[em.sym]);
      uint32_t ll = ll_base + bread(&b, LL_BITS[el.sym]);
      if (i + 1 < nbseq) {
        sl = el.base + bread(&b, el.nbits);
        sm = em.base + bread(&b, em.nbits);
        so = eo.base + bread(&b, eo.nbits);
      }
      if (b.bits < 0) return ERR;
      /* execute, zstd_decompress_block.c:896-1010 */
      if ((long)ll + (long)ml > cap - op) return ERR;
      if ((long)ll > litsize - litpos) return ERR;
      memcpy(dst + op, lit + litpos, ll);
      op += ll;
      litpos += ll;
      if ((long)offset > op) return ERR;
      for (uint32_t k = 0; k < ml; k++) dst[op + k] = dst[op - offset + k];
      op += ml;
    }
    if (b.bits != 0) return ERR; /* :1581 must reach the exact end */
  }
  long rest = litsize - litpos;
  if (rest > cap - op) return ERR;
  memcpy(dst + op, lit + litpos, (size_t)rest);
  return op + rest;
}

/* zstd_decompress.c:447-560 + 953-1060: one or more frames back to back */
long b2o_zstd_decompress(const uint8_t* src, long n, uint8_t* dst, long cap) {
  static uint8_t litbuf[128 * 1024 + 32]; /* oracle is single-threaded test code */
  long ip = 0, total = 0;
  while (ip < n) {
    if (n - ip < 4) return ERR;
    uint32_t magic = src[ip] | (src[ip + 1] << 8) | (src[ip + 2] << 16) | ((uint32_t)src[ip + 3] << 24);
    if ((magic & 0xFFFFFFF0u) == 0x184D2A50u) { /* skippable frame */
      if (n - ip < 8) return ERR;
      uint32_t sz = src[ip + 4] | (src[ip + 5] << 8) | (src[ip + 6] << 16) | ((uint32_t)src[ip + 7] << 24);
      if ((long)sz > n - ip - 8) return ERR;
      ip += 8 + sz;
      continue;
    }
    if (magic != 0xFD2FB528u) return ERR;
    ip += 4;
    if (ip >= n) return ERR;
    uint8_t fhd = src[ip++];
    int fcs_flag = fhd >> 6, single = (fhd >> 5) & 1, checksum = (fhd >> 2) & 1, did = fhd & 3;
    if (fhd & 0x08) return ERR; /* reserved bit */
    uint64_t window = 0, fcs = 0;
    int have_fcs = 0;
    if (!single) {
      if (ip >= n) return ERR;
      uint8_t wd = src[ip++];
      int wlog = (wd >> 3) + 10;
      if (wlog > 31) return ERR;
      window = 1ull << wlog;
      window += (window >> 3) * (wd & 7);
    }
    static const int did_size[4] = {0, 1, 2, 4};
    if (did) { /* a dictionary id is only legal when it is 0 here */
      uint32_t id = 0;
      if (ip + did_size[did] > n) return ERR;
      for (int i = 0; i < did_size[did]; i++) id |= (uint32_t)src[ip + i] << (8 * i);
      ip += did_size[did];
      if (id != 0) return ERR;
    }
    int fcs_size = fcs_flag == 0 ? single : (1 << fcs_flag);
    if (fcs_size) {
      if (ip + fcs_size > n) return ERR;
      for (int i = 0; i < fcs_size; i++) fcs |= (uint64_t)src[ip + i] << (8 * i);
      if (fcs_size == 2) fcs += 256;
      ip += fcs_size;
      have_fcs = 1;
    }
    if (single) window = fcs;
    long blockmax = window < 128 * 1024 ? (long)window : 128 * 1024;

    frame_state fs;
    memset(&fs, 0, sizeof(fs));
    fs.rep[0] = 1; fs.rep[1] = 4; fs.rep[2] = 8;
    uint8_t* fdst = dst + total;
    long fcap = cap - total, op = 0;
    for (;;) {
      if (ip + 3 > n) return ERR;
      uint32_t bh = src[ip] | (src[ip + 1] << 8) | (src[ip + 2] << 16);
      ip += 3;
      int last = bh & 1, type = (bh >> 1) & 3;
      long bsize = bh >> 3;
      if (type == 3) return ERR;
      if (type == 0) {
        if (bsize > n - ip || bsize > fcap - op) return ERR;
        memcpy(fdst + op, src + ip, (size_t)bsize);
        op += bsize;
        ip += bsize;
      } else if (type == 1) {
        if (ip >= n || bsize > fcap - op) return ERR;
        memset(fdst + op, src[ip], (size_t)bsize);
        op += bsize;
        ip += 1;
      } else {
        if (bsize > n - ip || bsize > blockmax) return ERR; /* zstd_decompress_block.c:78 */
        long r = decode_block(&fs, src + ip, bsize, fdst, op, fcap, litbuf, blockmax);
        if (r < 0) return ERR;
        op = r;
        ip += bsize;
      }
      if (last) break;
    }
    if (have_fcs && (uint64_t)op != fcs) return ERR;
    if (checksum) { if (ip + 4 > n) return ERR; ip += 4; }
    total += op;
  }
  return total;
}
