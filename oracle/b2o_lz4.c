/* oracle/b2o_lz4.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the LZ4 block codec as the reference
 * uses it: LZ4_compress_fast(acceleration = 1) (b2c:390-417 -> L4:1453, L4:1380-1403, L4:924-1330) and
 * LZ4_decompress_safe (b2c:433-450 -> L4:2451, L4:2016-2440).  L4 =
 * /root/reference/thirdparty/c-blosc2/internal-complibs/lz4-1.10.0/lz4.c.  Index-based, no dictionary modes. */
#include <string.h>
#include "b2o.h"

enum { MINMATCH = 4, LASTLITERALS = 5, MFLIMIT = 12, ML_MASK = 15, RUN_MASK = 15, SKIP_TRIGGER = 6 };
#define LZ4_64K_LIMIT (65536 + (MFLIMIT - 1))
#define LZ4_MAX_DISTANCE 65535

static uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
static uint64_t rd64(const uint8_t* p) { uint64_t v; memcpy(&v, p, 8); return v; }

/* L4:777-804.  byU16 tables hash 4 bytes into 13 bits; byU32 tables (64-bit host) hash 5 bytes into 12 bits. */
static uint32_t hash_pos(const uint8_t* p, int by_u16) {
  if (by_u16) return (rd32(p) * 2654435761U) >> (32 - 13);
  return (uint32_t)(((rd64(p) << 24) * 889523592379ULL) >> (64 - 12));
}

/* L4:689-707 LZ4_count: number of equal bytes, scanning pin up to limit */
static int count_eq(const uint8_t* s, int pin, int pmatch, int limit) {
  int start = pin;
  while (pin < limit && s[pin] == s[pmatch]) { pin++; pmatch++; }
  return pin - start;
}

static int lz4_compress_bound(int n) { return n + n / 255 + 16; }

/* L4:924-1330 LZ4_compress_generic_validated, specialised to noDict / acceleration 1.
 * Returns 0 when the output does not fit `cap` (limitedOutput), exactly like the reference. */
int b2o_lz4_compress(const uint8_t* src, int n, uint8_t* dst, int cap) {
  if (n <= 0 || n > 0x7E000000) return 0;
  const int limited = cap < lz4_compress_bound(n); /* L4:1389 */
  const int by_u16 = n < LZ4_64K_LIMIT;              /* L4:1390,1396 */
  uint32_t table[1 << 13];
  memset(table, 0, sizeof(table));

  int ip = 0, anchor = 0, op = 0;
  const int iend = n, mflimit_plus_one = iend - MFLIMIT + 1, matchlimit = iend - LASTLITERALS;
  const int olimit = cap;
  uint32_t forward_h;
  int match = 0, token;

  if (n < MFLIMIT + 1) goto last_literals; /* L4:1002 */

  table[hash_pos(src + ip, by_u16)] = 0; /* first byte, L4:1005-1011 */
  ip++;
  forward_h = hash_pos(src + ip, by_u16);

  for (;;) {
    /* find a match, L4:1040-1110 */
    {
      int forward_ip = ip, step = 1, search_nb = 1 << SKIP_TRIGGER;
      for (;;) {
        uint32_t h = forward_h;
        int current = forward_ip;
        int match_index = (int)table[h];
        ip = forward_ip;
        forward_ip += step;
        step = (search_nb++ >> SKIP_TRIGGER);
        if (forward_ip > mflimit_plus_one) goto last_literals;
        match = match_index;
        forward_h = hash_pos(src + forward_ip, by_u16);
        table[h] = (uint32_t)current;
        if (!by_u16 && match_index + LZ4_MAX_DISTANCE < current) continue; /* too far (byU32 only) */
        if (rd32(src + match) == rd32(src + ip)) break;
      }
    }
    /* catch up, L4:1113-1117 */
    while (ip > anchor && match > 0 && src[ip - 1] == src[match - 1]) { ip--; match--; }

    /* encode literals, L4:1120-1145 */
    {
      int lit = ip - anchor;
      token = op++;
      if (limited && op + lit + (2 + 1 + LASTLITERALS) + lit / 255 > olimit) return 0;
      if (lit >= RUN_MASK) {
        int len = lit - RUN_MASK;
        dst[token] = (uint8_t)(RUN_MASK << 4);
        for (; len >= 255; len -= 255) dst[op++] = 255;
        dst[op++] = (uint8_t)len;
      } else {
        dst[token] = (uint8_t)(lit << 4);
      }
      memcpy(dst + op, src + anchor, (size_t)lit);
      op += lit;
    }

  next_match:
    /* offset, L4:1166-1174 */
    dst[op++] = (uint8_t)((ip - match) & 255);
    dst[op++] = (uint8_t)((ip - match) >> 8);
    /* match length, L4:1177-1227 */
    {
      int code = count_eq(src, ip + MINMATCH, match + MINMATCH, matchlimit);
      ip += code + MINMATCH;
      if (limited && op + (1 + LASTLITERALS) + (code + 240) / 255 > olimit) return 0;
      if (code >= ML_MASK) {
        dst[token] += ML_MASK;
        code -= ML_MASK;
        while (code >= 255) { dst[op++] = 255; code -= 255; }
        dst[op++] = (uint8_t)code;
      } else {
        dst[token] += (uint8_t)code;
      }
    }
    anchor = ip;
    if (ip >= mflimit_plus_one) break; /* L4:1234 */

    table[hash_pos(src + ip - 2, by_u16)] = (uint32_t)(ip - 2); /* fill table, L4:1237-1243 */

    /* test next position, L4:1254-1294 */
    {
      uint32_t h = hash_pos(src + ip, by_u16);
      int current = ip, match_index = (int)table[h];
      match = match_index;
      table[h] = (uint32_t)current;
      if ((by_u16 || match_index + LZ4_MAX_DISTANCE >= current) && rd32(src + match) == rd32(src + ip)) {
        token = op++;
        dst[token] = 0;
        goto next_match;
      }
    }
    forward_h = hash_pos(src + (++ip), by_u16); /* L4:1297 */
  }

last_literals: /* L4:1301-1327 */
  {
    int last_run = iend - anchor;
    if (limited && op + last_run + 1 + (last_run + 255 - RUN_MASK) / 255 > olimit) return 0;
    if (last_run >= RUN_MASK) {
      int acc = last_run - RUN_MASK;
      dst[op++] = (uint8_t)(RUN_MASK << 4);
      for (; acc >= 255; acc -= 255) dst[op++] = 255;
      dst[op++] = (uint8_t)acc;
    } else {
      dst[op++] = (uint8_t)(last_run << 4);
    }
    memcpy(dst + op, src + anchor, (size_t)last_run);
    op += last_run;
  }
  return op;
}

/* L4:1976-2011 read_variable_length; returns -1 on error.  `ilimit` is an index. */
static long read_varlen(const uint8_t* s, int* ip, int ilimit, int initial_check) {
  long length = 0;
  unsigned b;
  if (initial_check && *ip >= ilimit) return -1;
  b = s[(*ip)++];
  length += b;
  if (*ip > ilimit) return -1;
  if (b != 255) return length;
  do {
    b = s[(*ip)++];
    length += b;
    if (*ip > ilimit) return -1;
  } while (b == 255);
  return length;
}

/* L4:2016-2440 LZ4_decompress_generic(endOnInputSize, decode_full_block, noDict), safe loop only (the fast
 * loop is an optimisation with identical results).  Returns decoded size or -1. */
int b2o_lz4_decompress(const uint8_t* src, int n, uint8_t* dst, int cap) {
  int ip = 0, op = 0;
  const int iend = n, oend = cap;
  if (cap < 0) return -1;
  if (cap == 0) return (n == 1 && src[0] == 0) ? 0 : -1; /* L4:2064-2068 */
  if (n == 0) return -1;
  for (;;) {
    unsigned token = src[ip++];
    long length = token >> 4;
    if (length == RUN_MASK) { /* L4:2241-2247 */
      long addl = read_varlen(src, &ip, iend - RUN_MASK, 1);
      if (addl < 0) return -1;
      length += addl;
    }
    long cpy = op + length;
    if (cpy > oend - MFLIMIT || ip + length > iend - (2 + 1 + LASTLITERALS)) { /* L4:2279 */
      if (ip + length != iend || cpy > oend) return -1;                         /* L4:2310 */
      memmove(dst + op, src + ip, (size_t)length);
      op += (int)length;
      break;
    }
    memcpy(dst + op, src + ip, (size_t)length);
    ip += (int)length;
    op = (int)cpy;

    int offset = src[ip] | (src[ip + 1] << 8); /* L4:2334 */
    ip += 2;
    int match = op - offset;
    length = token & ML_MASK;
    if (length == ML_MASK) { /* L4:2342-2347 */
      long addl = read_varlen(src, &ip, iend - LASTLITERALS + 1, 0);
      if (addl < 0) return -1;
      length += addl;
    }
    length += MINMATCH;
    if (match < 0) return -1; /* L4:2353 offset outside buffer */
    cpy = op + length;
    if (cpy > oend - MFLIMIT) { /* L4:2419-2423 */
      if (cpy > oend - LASTLITERALS) return -1;
    }
    if (offset == 0) { /* accepted by the reference: the match reads itself after L4:500 / L4:2407 wrote a zero word
                          at its start, so it comes out as zeros (never produced by a valid encoder; pinned against
                          the reference in tests/test_oracle.py) */
      memset(dst + op, 0, (size_t)length);
    } else
    for (long k = 0; k < length; k++) dst[op + k] = dst[match + k];
    op = (int)cpy;
  }
  return op;
}
