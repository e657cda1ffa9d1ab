/* oracle/b2o_blosclz.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the blosclz stream codec.
 * Follows /root/reference/thirdparty/c-blosc2/blosc/blosclz.c: get_cratio :320-419, blosclz_compress
 * :422-619, blosclz_decompress :685-795 (cited below as BL:line).  Index-based restatement. */
#include <string.h>
#include "b2o.h"

enum { MAX_COPY = 32, MAX_DISTANCE = 8191, MAX_FARDISTANCE = 65535 + 8191 - 1, HASH_LOG = 14 };

static uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
static uint32_t hashf(uint32_t seq, int hashlog) { return (seq * 2654435761U) >> (32 - hashlog); }

/* BL:120-145 get_run : ip advances while the byte at ref equals s[ip0-1]; no bound on ref other than ip */
static int run_end(const uint8_t* s, int ip, int ip_bound, int ref) {
  uint8_t x = s[ip - 1];
  while (ip < ip_bound && s[ref] == x) { ip++; ref++; }
  return ip;
}

/* BL:149-166 get_match : returns one past the first differing byte, or ip_bound if none differs */
static int match_end(const uint8_t* s, int ip, int ip_bound, int ref) {
  while (ip < ip_bound) {
    if (s[ref++] != s[ip++]) break;
  }
  return ip;
}

/* BL:320-419 entropy probe */
static double get_cratio(const uint8_t* ibase, int maxlen, int minlen, int ipshift, uint32_t* htab, int hashlog) {
  int ip = 0;
  int32_t oc = 0;
  const uint16_t hashlen = (uint16_t)(1U << hashlog);
  uint16_t limit = (uint16_t)((maxlen > hashlen) ? hashlen : maxlen);
  int ip_bound = (int)limit - 1, ip_limit = (int)limit - 12;
  uint8_t copy;
  memset(htab, 0, hashlen * sizeof(uint32_t));
  copy = 4;
  oc += 5;
  while (ip < ip_limit) {
    int anchor = ip;
    uint32_t seq = rd32(ibase + ip);
    uint32_t hval = hashf(seq, hashlog);
    int ref = (int)htab[hval];
    unsigned distance = (unsigned)(anchor - ref);
    htab[hval] = (uint32_t)anchor;
    int literal = 0;
    if (distance == 0 || distance >= MAX_FARDISTANCE) {
      literal = 1;
    } else if (rd32(ibase + ref) == rd32(ibase + ip)) {
      ref += 4;
      ip = anchor + 4;
      distance--;
      ip = distance ? match_end(ibase, ip, ip_bound, ref) : run_end(ibase, ip, ip_bound, ref);
      ip -= ipshift;
      if (ip - anchor < minlen) literal = 1;
    } else {
      literal = 1;
    }
    if (literal) { /* LITERAL2, BL:262-270 */
      oc++;
      anchor++;
      ip = anchor;
      copy++;
      if (copy == MAX_COPY) { copy = 0; oc++; }
      continue;
    }
    int len = ip - anchor;
    if (!copy) oc--;
    copy = 0;
    if (distance < MAX_DISTANCE) {
      if (len >= 7) oc += ((len - 7) / 255) + 1;
      oc += 2;
    } else {
      if (len >= 7) oc += ((len - 7) / 255) + 1;
      oc += 4;
    }
    seq = rd32(ibase + ip);
    hval = hashf(seq, hashlog);
    htab[hval] = (uint32_t)ip;
    ip += 2;
    oc++;
  }
  return (double)ip / (double)oc;
}

/* BL:422-619 */
int b2o_blosclz_compress(int clevel, const uint8_t* ibase, int length, uint8_t* out, int maxout) {
  static const uint8_t hashlog_[10] = {0, HASH_LOG - 2, HASH_LOG - 1, HASH_LOG, HASH_LOG, HASH_LOG,
                                       HASH_LOG, HASH_LOG, HASH_LOG, HASH_LOG};
  static const double cratio_[10] = {0, 2, 1.5, 1.2, 1.2, 1.2, 1.2, 1.15, 1.1, 1.0};
  uint32_t htab[1U << HASH_LOG];
  const int ipshift = 4, minlen = 4;
  int hashlog = hashlog_[clevel];
  int maxlen = length;
  if (clevel < 2) maxlen /= 8;
  else if (clevel < 4) maxlen /= 4;
  else if (clevel < 7) maxlen /= 2;
  int shift = length - maxlen;
  double cratio = get_cratio(ibase + shift, maxlen, minlen, ipshift, htab, hashlog);
  if (cratio < cratio_[clevel]) return 0;

  int ip = 0, ip_bound = length - 1, ip_limit = length - 12;
  int op = 0;
  const int op_limit = maxout;
  uint8_t copy;
  if (length < 16 || maxout < 66) return 0;
  memset(htab, 0, (1U << hashlog) * sizeof(uint32_t));

  copy = 4;
  out[op++] = MAX_COPY - 1;
  out[op++] = ibase[ip++];
  out[op++] = ibase[ip++];
  out[op++] = ibase[ip++];
  out[op++] = ibase[ip++];

  while (ip < ip_limit) {
    int anchor = ip;
    uint32_t seq = rd32(ibase + ip);
    uint32_t hval = hashf(seq, hashlog);
    int ref = (int)htab[hval];
    unsigned distance = (unsigned)(anchor - ref);
    unsigned len = 0;
    htab[hval] = (uint32_t)anchor;
    int literal = 0;
    if (distance == 0 || distance >= MAX_FARDISTANCE) {
      literal = 1;
    } else if (rd32(ibase + ref) == rd32(ibase + ip)) {
      ref += 4;
      ip = anchor + 4;
      distance--;
      ip = distance ? match_end(ibase, ip, ip_bound, ref) : run_end(ibase, ip, ip_bound, ref);
      ip -= ipshift;
      len = (unsigned)(ip - anchor);
      if ((int)len < minlen || (len <= 5 && distance >= MAX_DISTANCE)) literal = 1;
    } else {
      literal = 1;
    }
    if (literal) { /* LITERAL, BL:248-258 */
      if (op + 2 > op_limit) return 0;
      out[op++] = ibase[anchor++];
      ip = anchor;
      copy++;
      if (copy == MAX_COPY) { copy = 0; out[op++] = MAX_COPY - 1; }
      continue;
    }
    if (copy) out[op - copy - 1] = (uint8_t)(copy - 1);
    else op--;
    copy = 0;

    if (distance < MAX_DISTANCE) {
      if (len < 7) { /* MATCH_SHORT */
        if (op + 2 > op_limit) return 0;
        out[op++] = (uint8_t)((len << 5U) + (distance >> 8U));
        out[op++] = (uint8_t)(distance & 255U);
      } else { /* MATCH_LONG */
        if (op + 1 > op_limit) return 0;
        out[op++] = (uint8_t)((7U << 5U) + (distance >> 8U));
        for (len -= 7; len >= 255; len -= 255) {
          if (op + 1 > op_limit) return 0;
          out[op++] = 255;
        }
        if (op + 2 > op_limit) return 0;
        out[op++] = (uint8_t)len;
        out[op++] = (uint8_t)(distance & 255U);
      }
    } else {
      distance -= MAX_DISTANCE;
      if (len < 7) { /* MATCH_SHORT_FAR */
        if (op + 4 > op_limit) return 0;
        out[op++] = (uint8_t)((len << 5U) + 31);
        out[op++] = 255;
        out[op++] = (uint8_t)(distance >> 8U);
        out[op++] = (uint8_t)(distance & 255U);
      } else { /* MATCH_LONG_FAR */
        if (op + 1 > op_limit) return 0;
        out[op++] = (7U << 5U) + 31;
        for (len -= 7; len >= 255; len -= 255) {
          if (op + 1 > op_limit) return 0;
          out[op++] = 255;
        }
        if (op + 4 > op_limit) return 0;
        out[op++] = (uint8_t)len;
        out[op++] = 255;
        out[op++] = (uint8_t)(distance >> 8U);
        out[op++] = (uint8_t)(distance & 255U);
      }
    }
    /* update the hash at match boundary, BL:574-587 */
    seq = rd32(ibase + ip);
    hval = hashf(seq, hashlog);
    htab[hval] = (uint32_t)(ip++);
    if (clevel == 9) {
      seq >>= 8U;
      hval = hashf(seq, hashlog);
      htab[hval] = (uint32_t)(ip++);
    } else {
      ip++;
    }
    if (op + 1 > op_limit) return 0;
    out[op++] = MAX_COPY - 1;
  }
  while (ip <= ip_bound) { /* left-over as literal copy, BL:596-604 */
    if (op + 2 > op_limit) return 0;
    out[op++] = ibase[ip++];
    copy++;
    if (copy == MAX_COPY) { copy = 0; out[op++] = MAX_COPY - 1; }
  }
  if (copy) out[op - copy - 1] = (uint8_t)(copy - 1);
  else op--;
  out[0] |= (1U << 5U); /* marker, BL:613 */
  return op;
}

/* BL:685-795 */
int b2o_blosclz_decompress(const uint8_t* in, int length, uint8_t* out, int maxout) {
  int ip = 0, op = 0;
  const int ip_limit = length, op_limit = maxout;
  uint32_t ctrl;
  if (length == 0) return 0;
  ctrl = in[ip++] & 31U;
  for (;;) {
    if (ctrl >= 32) {
      int32_t len = (int32_t)(ctrl >> 5U) - 1;
      int32_t ofs = (int32_t)(ctrl & 31U) << 8U;
      uint8_t code;
      long ref = (long)op - ofs;
      if (len == 7 - 1) {
        do {
          if (ip + 1 >= ip_limit) return 0;
          code = in[ip++];
          len += code;
        } while (code == 255);
      } else {
        if (ip + 1 >= ip_limit) return 0;
      }
      code = in[ip++];
      len += 3;
      ref -= code;
      if (code == 255 && ofs == (31 << 8)) { /* far match, BL:723-732 */
        if (ip + 1 >= ip_limit) return 0;
        ofs = in[ip++] << 8;
        ofs += in[ip++];
        ref = (long)op - ofs - MAX_DISTANCE;
      }
      if (op + len > op_limit) return 0;
      if (ref - 1 < 0) return 0;
      if (ip >= ip_limit) break; /* BL:742: a trailing match is NOT copied */
      ctrl = in[ip++];
      ref--;
      for (int32_t k = 0; k < len; k++) out[op + k] = out[ref + k];
      op += len;
    } else {
      ctrl++;
      if (op + (int)ctrl > op_limit) return 0;
      if (ip + (int)ctrl > ip_limit) return 0;
      memcpy(out + op, in + ip, ctrl);
      op += (int)ctrl;
      ip += (int)ctrl;
      if (ip >= ip_limit) break;
      ctrl = in[ip++];
    }
  }
  return op;
}
