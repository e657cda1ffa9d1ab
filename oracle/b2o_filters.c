/* oracle/b2o_filters.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the shuffle / bitshuffle filters.
 * Follows blosc/shuffle.c:421-514, blosc/shuffle-generic.h:34-81, blosc/bitshuffle-generic.c:34-258
 * (reference = /root/reference/thirdparty/c-blosc2).  Written as direct index formulas, not as the
 * reference's three-pass transposes; the resulting byte layout is identical. */
#include <string.h>
#include "b2o.h"

/* shuffle-generic.h:34-53 : dest[j*ne + i] = src[i*ts + j]; trailing n%ts bytes copied as they are */
void b2o_shuffle(int ts, int32_t n, const uint8_t* src, uint8_t* dst) {
  int32_t ne = n / ts, rem = n % ts;
  for (int j = 0; j < ts; j++)
    for (int32_t i = 0; i < ne; i++) dst[(size_t)j * ne + i] = src[(size_t)i * ts + j];
  memcpy(dst + (n - rem), src + (n - rem), (size_t)rem);
}

/* shuffle-generic.h:62-81 */
void b2o_unshuffle(int ts, int32_t n, const uint8_t* src, uint8_t* dst) {
  int32_t ne = n / ts, rem = n % ts;
  for (int32_t i = 0; i < ne; i++)
    for (int j = 0; j < ts; j++) dst[(size_t)i * ts + j] = src[(size_t)j * ne + i];
  memcpy(dst + (n - rem), src + (n - rem), (size_t)rem);
}

/* bitshuffle-generic.c:147-167 (bshuf_trans_bit_elem_scal): for `size` elements (multiple of 8) of ts bytes,
 * output is 8*ts bit-rows of size/8 bytes; row (jj*8+kk), byte i, bit b  =  bit kk of byte jj of element 8i+b. */
static void bit_transpose(int ts, size_t size, const uint8_t* src, uint8_t* dst) {
  size_t nrow = size / 8;
  memset(dst, 0, size * (size_t)ts);
  for (size_t e = 0; e < size; e++)
    for (int jj = 0; jj < ts; jj++) {
      uint8_t v = src[e * ts + jj];
      for (int kk = 0; kk < 8; kk++)
        if ((v >> kk) & 1) dst[((size_t)jj * 8 + kk) * nrow + e / 8] |= (uint8_t)(1u << (e & 7));
    }
}

static void bit_untranspose(int ts, size_t size, const uint8_t* src, uint8_t* dst) {
  size_t nrow = size / 8;
  memset(dst, 0, size * (size_t)ts);
  for (size_t e = 0; e < size; e++)
    for (int jj = 0; jj < ts; jj++) {
      uint8_t v = 0;
      for (int kk = 0; kk < 8; kk++)
        v |= (uint8_t)(((src[((size_t)jj * 8 + kk) * nrow + e / 8] >> (e & 7)) & 1) << kk);
      dst[e * ts + jj] = v;
    }
}

/* shuffle.c:447-467 : only size - size%8 elements are bit-transposed, the rest is copied */
int32_t b2o_bitshuffle(int ts, int32_t n, const uint8_t* src, uint8_t* dst) {
  size_t size = (size_t)(n / ts);
  size -= size % 8;
  bit_transpose(ts, size, src, dst);
  size_t off = size * (size_t)ts;
  memcpy(dst + off, src + off, (size_t)n - off);
  return n;
}

/* shuffle.c:471-514 : format version 2 (Blosc1) = all-or-nothing rule; version >= 3 = rule above */
int32_t b2o_bitunshuffle(int ts, int32_t n, const uint8_t* src, uint8_t* dst, int format_version) {
  size_t size = (size_t)(n / ts);
  if (format_version == 2) {
    if (size % 8 == 0) {
      bit_untranspose(ts, size, src, dst);
      size_t off = size * (size_t)ts;
      memcpy(dst + off, src + off, (size_t)n - off);
    } else {
      memcpy(dst, src, (size_t)n);
    }
    return n;
  }
  size -= size % 8;
  bit_untranspose(ts, size, src, dst);
  size_t off = size * (size_t)ts;
  memcpy(dst + off, src + off, (size_t)n - off);
  return n;
}
