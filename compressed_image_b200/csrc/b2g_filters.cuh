// b2g_filters.cuh -- byte-shuffle / bit-shuffle filters fused into the codec kernels (rows K1-K4 of SURVEY.md 8a).
//
// Layouts (reference: thirdparty/c-blosc2/blosc/shuffle.c:421-514, shuffle-generic.h:34-81,
// bitshuffle-generic.c:147-167, 240-258):
//   shuffle     filtered[j*ne + i] = raw[i*ts + j],  ne = bsize/ts; the bsize%ts trailing bytes are copied.
//   bitshuffle  only size = ne - ne%8 elements are transposed: 8*ts bit-rows of size/8 bytes; row (jj*8+kk), byte i,
//               bit b = bit kk of byte jj of element 8i+b; everything after size*ts bytes is copied.
//               Chunk format version 2 (Blosc1) instead skips the transpose entirely when ne%8 != 0.
// The "window" (filtered block) lives in shared memory (or a global scratch slab for very large blocks); the raw
// block lives in global memory.  Fast paths move 16 raw bytes per thread with 128-bit global accesses; generic
// paths handle any typesize / alignment byte by byte.  `tid`/`nthr` = cooperating threads; the caller synchronises.
#pragma once
#include "b2g_copy.cuh"

namespace b2g {

// ---- plane <-> element byte transposes in registers -----------------------------------------------------------
// 4 words w[0..3] (4 elements of 4 bytes)  <->  4 planes (plane j = byte j of the 4 elements)
__device__ __forceinline__ void transpose4x4(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t& o0,
                                             uint32_t& o1, uint32_t& o2, uint32_t& o3) {
  uint32_t t0 = __byte_perm(a, b, 0x5140), t1 = __byte_perm(c, d, 0x5140);
  uint32_t t2 = __byte_perm(a, b, 0x7362), t3 = __byte_perm(c, d, 0x7362);
  o0 = __byte_perm(t0, t1, 0x5410);
  o1 = __byte_perm(t0, t1, 0x7632);
  o2 = __byte_perm(t2, t3, 0x5410);
  o3 = __byte_perm(t2, t3, 0x7632);
}

// ---- shuffle: 16 raw bytes per thread -----------------------------------------------------------------------------
template <int TS>
__device__ __forceinline__ void unshuffle16(const uint8_t* win, int ne, int k, uint4* dst4) {
  // raw bytes [16k, 16k+16) = elements e0 .. e0 + 16/TS
  const int e0 = k * (16 / TS);
  uint4 v;
  if (TS == 2) {
    uint2 a = *(const uint2*)(win + e0), b = *(const uint2*)(win + ne + e0);
    v.x = __byte_perm(a.x, b.x, 0x5140); v.y = __byte_perm(a.x, b.x, 0x7362);
    v.z = __byte_perm(a.y, b.y, 0x5140); v.w = __byte_perm(a.y, b.y, 0x7362);
  } else if (TS == 4) {
    uint32_t a = *(const uint32_t*)(win + e0), b = *(const uint32_t*)(win + ne + e0);
    uint32_t c = *(const uint32_t*)(win + 2 * ne + e0), d = *(const uint32_t*)(win + 3 * ne + e0);
    transpose4x4(a, b, c, d, v.x, v.y, v.z, v.w);
  } else if (TS == 8) {
    uint32_t p[8];
#pragma unroll
    for (int j = 0; j < 8; j++) p[j] = *(const uint16_t*)(win + (size_t)j * ne + e0);
    uint32_t x01 = __byte_perm(p[0], p[1], 0x5140), x23 = __byte_perm(p[2], p[3], 0x5140);
    uint32_t x45 = __byte_perm(p[4], p[5], 0x5140), x67 = __byte_perm(p[6], p[7], 0x5140);
    v.x = __byte_perm(x01, x23, 0x5410); v.y = __byte_perm(x45, x67, 0x5410);
    v.z = __byte_perm(x01, x23, 0x7632); v.w = __byte_perm(x45, x67, 0x7632);
  } else {  // 16
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
      uint32_t b0 = win[(size_t)(4 * q) * ne + e0], b1 = win[(size_t)(4 * q + 1) * ne + e0];
      uint32_t b2 = win[(size_t)(4 * q + 2) * ne + e0], b3 = win[(size_t)(4 * q + 3) * ne + e0];
      w[q] = b0 | (b1 << 8) | (b2 << 16) | (b3 << 24);
    }
    v = make_uint4(w[0], w[1], w[2], w[3]);
  }
  dst4[k] = v;
}

template <int TS>
__device__ __forceinline__ void shuffle16(const uint4* src4, int ne, int k, uint8_t* win) {
  const int e0 = k * (16 / TS);
  uint4 v = src4[k];
  if (TS == 2) {
    uint2 a, b;
    a.x = __byte_perm(v.x, v.y, 0x6420); b.x = __byte_perm(v.x, v.y, 0x7531);
    a.y = __byte_perm(v.z, v.w, 0x6420); b.y = __byte_perm(v.z, v.w, 0x7531);
    *(uint2*)(win + e0) = a;
    *(uint2*)(win + ne + e0) = b;
  } else if (TS == 4) {
    uint32_t a, b, c, d;
    transpose4x4(v.x, v.y, v.z, v.w, a, b, c, d);
    *(uint32_t*)(win + e0) = a;
    *(uint32_t*)(win + ne + e0) = b;
    *(uint32_t*)(win + 2 * ne + e0) = c;
    *(uint32_t*)(win + 3 * ne + e0) = d;
  } else if (TS == 8) {
    uint32_t x = __byte_perm(v.x, v.z, 0x5140), y = __byte_perm(v.x, v.z, 0x7362);
    uint32_t z = __byte_perm(v.y, v.w, 0x5140), u = __byte_perm(v.y, v.w, 0x7362);
    *(uint16_t*)(win + e0) = (uint16_t)x;
    *(uint16_t*)(win + (size_t)ne + e0) = (uint16_t)(x >> 16);
    *(uint16_t*)(win + (size_t)2 * ne + e0) = (uint16_t)y;
    *(uint16_t*)(win + (size_t)3 * ne + e0) = (uint16_t)(y >> 16);
    *(uint16_t*)(win + (size_t)4 * ne + e0) = (uint16_t)z;
    *(uint16_t*)(win + (size_t)5 * ne + e0) = (uint16_t)(z >> 16);
    *(uint16_t*)(win + (size_t)6 * ne + e0) = (uint16_t)u;
    *(uint16_t*)(win + (size_t)7 * ne + e0) = (uint16_t)(u >> 16);
  } else {  // 16
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int q = 0; q < 4; q++)
#pragma unroll
      for (int b = 0; b < 4; b++) win[(size_t)(4 * q + b) * ne + e0] = (uint8_t)(w[q] >> (8 * b));
  }
}

__device__ __forceinline__ bool shuffle_fast_ok(int ts, int bsize, const void* win, const void* raw) {
  if (!(ts == 2 || ts == 4 || ts == 8 || ts == 16)) return false;
  int ne = bsize / ts;
  return (bsize % 16) == 0 && (ne % (16 / ts)) == 0 && (((uintptr_t)win | (uintptr_t)raw) & 15) == 0;
}

// win (filtered) -> raw
__device__ __forceinline__ void unshuffle_block(int ts, int bsize, const uint8_t* win, uint8_t* raw, int tid,
                                                int nthr) {
  if (ts <= 1) { copy_bytes(raw, win, bsize, tid, nthr); return; }
  const int ne = bsize / ts;
  if (shuffle_fast_ok(ts, bsize, win, raw)) {
    const int groups = bsize >> 4;
    uint4* d4 = (uint4*)raw;
    if (ts == 2) for (int k = tid; k < groups; k += nthr) unshuffle16<2>(win, ne, k, d4);
    else if (ts == 4) for (int k = tid; k < groups; k += nthr) unshuffle16<4>(win, ne, k, d4);
    else if (ts == 8) for (int k = tid; k < groups; k += nthr) unshuffle16<8>(win, ne, k, d4);
    else for (int k = tid; k < groups; k += nthr) unshuffle16<16>(win, ne, k, d4);
    return;
  }
  const int body = ne * ts;
  for (int k = tid; k < body; k += nthr) {
    int i = k / ts, j = k - i * ts;
    raw[k] = win[(size_t)j * ne + i];
  }
  for (int k = body + tid; k < bsize; k += nthr) raw[k] = win[k];
}

// raw -> win (filtered)
__device__ __forceinline__ void shuffle_block(int ts, int bsize, const uint8_t* raw, uint8_t* win, int tid,
                                              int nthr) {
  if (ts <= 1) { copy_bytes(win, raw, bsize, tid, nthr); return; }
  const int ne = bsize / ts;
  if (shuffle_fast_ok(ts, bsize, win, raw)) {
    const int groups = bsize >> 4;
    const uint4* s4 = (const uint4*)raw;
    if (ts == 2) for (int k = tid; k < groups; k += nthr) shuffle16<2>(s4, ne, k, win);
    else if (ts == 4) for (int k = tid; k < groups; k += nthr) shuffle16<4>(s4, ne, k, win);
    else if (ts == 8) for (int k = tid; k < groups; k += nthr) shuffle16<8>(s4, ne, k, win);
    else for (int k = tid; k < groups; k += nthr) shuffle16<16>(s4, ne, k, win);
    return;
  }
  const int body = ne * ts;
  for (int k = tid; k < body; k += nthr) {
    int i = k / ts, j = k - i * ts;
    win[(size_t)j * ne + i] = raw[k];
  }
  for (int k = body + tid; k < bsize; k += nthr) win[k] = raw[k];
}

// ---- bitshuffle: one group of 8 elements per thread -----------------------------------------------------------
// plane jj of group i: u64 whose byte b = byte jj of element 8i+b.  transpose8x8(plane) -> byte kk = bit-row kk.
template <int TS>
__device__ __forceinline__ void bitunshuffle_group(const uint8_t* win, int nrow, int i, uint8_t* raw) {
  uint64_t P[TS];
#pragma unroll
  for (int jj = 0; jj < TS; jj++) {
    uint64_t r = 0;
#pragma unroll
    for (int kk = 0; kk < 8; kk++) r |= (uint64_t)win[(size_t)(jj * 8 + kk) * nrow + i] << (8 * kk);
    P[jj] = transpose8x8(r);
  }
  uint8_t* o = raw + (size_t)i * 8 * TS;
  if (TS == 1) {
    *(uint2*)o = make_uint2((uint32_t)P[0], (uint32_t)(P[0] >> 32));
  } else if (TS == 2) {
    uint32_t a0 = (uint32_t)P[0], a1 = (uint32_t)(P[0] >> 32), b0 = (uint32_t)P[1], b1 = (uint32_t)(P[1] >> 32);
    *(uint4*)o = make_uint4(__byte_perm(a0, b0, 0x5140), __byte_perm(a0, b0, 0x7362), __byte_perm(a1, b1, 0x5140),
                            __byte_perm(a1, b1, 0x7362));
  } else if (TS == 4) {
    uint4 lo, hi;
    transpose4x4((uint32_t)P[0], (uint32_t)P[1], (uint32_t)P[2], (uint32_t)P[3], lo.x, lo.y, lo.z, lo.w);
    transpose4x4((uint32_t)(P[0] >> 32), (uint32_t)(P[1] >> 32), (uint32_t)(P[2] >> 32), (uint32_t)(P[3] >> 32),
                 hi.x, hi.y, hi.z, hi.w);
    ((uint4*)o)[0] = lo;
    ((uint4*)o)[1] = hi;
  } else {  // 8
    uint32_t A[8], B[8];  // A: bytes 0-3 of elements 0..7, B: bytes 4-7
    transpose4x4((uint32_t)P[0], (uint32_t)P[1], (uint32_t)P[2], (uint32_t)P[3], A[0], A[1], A[2], A[3]);
    transpose4x4((uint32_t)(P[0] >> 32), (uint32_t)(P[1] >> 32), (uint32_t)(P[2] >> 32), (uint32_t)(P[3] >> 32),
                 A[4], A[5], A[6], A[7]);
    transpose4x4((uint32_t)P[4], (uint32_t)P[5], (uint32_t)P[6], (uint32_t)P[7], B[0], B[1], B[2], B[3]);
    transpose4x4((uint32_t)(P[4] >> 32), (uint32_t)(P[5] >> 32), (uint32_t)(P[6] >> 32), (uint32_t)(P[7] >> 32),
                 B[4], B[5], B[6], B[7]);
#pragma unroll
    for (int q = 0; q < 4; q++) ((uint4*)o)[q] = make_uint4(A[2 * q], B[2 * q], A[2 * q + 1], B[2 * q + 1]);
  }
}

template <int TS>
__device__ __forceinline__ void bitshuffle_group(const uint8_t* raw, int nrow, int i, uint8_t* win) {
  uint64_t P[TS];
  const uint8_t* s = raw + (size_t)i * 8 * TS;
  if (TS == 1) {
    uint2 v = *(const uint2*)s;
    P[0] = (uint64_t)v.x | ((uint64_t)v.y << 32);
  } else if (TS == 2) {
    uint4 v = *(const uint4*)s;
    uint32_t a0 = __byte_perm(v.x, v.y, 0x6420), b0 = __byte_perm(v.x, v.y, 0x7531);
    uint32_t a1 = __byte_perm(v.z, v.w, 0x6420), b1 = __byte_perm(v.z, v.w, 0x7531);
    P[0] = (uint64_t)a0 | ((uint64_t)a1 << 32);
    P[1] = (uint64_t)b0 | ((uint64_t)b1 << 32);
  } else if (TS == 4) {
    uint4 lo = ((const uint4*)s)[0], hi = ((const uint4*)s)[1];
    uint32_t a[4], b[4];
    transpose4x4(lo.x, lo.y, lo.z, lo.w, a[0], a[1], a[2], a[3]);
    transpose4x4(hi.x, hi.y, hi.z, hi.w, b[0], b[1], b[2], b[3]);
#pragma unroll
    for (int j = 0; j < 4; j++) P[j] = (uint64_t)a[j] | ((uint64_t)b[j] << 32);
  } else {  // 8
    uint4 v[4];
#pragma unroll
    for (int q = 0; q < 4; q++) v[q] = ((const uint4*)s)[q];
    uint32_t a[4], b[4], c[4], d[4];
    // elements 0..3: low words v0.x v0.z v1.x v1.z ; high words v0.y v0.w v1.y v1.w
    transpose4x4(v[0].x, v[0].z, v[1].x, v[1].z, a[0], a[1], a[2], a[3]);
    transpose4x4(v[2].x, v[2].z, v[3].x, v[3].z, b[0], b[1], b[2], b[3]);
    transpose4x4(v[0].y, v[0].w, v[1].y, v[1].w, c[0], c[1], c[2], c[3]);
    transpose4x4(v[2].y, v[2].w, v[3].y, v[3].w, d[0], d[1], d[2], d[3]);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      P[j] = (uint64_t)a[j] | ((uint64_t)b[j] << 32);
      P[4 + j] = (uint64_t)c[j] | ((uint64_t)d[j] << 32);
    }
  }
#pragma unroll
  for (int jj = 0; jj < TS; jj++) {
    uint64_t r = transpose8x8(P[jj]);
#pragma unroll
    for (int kk = 0; kk < 8; kk++) win[(size_t)(jj * 8 + kk) * nrow + i] = (uint8_t)(r >> (8 * kk));
  }
}

__device__ __forceinline__ bool bitshuffle_fast_ok(int ts, const void* raw) {
  if (!(ts == 1 || ts == 2 || ts == 4 || ts == 8)) return false;
  return ((uintptr_t)raw & 15) == 0;
}

// win (filtered) -> raw.  format_version = chunk header byte 0.
__device__ __forceinline__ void bitunshuffle_block(int ts, int bsize, const uint8_t* win, uint8_t* raw,
                                                   int format_version, int tid, int nthr) {
  int ne = bsize / ts;
  if (format_version == 2 && (ne % 8) != 0) { copy_bytes(raw, win, bsize, tid, nthr); return; }  // shuffle.c:478-497
  const int size = ne - (ne % 8), nrow = size / 8;
  if (bitshuffle_fast_ok(ts, raw)) {
    if (ts == 1) for (int i = tid; i < nrow; i += nthr) bitunshuffle_group<1>(win, nrow, i, raw);
    else if (ts == 2) for (int i = tid; i < nrow; i += nthr) bitunshuffle_group<2>(win, nrow, i, raw);
    else if (ts == 4) for (int i = tid; i < nrow; i += nthr) bitunshuffle_group<4>(win, nrow, i, raw);
    else for (int i = tid; i < nrow; i += nthr) bitunshuffle_group<8>(win, nrow, i, raw);
  } else {
    const int pairs = nrow * ts;
    for (int q = tid; q < pairs; q += nthr) {
      int i = q / ts, jj = q - i * ts;
      uint64_t r = 0;
      for (int kk = 0; kk < 8; kk++) r |= (uint64_t)win[(size_t)(jj * 8 + kk) * nrow + i] << (8 * kk);
      r = transpose8x8(r);
      for (int b = 0; b < 8; b++) raw[((size_t)8 * i + b) * ts + jj] = (uint8_t)(r >> (8 * b));
    }
  }
  const int off = size * ts;
  for (int k = off + tid; k < bsize; k += nthr) raw[k] = win[k];
}

// raw -> win (filtered); always the format-version >= 3 rule (we only write version-5 chunks).
__device__ __forceinline__ void bitshuffle_block(int ts, int bsize, const uint8_t* raw, uint8_t* win, int tid,
                                                 int nthr) {
  int ne = bsize / ts;
  const int size = ne - (ne % 8), nrow = size / 8;
  if (bitshuffle_fast_ok(ts, raw)) {
    if (ts == 1) for (int i = tid; i < nrow; i += nthr) bitshuffle_group<1>(raw, nrow, i, win);
    else if (ts == 2) for (int i = tid; i < nrow; i += nthr) bitshuffle_group<2>(raw, nrow, i, win);
    else if (ts == 4) for (int i = tid; i < nrow; i += nthr) bitshuffle_group<4>(raw, nrow, i, win);
    else for (int i = tid; i < nrow; i += nthr) bitshuffle_group<8>(raw, nrow, i, win);
  } else {
    const int pairs = nrow * ts;
    for (int q = tid; q < pairs; q += nthr) {
      int i = q / ts, jj = q - i * ts;
      uint64_t r = 0;
      for (int b = 0; b < 8; b++) r |= (uint64_t)raw[((size_t)8 * i + b) * ts + jj] << (8 * b);
      r = transpose8x8(r);
      for (int kk = 0; kk < 8; kk++) win[(size_t)(jj * 8 + kk) * nrow + i] = (uint8_t)(r >> (8 * kk));
    }
  }
  const int off = size * ts;
  for (int k = off + tid; k < bsize; k += nthr) win[k] = raw[k];
}

__device__ __forceinline__ void filter_backward(int filter, int ts, int bsize, const uint8_t* win, uint8_t* raw,
                                                int format_version, int tid, int nthr) {
  if (filter == kFilterShuffle) unshuffle_block(ts, bsize, win, raw, tid, nthr);
  else if (filter == kFilterBitshuffle) bitunshuffle_block(ts, bsize, win, raw, format_version, tid, nthr);
  else copy_bytes(raw, win, bsize, tid, nthr);
}

__device__ __forceinline__ void filter_forward(int filter, int ts, int bsize, const uint8_t* raw, uint8_t* win,
                                               int tid, int nthr) {
  if (filter == kFilterShuffle) shuffle_block(ts, bsize, raw, win, tid, nthr);
  else if (filter == kFilterBitshuffle) bitshuffle_block(ts, bsize, raw, win, tid, nthr);
  else copy_bytes(win, raw, bsize, tid, nthr);
}

}  // namespace b2g
