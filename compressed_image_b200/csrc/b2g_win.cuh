// b2g_win.cuh -- byte-addressed views of a decode window for the hot loops of the stream decoders.
//
// A window lives either in shared memory (the usual case) or in global memory (streams too large for shared memory,
// unfiltered blocks decoded straight into the destination).  MemShared holds the window as a 32-bit shared-space
// address and goes through ld.shared / st.shared directly: left to itself the compiler re-derives the shared address
// of a generic pointer (S2R SR_CgaCtaId + LEA) in front of every access of a loop, which puts a long-latency special
// register read on the critical path of the token chain.  MemGeneric is the same interface over a plain pointer (and
// the only one in the CPU simulator build).
//
// Accesses through one view are never reordered against each other (volatile asm / may-alias byte pointers).
#pragma once
#include "b2g_dev.cuh"

namespace b2g {

struct MemGeneric {
  uint8_t* p;
  __device__ __forceinline__ explicit MemGeneric(const void* q) : p((uint8_t*)q) {}
  __device__ __forceinline__ MemGeneric offset(int o) const { return MemGeneric(p + o); }
  __device__ __forceinline__ uint32_t ld8(int i) const { return p[i]; }
  __device__ __forceinline__ void st8(int i, uint32_t v) const { p[i] = (uint8_t)v; }
  __device__ __forceinline__ uint32_t ld32(int i) const { return *(const uint32_t*)(p + i); }   // p + i word aligned
  __device__ __forceinline__ void st32(int i, uint32_t v) const { *(uint32_t*)(p + i) = v; }    // p + i word aligned
  __device__ __forceinline__ unsigned low2(int i) const { return (unsigned)((uintptr_t)(p + i) & 3); }
  // a view whose index 0 is as cheap to address as it gets, and the index of this view's byte 0 in it
  __device__ __forceinline__ MemGeneric origin(int& base) const { base = 0; return *this; }
};

#ifdef B2G_SIM
using MemShared = MemGeneric;
#else
struct MemShared {
  uint32_t b;
  __device__ __forceinline__ explicit MemShared(const void* q) : b((uint32_t)__cvta_generic_to_shared(q)) {}
  __device__ __forceinline__ MemShared offset(int o) const { MemShared m(*this); m.b += (uint32_t)o; return m; }
  __device__ __forceinline__ uint32_t ld8(int i) const {
    uint32_t v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(b + (uint32_t)i));
    return v;
  }
  __device__ __forceinline__ void st8(int i, uint32_t v) const {
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(b + (uint32_t)i), "r"(v) : "memory");
  }
  __device__ __forceinline__ uint32_t ld32(int i) const {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(b + (uint32_t)i));
    return v;
  }
  __device__ __forceinline__ void st32(int i, uint32_t v) const {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(b + (uint32_t)i), "r"(v) : "memory");
  }
  __device__ __forceinline__ unsigned low2(int i) const { return (b + (uint32_t)i) & 3u; }
  // indices of the returned view ARE shared-memory addresses: no add in front of a load whose index is a loop variable
  __device__ __forceinline__ MemShared origin(int& base) const { base = (int)b; MemShared m(*this); m.b = 0; return m; }
};
#endif

template <bool SMEM> struct MemSel { using type = MemGeneric; };
template <> struct MemSel<true> { using type = MemShared; };

// Lane-private copy: every lane moves ITS n bytes from I[s..) to O[d..) (n, s, d differ per lane; n <= a few dozen).
// Ranges of different lanes and source / destination of one lane do not overlap.  The destination is brought to a word
// boundary with up to three byte moves, the body moves words (the source word assembled from two aligned loads with a
// funnel shift, one new load per word), up to three bytes finish.  Loop bounds are warp maxima: all lanes call.
// The body may read up to 4 bytes beyond the source's last byte (the caller keeps that inside the stream).
template <class IN, class OUT>
__device__ __forceinline__ void lane_copy(const IN& I, int s, const OUT& O, int d, int n) {
  {
    const int h = min((int)((4u - O.low2(d)) & 3u), n);
    uint32_t v0 = 0, v1 = 0, v2 = 0;
    if (h > 0) v0 = I.ld8(s);
    if (h > 1) v1 = I.ld8(s + 1);
    if (h > 2) v2 = I.ld8(s + 2);
    if (h > 0) O.st8(d, v0);
    if (h > 1) O.st8(d + 1, v1);
    if (h > 2) O.st8(d + 2, v2);
    s += h; d += h; n -= h;
  }
  const int nw = n >> 2;
  const int maxw = (int)__reduce_max_sync(kFull, (unsigned)nw);
  const unsigned sh = I.low2(s) * 8u;
  const int sa = s - (int)(sh >> 3);
  uint32_t w0 = 0;
  if (nw > 0) w0 = I.ld32(sa);
  int i = 0;
  for (; i + 1 < maxw; i += 2) {   // two words per step, both loads first
    uint32_t w1 = 0, w2 = 0;
    if (i < nw) w1 = I.ld32(sa + 4 * i + 4);
    if (i + 1 < nw) w2 = I.ld32(sa + 4 * i + 8);
    if (i < nw) O.st32(d + 4 * i, __funnelshift_r(w0, w1, sh));
    if (i + 1 < nw) O.st32(d + 4 * i + 4, __funnelshift_r(w1, w2, sh));
    w0 = w2;
  }
  if (i < maxw) {
    uint32_t w1 = 0;
    if (i < nw) w1 = I.ld32(sa + 4 * i + 4);
    if (i < nw) O.st32(d + 4 * i, __funnelshift_r(w0, w1, sh));
  }
  {
    const int t = n & 3, ts = s + 4 * nw, td = d + 4 * nw;
    uint32_t v0 = 0, v1 = 0, v2 = 0;
    if (t > 0) v0 = I.ld8(ts);
    if (t > 1) v1 = I.ld8(ts + 1);
    if (t > 2) v2 = I.ld8(ts + 2);
    if (t > 0) O.st8(td, v0);
    if (t > 1) O.st8(td + 1, v1);
    if (t > 2) O.st8(td + 2, v2);
  }
}

// Warp-wide copy of n bytes from I[s..) to O[d..) (uniform arguments, ranges do not overlap): destination words spread
// over the lanes.  May read up to 7 bytes beyond the source's last byte.
template <class IN, class OUT>
__device__ __forceinline__ void warp_copy(const IN& I, int s, const OUT& O, int d, int n, int lane) {
  const int h = min((int)((4u - O.low2(d)) & 3u), n);
  if (lane < h) O.st8(d + lane, I.ld8(s + lane));
  s += h; d += h; n -= h;
  const int nw = n >> 2;
  const unsigned sh = I.low2(s) * 8u;
  const int sa = s - (int)(sh >> 3);
  for (int i = lane; i < nw; i += 32) {
    const uint32_t w0 = I.ld32(sa + 4 * i), w1 = I.ld32(sa + 4 * i + 4);
    O.st32(d + 4 * i, __funnelshift_r(w0, w1, sh));
  }
  const int t = n & 3;
  if (lane < t) O.st8(d + 4 * nw + lane, I.ld8(s + 4 * nw + lane));
}

// O[d .. d+n) = value, by one warp (uniform arguments): head bytes up to a word boundary, words, tail bytes.
template <class OUT>
__device__ __forceinline__ void warp_fill(const OUT& O, int d, uint32_t value, int n, int lane) {
  const int h = min((int)((4u - O.low2(d)) & 3u), n);
  if (lane < h) O.st8(d + lane, value);
  d += h; n -= h;
  const int nw = n >> 2;
  const uint32_t w = 0x01010101u * (value & 0xffu);
  for (int i = lane; i < nw; i += 32) O.st32(d + 4 * i, w);
  const int t = n & 3;
  if (lane < t) O.st8(d + 4 * nw + lane, value);
}

// LZ77 back-reference by one warp through a view: O[op .. op+len) = O[op-dist ..) with byte-serial (overlapping)
// semantics, dist >= 1 (lz_match_copy of b2g_copy.cuh for a window behind a view: in a shared-memory window every
// access is an ld.shared / st.shared instead of a generic load / store).  The caller has made earlier writes visible.
template <class OUT>
__device__ __forceinline__ void warp_match_copy(const OUT& O, int op, int dist, int len, int lane) {
  const int match = op - dist;
  if (dist >= len) {
    if (len >= 64) {
      // word-granular; the 4-byte slack warp_copy reads past the source's end lies in [op, op + 8): written by this very
      // copy, never used (the funnel shift only keeps source bytes)
      warp_copy(O, match, O, op, len, lane);
    } else {
      for (int i = lane; i < len; i += 32) O.st8(op + i, O.ld8(match + i));
    }
  } else if (dist == 1) {
    warp_fill(O, op, O.ld8(match), len, lane);
  } else if (dist >= 32) {
    for (int i = 0; i < len; i += 32) {  // each 32-byte step only reads bytes finished by earlier steps
      if (i + lane < len) O.st8(op + i + lane, O.ld8(match + i + lane));
      __syncwarp();
    }
  } else {  // short period: replicate the `dist` bytes before op
    int r = lane % dist;
    const int step = 32 % dist;
    for (int i = lane; i < len; i += 32) {
      O.st8(op + i, O.ld8(match + r));
      r += step;
      if (r >= dist) r -= dist;
    }
  }
}

}  // namespace b2g
