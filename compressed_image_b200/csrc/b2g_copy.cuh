// b2g_copy.cuh -- cooperative byte copies / fills for arbitrary alignments (used for stored streams, memcpyed
// chunks, zero / run streams and the scratch -> chunk compaction of the encoder).  `tid`/`nthr` describe the
// cooperating thread group (a warp: lane, 32; a CTA: threadIdx.x, blockDim.x).  No synchronisation inside.
#pragma once
#include "b2g_dev.cuh"

namespace b2g {

// 16 bytes starting at an arbitrarily aligned address, assembled from aligned 32-bit words.  Every word that is
// loaded contains at least one requested byte.
__device__ __forceinline__ uint4 ld16_unaligned(const uint8_t* s) {
  const unsigned sh = (unsigned)((uintptr_t)s & 3) * 8;
  const uint32_t* b = (const uint32_t*)((uintptr_t)s & ~(uintptr_t)3);
  uint32_t w0 = b[0], w1 = b[1], w2 = b[2], w3 = b[3];
  uint32_t w4 = sh ? b[4] : 0u;
  uint4 v;
  v.x = __funnelshift_r(w0, w1, sh);
  v.y = __funnelshift_r(w1, w2, sh);
  v.z = __funnelshift_r(w2, w3, sh);
  v.w = __funnelshift_r(w3, w4, sh);
  return v;
}

// 16 bytes global -> shared without a register in between (LDGSTS); both addresses 16-byte aligned.
__device__ __forceinline__ void async_copy16(void* smem_dst, const void* gmem_src) {
#ifdef B2G_SIM
  *(uint4*)smem_dst = *(const uint4*)gmem_src;
#else
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
#endif
}
__device__ __forceinline__ void async_copy_wait() {
#ifndef B2G_SIM
  asm volatile("cp.async.wait_all;" ::: "memory");
#endif
}

// dst[0..n) = src[0..n); the ranges must not overlap.
__device__ __forceinline__ void copy_bytes(uint8_t* dst, const uint8_t* src, int n, int tid, int nthr) {
  if (n <= 0) return;
  int head = (int)((16 - ((uintptr_t)dst & 15)) & 15);
  if (head > n) head = n;
  for (int i = tid; i < head; i += nthr) dst[i] = src[i];
  const uint8_t* s = src + head;
  uint8_t* d = dst + head;
  int rem = n - head;
  int groups = rem >> 4;
  if (groups > 0 && ((uintptr_t)s & 3) != 0 && rem - (groups << 4) < 3) groups--;  // keep the last word in bounds
  if (((uintptr_t)s & 15) == 0) {
    const uint4* s4 = (const uint4*)s;
    uint4* d4 = (uint4*)d;
    for (int g = tid; g < groups; g += nthr) d4[g] = s4[g];
  } else {
    uint4* d4 = (uint4*)d;
    for (int g = tid; g < groups; g += nthr) d4[g] = ld16_unaligned(s + ((size_t)g << 4));
  }
  for (int i = (groups << 4) + tid; i < rem; i += nthr) d[i] = s[i];
}

// dst[0..n) = value
__device__ __forceinline__ void fill_bytes(uint8_t* dst, uint8_t value, int n, int tid, int nthr) {
  if (n <= 0) return;
  int head = (int)((16 - ((uintptr_t)dst & 15)) & 15);
  if (head > n) head = n;
  for (int i = tid; i < head; i += nthr) dst[i] = value;
  uint8_t* d = dst + head;
  int rem = n - head;
  int groups = rem >> 4;
  uint32_t w = 0x01010101u * value;
  uint4 v = make_uint4(w, w, w, w);
  uint4* d4 = (uint4*)d;
  for (int g = tid; g < groups; g += nthr) d4[g] = v;
  for (int i = (groups << 4) + tid; i < rem; i += nthr) d[i] = value;
}

// lane % d for 0 <= lane < 32, 1 <= d < 32 (warp-uniform d) without an integer division: (lane + 0.5) / d is never
// within rounding distance of an integer.
__device__ __forceinline__ int lane_mod(int lane, int d) {
  const int q = (int)(((float)lane + 0.5f) * (1.0f / (float)d));
  return lane - q * d;
}

// LZ77 back-reference copy by one warp: out[op .. op+len) = out[op-dist ..] with byte-serial (overlapping)
// semantics.  dist >= 1.  The caller has made earlier writes visible (__syncwarp) before calling.
__device__ __forceinline__ void lz_match_copy(uint8_t* out, int op, int dist, int len, int lane) {
  const int match = op - dist;
  if (dist >= len) {
    if (len >= 64) copy_bytes(out + op, out + match, len, lane, 32);
    else for (int i = lane; i < len; i += 32) out[op + i] = out[match + i];
  } else if (dist == 1) {
    fill_bytes(out + op, out[match], len, lane, 32);
  } else if (dist >= 32) {
    for (int i = 0; i < len; i += 32) {  // each 32-byte step only reads bytes finished by earlier steps
      if (i + lane < len) out[op + i + lane] = out[match + i + lane];
      __syncwarp();
    }
  } else {  // short period: replicate the `dist` bytes before op
    int r = lane % dist;
    const int step = 32 % dist;
    for (int i = lane; i < len; i += 32) {
      out[op + i] = out[match + r];
      r += step;
      if (r >= dist) r -= dist;
    }
  }
}

}  // namespace b2g
