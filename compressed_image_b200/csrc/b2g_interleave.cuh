// Interleave / deinterleave of planar channels <-> pixel-interleaved buffers.
//
// Replaces compressed::image_algo::interleave<T> / deinterleave<T>
// (compressed_image/include/compressed/image_algo.h:34-65 and :84-111; call site image.h:1880, the step right before
// the compressor on the read pipeline):   interleaved[i * nch + c] = plane[c][i]   for elements of `ts` bytes.
//
// Pure bandwidth: every byte is read once and written once.  The vector kernels (2..4 channels of 1/2/4/8-byte
// elements, all pointers 16-byte aligned) give each thread 16 bytes of every plane: NCH 128-bit accesses on the planar
// side, NCH consecutive 128-bit accesses on the interleaved side, and a compile-time byte permutation in registers
// between them.  A warp therefore touches 512 contiguous bytes per plane and 512 * NCH contiguous bytes of the
// interleaved buffer per step.  Everything else (other channel counts, odd alignments, the last < 16 bytes of each
// plane) goes through the element kernel, which keeps its stores coalesced.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b2g {

constexpr int kIlvMaxPlanes = 64;    // planes per launch of the element kernel (more channels: several launches)
constexpr int kIlvThreads = 256;

struct IlvPlanes {
  void* p[kIlvMaxPlanes];
};
struct IlvPlanes4 {
  void* p[4];
};

// Byte `idx` of a register array of little-endian words (idx is a compile-time constant after unrolling).
__device__ __forceinline__ uint32_t ilv_byte(const uint32_t* w, int idx) { return (w[idx >> 2] >> ((idx & 3) * 8)) & 0xffu; }

// planar word `ow` (0..3) of channel c from the 4 * NCH interleaved words of one group (16 / TS pixels)
template <int TS, int NCH>
__device__ __forceinline__ uint32_t ilv_planar_word(const uint32_t* iw, int c, int ow) {
  if constexpr (TS >= 4) {
    constexpr int WE = TS / 4;
    return iw[((ow / WE) * NCH + c) * WE + (ow % WE)];
  } else if constexpr (TS == 2) {
    const int h0 = (2 * ow) * NCH + c, h1 = (2 * ow + 1) * NCH + c;    // halfword indices in the interleaved group
    const uint32_t lo = (iw[h0 >> 1] >> ((h0 & 1) * 16)) & 0xffffu;
    const uint32_t hi = (iw[h1 >> 1] >> ((h1 & 1) * 16)) & 0xffffu;
    return lo | (hi << 16);
  } else {
    uint32_t v = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) v |= ilv_byte(iw, (4 * ow + b) * NCH + c) << (8 * b);
    return v;
  }
}

// interleaved word `j` (0 .. 4 * NCH - 1) of one group from the planar words pw[c * 4 + ow]
template <int TS, int NCH>
__device__ __forceinline__ uint32_t ilv_inter_word(const uint32_t* pw, int j) {
  if constexpr (TS >= 4) {
    constexpr int WE = TS / 4;
    const int e = j / WE, k = j % WE;       // element index in the group, word of the element
    return pw[(e % NCH) * 4 + (e / NCH) * WE + k];
  } else if constexpr (TS == 2) {
    const int e0 = 2 * j, e1 = 2 * j + 1;
    const int s0 = (e0 % NCH) * 8 + (e0 / NCH), s1 = (e1 % NCH) * 8 + (e1 / NCH);   // halfword indices in pw
    const uint32_t lo = (pw[s0 >> 1] >> ((s0 & 1) * 16)) & 0xffffu;
    const uint32_t hi = (pw[s1 >> 1] >> ((s1 & 1) * 16)) & 0xffffu;
    return lo | (hi << 16);
  } else {
    uint32_t v = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const int e = 4 * j + b;
      v |= ilv_byte(pw, (e % NCH) * 16 + (e / NCH)) << (8 * b);
    }
    return v;
  }
}

// groups = number of 16-byte vectors per plane.  Grid-stride over groups.
template <int TS, int NCH>
__global__ void __launch_bounds__(kIlvThreads) k_deinterleave_vec(const uint4* __restrict__ inter, IlvPlanes4 pl, int64_t groups) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
    uint32_t iw[4 * NCH];
#pragma unroll
    for (int k = 0; k < NCH; k++) {
      const uint4 v = __ldcs(inter + g * NCH + k);
      iw[4 * k + 0] = v.x; iw[4 * k + 1] = v.y; iw[4 * k + 2] = v.z; iw[4 * k + 3] = v.w;
    }
#pragma unroll
    for (int c = 0; c < NCH; c++) {
      uint4 o;
      o.x = ilv_planar_word<TS, NCH>(iw, c, 0);
      o.y = ilv_planar_word<TS, NCH>(iw, c, 1);
      o.z = ilv_planar_word<TS, NCH>(iw, c, 2);
      o.w = ilv_planar_word<TS, NCH>(iw, c, 3);
      __stcs((uint4*)pl.p[c] + g, o);
    }
  }
}

template <int TS, int NCH>
__global__ void __launch_bounds__(kIlvThreads) k_interleave_vec(uint4* __restrict__ inter, IlvPlanes4 pl, int64_t groups) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += stride) {
    uint32_t pw[4 * NCH];
#pragma unroll
    for (int c = 0; c < NCH; c++) {
      const uint4 v = __ldcs((const uint4*)pl.p[c] + g);
      pw[4 * c + 0] = v.x; pw[4 * c + 1] = v.y; pw[4 * c + 2] = v.z; pw[4 * c + 3] = v.w;
    }
#pragma unroll
    for (int k = 0; k < NCH; k++) {
      uint4 o;
      o.x = ilv_inter_word<TS, NCH>(pw, 4 * k + 0);
      o.y = ilv_inter_word<TS, NCH>(pw, 4 * k + 1);
      o.z = ilv_inter_word<TS, NCH>(pw, 4 * k + 2);
      o.w = ilv_inter_word<TS, NCH>(pw, 4 * k + 3);
      __stcs(inter + g * NCH + k, o);
    }
  }
}

// Tile kernels: any channel count up to kIlvMaxPlanes whose pixel is a whole number of 32-bit words (nch * TS % 4 == 0),
// TS in {1, 2, 4, 8}, pointers 4-byte aligned -- e.g. the 16-channel f32 compositing stack of cfg 3.  A CTA moves a tile
// of P pixels through shared memory: the interleaved side is read / written as consecutive words, the planar side as
// consecutive words of one plane per warp step, and the pixel stride in shared memory is padded to an odd number of
// words so that walking along a plane hits 32 different banks.
constexpr int kIlvTileWords = 8192;   // 32 KiB of shared memory per CTA (24 KiB / 8 CTAs per SM measured no faster)

// 4-byte asynchronous global -> shared copy (LDGSTS): a whole tile is in flight per CTA without staging in registers
__device__ __forceinline__ void ilv_async4(uint32_t* smem_dst, const uint32_t* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void ilv_async_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <int TS>
__device__ __forceinline__ uint32_t ilv_tile_gather(const uint32_t* sm, int psw, int c, int q) {
  if constexpr (TS == 4) return sm[q * psw + c];
  else if constexpr (TS == 8) return sm[(q >> 1) * psw + 2 * c + (q & 1)];
  else if constexpr (TS == 2) {
    const uint16_t* h = (const uint16_t*)sm;
    return (uint32_t)h[(2 * q) * psw * 2 + c] | ((uint32_t)h[(2 * q + 1) * psw * 2 + c] << 16);
  } else {
    const uint8_t* b = (const uint8_t*)sm;
    return (uint32_t)b[(4 * q) * psw * 4 + c] | ((uint32_t)b[(4 * q + 1) * psw * 4 + c] << 8) |
           ((uint32_t)b[(4 * q + 2) * psw * 4 + c] << 16) | ((uint32_t)b[(4 * q + 3) * psw * 4 + c] << 24);
  }
}

template <int TS>
__device__ __forceinline__ void ilv_tile_scatter(uint32_t* sm, int psw, int c, int q, uint32_t v) {
  if constexpr (TS == 4) sm[q * psw + c] = v;
  else if constexpr (TS == 8) sm[(q >> 1) * psw + 2 * c + (q & 1)] = v;
  else if constexpr (TS == 2) {
    uint16_t* h = (uint16_t*)sm;
    h[(2 * q) * psw * 2 + c] = (uint16_t)v;
    h[(2 * q + 1) * psw * 2 + c] = (uint16_t)(v >> 16);
  } else {
    uint8_t* b = (uint8_t*)sm;
    b[(4 * q) * psw * 4 + c] = (uint8_t)v;
    b[(4 * q + 1) * psw * 4 + c] = (uint8_t)(v >> 8);
    b[(4 * q + 2) * psw * 4 + c] = (uint8_t)(v >> 16);
    b[(4 * q + 3) * psw * 4 + c] = (uint8_t)(v >> 24);
  }
}

// n = pixels handled here (n * TS % 4 == 0), P = pixels per tile (multiple of 32), W = nch * TS / 4, psw = W | 1.
template <int TS>
__global__ void __launch_bounds__(kIlvThreads) k_deinterleave_tile(const uint32_t* __restrict__ inter, IlvPlanes pl, int nch, int W,
                                                                   int psw, int P, int64_t n) {
  __shared__ uint32_t sm[kIlvTileWords];
  const int64_t tiles = (n + P - 1) / P;
  const int tid = threadIdx.x, dp = kIlvThreads / W, dk = kIlvThreads % W;
  for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int64_t p0 = t * P;
    const int np = (int)min((int64_t)P, n - p0);
    const uint32_t* src = inter + p0 * W;
    const int words = np * W;
    int p = tid / W, k = tid % W;
    for (int w = tid; w < words; w += kIlvThreads) {
      ilv_async4(sm + p * psw + k, src + w);
      p += dp; k += dk;
      if (k >= W) { k -= W; p++; }
    }
    ilv_async_wait();
    __syncthreads();
    const int qw = np * TS / 4;             // plane words of this tile
    for (int c = 0; c < nch; c++) {
      uint32_t* dst = (uint32_t*)pl.p[c] + p0 * TS / 4;
      for (int q = tid; q < qw; q += kIlvThreads) __stcs(dst + q, ilv_tile_gather<TS>(sm, psw, c, q));
    }
    __syncthreads();
  }
}

template <int TS>
__global__ void __launch_bounds__(kIlvThreads) k_interleave_tile(uint32_t* __restrict__ inter, IlvPlanes pl, int nch, int W, int psw,
                                                                 int P, int64_t n) {
  __shared__ uint32_t sm[kIlvTileWords];
  const int64_t tiles = (n + P - 1) / P;
  const int tid = threadIdx.x, dp = kIlvThreads / W, dk = kIlvThreads % W;
  for (int64_t t = blockIdx.x; t < tiles; t += gridDim.x) {
    const int64_t p0 = t * P;
    const int np = (int)min((int64_t)P, n - p0);
    const int qw = np * TS / 4;
    if constexpr (TS >= 4) {
      for (int c = 0; c < nch; c++) {
        const uint32_t* src = (const uint32_t*)pl.p[c] + p0 * TS / 4;
        for (int q = tid; q < qw; q += kIlvThreads)
          ilv_async4(sm + (TS == 4 ? q * psw + c : (q >> 1) * psw + 2 * c + (q & 1)), src + q);
      }
    } else {                                // sub-word elements are split on the way in: four planes' loads ahead of the stores
      for (int c0 = 0; c0 < nch; c0 += 4) {
        for (int q = tid; q < qw; q += kIlvThreads) {
          uint32_t v[4];
#pragma unroll
          for (int u = 0; u < 4; u++)
            if (c0 + u < nch) v[u] = __ldcs((const uint32_t*)pl.p[c0 + u] + p0 * TS / 4 + q);
#pragma unroll
          for (int u = 0; u < 4; u++)
            if (c0 + u < nch) ilv_tile_scatter<TS>(sm, psw, c0 + u, q, v[u]);
        }
      }
    }
    if constexpr (TS >= 4) ilv_async_wait();
    __syncthreads();
    uint32_t* dst = inter + p0 * W;
    const int words = np * W;
    int p = tid / W, k = tid % W;
    for (int w = tid; w < words; w += kIlvThreads) {
      __stcs(dst + w, sm[p * psw + k]);
      p += dp; k += dk;
      if (k >= W) { k -= W; p++; }
    }
    __syncthreads();
  }
}

// Element kernel: elements [first, n) of `count` planes that are channels c0 .. c0 + count - 1 of an nch-channel pixel.
// One thread per element, ordered so that the STORES of a warp are consecutive (interleave: along the pixel,
// deinterleave: along the plane, blockIdx.y = plane).  Bytewise, so no alignment is assumed.
__global__ void __launch_bounds__(kIlvThreads) k_interleave_elems(uint8_t* __restrict__ inter, IlvPlanes pl, int ts, int nch, int c0,
                                                                  int count, int64_t first, int64_t n) {
  const int64_t total = (n - first) * count;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int64_t i = first + t / count;
    const int c = (int)(t % count);
    const uint8_t* s = (const uint8_t*)pl.p[c] + i * ts;
    uint8_t* d = inter + (i * nch + c0 + c) * ts;
    for (int b = 0; b < ts; b++) d[b] = s[b];
  }
}

__global__ void __launch_bounds__(kIlvThreads) k_deinterleave_elems(const uint8_t* __restrict__ inter, IlvPlanes pl, int ts, int nch,
                                                                    int c0, int count, int64_t first, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int c = blockIdx.y; c < count; c += gridDim.y) {
    uint8_t* plane = (uint8_t*)pl.p[c];
    for (int64_t i = first + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      const uint8_t* s = inter + (i * nch + c0 + c) * ts;
      uint8_t* d = plane + i * ts;
      for (int b = 0; b < ts; b++) d[b] = s[b];
    }
  }
}

}  // namespace b2g
