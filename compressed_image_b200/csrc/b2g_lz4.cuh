// b2g_lz4.cuh -- warp-cooperative LZ4 *block* decoder and encoder (rows C1/C2 of SURVEY.md 8a).
//
// Format: token (hi4 literal length, lo4 match length - 4; 15 => 255-run extension bytes), literals, 2-byte LE
// offset, match-length extension.  Behavioural reference: LZ4_decompress_safe / LZ4_compress_fast of the vendored
// LZ4 1.10.0 (thirdparty/c-blosc2/internal-complibs/lz4-1.10.0/lz4.c:2016-2440, 924-1330) as c-blosc2 calls them
// (blosc/blosc2.c:390-450).  One warp owns one stream: the 32 lanes parse a token together (a 32-byte register
// window over the input replaces dependent byte loads), copy literals and matches 32 bytes (or 16 bytes x 32
// lanes) at a time.  The encoder lives in b2g_lzenc.cuh.
#pragma once
#include "b2g_copy.cuh"

namespace b2g {

// 255-run length extension starting at in[pos]: adds to `len`, advances `pos`.  Fails when pos would pass `ilimit`.
__device__ __forceinline__ bool lz4_varlen(const uint8_t* in, int n, int& pos, int ilimit, int& len, int lane) {
  for (;;) {
    uint32_t b = (pos + lane < n) ? in[pos + lane] : 0u;
    unsigned m = __ballot_sync(kFull, b != 255u);
    if (m == 0) {
      len += 255 * 32;
      pos += 32;
      if (pos > ilimit || len < 0) return false;
      continue;
    }
    int k = __ffs((int)m) - 1;
    len += 255 * k + (int)__shfl_sync(kFull, b, k);
    pos += k + 1;
    return pos <= ilimit;
  }
}

// Decodes in[0..n) into out[0..cap).  Returns the decoded size (must equal cap for blosc) or -1.
// All 32 lanes must call with identical arguments; `out` may be shared or global memory.
__device__ __forceinline__ int lz4_decode_warp(const uint8_t* in, int n, uint8_t* out, int cap) {
  const int lane = lane_id();
  int ip = 0, op = 0;
  if (n <= 0 || cap <= 0) return -1;
  for (;;) {
    if (ip >= n) return -1;
    // 32-byte register window over the input: lane l holds in[ip + l]
    const uint32_t wb = (ip + lane < n) ? in[ip + lane] : 0u;
    const uint32_t token = __shfl_sync(kFull, wb, 0);
    int lit = (int)(token >> 4);
    int pos = ip + 1;
    if (lit == 15) {
      if (pos >= n - 15) return -1;
      if (!lz4_varlen(in, n, pos, n - 15, lit, lane)) return -1;
    }
    const int hdr = pos - ip;
    const int cpy = op + lit;
    if (cpy > cap - 12 || pos + lit > n - 8) {
      // must be the last sequence: literals only, consuming the input exactly (lz4.c:2279-2320)
      if (pos + lit != n || cpy > cap) return -1;
      if (hdr + lit <= 32) {
        if (lane >= hdr && lane < hdr + lit) out[op + lane - hdr] = (uint8_t)wb;
      } else {
        copy_bytes(out + op, in + pos, lit, lane, 32);
      }
      op = cpy;
      break;
    }
    uint32_t offset;
    if (hdr + lit + 2 <= 32) {
      if (lane >= hdr && lane < hdr + lit) out[op + lane - hdr] = (uint8_t)wb;
      offset = __shfl_sync(kFull, wb, hdr + lit) | (__shfl_sync(kFull, wb, hdr + lit + 1) << 8);
    } else {
      if (lit >= 64) copy_bytes(out + op, in + pos, lit, lane, 32);
      else for (int i = lane; i < lit; i += 32) out[op + i] = in[pos + i];
      offset = (uint32_t)in[pos + lit] | ((uint32_t)in[pos + lit + 1] << 8);
    }
    ip = pos + lit + 2;
    op = cpy;
    int ml = (int)(token & 15u);
    if (ml == 15) {
      if (!lz4_varlen(in, n, ip, n - 4, ml, lane)) return -1;
    }
    ml += 4;
    const int match = op - (int)offset;
    if (match < 0 || offset == 0) return -1;
    const int mend = op + ml;
    if (mend > cap - 5) return -1;  // lz4.c:2419-2423: the last 5 bytes are always literals
    __syncwarp();                   // literals (and every earlier match) visible before the match is read
    lz_match_copy(out, op, (int)offset, ml, lane);
    op = mend;
  }
  __syncwarp();
  return op;
}

// ---- helpers shared with the encoder (b2g_lzenc.cuh) ---------------------------------------------------------------

__device__ __forceinline__ uint32_t ld32_any(const uint8_t* p) {  // 4 bytes at any alignment (shared or global)
  const unsigned sh = (unsigned)((uintptr_t)p & 3) * 8;
  const uint32_t* b = (const uint32_t*)((uintptr_t)p & ~(uintptr_t)3);
  uint32_t lo = b[0];
  uint32_t hi = sh ? b[1] : 0u;
  return __funnelshift_r(lo, hi, sh);
}

// Number of equal bytes of a[0..max) and b[0..max), compared warp-wide 32 x 4 bytes per round.
__device__ __forceinline__ int warp_match_len(const uint8_t* a, const uint8_t* b, int maxlen, int lane) {
  int done = 0;
  while (done < maxlen) {
    int i = done + lane * 4;
    int rem = maxlen - i;
    unsigned neq;
    if (rem >= 4) {
      neq = ld32_any(a + i) ^ ld32_any(b + i);
    } else {
      neq = 0;
      for (int k = 0; k < 4; k++)
        if (k >= rem || a[i + k] != b[i + k]) { neq |= 0xffu << (8 * k); }
    }
    unsigned m = __ballot_sync(kFull, neq != 0);
    if (m) {
      int l = __ffs((int)m) - 1;
      unsigned x = __shfl_sync(kFull, neq, l);
      int n = done + l * 4 + ((__ffs((int)x) - 1) >> 3);
      return n < maxlen ? n : maxlen;
    }
    done += 128;
  }
  return maxlen;
}

}  // namespace b2g
