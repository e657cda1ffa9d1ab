// b2g_blosclz.cuh -- warp-cooperative blosclz decoder and encoder (rows C3/C4 of SURVEY.md 8a).
//
// Format (thirdparty/c-blosc2/blosc/blosclz.c:685-795 is the behavioural reference for the decoder, :422-619 for
// what an encoder may emit): control byte c < 32 => literal run of c+1 bytes; c >= 32 => match with
// len = (c>>5)-1 (+255-chain when that is 6) + 3, biased distance d = ((c&31)<<8) + code, source = op - d - 1;
// (c&31)==31 && code==255 => far match, next two bytes big-endian D, d = D + 8191.  The first control byte carries
// bit 5 as a codec marker and is therefore always a literal run; the stream must end with a literal run (a trailing
// match is not copied by the reference decoder, :742).
#pragma once
#include "b2g_lz4.cuh"

namespace b2g {

constexpr int kBlzMaxDistance = 8191;

// Decodes in[0..n) into out[0..cap).  Returns the decoded size, 0 on any error (as the reference does).
// INPLACE: in == out + a (see lz4_decode_warp_t); -2 = the output would pass unread input, decode from the original.
template <bool INPLACE>
__device__ __forceinline__ int blosclz_decode_warp_t(const uint8_t* in, int n, uint8_t* out, int cap, int a) {
  const int lane = lane_id();
  if (n <= 0) return 0;
  int ip = 0, op = 0;
  uint32_t ctrl = in[ip++] & 31u;
  for (;;) {
    if (ctrl >= 32) {
      int len = (int)(ctrl >> 5) - 1;
      int ofs = (int)(ctrl & 31u) << 8;
      if (len == 6) {
        // do { if (ip + 1 >= n) fail; code = in[ip++]; len += code; } while (code == 255)
        if (!lz4_varlen(in, n, ip, n - 1, len, lane)) return 0;
      } else {
        if (ip + 1 >= n) return 0;
      }
      uint32_t code = in[ip++];
      len += 3;
      int dist = ofs + (int)code;  // biased: source = op - dist - 1
      if (code == 255 && ofs == (31 << 8)) {
        if (ip + 1 >= n) return 0;
        dist = (((int)in[ip] << 8) | (int)in[ip + 1]) + kBlzMaxDistance;
        ip += 2;
      }
      if (op + len > cap) return 0;
      if (op - dist - 1 < 0) return 0;
      if (ip >= n) break;  // a trailing match is NOT copied (blosclz.c:742)
      ctrl = in[ip++];
      if (INPLACE && op + len > a + ip) return -2;
      __syncwarp();
      lz_match_copy(out, op, dist + 1, len, lane);
      op += len;
    } else {
      const int run = (int)ctrl + 1;
      if (op + run > cap) return 0;
      if (ip + run > n) return 0;
      const uint32_t nctrl = ip + run < n ? in[ip + run] : 0u;  // read before the run is written: in place, the run
      if (INPLACE) __syncwarp();  // may end right below it, and the NEXT run may start on it: every lane has read it
      lz_literal_copy<INPLACE>(out, op, in, ip, run, a + ip - op, lane);
      op += run;
      ip += run;
      if (ip >= n) break;
      ctrl = nctrl;
      ip++;
    }
  }
  __syncwarp();
  return op;
}

__device__ __forceinline__ int blosclz_decode_warp(const uint8_t* in, int n, uint8_t* out, int cap) {
  return blosclz_decode_warp_t<false>(in, n, out, cap, 0);
}

// Literal runs: `lit` bytes from in[anchor..] as chained groups of (control byte, <= 32 literals).
__device__ __forceinline__ int blosclz_put_literals(const uint8_t* in, int anchor, int lit, uint8_t* out, int op,
                                                    int lane) {
  const int groups = (lit + 31) >> 5;
  const int total = lit + groups;
  for (int k = lane; k < total; k += 32) {
    int g = k / 33, w = k - g * 33;
    uint8_t v;
    if (w == 0) {
      int left = lit - (g << 5);
      v = (uint8_t)((left > 32 ? 32 : left) - 1);
    } else {
      v = in[anchor + (g << 5) + w - 1];
    }
    out[op + k] = v;
  }
  return op + total;
}

}  // namespace b2g
