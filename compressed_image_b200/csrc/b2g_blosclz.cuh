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
__device__ __forceinline__ int blosclz_decode_warp(const uint8_t* in, int n, uint8_t* out, int cap) {
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
      __syncwarp();
      lz_match_copy(out, op, dist + 1, len, lane);
      op += len;
    } else {
      const int run = (int)ctrl + 1;
      if (op + run > cap) return 0;
      if (ip + run > n) return 0;
      if (lane < run) out[op + lane] = in[ip + lane];
      op += run;
      ip += run;
      if (ip >= n) break;
      ctrl = in[ip++];
    }
  }
  __syncwarp();
  return op;
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

// Compresses in[0..n) (shared memory) into out[0..cap).  Returns the size or 0 (does not fit / too small: the
// caller stores the stream raw).  Same search structure as lz4_encode_warp, blosclz emission.
__device__ __forceinline__ int blosclz_encode_warp(const uint8_t* in, int n, uint8_t* out, int cap, uint16_t* htab,
                                                   int hlog) {
  const int lane = lane_id();
  if (n < 16 || cap < 66) return 0;  // blosclz.c:480-482
  const int iplimit = n - 13;    // last position probed (blosclz.c ip_limit = length - 12, exclusive)
  const int matchlimit = n - 4;  // matches end at or before this; the rest is the closing literal run
  int ip = 1, anchor = 0, op = 0;
  for (int i = lane; i < (1 << hlog); i += 32) htab[i] = 0;
  __syncwarp();
  int miss = 0;
  while (ip <= iplimit) {
    const int p = ip + lane;
    const bool live = p <= iplimit;
    uint32_t seq = 0, h = 0;
    int cand = -1;
    if (live) {
      seq = ld32_any(in + p);
      h = (seq * 2654435761u) >> (32 - hlog);
      cand = p - (int)((uint32_t)(p - (int)htab[h]) & 0xffffu);
    }
    bool hit = live && cand >= 0 && cand < p && ld32_any(in + cand) == seq;
    unsigned m = __ballot_sync(kFull, hit);
    // insert only the positions up to the winning one: later positions are probed again and must still
    // find their older candidates (every lane has read the table before any lane writes it)
    if (live && (m == 0 || lane <= __ffs((int)m) - 1)) htab[h] = (uint16_t)p;
    if (m == 0) {
      miss++;
      ip += 32 + ((miss >> 3) << 5);
      __syncwarp();
      continue;
    }
    miss = 0;
    const int w = __ffs((int)m) - 1;
    int mp = ip + w;
    int mc = __shfl_sync(kFull, cand, w);
    {
      int back = 0;
      const int maxback = min(mp - anchor, mc);
      while (back < maxback) {
        int i = back + lane;
        bool eq = i < maxback && in[mp - 1 - i] == in[mc - 1 - i];
        unsigned e = __ballot_sync(kFull, !eq);
        if (e) { back += __ffs((int)e) - 1; break; }
        back += 32;
      }
      if (back > maxback) back = maxback;
      mp -= back;
      mc -= back;
    }
    const int ml = 4 + warp_match_len(in + mp + 4, in + mc + 4, matchlimit - (mp + 4), lane);
    const int db = mp - mc - 1;  // biased distance
    const bool far = db >= kBlzMaxDistance;
    if (far && ml < 8) {  // a 4-byte far token does not pay for a short match: treat the position as a literal
      ip = ip + w + 1;
      __syncwarp();
      continue;
    }
    const int lit = mp - anchor;
    const int L = ml - 2;
    {
      int need = lit + ((lit + 31) >> 5) + 1 + (L >= 7 ? (L - 7) / 255 + 1 : 0) + (far ? 3 : 1);
      if (op + need + (n - (mp + ml)) / 32 + 6 > cap) return 0;
    }
    if (lit > 0) op = blosclz_put_literals(in, anchor, lit, out, op, lane);
    const int dd = far ? db - kBlzMaxDistance : db;
    const int hi5 = far ? 31 : (db >> 8);
    if (lane == 0) out[op] = (uint8_t)(((L >= 7 ? 7 : L) << 5) | hi5);
    op++;
    if (L >= 7) op = lz4_put_varlen(out, op, L - 7, lane);
    if (lane == 0) {
      if (far) {
        out[op] = 255;
        out[op + 1] = (uint8_t)(dd >> 8);
        out[op + 2] = (uint8_t)dd;
      } else {
        out[op] = (uint8_t)db;
      }
    }
    op += far ? 3 : 1;
    ip = mp + ml;
    anchor = ip;
    if (lane < 2 && ip - 2 + lane <= iplimit && ip - 2 + lane > 0) {
      int q = ip - 2 + lane;
      uint32_t s2 = ld32_any(in + q);
      htab[(s2 * 2654435761u) >> (32 - hlog)] = (uint16_t)q;
    }
    __syncwarp();
  }
  const int lit = n - anchor;  // >= 4: the closing literal run
  if (op + lit + ((lit + 31) >> 5) > cap) return 0;
  op = blosclz_put_literals(in, anchor, lit, out, op, lane);
  __syncwarp();
  if (lane == 0) out[0] |= 32;  // codec marker (blosclz.c:613)
  __syncwarp();
  return op;
}

}  // namespace b2g
