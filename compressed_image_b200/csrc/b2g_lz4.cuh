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

// Literal run of a sequence: out[op .. op+lit) = in[pos .. pos+lit).  When the stream is decoded in place (the
// compressed bytes sit further up in the very window that is being written, see k_decode) the two ranges may overlap
// with the destination below the source; then every 32-byte step loads before it stores.
template <bool INPLACE>
__device__ __forceinline__ void lz_literal_copy(uint8_t* out, int op, const uint8_t* in, int pos, int lit, int delta,
                                                int lane) {
  if (!INPLACE || delta >= lit) {
    if (lit <= 32) { if (lane < lit) out[op + lane] = in[pos + lane]; }
    else if (lit <= 64) {  // both loads in flight before the stores
      const uint8_t v0 = in[pos + lane];
      uint8_t v1 = 0;
      if (lane + 32 < lit) v1 = in[pos + 32 + lane];
      out[op + lane] = v0;
      if (lane + 32 < lit) out[op + 32 + lane] = v1;
    }
    else if (lit < 128) for (int i = lane; i < lit; i += 32) out[op + i] = in[pos + i];
    else copy_bytes(out + op, in + pos, lit, lane, 32);
  } else {
    for (int i = 0; i < lit; i += 32) {
      uint8_t v = 0;
      if (i + lane < lit) v = in[pos + i + lane];
      __syncwarp();
      if (i + lane < lit) out[op + i + lane] = v;
    }
  }
}

// Decodes in[0..n) into out[0..cap).  Returns the decoded size (must equal cap for blosc) or -1.
// All 32 lanes must call with identical arguments; `out` may be shared or global memory.
// INPLACE: in == out + a (the input is staged at the top of the output window).  Output never passes unread input
// in a stream that compresses; when a sequence would, -2 is returned and the caller decodes from the original bytes.
// Header bytes are fetched with warp-uniform loads (one broadcast each), and the next token is fetched before the
// copies of the current sequence are issued, so its latency overlaps them.
constexpr int kLz4Zone = 600;

template <bool INPLACE>
__device__ __forceinline__ int lz4_decode_warp_t(const uint8_t* in, int n, uint8_t* out, int cap, int a) {
  const int lane = lane_id();
  if (n <= 0 || cap <= 0) return -1;
  int ip = 0, op = 0;
  uint32_t tok = in[0];
  uint32_t nxt = n > 1 ? in[1] : 0u;
  for (;;) {
    // ---- the zone: far from both ends of the stream ------------------------------------------------------------------
    // A sequence whose length fields take at most one extension byte each produces <= 269 + 273 bytes from <= 276, so
    // kLz4Zone bytes away from the end of the input and of the output none of the end-of-stream rules (lz4.c:2279-2320,
    // 2419-2423) and none of the input bounds can fire: they are not evaluated.  Anything else (a 255 extension byte,
    // the zone's edge) drops to the general sequence below, which starts again from the same token.
    while (op <= cap - kLz4Zone && ip <= n - kLz4Zone) {
      int lit = (int)(tok >> 4);
      int pos = ip + 1;
      if (lit == 15) {
        if (nxt == 255u) break;
        lit += (int)nxt;
        pos++;
      }
      const int opos = pos + lit;
      const uint32_t offset = (uint32_t)in[opos] | ((uint32_t)in[opos + 1] << 8);
      int ipn = opos + 2;
      uint32_t t0 = in[ipn], t1 = in[ipn + 1];
      int ml = (int)(tok & 15u);
      if (ml == 15 && t0 == 255u) break;  // nothing has been written yet
      const int cpy = op + lit;
      lz_literal_copy<INPLACE>(out, op, in, pos, lit, a + pos - op, lane);
      if (ml == 15) {
        ml += (int)t0;
        ipn++;
        t0 = t1;
        t1 = in[ipn + 1];
      }
      ml += 4;
      if ((int)offset > cpy) return -1;
      const int mend = cpy + ml;
      if (INPLACE && mend > a + ipn) return -2;
      __syncwarp();
      if (offset == 0) fill_bytes(out + cpy, 0, ml, lane, 32);
      else if (ml <= 32 && (int)offset >= ml) { if (lane < ml) out[cpy + lane] = out[cpy - (int)offset + lane]; }
      else lz_match_copy(out, cpy, (int)offset, ml, lane);
      op = mend;
      ip = ipn;
      tok = t0;
      nxt = t1;
    }
    // ---- one general sequence ---------------------------------------------------------------------------------------
    if (ip >= n) return -1;
    int lit = (int)(tok >> 4);
    int pos = ip + 1;
    if (lit == 15) {
      if (pos >= n - 15) return -1;
      if (nxt != 255u) { lit += (int)nxt; pos++; }
      else if (!lz4_varlen(in, n, pos, n - 15, lit, lane)) return -1;
    }
    const int cpy = op + lit;
    const int delta = a + pos - op;
    if (cpy > cap - 12 || pos + lit > n - 8) {
      // must be the last sequence: literals only, consuming the input exactly (lz4.c:2279-2320)
      if (pos + lit != n || cpy > cap) return -1;
      lz_literal_copy<INPLACE>(out, op, in, pos, lit, delta, lane);
      op = cpy;
      break;
    }
    const int opos = pos + lit;
    const uint32_t offset = (uint32_t)in[opos] | ((uint32_t)in[opos + 1] << 8);
    int ipn = opos + 2;                       // <= n - 6
    uint32_t t0 = in[ipn], t1 = in[ipn + 1];  // next token (or the match-length extension)
    lz_literal_copy<INPLACE>(out, op, in, pos, lit, delta, lane);
    int ml = (int)(tok & 15u);
    if (ml == 15) {
      if (t0 != 255u) { ml += (int)t0; ipn++; }
      else if (!lz4_varlen(in, n, ipn, n - 4, ml, lane)) return -1;
      t0 = ipn < n ? in[ipn] : 0u;
      t1 = ipn + 1 < n ? in[ipn + 1] : 0u;
    }
    ml += 4;
    if ((int)offset > cpy) return -1;
    const int mend = cpy + ml;
    if (mend > cap - 5) return -1;  // lz4.c:2419-2423: the last 5 bytes are always literals
    if (INPLACE && mend > a + ipn) return -2;
    __syncwarp();                   // literals (and every earlier match) visible before the match is read
    if (offset == 0) fill_bytes(out + cpy, 0, ml, lane, 32);  // accepted by LZ4_decompress_safe: the match reads itself
                                                              // after lz4.c:500 / :2407 zeroed its first word -> zeros
    else if (ml <= 32 && (int)offset >= ml) { if (lane < ml) out[cpy + lane] = out[cpy - (int)offset + lane]; }
    else lz_match_copy(out, cpy, (int)offset, ml, lane);
    op = mend;
    ip = ipn;
    tok = t0;
    nxt = t1;
  }
  __syncwarp();
  return op;
}

__device__ __forceinline__ int lz4_decode_warp(const uint8_t* in, int n, uint8_t* out, int cap) {
  return lz4_decode_warp_t<false>(in, n, out, cap, 0);
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
