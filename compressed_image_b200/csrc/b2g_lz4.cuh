// b2g_lz4.cuh -- warp-cooperative LZ4 *block* decoder and encoder (rows C1/C2 of SURVEY.md 8a).
//
// Format: token (hi4 literal length, lo4 match length - 4; 15 => 255-run extension bytes), literals, 2-byte LE
// offset, match-length extension.  Behavioural reference: LZ4_decompress_safe / LZ4_compress_fast of the vendored
// LZ4 1.10.0 (thirdparty/c-blosc2/internal-complibs/lz4-1.10.0/lz4.c:2016-2440, 924-1330) as c-blosc2 calls them
// (blosc/blosc2.c:390-450).  One warp owns one stream: the 32 lanes parse a token together (a 32-byte register
// window over the input replaces dependent byte loads), copy literals and matches 32 bytes (or 16 bytes x 32
// lanes) at a time, and - in the encoder - probe 32 consecutive positions of the hash table per step.
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

// ---- encoder ----------------------------------------------------------------------------------------------------
// Greedy LZ4 with a 16-bit position hash table in shared memory.  Per step the 32 lanes hash the 32 positions
// ip..ip+31, look their candidates up, and the first lane whose candidate really matches wins; the pending
// literals and the match are then emitted cooperatively.  The table stores positions modulo 65536, so a
// reconstructed candidate is always within LZ4's 65535 offset limit.  Output is any valid LZ4 block that obeys the
// end-of-block rules the reference decoder enforces (last 5 bytes literal, last match starts >= 12 bytes before
// the end: lz4.c:245-249, 2279, 2423); match choices differ from the CPU encoder, the format does not.

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

// Emits `len - 15`-style 255-run extension bytes for a length field (value already reduced by 15).
__device__ __forceinline__ int lz4_put_varlen(uint8_t* out, int op, int v, int lane) {
  int n255 = v / 255;
  for (int i = lane; i < n255; i += 32) out[op + i] = 255;
  if (lane == 0) out[op + n255] = (uint8_t)(v - n255 * 255);
  return op + n255 + 1;
}

// Compresses in[0..n) (shared memory) into out[0..cap) (global or shared).  Returns the compressed size, or 0 when
// it does not fit in cap (the caller then stores the stream raw, as blosc_c does: blosc2.c:1288-1324).
// htab: (1 << hlog) uint16 entries of shared memory owned by this warp.
__device__ __forceinline__ int lz4_encode_warp(const uint8_t* in, int n, uint8_t* out, int cap, uint16_t* htab,
                                               int hlog) {
  const int lane = lane_id();
  const int mflimit = n - 12;     // last position where a match may start
  const int matchlimit = n - 5;   // matches end at or before this
  int ip = 1, anchor = 0, op = 0;
  for (int i = lane; i < (1 << hlog); i += 32) htab[i] = 0;
  __syncwarp();
  if (n >= 13) {
    int miss = 0;
    while (ip <= mflimit) {
      const int p = ip + lane;
      const bool live = p <= mflimit;
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
        // acceleration as in LZ4's skip strength: after repeated misses advance faster through noise
        miss++;
        ip += 32 + ((miss >> 3) << 5);
        __syncwarp();
        continue;
      }
      miss = 0;
      const int w = __ffs((int)m) - 1;
      int mp = ip + w;
      int mc = __shfl_sync(kFull, cand, w);
      // extend backwards over pending literals (lz4.c:1113-1117)
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
      const int lit = mp - anchor;
      // worst-case size of this sequence plus the mandatory tail; bail out if it cannot fit (limited output)
      {
        int need = 1 + (lit >= 15 ? (lit - 15) / 255 + 1 : 0) + lit + 2 + (ml - 4 >= 15 ? (ml - 19) / 255 + 1 : 0);
        if (op + need + (1 + 5) > cap) return 0;
      }
      const int token_pos = op++;
      if (lit >= 15) op = lz4_put_varlen(out, op, lit - 15, lane);
      if (lit >= 64) copy_bytes(out + op, in + anchor, lit, lane, 32);
      else for (int i = lane; i < lit; i += 32) out[op + i] = in[anchor + i];
      op += lit;
      const int off = mp - mc;
      if (lane == 0) {
        out[token_pos] = (uint8_t)((lit >= 15 ? 15 : lit) << 4 | (ml - 4 >= 15 ? 15 : ml - 4));
        out[op] = (uint8_t)off;
        out[op + 1] = (uint8_t)(off >> 8);
      }
      op += 2;
      if (ml - 4 >= 15) op = lz4_put_varlen(out, op, ml - 19, lane);
      ip = mp + ml;
      anchor = ip;
      // seed the table with two positions inside the match tail so the next step can chain (lz4.c:1237-1243)
      if (lane < 2 && ip - 2 + lane <= mflimit && ip - 2 + lane > 0) {
        int q = ip - 2 + lane;
        uint32_t s2 = ld32_any(in + q);
        htab[(s2 * 2654435761u) >> (32 - hlog)] = (uint16_t)q;
      }
      __syncwarp();
    }
  }
  // last literals
  const int lit = n - anchor;
  {
    int need = 1 + (lit >= 15 ? (lit - 15) / 255 + 1 : 0) + lit;
    if (op + need > cap) return 0;
  }
  if (lane == 0) out[op] = (uint8_t)((lit >= 15 ? 15 : lit) << 4);
  op++;
  if (lit >= 15) op = lz4_put_varlen(out, op, lit - 15, lane);
  if (lit >= 64) copy_bytes(out + op, in + anchor, lit, lane, 32);
  else for (int i = lane; i < lit; i += 32) out[op + i] = in[anchor + i];
  op += lit;
  __syncwarp();
  return op;
}

}  // namespace b2g
