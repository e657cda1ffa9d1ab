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
#include "b2g_win.cuh"

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

// Inclusive prefix sum over the lanes of a warp.
__device__ __forceinline__ int warp_incl_scan(int v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(kFull, v, o);
    if (lane >= o) v += t;
  }
  return v;
}

// Decodes in[0..n) into out[0..cap).  Returns the decoded size (must equal cap for blosc) or -1.
// All 32 lanes must call with identical arguments; `out` may be shared or global memory.
// INPLACE: in == out + a (the input is staged at the top of the output window).  Output never passes unread input
// in a stream that compresses; when a sequence would, -2 is returned and the caller decodes from the original bytes.
//
// The stream is decoded in BATCHES of up to 32 sequences:
//   parse    the token chain is the only serial part of LZ4: where a sequence starts depends on the literal length of
//            the one before.  The warp follows it with warp-uniform byte loads (nothing but the next token position is
//            computed) and leaves the position of sequence k in lane k.  Sequences whose length fields run into a second
//            extension byte, and the last few of the stream, stop the parser; they take the general path below.
//   decode   every lane decodes ITS sequence (literal length, offset, match length); a prefix sum over the lanes gives
//            the output positions.  The batch is cut where the output would reach the end-of-stream zone or (in place)
//            the batch's own input: inside the cut every destination lies below every source of the batch.
//   literals every lane copies its own literal run (the first kLitLane bytes; the whole warp finishes longer runs).
//   matches  a match whose source lies in bytes that no match of the same batch writes (literals of the batch or anything
//            before it; found with a binary search over the lanes' output positions) is copied by its own lane, all of
//            them at once; the others (and long ones) follow in stream order, copied by the whole warp.
#ifdef B2G_SIM
struct SimDecodeStats { unsigned long long batches, batch_seqs, lane_matches, ordered_matches, general_seqs; };
inline SimDecodeStats& sim_decode_stats() { static SimDecodeStats s = {0, 0, 0, 0, 0}; return s; }  // test visibility only
#define B2G_SIM_STAT(field, v) do { if (lane == 0) sim_decode_stats().field += (unsigned long long)(v); } while (0)
#else
#define B2G_SIM_STAT(field, v) do { } while (0)
#endif

constexpr int kLitLane = 64;   // literal bytes a sequence's own lane copies
constexpr int kMatLane = 32;   // longest match a lane copies on its own

// Executes `cnt` decoded sequences at once, lane k < cnt holding sequence k: its literal run I[lpos .. lpos + lit) goes
// to O[myop ..), its match of `ml` bytes from `off` back to O[cpy ..) (cpy = myop + lit; the next sequence's myop is
// cpy + ml).  The caller has checked the bounds, that no destination of the batch reaches a literal source of the batch
// (in-place windows), and that off <= cpy.  Shared by the LZ4 decoder and the zstd sequence executor.
//   literals  every lane copies its own run (the first kLitLane bytes; the whole warp finishes longer runs).
//   matches   a match whose source lies in bytes that no match of the same batch writes (literals of the batch or
//             anything before it; found with a binary search over the lanes' output positions) is copied by its own
//             lane, all of them at once; the others (and long ones) follow in stream order, copied by the whole warp.
// off == 0 (LZ4 only: accepted by LZ4_decompress_safe, the match reads itself after lz4.c:500 / :2407 zeroed its first
// word) produces zeros.
template <class IN, class OUT>
__device__ __forceinline__ void lz_batch_execute(const IN& I, const OUT& O, uint8_t* out, int lane, int cnt, int lpos,
                                                 int myop, int lit, int cpy, int off, int ml) {
  const bool exec = lane < cnt;
  // ---- literals --------------------------------------------------------------------------------------------------------
  lane_copy(I, lpos, O, myop, exec ? min(lit, kLitLane) : 0);
  unsigned lm = __ballot_sync(kFull, exec && lit > kLitLane);
  while (lm) {
    const int j = __ffs((int)lm) - 1;
    lm &= lm - 1;
    const int sj = __shfl_sync(kFull, lpos, j) + kLitLane, dj = __shfl_sync(kFull, myop, j) + kLitLane;
    warp_copy(I, sj, O, dj, __shfl_sync(kFull, lit, j) - kLitLane, lane);
  }
  __syncwarp();
  // ---- matches ---------------------------------------------------------------------------------------------------------
  const int s = cpy - off;                 // source start; the source ends at se or runs into the match itself
  const int se = min(s + ml, cpy);
  int j = 0;                               // the sequence whose output holds byte s (0: before the batch)
#pragma unroll
  for (int b = 16; b; b >>= 1) {
    const int c = j + b;
    const int v = __shfl_sync(kFull, myop, c & 31);
    if (c < cnt && v <= s) j = c;
  }
  const int cj = __shfl_sync(kFull, cpy, j);   // [s, se) ends before that sequence's match: literals only
  const bool par = exec && ml <= kMatLane && off != 0 && se <= cj;
  // sources at least 4 bytes back: four bytes per step (loads first); closer ones: byte by byte, so that a source
  // overlapping the match itself reads what the lane has just written
  const int nm4 = par && off >= 4 ? ml : 0;
  const int max4 = (int)__reduce_max_sync(kFull, (unsigned)nm4);
  for (int i = 0; i < max4; i += 4) {
    uint32_t v0 = 0, v1 = 0, v2 = 0, v3 = 0;
    if (i < nm4) v0 = O.ld8(s + i);
    if (i + 1 < nm4) v1 = O.ld8(s + i + 1);
    if (i + 2 < nm4) v2 = O.ld8(s + i + 2);
    if (i + 3 < nm4) v3 = O.ld8(s + i + 3);
    if (i < nm4) O.st8(cpy + i, v0);
    if (i + 1 < nm4) O.st8(cpy + i + 1, v1);
    if (i + 2 < nm4) O.st8(cpy + i + 2, v2);
    if (i + 3 < nm4) O.st8(cpy + i + 3, v3);
  }
  const int nm1 = par && off < 4 ? ml : 0;
  const int max1 = (int)__reduce_max_sync(kFull, (unsigned)nm1);
  for (int i = 0; i < max1; i++)
    if (i < nm1) O.st8(cpy + i, O.ld8(s + i));
  __syncwarp();
  unsigned rest = __ballot_sync(kFull, exec && !par && ml > 0);   // (ml == 0: a blosclz literal-run token)
  B2G_SIM_STAT(batches, 1);
  B2G_SIM_STAT(batch_seqs, cnt);
  B2G_SIM_STAT(ordered_matches, __popc(rest));
  while (rest) {
    const int k = __ffs((int)rest) - 1;
    rest &= rest - 1;
    const int c = __shfl_sync(kFull, cpy, k), o = __shfl_sync(kFull, off, k), m = __shfl_sync(kFull, ml, k);
    if (o == 0) warp_fill(O, c, 0u, m, lane);
    else warp_match_copy(O, c, o, m, lane);
    __syncwarp();
  }
}

template <bool INPLACE, bool SMEM>
__device__ __forceinline__ int lz4_decode_warp_t(const uint8_t* in, int n, uint8_t* out, int cap, int a) {
  const int lane = lane_id();
  if (n <= 0 || cap <= 0) return -1;
  const typename MemSel<SMEM>::type O(out);                // the hot loops go through these views
  const typename MemSel<SMEM && INPLACE>::type I(in);
  int ip = 0, op = 0;
  int tp = 0;            // lane k: token position of the k-th parsed, not yet executed sequence
  int npend = 0;         // how many of those there are
  int pip = 0;           // where the parser goes on (== ip when nothing is pending)
  bool pstop = n < 32;   // the sequence at pip is not one the parser takes
  const int nlim = n - 9;
  for (;;) {
    // ---- parse ahead ---------------------------------------------------------------------------------------------------
    while (!pstop) {
      // the usual sequence (at most one extension byte, on the literal length): nothing but the next token position
      // (the chain token load -> next position -> token load is what a noisy stream's time is made of: positions are
      // absolute addresses of the view, both candidates for the next position hang off the loads side by side)
      {
        int base;
        const auto A = I.origin(base);
        int pa = base + pip;
        const int alim = base + nlim;
        while (npend < 32) {
          const uint32_t t = A.ld8(pa), e = A.ld8(pa + 1);
          const int q1 = pa + 3 + (int)(t >> 4);              // token, literals, offset
          const int q2 = pa + 19 + (int)e;                    // token, extension byte, 15 + e literals, offset
          const bool l15 = t >= 0xf0u;
          const int q = l15 ? q2 : q1;
          if ((t & 15u) == 15u || (l15 && e == 255u) || q > alim) break;
          if (lane == npend) tp = pa - base;
          npend++;
          pa = q;
        }
        pip = pa - base;
      }
      if (npend == 32) break;
      // anything else: lengths with longer extensions; the end-of-stream rules (lz4.c:2279-2320) belong to the general path
      const uint32_t t = I.ld8(pip);
      int x = pip + 1, lit = (int)(t >> 4);          // x: where the literals start
      if (lit == 15) {
        uint32_t b;
        do {
          if (x > nlim) { pstop = true; break; }
          b = I.ld8(x++);
          lit = min(lit + (int)b, 1 << 30);          // saturate: a length beyond any stream stops the parser below
        } while (b == 255u);
      }
      int q = x + lit + 2;
      if (q > nlim) pstop = true;
      if (!pstop && (t & 15u) == 15u) {
        uint32_t b;
        do {
          b = I.ld8(q++);
          if (q > nlim) { pstop = true; break; }
        } while (b == 255u);
      }
      if (pstop) break;
      if (lane == npend) tp = pip;
      npend++;
      pip = q;
    }
    if (npend > 0) {
      // ---- every lane decodes its own sequence ---------------------------------------------------------------------------
      const bool have = lane < npend;
      const int tp0 = __shfl_sync(kFull, tp, 0);
      const int tq = have ? tp : tp0;
      const uint32_t t = I.ld8(tq);
      int lit = (int)(t >> 4), lpos = tq + 1;
      if (lit == 15) {                           // the parser has been over these bytes: they lie inside the stream
        uint32_t b;
        do { b = I.ld8(lpos++); lit = min(lit + (int)b, 1 << 30); } while (b == 255u);
      }
      const int opos = lpos + lit;
      const int off = (int)(I.ld8(opos) | (I.ld8(opos + 1) << 8));
      int ml = (int)(t & 15u), nip = opos + 2;
      if (ml == 15) {
        uint32_t b;
        do { b = I.ld8(nip++); ml = min(ml + (int)b, (1 << 30) - 8); } while (b == 255u);
      }
      ml += 4;
      if (!have) { lit = 0; ml = 0; }
      const int len = lit + ml;
      const int incl = warp_incl_scan(len, lane);
      const int myop = op + incl - len, cpy = myop + lit, mend = cpy + ml;
      bool ok = have && cpy <= cap - 12 && mend <= cap - 5 && cpy >= myop && mend > cpy;   // lz4.c:2279, 2419-2423: the
      if (INPLACE) ok = ok && mend <= a + tp0;                // stream's end is not batch work; in place, every destination
      const unsigned okm = __ballot_sync(kFull, ok);          // stays below every source of the batch
      const int cnt = okm == kFull ? 32 : __ffs((int)~okm) - 1;
      if (cnt > 0) {
        const bool exec = lane < cnt;
        if (__any_sync(kFull, exec && off > cpy)) return -1;  // lz4.c:2356 offset beyond the start of the output
        lz_batch_execute(I, O, out, lane, cnt, lpos, myop, lit, cpy, off, ml);
        op = __shfl_sync(kFull, mend, cnt - 1);
        ip = __shfl_sync(kFull, nip, cnt - 1);
        tp = __shfl_down_sync(kFull, tp, (unsigned)cnt);
        npend -= cnt;
        continue;
      }
    }
    // ---- one general sequence -------------------------------------------------------------------------------------------
    B2G_SIM_STAT(general_seqs, 1);
    if (ip >= n) return -1;
    const uint32_t tok = in[ip];
    int lit = (int)(tok >> 4);
    int pos = ip + 1;
    if (lit == 15) {
      if (pos >= n - 15) return -1;
      if (!lz4_varlen(in, n, pos, n - 15, lit, lane)) return -1;
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
    lz_literal_copy<INPLACE>(out, op, in, pos, lit, delta, lane);
    int ml = (int)(tok & 15u);
    if (ml == 15 && !lz4_varlen(in, n, ipn, n - 4, ml, lane)) return -1;
    ml += 4;
    if ((int)offset > cpy) return -1;
    const int mend = cpy + ml;
    if (mend > cap - 5) return -1;  // lz4.c:2419-2423: the last 5 bytes are always literals
    if (INPLACE && mend > a + ipn) return -2;
    __syncwarp();                   // literals (and every earlier match) visible before the match is read
    if (offset == 0) warp_fill(O, cpy, 0u, ml, lane);
    else warp_match_copy(O, cpy, (int)offset, ml, lane);
    __syncwarp();
    op = mend;
    ip = ipn;
    // the parser's view: the sequence it had first is done (it parsed the same bytes), or it starts again from here
    if (npend > 0) {
      tp = __shfl_down_sync(kFull, tp, 1u);
      npend--;
      if (npend > 0 && __shfl_sync(kFull, tp, 0) != ip) npend = 0;
    }
    if (npend == 0) { pip = ip; pstop = ip > n - 16; }
  }
  __syncwarp();
  return op;
}

__device__ __forceinline__ int lz4_decode_warp(const uint8_t* in, int n, uint8_t* out, int cap) {
  return lz4_decode_warp_t<false, false>(in, n, out, cap, 0);
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
