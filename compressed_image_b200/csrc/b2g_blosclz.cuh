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
// ip0 / op0: where to take the stream up (0, 0 = from its start; the batched decoder below hands over near the end).
template <bool INPLACE>
__device__ __forceinline__ int blosclz_decode_warp_t(const uint8_t* in, int n, uint8_t* out, int cap, int a, int ip0 = 0,
                                                     int op0 = 0) {
  const int lane = lane_id();
  if (n <= 0) return 0;
  int ip = ip0, op = op0;
  if (ip >= n) return op;
  uint32_t ctrl = in[ip++];
  if (ip0 == 0) ctrl &= 31u;
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

// The same stream decoded in BATCHES of up to 32 tokens, like LZ4 (b2g_lz4.cuh): the warp follows the token chain with
// warp-uniform byte loads and leaves the position of token k in lane k; every lane decodes ITS token -- a literal run
// (lit bytes from the stream, no match) or a match (no literals) --, prefix sums give the output positions, and the
// copies run lane-parallel through lz_batch_execute.  Bit-shuffled float data is a token every few bytes: one by one
// with the whole warp per token, such a stream cost ten times its size in instructions.  The tokens whose bytes come
// within kBlzTail of the stream's end, and everything after a batch that could not run (output about to reach the
// staged input), go to the one-by-one decoder above, which owns the end-of-stream rules and every error return.
constexpr int kBlzTail = 8;

template <bool INPLACE, bool SMEM>
__device__ __forceinline__ int blosclz_decode_batched(const uint8_t* in, int n, uint8_t* out, int cap, int a) {
  const int lane = lane_id();
  if (n <= 0) return 0;
  const typename MemSel<SMEM>::type O(out);
  const typename MemSel<SMEM && INPLACE>::type I(in);
  const int nlim = n - kBlzTail;
  int ip = 0, op = 0;      // next token / output position (all tokens before ip are executed)
  int pip = 0;             // where the parser goes on
  int tp = 0, npend = 0;   // lane k: position of the k-th pending token
  bool pstop = false;
  for (;;) {
    // ---- parse ahead: nothing but the next token position ------------------------------------------------------------------
    while (!pstop && npend < 32) {
      // the usual tokens -- literal runs, matches without length extension (near or far) -- with nothing but the next
      // position on the chain: control byte and the byte after it are loaded together, both candidates for the next
      // position hang off them side by side, positions are absolute addresses of the view
      if (pip > 0) {                                            // (the first token carries the marker bit: general path)
        int base;
        const auto A = I.origin(base);
        int pa = base + pip;
        const int alim = base + nlim;
        while (npend < 32 && pa < alim) {
          const uint32_t c = A.ld8(pa), code = A.ld8(pa + 1);
          const int ql = pa + 2 + (int)c;                       // control byte, c + 1 literals
          const int qm = pa + ((code == 255u && (c & 31u) == 31u) ? 4 : 2);   // control byte, code [, 16-bit distance]
          const int q = c < 32u ? ql : qm;
          if (c >= 224u || q > alim) break;
          if (lane == npend) tp = pa - base;
          npend++;
          pa = q;
        }
        pip = pa - base;
        if (npend == 32) break;
      }
      // anything else, one token: the first one, length extensions, the end of the stream
      if (pip >= nlim) { pstop = true; break; }
      uint32_t c = I.ld8(pip);
      if (pip == 0) c &= 31u;
      int q;
      if (c < 32u) {
        q = pip + 2 + (int)c;                                   // control byte, c + 1 literals
      } else {
        q = pip + 1;
        if ((c >> 5) == 7u) {                                   // length extension bytes
          uint32_t b;
          do {
            if (q >= nlim) { pstop = true; break; }
            b = I.ld8(q++);
          } while (b == 255u);
          if (pstop) break;
        }
        const uint32_t code = I.ld8(q++);
        if (code == 255u && (c & 31u) == 31u) q += 2;           // far match: 16-bit distance follows
      }
      if (q > nlim) { pstop = true; break; }
      if (lane == npend) tp = pip;
      npend++;
      pip = q;
    }
    if (npend == 0) break;
    // ---- every lane decodes its own token ----------------------------------------------------------------------------------
    const bool have = lane < npend;
    const int tp0 = __shfl_sync(kFull, tp, 0);
    const int tq = have ? tp : tp0;
    uint32_t c = I.ld8(tq);
    if (tq == 0) c &= 31u;
    int lit = 0, ml = 0, off = 0, lpos = tq + 1, nip;
    if (c < 32u) {
      lit = (int)c + 1;
      nip = tq + 1 + lit;
    } else {
      int len = (int)(c >> 5) - 1, x = tq + 1;
      if (len == 6) {
        uint32_t b;
        do { b = I.ld8(x++); len = min(len + (int)b, 1 << 30); } while (b == 255u);
      }
      const uint32_t code = I.ld8(x++);
      int dist = (int)((c & 31u) << 8) + (int)code;
      if (code == 255u && (c & 31u) == 31u) {
        dist = (int)((I.ld8(x) << 8) | I.ld8(x + 1)) + kBlzMaxDistance;
        x += 2;
      }
      ml = len + 3;
      off = dist + 1;                                           // source = op - dist - 1 (the distance is biased)
      nip = x;
    }
    if (!have) { lit = 0; ml = 0; }
    const int len = lit + ml;
    const int incl = warp_incl_scan(len, lane);
    const int myop = op + incl - len, cpy = myop + lit, mend = cpy + ml;
    // a token that the one-by-one decoder would refuse (output beyond cap, source before the start) ends the batch there:
    // that decoder then meets it again and returns what the reference returns
    bool ok = have && mend <= cap && mend >= myop && off <= cpy;
    if (INPLACE) ok = ok && mend <= a + tp0;   // every destination below every source of the batch
    const unsigned okm = __ballot_sync(kFull, ok);
    const int cnt = okm == kFull ? 32 : __ffs((int)~okm) - 1;
    if (cnt == 0) break;
    lz_batch_execute(I, O, out, lane, cnt, lpos, myop, lit, cpy, off, ml);
    op = __shfl_sync(kFull, mend, cnt - 1);
    ip = __shfl_sync(kFull, nip, cnt - 1);
    tp = __shfl_down_sync(kFull, tp, (unsigned)cnt);
    npend -= cnt;
  }
  __syncwarp();
  return blosclz_decode_warp_t<INPLACE>(in, n, out, cap, a, ip, op);
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
