// b2g_seq.cuh -- two-phase LZ decode: a lane-per-stream PARSE that turns an LZ4 / blosclz stream into sequence
// records, and a warp-per-stream EXECUTOR that replays the records into the warp's shared-memory window.
//
// Why: decoding one stream with one warp is bound by the byte-serial token chain (token -> literal length -> position
// of the next token -> load -> ...), ~60 instructions and ~500 cycles per sequence with every lane doing the same
// scalar work.  The chain needs no data but the compressed bytes, so it is run here by ONE LANE per stream (32 streams
// per warp, k_parse), validating exactly what LZ4_decompress_safe (lz4-1.10.0/lz4.c:2016-2440) and blosclz_decompress
// (blosc/blosclz.c:685-795) validate.  What is left for the executor has no serial chain: 32 records are fetched with
// one coalesced load, output positions are a warp prefix sum, all literal runs of the group are copied first (they only
// depend on the input) and the matches are replayed in order.
//
// Record = 8 bytes {lit u16 | ml u16, off u16 | lpos u16}: `lit` literal bytes from payload[lpos..], then `ml` match
// bytes from `off` bytes back (ml == 0: no match).  Streams on this path are <= 65535 bytes, so every field fits.
#pragma once
#include "b2g_blosclz.cuh"

namespace b2g {

// What k_parse leaves for the executor, one per work item.
struct ItemRec {
  int32_t pos;     // payload offset inside the chunk (right after the stream's csize field)
  int32_t csize;   // the stream's csize field
  int32_t nseq;    // > 0: records are valid; 0: not parsed (special stream, not eligible, or any anomaly: the
                   // one-pass decoder then reproduces the reference's accept / reject decision)
  int32_t slack;   // max over sequences of (output end - input end): the in-place window needs a >= slack
};

constexpr int kSeqMaxStream = 65535;       // u16 record fields
constexpr int kSeqGroupSmem = 32 * 16;     // executor scratch per warp: 32 x uint4

__device__ __forceinline__ uint2 seq_pack(int lit, int ml, int off, int lpos) {
  return make_uint2((uint32_t)lit | ((uint32_t)ml << 16), (uint32_t)off | ((uint32_t)lpos << 16));
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

// ---- lane parsers -------------------------------------------------------------------------------------------------
// A lane reads its stream through a PIECE: kPieceBytes of the payload staged in shared memory by the whole warp with
// coalesced 16-byte loads (32 lanes reading 32 different streams straight from global memory cost one L1 wavefront per
// lane and instruction).  A step that needs bytes beyond the piece returns kStepRefill with the position it wants the
// next piece to start at; it has changed nothing and is simply run again.
constexpr int kPieceBlocks = 65;                    // 16-byte blocks per piece: >= 1025 payload bytes from any position
constexpr int kPieceBytes = kPieceBlocks * 16;
enum { kStepGo = 0, kStepRefill = 1, kStepDone = 2, kStepFail = 3 };

struct LaneParse {
  const uint8_t* in;     // payload in global memory
  uint2* rec;            // the stream's record region
  int n, cap, maxrec;    // payload bytes, stream bytes, record slots
  int ip, op, nrec, slack;
  int phase, lit, pos, mlc;   // lz4: between the literal-length half and the offset half of a sequence
  int plit, ppos;             // blosclz: pending literal run
  int want;              // position the next piece must start at
  int lo, hi;            // payload positions [lo, hi) are staged at sm[p - lo]
  const uint8_t* sm;     // this lane's piece (shared memory)
};

__device__ __forceinline__ uint32_t pc_byte(const LaneParse& L, int p) {
#ifdef B2G_SIM
  return L.sm[p - L.lo];
#else
  uint32_t v;  // the piece is shared memory: say so (through the struct the pointer would be a generic one)
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(L.sm) + (unsigned)(p - L.lo)));
  return v;
#endif
}
// 4 payload bytes from position p (any alignment) as a little-endian word; p + 4 <= L.hi.
__device__ __forceinline__ uint32_t pc_word(const LaneParse& L, int p) {
#ifdef B2G_SIM
  const uint8_t* q = L.sm + (p - L.lo);
  return (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
#else
  const unsigned a = (unsigned)__cvta_generic_to_shared(L.sm) + (unsigned)(p - L.lo);
  uint32_t w0, w1;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(a & ~3u));
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w1) : "r"((a & ~3u) + 4u));
  return __funnelshift_r(w0, w1, (a & 3u) * 8u);
#endif
}

// 255-run length extension: same accept / reject and same result as lz4_varlen (b2g_lz4.cuh), which reads bytes past
// the end of the input as 0 and checks the position against `ilimit` once the run has ended.  Returns kStepRefill when
// the run leaves the piece.
__device__ __forceinline__ int pc_varlen(const LaneParse& L, int& pos, int ilimit, int& len) {
  for (;;) {
    if (pos >= L.hi && pos < L.n) return kStepRefill;
    const uint32_t b = pos < L.n ? pc_byte(L, pos) : 0u;
    pos++;
    len += (int)b;
    if (b != 255u) break;
  }
  return pos <= ilimit ? kStepGo : kStepFail;
}

// The LZ4 token chain of one lane, as far as its piece reaches.  Mirrors lz4_decode_warp_t check for check.  The common
// sequence costs ONE shared-memory access: the 4-byte window at the offset also holds the first match-length extension
// byte or the next token and its first literal-length extension byte.  Returns kStepRefill / kStepDone / kStepFail.
__device__ __forceinline__ int lz4_parse_run(LaneParse& L) {
  const int n = L.n, cap = L.cap, hi = L.hi, maxrec = L.maxrec;
  int ip = L.ip, op = L.op, nrec = L.nrec, slack = L.slack;
  int lit = L.lit, pos = L.pos, mlc = L.mlc;
  int tokc = -1, b1 = -1;  // next token / the byte after it, when already seen (-1: not)
  int ret;
  if (L.phase == 1) goto second_half;
  for (;;) {
    if (ip >= n || nrec >= maxrec) { ret = kStepFail; break; }
    if (tokc < 0) {
      if (ip >= hi) { L.phase = 0; L.want = ip; ret = kStepRefill; break; }
      tokc = (int)pc_byte(L, ip);
      b1 = -1;
    }
    lit = tokc >> 4;
    mlc = tokc & 15;
    pos = ip + 1;
    if (lit == 15) {
      if (pos >= n - 15) { ret = kStepFail; break; }
      if (b1 >= 0 && b1 != 255) {
        lit += b1;  // pos + 1 <= n - 15 holds
        pos++;
      } else {
        const int r = pc_varlen(L, pos, n - 15, lit);
        if (r == kStepRefill) { L.phase = 0; L.want = ip; ret = r; break; }  // the token is read again
        if (r != kStepGo) { ret = r; break; }
      }
    }
    if (op + lit > cap - 12 || pos + lit > n - 8) {  // last sequence: literals only, consuming the input exactly
      if (pos + lit != n || op + lit > cap) { ret = kStepFail; break; }
      L.rec[nrec++] = seq_pack(lit, 0, 0, pos);
      op += lit;
      ret = kStepDone;
      break;
    }
  second_half:
    {
      const int opos = pos + lit;  // <= n - 8: offset, then a match-length extension or the next token
      if (opos + 4 > hi) { L.phase = 1; L.want = opos; ret = kStepRefill; break; }
      const uint32_t w = pc_word(L, opos);
      const int offset = (int)(w & 0xffffu);
      int ipn = opos + 2;
      int ml = mlc;
      if (ml == 15) {
        const int e = (int)((w >> 16) & 0xffu);
        if (e != 255) {
          ml += e;  // ipn + 1 <= n - 4 holds
          ipn++;
          tokc = (int)(w >> 24);
          b1 = -1;
        } else {
          const int r = pc_varlen(L, ipn, n - 4, ml);
          if (r == kStepRefill) { L.phase = 1; L.want = opos; ret = r; break; }
          if (r != kStepGo) { ret = r; break; }
          tokc = -1;
        }
      } else {
        tokc = (int)((w >> 16) & 0xffu);
        b1 = (int)(w >> 24);
      }
      ml += 4;
      const int cpy = op + lit;
      if (offset > cpy) { ret = kStepFail; break; }  // offset 0 is accepted (zeros), as in lz4_decode_warp_t
      const int mend = cpy + ml;
      if (mend > cap - 5) { ret = kStepFail; break; }  // lz4.c:2419-2423: the last 5 bytes are always literals
      L.rec[nrec++] = seq_pack(lit, ml, offset, pos);
      slack = max(slack, mend - ipn);
      op = mend;
      ip = ipn;
    }
  }
  L.ip = ip; L.op = op; L.nrec = nrec; L.slack = slack; L.lit = lit; L.pos = pos; L.mlc = mlc;
  return ret;
}

// One control byte of a blosclz stream (L.ip = position of the control byte; the first one is masked by the caller via
// phase == 0).  Mirrors blosclz_decode_warp_t.  A literal run and the match that follows it share one record.
__device__ __forceinline__ int blosclz_parse_step(LaneParse& L) {
  if (L.nrec + 2 > L.maxrec) return kStepFail;
  int ip = L.ip;
  L.want = ip;
  if (ip >= L.hi) return kStepRefill;
  uint32_t ctrl = pc_byte(L, ip++);
  if (L.phase == 0) ctrl &= 31u;  // the first control byte carries the codec marker in bit 5 (blosclz.c:694)
  if (ctrl >= 32) {
    int len = (int)(ctrl >> 5) - 1;
    const int ofs = (int)(ctrl & 31u) << 8;
    if (len == 6) {
      const int r = pc_varlen(L, ip, L.n - 1, len);
      if (r != kStepGo) return r;
    } else {
      if (ip + 1 >= L.n) return kStepFail;
    }
    if (min(ip + 3, L.n) > L.hi) return kStepRefill;  // code + a far distance
    const uint32_t code = pc_byte(L, ip++);
    len += 3;
    int dist = ofs + (int)code;  // biased: source = op - dist - 1
    if (code == 255 && ofs == (31 << 8)) {
      if (ip + 1 >= L.n) return kStepFail;
      dist = (((int)pc_byte(L, ip) << 8) | (int)pc_byte(L, ip + 1)) + kBlzMaxDistance;
      ip += 2;
    }
    if (L.op + len > L.cap) return kStepFail;
    if (L.op - dist - 1 < 0) return kStepFail;
    if (ip >= L.n) {  // a trailing match is NOT copied (blosclz.c:742)
      if (L.plit) L.rec[L.nrec++] = seq_pack(L.plit, 0, 0, L.ppos);
      return kStepDone;
    }
    L.rec[L.nrec++] = seq_pack(L.plit, len, dist + 1, L.ppos);
    L.plit = 0;
    L.op += len;
    L.slack = max(L.slack, L.op - (ip + 1));  // the one-pass decoder has consumed the next control byte by now
    L.ip = ip;
  } else {
    const int run = (int)ctrl + 1;
    if (L.op + run > L.cap) return kStepFail;
    if (ip + run > L.n) return kStepFail;
    if (L.plit) L.rec[L.nrec++] = seq_pack(L.plit, 0, 0, L.ppos);
    L.plit = run;
    L.ppos = ip;
    L.op += run;
    ip += run;
    if (ip >= L.n) {
      L.rec[L.nrec++] = seq_pack(L.plit, 0, 0, L.ppos);
      return kStepDone;
    }
    L.ip = ip;
  }
  L.phase = 1;
  return kStepGo;
}

// 32 streams, one per lane (lanes without a stream pass active = false): pieces are staged by the whole warp, then
// every lane steps through its piece.  `sm` = 32 x kPieceBytes of the warp's shared memory.  Returns the record count
// when the stream decodes to exactly L.cap bytes, else 0.
__device__ __forceinline__ int seq_parse_warp(LaneParse& L, bool active, bool lz4, uint8_t* sm) {
  const int lane = lane_id();
  L.ip = 0; L.op = 0; L.nrec = 0; L.slack = 0; L.phase = 0; L.plit = 0; L.ppos = 0; L.want = 0; L.lo = 0; L.hi = 0;
  L.lit = 0; L.pos = 0; L.mlc = 0;
  L.sm = sm + lane * kPieceBytes;
  int state = (active && L.n > 0 && (!lz4 || L.cap > 0)) ? kStepRefill : kStepFail;
  int stuck = -1;  // position of the last refill request: asking for the same piece twice means it can never suffice
  for (;;) {
    const unsigned need = __ballot_sync(kFull, state == kStepRefill);
    if (need == 0) break;
    // ---- stage the pieces: the warp fetches the 16-byte blocks of one stream after the other ---------------------------
    const unsigned long long gaddr = (unsigned long long)(uintptr_t)(L.in + L.want);
    const unsigned long long gend = (unsigned long long)(uintptr_t)(L.in + L.n);
    for (unsigned m = need; m; m &= m - 1) {  // asynchronous copies: every block of the round is in flight at once
      const int s = __ffs((int)m) - 1;
      const unsigned long long g = __shfl_sync(kFull, gaddr, s) & ~15ull;
      const unsigned long long e = __shfl_sync(kFull, gend, s);
#pragma unroll
      for (int b = lane; b < kPieceBlocks; b += 32)
        if (g + 16ull * b < e) async_copy16(sm + s * kPieceBytes + 16 * b, (const void*)(uintptr_t)(g + 16ull * b));
    }
    async_copy_wait();
    if (state == kStepRefill) {
      L.lo = L.want - (int)(gaddr & 15ull);
      L.hi = min(L.n, L.lo + kPieceBytes);
      if (L.want == stuck) state = kStepFail;  // one header longer than a piece: leave it to the one-pass decoder
      else { stuck = L.want; state = kStepGo; }
    }
    __syncwarp();
    if (state == kStepGo) {
      if (lz4) state = lz4_parse_run(L);
      else { do { state = blosclz_parse_step(L); } while (state == kStepGo); }
    }
    __syncwarp();
  }
  return (state == kStepDone && L.op == L.cap) ? L.nrec : 0;
}

// Replays `nseq` records into win[0..cap); the stream's payload is staged at win + a (in place: a >= slack has been
// checked by the caller, so no write ever reaches input that is still to be read).  `sq` = 32 x uint4 of the warp's
// shared memory.  Every record is bounds-checked again here (per lane, 32 at a time): the records live in global memory
// and a malformed chunk may alias two streams onto one record region.  Returns the decoded size, or -1.
__device__ __forceinline__ int seq_execute_warp(const uint2* rec, int nseq, uint8_t* win, int a, int csize, int cap,
                                                uint4* sq) {
  const int lane = lane_id();
  int op_base = 0;
  uint2 rn = make_uint2(0u, 0u);
  if (lane < nseq) rn = rec[lane];
  for (int g = 0; g < nseq; g += 32) {
    const uint2 r = rn;
    rn = make_uint2(0u, 0u);
    if (g + 32 + lane < nseq) rn = rec[g + 32 + lane];  // next group's records are in flight during this group
    const int lit = (int)(r.x & 0xffffu), ml = (int)(r.x >> 16), off = (int)(r.y & 0xffffu), lpos = (int)(r.y >> 16);
    const int tot = lit + ml;
    int incl = tot;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int v = __shfl_up_sync(kFull, incl, d);
      if (lane >= d) incl += v;
    }
    const int op = op_base + incl - tot;
    const int cpy = op + lit;
    const bool bad = (op + tot > cap) || (lpos + lit > csize) || (off > cpy);
    if (__any_sync(kFull, bad)) return -1;
    __syncwarp();  // the previous group has been read out of sq
    sq[lane] = make_uint4(r.x, (uint32_t)(a + lpos), (uint32_t)op, (uint32_t)off | ((uint32_t)cpy << 16));
    op_base = __shfl_sync(kFull, op + tot, 31);
    __syncwarp();
    const int cnt = min(32, nseq - g);
    // ---- literal runs of the whole group: independent of every match ------------------------------------------
    for (int k0 = 0; k0 < cnt; k0 += 8) {
      uint8_t v0[8], v1[8];  // a lane carries bytes `lane` and `32 + lane` of each run: runs of up to 64 bytes
      bool longrun = false;
#pragma unroll
      for (int u = 0; u < 8; u++) {
        const uint4 q = sq[k0 + u];  // lanes past the last record hold zero records
        const int L = (int)(q.x & 0xffffu);
        v0[u] = 0; v1[u] = 0;
        if (lane < L) v0[u] = win[q.y + lane];
        if (lane + 32 < L) v1[u] = win[q.y + 32 + lane];
        longrun |= L > 64;
      }
      __syncwarp();  // in place, a run's destination may overlap the source of an earlier run of this batch
      if (!longrun) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
          const uint4 q = sq[k0 + u];
          const int L = (int)(q.x & 0xffffu);
          if (lane < L) win[q.z + lane] = v0[u];
          if (lane + 32 < L) win[q.z + 32 + lane] = v1[u];
        }
      } else {
        // A run longer than 64 bytes: its tail is still unread, and in place the destinations of the LATER runs of this
        // batch may lie on it.  Copy the batch run by run instead, each a forward memmove in 32-byte steps (a run's
        // destination never reaches the source of a later run, and never passes its own source).
        for (int u = 0; u < 8; u++) {
          const uint4 q = sq[k0 + u];
          const int L = (int)(q.x & 0xffffu);
          for (int i = 0; i < L; i += 32) {
            uint8_t t = 0;
            if (i + lane < L) t = win[q.y + i + lane];
            __syncwarp();
            if (i + lane < L) win[q.z + i + lane] = t;
          }
          __syncwarp();
        }
      }
    }
    __syncwarp();
    // ---- matches, in order ------------------------------------------------------------------------------------------
    for (int k = 0; k < cnt; k++) {
      const uint4 q = sq[k];
      const int m = (int)(q.x >> 16);
      if (m == 0) continue;
      const int o = (int)(q.w & 0xffffu), c = (int)(q.w >> 16);
      if (o == 0) {  // lz4 offset 0: zeros (see lz4_decode_warp_t)
        fill_bytes(win + c, 0, m, lane, 32);
      } else if (m <= 32) {
        if (o >= m) {
          if (lane < m) win[c + lane] = win[c - o + lane];
        } else if (o == 1) {
          if (lane < m) win[c + lane] = win[c - 1];
        } else {
          if (lane < m) win[c + lane] = win[c - o + lane_mod(lane, o)];
        }
      } else {
        lz_match_copy(win, c, o, m, lane);
      }
      __syncwarp();
    }
  }
  return op_base;
}

}  // namespace b2g
