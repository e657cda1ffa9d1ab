// b2g_lzenc.cuh -- tile-phased warp encoder for LZ4 block streams and blosclz streams (rows C1/C3 of SURVEY.md 8a).
//
// One warp owns one stream (a byte plane of a shuffled block, resident in shared memory).  The stream is walked in
// tiles of 512 or 1024 positions, each tile in three phases, so that the only byte-serial part of LZ77 -- the greedy
// parse -- touches nothing but a few registers and broadcast shared-memory reads:
//
//   1. candidates  (position-parallel)  32 positions per step: hash the 4 bytes at p, look the most recent earlier
//                  position with that hash up (__match_any_sync resolves the 32 positions of the step among
//                  themselves, the shared-memory table everything older), verify the 4 bytes, and record the match
//                  distance in off[] plus one flag bit per position (the flag word of group g lives in lane g).
//   2. parse       (serial, warp-uniform)  jump from flag bit to flag bit with ffs over the flag words, extend the
//                  match 128 bytes per step with the whole warp, append (start, length, distance) to a sequence list.
//   3. emit        (sequence-parallel)  one lane per sequence: encoded sizes -> warp prefix sum -> every lane writes its
//                  own token / length / offset bytes and short literal runs; long literal runs are copied warp-wide.
//
// The output is any valid stream of the codec's format that obeys what the reference decoders enforce
// (LZ4: last 5 bytes literal, last match starts >= 12 bytes before the end, lz4.c:245-249, 2279, 2423;
// blosclz: first token a literal run with the marker bit, stream ends with a literal run, blosclz.c:613, 694, 742).
// Match choices differ from the CPU encoders (every position is a candidate, most recent occurrence wins), the
// formats do not.  Incompressible data is abandoned gradually like LZ4's skip strength: a tile without matches
// halves the density at which the next tile is probed.
#pragma once
#include "b2g_blosclz.cuh"

namespace b2g {

// Positions per tile = groups of 32 (the flag word of group g lives in lane g).  Streams above 8 KiB take tiles of 512:
// their window is what limits the warps per SM, and the smaller work area lets a tenth warp in; smaller streams fit
// sixteen warps either way and run faster with fewer, larger tiles.
constexpr int kEncTileMax = 1024;
__host__ __device__ __forceinline__ int enc_tile_for(int max_stream) { return max_stream > 8192 ? 512 : 1024; }
// per-warp work area after the hash table: off[tile] u16, sp[tile / 4] u32, sml[tile / 4] u32, so[tile / 4] u16
// (every sequence consumes >= 4 positions of its tile)
__host__ __device__ __forceinline__ int enc_work_bytes(int tile) { return tile * 2 + (tile / 4) * 4 * 2 + (tile / 4) * 2; }

#ifdef B2G_SIM
__device__ __forceinline__ unsigned match_any_u32(unsigned v) { return cusim_match_any(v); }
#else
__device__ __forceinline__ unsigned match_any_u32(unsigned v) { return __match_any_sync(kFull, v); }
#endif

struct EncWork {
  uint16_t* htab;   // [1 << hlog]
  uint16_t* off;    // [tile] match distance of a flagged position of the current tile
  uint32_t* sp;     // [tile / 4] match start
  uint32_t* sml;    // [tile / 4] match length
  uint16_t* so;     // [tile / 4] match distance
  int hlog, tile;
  int minml;        // shortest match worth a sequence (4 = every match the format allows)
};

__device__ __forceinline__ EncWork enc_work(uint8_t* base, int htab_bytes, int hlog, int tile, int minml = 4) {
  EncWork w;
  w.htab = (uint16_t*)base;
  w.off = (uint16_t*)(base + htab_bytes);
  w.sp = (uint32_t*)(base + htab_bytes + tile * 2);
  w.sml = w.sp + tile / 4;
  w.so = (uint16_t*)(w.sml + tile / 4);
  w.hlog = hlog;
  w.tile = tile;
  w.minml = minml;
  return w;
}

// bytes a literal-length / match-length field adds beyond the token (LZ4 255-run extension)
__device__ __forceinline__ int lz4_ext_bytes(int v) { return v >= 15 ? (v - 15) / 255 + 1 : 0; }
__device__ __forceinline__ int lz4_put_ext_lane(uint8_t* out, int op, int v) {  // one lane writes the whole field
  v -= 15;
  while (v >= 255) { out[op++] = 255; v -= 255; }
  out[op++] = (uint8_t)v;
  return op;
}

// The encoder reads its input through a shared-space view (MemShared: ld.shared with 32-bit addresses; through a plain
// generic pointer every access is a 64-bit generic load -- LD, not LDS -- with the address arithmetic to match).
using EncIn = MemSel<true>::type;
// 4 bytes at any position (the view's byte 0 is word aligned; reads up to 4 bytes beyond: the window's slack)
__device__ __forceinline__ uint32_t enc_ld32(const EncIn& I, int i) {
  const unsigned sh = (unsigned)(i & 3) * 8u;
  const int a = i & ~3;
  return __funnelshift_r(I.ld32(a), I.ld32(a + 4), sh);
}
// Number of equal bytes of I[a..a+maxlen) and I[b..b+maxlen), compared warp-wide 32 x 4 bytes per round.
__device__ __forceinline__ int enc_match_len(const EncIn& I, int a, int b, int maxlen, int lane) {
  int done = 0;
  while (done < maxlen) {
    const int i = done + lane * 4;
    const int rem = maxlen - i;
    unsigned neq = enc_ld32(I, a + i) ^ enc_ld32(I, b + i);
    if (rem < 4) neq = rem <= 0 ? 0xffffffffu : (neq | (0xffffffffu << (8 * rem)));
    const unsigned m = __ballot_sync(kFull, neq != 0);
    if (m) {
      const int l = __ffs((int)m) - 1;
      const unsigned x = __shfl_sync(kFull, neq, l);
      const int n = done + l * 4 + ((__ffs((int)x) - 1) >> 3);
      return n < maxlen ? n : maxlen;
    }
    done += 128;
  }
  return maxlen;
}
// blosclz literal runs (blosclz_put_literals) with the input behind a view
__device__ __forceinline__ int enc_blz_literals(const EncIn& I, int anchor, int lit, uint8_t* out, int op, int lane) {
  const int groups = (lit + 31) >> 5;
  const int total = lit + groups;
  for (int k = lane; k < total; k += 32) {
    const int g = k / 33, w = k - g * 33;
    uint32_t v;
    if (w == 0) {
      const int left = lit - (g << 5);
      v = (uint32_t)((left > 32 ? 32 : left) - 1);
    } else {
      v = I.ld8(anchor + (g << 5) + w - 1);
    }
    out[op + k] = (uint8_t)v;
  }
  return op + total;
}

constexpr int kEncSelfLit = 16;  // literal runs up to this length are copied by the sequence's own lane

// What one warp of a TEAM reports about its segment of a stream (k_compress_team: several warps encode consecutive
// segments of one stream side by side).  A segment's output is a BODY: its sequences without the literal run in front
// of the first match (LZ4: without that sequence's token either, the body starts with its offset field), because where
// that run starts is the end of the previous segment's last match.  lz_join_team() writes those pieces.
struct EncSeg {
  int32_t p0;     // start of the segment's first match, -1 = the segment has no sequence
  int32_t ml0;    // length of that match (LZ4: goes into the token the join writes)
  int32_t end;    // end of the segment's last match
  int32_t bytes;  // body bytes, -1 = the body did not fit its slot
};

// Compresses in[0..n) (shared memory) into out[0..cap) (global or shared).  Returns the compressed size, or 0 when
// the result does not fit in cap / the stream is too small for the codec (the caller then stores the stream raw, as
// blosc_c does: blosc2.c:1288-1324).  All 32 lanes call with identical arguments.
//
// SEG = the warp encodes the SEGMENT [seg_begin, seg_end) of the stream (seg_begin a multiple of the tile; matches start
// and end inside the segment, their sources lie anywhere before them) and writes a body (EncSeg) instead of a stream:
// the hash table is primed with the `prime` positions in front of the segment (a multiple of 128), the first sequence
// loses its literal run, there is no closing run.  Returns the body bytes, -1 when they do not fit in cap.
template <int CODEC, int TILE, bool SEG = false>
__device__ __forceinline__ int lz_encode_warp(const uint8_t* in, int n, uint8_t* out, int cap, const EncWork& w,
                                              int seg_begin = 0, int seg_end = 0, int prime = 0, EncSeg* seg = nullptr,
                                              int prime_step = 1) {
  constexpr bool LZ4 = CODEC == kCodecLz4;
  const int lane = lane_id();
  const EncIn I(in);
  const MemGeneric O(out);
  int plimit, mlimit;  // last position a match may start at; matches end at or before mlimit
  if (LZ4) {
    plimit = n - 12;   // lz4.c:245-249 (MFLIMIT 12, LASTLITERALS 5)
    mlimit = n - 5;
  } else {
    if (!SEG && (n < 16 || cap < 66)) return 0;  // blosclz.c:480-482 (a team checks the whole stream)
    plimit = n - 13;
    mlimit = n - 4;
  }
  if (SEG) {
    plimit = min(plimit, seg_end - 4);
    mlimit = min(mlimit, seg_end);
    if (lane == 0) seg->p0 = -1;
  }
  for (int i = lane; i < (1 << w.hlog); i += 32) w.htab[i] = 0;
  __syncwarp();
  if (SEG && prime > 0) {  // positions in front of the segment: inserted, never probed (whichever lane wins a slot is a
                           // valid candidate).  prime_step 4: every fourth position only -- a quarter of the work; a match
                           // whose source starts between two of them is found up to three bytes late and grown backwards
    if (prime_step == 4) {
      for (int pb = max(0, seg_begin - prime); pb < seg_begin; pb += 128) {
        const uint32_t s4 = I.ld32(pb + 4 * lane);
        w.htab[(s4 * 2654435761u) >> (32 - w.hlog)] = (uint16_t)(pb + 4 * lane);
        __syncwarp();
      }
    } else {
      for (int pb = max(0, seg_begin - prime); pb < seg_begin; pb += 128) {
        const int wb = pb + 4 * (lane >> 2);
        const unsigned sh = (unsigned)(lane & 3) * 8u;
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const uint32_t s4 = __funnelshift_r(I.ld32(wb + 32 * j), I.ld32(wb + 32 * j + 4), sh);
          w.htab[(s4 * 2654435761u) >> (32 - w.hlog)] = (uint16_t)(pb + 32 * j + lane);
          __syncwarp();  // later groups overwrite earlier ones
        }
      }
    }
  }
  int anchor = SEG ? seg_begin : 0, ip = anchor, op = 0, step = 1;
  bool head_pending = SEG;  // the segment's first sequence is still to come
  const unsigned lt_mask = (1u << lane) - 1u;
  constexpr int kEncTile = TILE, kEncGroups = TILE / 32;   // == w.tile (the work area is laid out for it)
  for (int tbase = SEG ? seg_begin : 0; tbase <= plimit; tbase += kEncTile) {
    if (ip >= tbase + kEncTile) continue;  // the tile lies inside an earlier match
    // ---- phase 1: candidates ----------------------------------------------------------------------------------
    // Four groups of 32 positions per step, as four independent dependency chains (the step is latency-bound, not
    // issue-bound).  Within a step every position sees the table as it was before the step plus the earlier positions
    // of its own group; a first occurrence 32..127 bytes back inside the same step is the only candidate missed.
    // The equal-hash resolution (__match_any_sync) has a long latency and depends only on the input bytes, not on the
    // table: it is issued one step ahead, so that it completes under the previous step's verification work.
    unsigned myflags = 0;
    auto next_step = [&](int gb) {  // first step at or after gb that has to be probed, kEncGroups = none
      for (; gb < kEncGroups; gb += 4) {
        const int bbase = tbase + gb * 32;
        if (bbase > plimit) return kEncGroups;
        if (bbase + 127 < ip) continue;
        if (step > 1 && ((gb >> 2) & (step - 1))) continue;
        return gb;
      }
      return kEncGroups;
    };
    uint32_t seq[4], h[4], nseqv[4], nh[4];
    unsigned peers[4], npeers[4];
    bool live[4], nlive[4];
    auto fetch = [&](int gb, uint32_t* fs, uint32_t* fh, unsigned* fp, bool* fl) {
      const int bbase = tbase + gb * 32;
      // the 4 bytes at bbase + 32 j + lane: `in` is word aligned and bbase a multiple of 32, so every lane's two words and
      // its shift are known without address arithmetic (positions beyond plimit read the window's slack, masked by fl)
      const int wb = bbase + 4 * (lane >> 2);
      const unsigned sh = (unsigned)(lane & 3) * 8u;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        fl[j] = bbase + 32 * j + lane <= plimit;
        fs[j] = __funnelshift_r(I.ld32(wb + 32 * j), I.ld32(wb + 32 * j + 4), sh);
        fh[j] = (fs[j] * 2654435761u) >> (32 - w.hlog);
      }
#pragma unroll
      for (int j = 0; j < 4; j++) fp[j] = match_any_u32(fl[j] ? fh[j] : (0x80000000u | (uint32_t)lane));
    };
    int gb = next_step(0);
    if (gb < kEncGroups) fetch(gb, nseqv, nh, npeers, nlive);
    while (gb < kEncGroups) {
      const int bbase = tbase + gb * 32;
#pragma unroll
      for (int j = 0; j < 4; j++) { seq[j] = nseqv[j]; h[j] = nh[j]; peers[j] = npeers[j]; live[j] = nlive[j]; }
      const int gbn = next_step(gb + 4);
      if (gbn < kEncGroups) fetch(gbn, nseqv, nh, npeers, nlive);
      uint32_t tab[4];
#pragma unroll
      for (int j = 0; j < 4; j++) tab[j] = w.htab[h[j]];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int p = bbase + 32 * j + lane;
        const unsigned prev = peers[j] & lt_mask;
        const int src = prev ? 31 - __clz((int)prev) : lane;
        const uint32_t pseq = __shfl_sync(kFull, seq[j], src);
        int c = p - (int)((uint32_t)(p - (int)tab[j]) & 0xffffu);   // table candidate, within 65535 of p
        const bool tab_ok = c >= 0 && c < p;
        const uint32_t cseq = enc_ld32(I, tab_ok ? c : 0);
        bool ok;
        if (prev) {  // the most recent earlier position with this hash is in this very group
          c = bbase + 32 * j + src;
          ok = live[j] && pseq == seq[j];
        } else {
          ok = live[j] && tab_ok && cseq == seq[j];
        }
        w.off[(gb + j) * 32 + lane] = ok ? (uint16_t)(p - c) : (uint16_t)0;
        const unsigned fl = __ballot_sync(kFull, ok);
        if (lane == gb + j) myflags = fl;
      }
      __syncwarp();  // every table read of the step happens before any table write
#pragma unroll
      for (int j = 0; j < 4; j++) {
        if (live[j] && (peers[j] >> lane) == 1u) w.htab[h[j]] = (uint16_t)(bbase + 32 * j + lane);  // last of its hash group
        __syncwarp();  // group j + 1 overwrites group j where their hashes meet (most recent wins), and every write
      }                // is done before the next step's reads
      gb = gbn;
    }
    __syncwarp();
    // ---- phase 2: greedy parse ------------------------------------------------------------------------------------
    int nseq = 0;
    int lastend = anchor;
    const unsigned gmask = __ballot_sync(kFull, myflags != 0);
    const int tend = min(tbase + kEncTile, plimit + 1);
    if (ip < tbase) ip = tbase;
    while (ip < tend) {
      const int rel = ip - tbase;
      int g = rel >> 5;
      unsigned wd = __shfl_sync(kFull, myflags, g) & (~0u << (rel & 31));
      if (!wd) {
        const unsigned gm = g < 31 ? (gmask & (~0u << (g + 1))) : 0u;
        if (!gm) break;
        g = __ffs((int)gm) - 1;
        wd = __shfl_sync(kFull, myflags, g);
      }
      const int b = __ffs((int)wd) - 1;
      int p = tbase + g * 32 + b;
      const int dist = (int)w.off[g * 32 + b];
      int c = p - dist;
      // the bytes in front of the match and of its source, for the backward extension below: fetched here so that their
      // latency overlaps the forward compare
      const bool can_back = p > lastend && c > 0;
      const uint32_t pb = I.ld8(can_back ? p - 1 : 0), cb = I.ld8(can_back ? c - 1 : 0);
      // Match length beyond the verified 4 bytes.  Most matches of a noisy plane end within the next 4 bytes: one
      // warp-uniform word compare settles those, the warp-wide compare only runs for the longer ones (same result).
      int ml;
      {
        const int room = mlimit - (p + 4);
        const uint32_t x = enc_ld32(I, p + 4) ^ enc_ld32(I, c + 4);  // p + 7 <= n - 5: inside the stream
        if (x != 0u || room <= 4) {
          const int e = x ? ((__ffs((int)x) - 1) >> 3) : 4;
          ml = 4 + max(0, min(e, room));
        } else {
          ml = 8 + enc_match_len(I, p + 8, c + 8, room - 4, lane);
        }
      }
      if (!LZ4 && dist - 1 >= kBlzMaxDistance && ml < 8) {  // a 4-byte far token does not pay for a short match
        ip = p + 1;
        continue;
      }
      // the faster levels leave the matches that save a byte or less in the literals (a 4-byte LZ4 match costs a token and
      // an offset: noisy byte planes are full of them, and every one is a sequence for the decoder)
      if (ml < w.minml && !(can_back && pb == cb)) {
        ip = p + 1;
        continue;
      }
      if (can_back && pb == cb) {  // extend over pending literals
        p--; c--; ml++;
        while (p > lastend && c > 0 && I.ld8(p - 1) == I.ld8(c - 1)) { p--; c--; ml++; }
      }
      if (lane == 0) { w.sp[nseq] = (uint32_t)p; w.sml[nseq] = (uint32_t)ml; w.so[nseq] = (uint16_t)dist; }
      nseq++;
      lastend = ip = p + ml;
    }
    step = nseq ? 1 : min(step * 2, 8);
    __syncwarp();
    // ---- phase 3: emit the tile's sequences, 32 at a time -----------------------------------------------------------
    for (int sb = 0; sb < nseq; sb += 32) {
      const int i = sb + lane;
      const bool have = i < nseq;
      int p = 0, ml = 0, dist = 0, start = 0;
      if (have) {
        p = (int)w.sp[i];
        ml = (int)w.sml[i];
        dist = (int)w.so[i];
        start = i == 0 ? anchor : (int)(w.sp[i - 1] + w.sml[i - 1]);
      }
      const bool head = SEG && head_pending && i == 0;       // its literal run (LZ4: and token) belong to the join
      const int lit = head ? 0 : p - start;
      const int L = ml - 2;                                  // blosclz length code
      const bool far = dist - 1 >= kBlzMaxDistance;          // blosclz far match
      int size = 0;
      if (have) {
        if (LZ4) size = (head ? 0 : 1 + lz4_ext_bytes(lit) + lit) + 2 + lz4_ext_bytes(ml - 4);
        else size = lit + ((lit + 31) >> 5) + 1 + (L >= 7 ? (L - 7) / 255 + 1 : 0) + (far ? 3 : 1);
      }
      int incl = size;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += t;
      }
      const int total = __shfl_sync(kFull, incl, 31);
      if (op + total > cap) {          // limited output, like LZ4_compress_fast with maxOutputSize (lz4.c:1453)
        if (SEG && lane == 0) seg->bytes = -1;
        return SEG ? -1 : 0;
      }
      int q = op + incl - size;        // where this lane's sequence starts
      int litpos = 0;
      if (have) {
        if (head) { seg->p0 = p; seg->ml0 = ml; }
        if (LZ4) {
          if (!head) {
            out[q++] = (uint8_t)(((lit >= 15 ? 15 : lit) << 4) | (ml - 4 >= 15 ? 15 : ml - 4));
            if (lit >= 15) q = lz4_put_ext_lane(out, q, lit);
            litpos = q;
            if (lit <= kEncSelfLit)
              for (int k = 0; k < lit; k++) out[q + k] = (uint8_t)I.ld8(start + k);
            q += lit;
          }
          out[q] = (uint8_t)dist;
          out[q + 1] = (uint8_t)(dist >> 8);
          q += 2;
          if (ml - 4 >= 15) q = lz4_put_ext_lane(out, q, ml - 4);
        } else {
          litpos = q;
          if (lit > 0 && lit <= kEncSelfLit) {
            out[q] = (uint8_t)(lit - 1);
            for (int k = 0; k < lit; k++) out[q + 1 + k] = (uint8_t)I.ld8(start + k);
          }
          q += lit + ((lit + 31) >> 5);
          const int db = dist - 1;
          out[q++] = (uint8_t)(((L >= 7 ? 7 : L) << 5) | (far ? 31 : (db >> 8)));
          if (L >= 7) {
            int v = L - 7;
            while (v >= 255) { out[q++] = 255; v -= 255; }
            out[q++] = (uint8_t)v;
          }
          if (far) {
            const int dd = db - kBlzMaxDistance;
            out[q] = 255;
            out[q + 1] = (uint8_t)(dd >> 8);
            out[q + 2] = (uint8_t)dd;
          } else {
            out[q] = (uint8_t)db;
          }
        }
      }
      unsigned lm = __ballot_sync(kFull, have && lit > kEncSelfLit);  // long literal runs: the whole warp copies
      while (lm) {
        const int j = __ffs((int)lm) - 1;
        lm &= lm - 1;
        const int s = __shfl_sync(kFull, start, j), d = __shfl_sync(kFull, litpos, j), l = __shfl_sync(kFull, lit, j);
        if (LZ4) {
          if (l <= 64) {  // the usual case: one or two bytes per lane, no alignment bookkeeping
            if (lane < l) out[d + lane] = (uint8_t)I.ld8(s + lane);
            if (lane + 32 < l) out[d + 32 + lane] = (uint8_t)I.ld8(s + 32 + lane);
          } else {
            warp_copy(I, s, O, d, l, lane);
          }
        } else {
          enc_blz_literals(I, s, l, out, d, lane);
        }
      }
      op += total;
    }
    if (nseq) { anchor = lastend; head_pending = false; }
    __syncwarp();
  }
  if (SEG) {
    if (lane == 0) { seg->end = anchor; seg->bytes = op; }
    __syncwarp();
    return op;
  }
  // ---- closing literal run --------------------------------------------------------------------------------------------
  const int lit = n - anchor;
  if (LZ4) {
    if (op + 1 + lz4_ext_bytes(lit) + lit > cap) return 0;
    if (lane == 0) {
      out[op] = (uint8_t)((lit >= 15 ? 15 : lit) << 4);
      if (lit >= 15) lz4_put_ext_lane(out, op + 1, lit);
    }
    op += 1 + lz4_ext_bytes(lit);
    warp_copy(I, anchor, O, op, lit, lane);
    op += lit;
  } else {
    if (op + lit + ((lit + 31) >> 5) > cap) return 0;
    op = enc_blz_literals(I, anchor, lit, out, op, lane);
    __syncwarp();
    if (lane == 0) out[0] |= 32;  // codec marker (blosclz.c:613)
  }
  __syncwarp();
  return op;
}

// Segment of warp `warp` of a team of S for a stream of n bytes: [begin, begin + len) with len a multiple of the tile.
__device__ __forceinline__ int team_segment_len(int n, int S, int tile) { return ((n + S - 1) / S + tile - 1) / tile * tile; }
__device__ __forceinline__ int team_body_stride(int seglen) { return seglen + 64; }

// The team's join: every thread of the S warps calls it with the same arguments (but its own warp / lane) after the team
// has synchronised on the segment records.  Lays the stream out -- for every segment with a sequence: the literal run
// from the end of the previous segment's last match to its first match (LZ4: under the token that also carries that
// match's length), then its body; the closing literal run -- and, when it fits in cap, writes it: each warp its own
// segment's run and body (body k lies at bodies + k * body_stride, written by this very warp), the last warp the
// closing run.  Returns the stream's size, 0 when it does not fit or a body overflowed.  (blosclz: the caller sets the
// marker bit of out[0] once the team has synchronised again.)
template <int CODEC>
__device__ __forceinline__ int lz_join_team(const uint8_t* in, int n, const EncSeg* segs, int S, const uint8_t* bodies,
                                            int body_stride, uint8_t* out, int cap, int warp, int lane) {
  constexpr bool LZ4 = CODEC == kCodecLz4;
  const EncIn I(in);
  const MemGeneric O(out);
  int prev = 0, pos = 0, my_start = 0, my_lit = -1, my_pos = 0;
  for (int k = 0; k < S; k++) {
    const int p0 = segs[k].p0, bytes = segs[k].bytes;
    if (bytes < 0) return 0;
    if (p0 < 0) continue;
    const int lit = p0 - prev;
    if (k == warp) { my_start = prev; my_lit = lit; my_pos = pos; }
    pos += (LZ4 ? 1 + lz4_ext_bytes(lit) + lit : lit + ((lit + 31) >> 5)) + bytes;
    prev = segs[k].end;
  }
  const int clit = n - prev;
  const int total = pos + (LZ4 ? 1 + lz4_ext_bytes(clit) + clit : clit + ((clit + 31) >> 5));
  if (total > cap) return 0;
  if (my_lit >= 0) {
    int q = my_pos;
    if (LZ4) {
      const int m = segs[warp].ml0 - 4;
      if (lane == 0) {
        out[q] = (uint8_t)(((my_lit >= 15 ? 15 : my_lit) << 4) | (m >= 15 ? 15 : m));
        if (my_lit >= 15) lz4_put_ext_lane(out, q + 1, my_lit);
      }
      q += 1 + lz4_ext_bytes(my_lit);
      warp_copy(I, my_start, O, q, my_lit, lane);
      q += my_lit;
    } else {
      q = enc_blz_literals(I, my_start, my_lit, out, q, lane);
    }
    const uint8_t* body = bodies + (size_t)warp * (size_t)body_stride;
    copy_bytes(out + q, body, segs[warp].bytes, lane, 32);
  }
  if (warp == S - 1) {
    if (LZ4) {
      if (lane == 0) {
        out[pos] = (uint8_t)((clit >= 15 ? 15 : clit) << 4);
        if (clit >= 15) lz4_put_ext_lane(out, pos + 1, clit);
      }
      warp_copy(I, prev, O, pos + 1 + lz4_ext_bytes(clit), clit, lane);
    } else {
      enc_blz_literals(I, prev, clit, out, pos, lane);
    }
  }
  return total;
}

}  // namespace b2g
