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

constexpr int kEncSelfLit = 16;  // literal runs up to this length are copied by the sequence's own lane

// Compresses in[0..n) (shared memory) into out[0..cap) (global or shared).  Returns the compressed size, or 0 when
// the result does not fit in cap / the stream is too small for the codec (the caller then stores the stream raw, as
// blosc_c does: blosc2.c:1288-1324).  All 32 lanes call with identical arguments.
template <int CODEC, int TILE>
__device__ __forceinline__ int lz_encode_warp(const uint8_t* in, int n, uint8_t* out, int cap, const EncWork& w) {
  constexpr bool LZ4 = CODEC == kCodecLz4;
  const int lane = lane_id();
  int plimit, mlimit;  // last position a match may start at; matches end at or before mlimit
  if (LZ4) {
    plimit = n - 12;   // lz4.c:245-249 (MFLIMIT 12, LASTLITERALS 5)
    mlimit = n - 5;
  } else {
    if (n < 16 || cap < 66) return 0;  // blosclz.c:480-482
    plimit = n - 13;
    mlimit = n - 4;
  }
  for (int i = lane; i < (1 << w.hlog); i += 32) w.htab[i] = 0;
  __syncwarp();
  int anchor = 0, ip = 0, op = 0, step = 1;
  const unsigned lt_mask = (1u << lane) - 1u;
  constexpr int kEncTile = TILE, kEncGroups = TILE / 32;   // == w.tile (the work area is laid out for it)
  for (int tbase = 0; tbase <= plimit; tbase += kEncTile) {
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
      const uint32_t* wb = (const uint32_t*)(in + bbase) + (lane >> 2);
      const unsigned sh = (unsigned)(lane & 3) * 8u;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        fl[j] = bbase + 32 * j + lane <= plimit;
        fs[j] = __funnelshift_r(wb[8 * j], wb[8 * j + 1], sh);
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
        const uint32_t cseq = ld32_any(in + (tab_ok ? c : 0));
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
      const uint32_t pb = in[can_back ? p - 1 : 0], cb = in[can_back ? c - 1 : 0];
      // Match length beyond the verified 4 bytes.  Most matches of a noisy plane end within the next 4 bytes: one
      // warp-uniform word compare settles those, the warp-wide compare only runs for the longer ones (same result).
      int ml;
      {
        const int room = mlimit - (p + 4);
        const uint32_t x = ld32_any(in + p + 4) ^ ld32_any(in + c + 4);  // p + 7 <= n - 5: inside the stream
        if (x != 0u || room <= 4) {
          const int e = x ? ((__ffs((int)x) - 1) >> 3) : 4;
          ml = 4 + max(0, min(e, room));
        } else {
          ml = 8 + warp_match_len(in + p + 8, in + c + 8, room - 4, lane);
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
        while (p > lastend && c > 0 && in[p - 1] == in[c - 1]) { p--; c--; ml++; }
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
      const int lit = p - start;
      const int L = ml - 2;                                  // blosclz length code
      const bool far = dist - 1 >= kBlzMaxDistance;          // blosclz far match
      int size = 0;
      if (have) {
        if (LZ4) size = 1 + lz4_ext_bytes(lit) + lit + 2 + lz4_ext_bytes(ml - 4);
        else size = lit + ((lit + 31) >> 5) + 1 + (L >= 7 ? (L - 7) / 255 + 1 : 0) + (far ? 3 : 1);
      }
      int incl = size;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(kFull, incl, o);
        if (lane >= o) incl += t;
      }
      const int total = __shfl_sync(kFull, incl, 31);
      if (op + total > cap) return 0;  // limited output, like LZ4_compress_fast with maxOutputSize (lz4.c:1453)
      int q = op + incl - size;        // where this lane's sequence starts
      int litpos = 0;
      if (have) {
        if (LZ4) {
          out[q++] = (uint8_t)(((lit >= 15 ? 15 : lit) << 4) | (ml - 4 >= 15 ? 15 : ml - 4));
          if (lit >= 15) q = lz4_put_ext_lane(out, q, lit);
          litpos = q;
          if (lit <= kEncSelfLit)
            for (int k = 0; k < lit; k++) out[q + k] = in[start + k];
          q += lit;
          out[q] = (uint8_t)dist;
          out[q + 1] = (uint8_t)(dist >> 8);
          q += 2;
          if (ml - 4 >= 15) q = lz4_put_ext_lane(out, q, ml - 4);
        } else {
          litpos = q;
          if (lit > 0 && lit <= kEncSelfLit) {
            out[q] = (uint8_t)(lit - 1);
            for (int k = 0; k < lit; k++) out[q + 1 + k] = in[start + k];
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
            if (lane < l) out[d + lane] = in[s + lane];
            if (lane + 32 < l) out[d + 32 + lane] = in[s + 32 + lane];
          } else {
            copy_bytes(out + d, in + s, l, lane, 32);
          }
        } else {
          blosclz_put_literals(in, s, l, out, d, lane);
        }
      }
      op += total;
    }
    if (nseq) anchor = lastend;
    __syncwarp();
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
    copy_bytes(out + op, in + anchor, lit, lane, 32);
    op += lit;
  } else {
    if (op + lit + ((lit + 31) >> 5) > cap) return 0;
    op = blosclz_put_literals(in, anchor, lit, out, op, lane);
    __syncwarp();
    if (lane == 0) out[0] |= 32;  // codec marker (blosclz.c:613)
  }
  __syncwarp();
  return op;
}

}  // namespace b2g
