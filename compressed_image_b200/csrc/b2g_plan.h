// b2g_plan.h -- host-side compress planning (pure C++, shared by b2g_api.cu and the test simulator).
// Restates the geometry rules of c-blosc2 that the reference reaches through blosc2_compress_ctx:
//   blocksize   blosc/stune.c:46-164  (a user blocksize wins; clamp to nbytes; multiple of typesize)
//   split rule  blosc/stune.c:185-210 (BLOSC_AUTO_SPLIT, the only mode compressed-image uses: wrapper.h:352)
//   memcpyed    blosc/blosc2.c:2400-2407 (clevel 0 or nbytes < 32)
#pragma once
#include <algorithm>
#include <climits>
#include <cstring>

#include "b2g_dev.cuh"

namespace b2g {

inline bool plan_split_block(int codec, int clevel, int filter, int typesize, int blocksize) {
  return ((codec == kCodecBlosclz || codec == kCodecLz4 || (codec == kCodecZstd && clevel <= 5)) &&
          filter == kFilterShuffle && typesize <= kMaxStreams && (blocksize / typesize) >= kMinBuffer);
}

inline int plan_blocksize(int user, int typesize, int nbytes) {
  int bs = user > 0 ? user : 32768;
  if (nbytes < typesize) return 1;
  if (bs > nbytes) bs = nbytes;
  if (bs > typesize) bs = bs / typesize * typesize;
  return std::max(bs, 1);
}

// Fills one CInfo; returns the number of item numbers the chunk occupies (>= 1 so that every chunk gets a header).
// win_bytes = largest stream the encoder window can take (<= 0: no limit, the caller sizes the window afterwards).
inline int plan_compress_chunk(CInfo& d, const CParams& p, int user_blocksize, int32_t nbytes, int32_t dst_cap,
                               int win_bytes) {
  memset(&d, 0, sizeof(d));
  d.nbytes = nbytes;
  if (nbytes < 0 || nbytes > INT32_MAX - kHdrExt) d.status = kErrMaxBuf;
  else if (dst_cap < kHdrExt) d.status = kErrMaxBuf;
  if (d.status == 0) {
    d.blocksize = plan_blocksize(user_blocksize, p.typesize, nbytes);
    d.nblocks = nbytes / d.blocksize;
    d.leftover = nbytes % d.blocksize;
    if (d.leftover > 0) d.nblocks++;
    d.memcpyed = (p.clevel == 0 || nbytes < kMinBuffer) ? 1 : 0;
    d.dont_split = plan_split_block(p.codec, p.clevel, p.filter, p.typesize, d.blocksize) ? 0 : 1;
    if (!d.memcpyed) {
      const int full = d.nblocks - (d.leftover > 0 ? 1 : 0);
      const int ns = d.dont_split ? 1 : p.typesize;
      d.nitems = full * ns + (d.leftover > 0 ? 1 : 0);
      d.max_stream = std::max(full > 0 ? d.blocksize / ns : 0, d.leftover);
      if (win_bytes > 0 && d.max_stream > win_bytes) d.status = kErrBlocksize;  // the encoder window is shared memory
    }
  }
  return std::max(std::max(d.nitems, d.nblocks), 1);
}

// Streams a block of this geometry is stored as (stune.c:185-210), and the warps a CTA gets for it.
inline int plan_nstreams(int codec, int clevel, int filter, int typesize, int blocksize) {
  return plan_split_block(codec, clevel, filter, typesize, blocksize) ? typesize : 1;
}
inline int plan_cta_warps(int nstreams) { return std::min(std::max(nstreams, 1), kCtaWarps); }

// Effort by compression level.  The reference's LZ4 path ignores clevel (accel is forced to 1, blosc2.c:413) and its
// blosclz path sizes the hash table by it (blosclz.c:57-62, 442-444); here clevel 9 -- what compressed-image uses by
// default -- means every match the format allows through the largest table, and the lower levels trade ratio for speed
// on both sides of the codec: a smaller table (one more warp per SM, fewer candidates to verify) and, for LZ4 from level
// 5 down, no sequences for matches that save a byte or less (two bytes or less at levels 1 and 2).
inline int plan_htab_shrink(int clevel) { return clevel >= 9 ? 0 : (clevel >= 6 ? 1 : 2); }
inline int plan_min_match(int codec, int clevel) {
  if (codec == kCodecBlosclz || clevel > 5) return 4;
  return clevel > 2 ? 5 : 6;
}

// Per-warp hash table of the encoder for streams of `neblock` bytes: 2^hlog uint16 entries with
// hlog = floor(log2(neblock)) - 3 - shrink clamped to [6, 12] (table bytes <= neblock / 4; k_compress derives the same
// hlog from the table size it is given).
inline int plan_htab_bytes(int neblock, int clevel = 9) {
  int lg = 0;
  while ((2 << lg) <= std::max(neblock, 1)) lg++;   // floor(log2(neblock))
  const int hlog = std::min(std::max(lg - 3 - plan_htab_shrink(clevel), 6), 12);
  return 2 << hlog;
}

// Encoder teams (k_compress_team): streams above one segment are encoded by several warps of one CTA side by side, each
// warp a segment of the stream in the CTA's one shared-memory window.  Warps per team for a batch whose largest stream
// is `max_stream` bytes (1 = every warp its own stream, k_compress), the segment of a `neblock`-byte stream (a multiple
// of the encoder tile), the positions in front of a segment its hash table is primed with, and the per-CTA area that
// holds the segments' bodies (segment + 64 bytes each) of the item the CTA is working on.
inline int plan_team_warps(int max_stream) {
  if (max_stream <= kTeamSegment) return 1;
  return std::min(kTeamMaxWarps, (max_stream + kTeamSegment - 1) / kTeamSegment);
}
// Two-warp teams (16 KiB planes) prime with every fourth position only: there the full insert costs as much as the
// second warp gains (measured: 1.60 ms against 1.59 ms with one warp per stream; 1.51 ms primed sparsely, ratio -0.08 %).
inline int plan_team_prime_step(int max_stream) { return max_stream <= 2 * kTeamSegment ? 4 : 1; }
inline int plan_team_body_bytes(int win) { return win + kTeamMaxWarps * (kTeamTile + 64); }   // per CTA

}  // namespace b2g
