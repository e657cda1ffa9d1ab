// b2g_kernels.cuh -- the batch kernels of the b2g engine (sm_100a).
//
//   k_plan_decode   one thread per chunk: parse + validate the blosc2 header (F1/F2/F4 of SURVEY.md 8a;
//                   blosc/blosc2.c:672-766, 2298-2391, 2922-2951), lay the batch out as work items (= streams).
//   k_decode<Z>     persistent independent warps pull stream items from a ticket; per stream: directory walk, the
//                   compressed bytes staged at the top of the warp's shared-memory window and decoded in place
//                   (lz4 / blosclz / zstd / stored / run / zero: blosc_d, blosc2.c:1568-1937), then the fused un-filter:
//                   a byte plane is scattered into the raw block, whole-block streams are un(bit)shuffled
//                   (pipeline_backward, blosc2.c:1336-1461).  <Z = true> adds the zstd work area; zstd chunks only.
//   k_compress      persistent independent warps, stream items in ticket order: fused filter into the warp's shared
//                   window (pipeline_forward, blosc2.c:940-1050; a byte plane is gathered straight from the raw block),
//                   run detection (blosc2.c:1054-1076), lz4 / blosclz, raw fallback (blosc_c, blosc2.c:1080-1332) into
//                   an L2-resident per-warp slab, decoupled look-back over the chunk's streams for the output offset,
//                   compaction into the chunk with bstarts.
//   k_finalize      chunk headers, SPECIAL_ZERO and memcpyed fallbacks (blosc2.c:2393-2529).
//   k_filter        standalone filters (K1-K4).
#pragma once
#include "b2g_lzenc.cuh"
#include "b2g_filters.cuh"
#include "b2g_zstd.cuh"

namespace b2g {

// ---- decode plan ---------------------------------------------------------------------------------------------------
struct PlanArgs {
  const DChunk* chunks;
  DInfo* info;
  DResult* res;
  int32_t* base;       // [2][n + 1] exclusive item offsets per class (0 = lz4/blosclz/stored/special, 1 = zstd)
  BatchCtl* ctl;       // [2]
  int32_t n;
  int32_t max_block;   // largest blocksize the decode kernels can take (window or scratch slab)
  int32_t zstd_ok;     // the zstd kernel will be launched
};

__device__ __forceinline__ int plan_one(const DChunk& ch, DInfo& d, int max_block, int zstd_ok) {
  const uint8_t* src = ch.src;
  const int32_t srcsize = ch.src_bytes;
  d.nitems = 0; d.nblocks = 0; d.nbytes = 0; d.vts = 0; d.nstreams = 1; d.serial = 0;
  if (srcsize < kHdrMin) return kErrRead;
  d.version = src[0];
  d.flags = src[2];
  d.typesize = src[3];
  const int32_t nbytes = ld_le32(src + 4), blocksize = ld_le32(src + 8), cbytes = ld_le32(src + 12);
  d.blocksize = blocksize; d.cbytes = cbytes;
  if (d.version > 5) return kErrVersion;
  if (cbytes < kHdrMin) return kErrHeader;
  if (blocksize <= 0 || (nbytes > 0 && blocksize > nbytes)) return kErrHeader;
  if (blocksize > kMaxBlocksize) return kErrHeader;
  // a negative size would make this chunk's item count negative and shift every later chunk of the batch (the reference
  // decodes one chunk per call and fails it further down: blosc2.c:2937 "nbytes > destsize" never fires, the block loop
  // does not run, and it returns the negative size)
  if (nbytes < 0) return kErrHeader;
  if (d.typesize == 0) return kErrHeader;
  const int ts = d.typesize;
  const bool ext = (d.flags & kFlagShuffle) && (d.flags & kFlagBitshuffle);
  d.hdr = ext ? kHdrExt : kHdrMin;
  d.filter = kFilterNone;
  d.special = 0;
  uint8_t b2flags = 0;
  if (ext) {
    if (cbytes < kHdrExt) return kErrHeader;
    if (srcsize < kHdrExt) return kErrRead;
    b2flags = src[31];
    d.special = (b2flags >> 4) & 7;
    if (d.special == 3) {
      int32_t vts = cbytes - kHdrExt;
      if (vts <= 0 || vts > nbytes || nbytes % vts) return kErrHeader;
      d.vts = vts;
    } else if (d.special != 0) {
      if (nbytes % ts) return kErrHeader;
    }
    for (int i = 0; i < 6; i++) {
      uint8_t f = src[16 + i];
      if (d.version == 4 && i == 5) f = 0;  // alpha series (blosc2.c:758-761)
      if (f == 0) continue;
      if ((f == kFilterShuffle || f == kFilterBitshuffle) && d.filter == kFilterNone) d.filter = f;
      else return kErrFilter;
    }
  } else {
    if (d.flags & kFlagDelta) return kErrFilter;
    if (d.flags & kFlagShuffle) d.filter = kFilterShuffle;
    if (d.flags & kFlagBitshuffle) d.filter = kFilterBitshuffle;
  }
  if (nbytes > ch.dst_cap) return kErrWrite;
  if (cbytes > srcsize) return kErrHeader;
  if (d.special > 4) return kErrData;
  d.memcpyed = (d.flags & kFlagMemcpyed) ? 1 : 0;
  if (d.memcpyed && cbytes != nbytes + d.hdr) return kErrData;
  d.nbytes = nbytes;
  if (nbytes == 0 && cbytes == d.hdr && !d.special) return 0;
  if (ext && (b2flags & 0x08) && !d.special) return kErrParam;
  if (ext && (b2flags & 0x01)) return kErrCodec;
  if (d.special == 2 && ts != 4 && ts != 8) return kErrData;
  int32_t nblocks = nbytes / blocksize;
  d.leftover = nbytes % blocksize;
  if (d.leftover > 0) nblocks++;
  d.nblocks = nblocks;
  d.fmt = d.flags >> 5;
  d.dont_split = (d.flags & kFlagDontSplit) >> 4;
  if (!d.special && !d.memcpyed) {
    if (srcsize < d.hdr + 4 * nblocks) return kErrRead;
    if (d.fmt != kFmtBlosclz && d.fmt != kFmtLz4 && d.fmt != kFmtZstd) return kErrCodec;
    if (blocksize > max_block) return kErrBlocksize;
    if (d.fmt == kFmtZstd && !zstd_ok) return kErrCtxCodec;
    // work items: one per stream (blosc2.c:1771-1776: a leftover block is never split)
    const int32_t full = nblocks - (d.leftover > 0 ? 1 : 0);
    if (!d.dont_split) {
      if (ts <= kMaxStreams) d.nstreams = ts;
      else if (full > 0) return kErrData;
    }
    // planes of a byte-shuffled block are independent work items; a split block under any other filter (Blosc1-era
    // bitshuffle chunks) needs all its streams in one window before the un-filter: one item, streams decoded in turn
    d.serial = (d.nstreams > 1 && d.filter != kFilterShuffle) ? 1 : 0;
    const int per_block = d.serial ? 1 : d.nstreams;
    if ((long long)full * per_block > (1 << 28)) return kErrMaxBuf;
    d.nitems = full * per_block + (d.leftover > 0 ? 1 : 0);
  } else {
    d.nitems = nblocks;  // special-value / memcpyed chunks: one item per block
  }
  return 0;
}

constexpr int kPlanThreads = 1024;
constexpr size_t kPlanSmemBytes = kPlanThreads * 2 * sizeof(int32_t);

__global__ void __launch_bounds__(1024) k_plan_decode(PlanArgs a) {
  int32_t* cnt = (int32_t*)dyn_smem();  // [2][blockDim.x] partial sums
  const int t = (int)threadIdx.x, nt = (int)blockDim.x;
  const int per = (a.n + nt - 1) / nt;
  const int lo = min(a.n, t * per), hi = min(a.n, lo + per);
  int32_t sum[2] = {0, 0};
  for (int i = lo; i < hi; i++) {
    DInfo d;
    int rc = plan_one(a.chunks[i], d, a.max_block, a.zstd_ok);
    if (rc < 0) d.nitems = 0;
    a.res[i].status = rc;
    a.res[i].out_bytes = rc < 0 ? 0 : d.nbytes;
    const int cls = (d.nitems > 0 && !d.special && !d.memcpyed && d.fmt == kFmtZstd) ? 1 : 0;
    d.cls = (uint8_t)cls;
    d.item_base = sum[cls];  // provisional: offset inside this thread's range
    sum[cls] += d.nitems;
    a.info[i] = d;
  }
  cnt[t] = sum[0];
  cnt[nt + t] = sum[1];
  __syncthreads();
  if (t == 0) {  // serial scan over <= 1024 partials
    for (int k = 0; k < 2; k++) {
      int32_t run = 0;
      for (int i = 0; i < nt; i++) { int32_t v = cnt[k * nt + i]; cnt[k * nt + i] = run; run += v; }
      a.ctl[k].total_items = run;
      a.ctl[k].ticket = 0;
      a.ctl[k].ring = 0;
      a.ctl[k].paired = 0;
      a.base[k * (a.n + 1) + a.n] = run;
    }
  }
  __syncthreads();
  int32_t off[2] = {cnt[t], cnt[nt + t]};
  for (int i = lo; i < hi; i++) {
    const int cls = a.info[i].cls;
    const int32_t nit = a.info[i].nitems;
    a.base[i] = off[0];
    a.base[(a.n + 1) + i] = off[1];
    a.info[i].item_base = off[cls];
    off[cls] += nit;
  }
}

// Work item -> chunk: largest c with base[c] <= item (chunks without items have empty ranges).  Tickets grow
// monotonically per CTA, so gallop forward from the previous answer, then bisect.
__device__ __forceinline__ int find_chunk(const int32_t* base, int n, int item, int hint) {
  int lo = hint, step = 1;
  int hi = lo + 1;
  while (hi < n && base[hi] <= item) { lo = hi; hi += step; step <<= 1; }
  if (hi > n) hi = n;
  // invariant: base[lo] <= item, and (hi == n or base[hi] > item)
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (base[mid] <= item) lo = mid; else hi = mid;
  }
  return lo;
}


// ---- decode ----------------------------------------------------------------------------------------------------------
// Work item = one STREAM of one block (a byte plane of a split shuffled block, or the whole block when it is stored as
// a single stream); special-value and memcpyed chunks are laid out as one item per block.  Every warp of the grid is an
// independent worker: it draws items from the class ticket, owns `wcap` bytes of shared memory and never waits for
// another warp (streams of one block take very different times: a noisy low byte plane has ten times the sequences of
// the smooth high plane).
struct DecodeArgs {
  const DChunk* chunks;
  const DInfo* info;
  DResult* res;
  const int32_t* base;   // this class's [n + 1]
  BatchCtl* ctl;         // this class's
  int32_t n;
  int32_t wcap;          // shared-memory window of one warp (bytes, multiple of 16)
  int32_t wstride;       // bytes of shared memory per warp: wcap (+ the zstd work area)
  uint8_t* slab;         // global scratch window for streams that do not fit in wcap: one per warp of the grid
  int64_t slab_stride;
  // Plane pairing: the byte planes of a split block are decoded by different warps at different times.  Instead of each
  // scattering its bytes into the raw block (partial sectors, read-modify-write in L2 / HBM), the planes that finish
  // first are parked in an L2-resident ring of staging slots and the LAST plane's warp writes the whole block
  // interleaved with 128-bit stores.  nullptr = off.
  uint8_t* stage;        // [nslots][slot_bytes]
  int32_t* slotflag;     // [nslots] 0 free, 1 in use
  int32_t* mail;         // [pair_cap] per item: 0 nothing yet, slot + 1 = parked there, -1 = scattered directly
  int32_t* blkcnt;       // [pair_cap] per block (index = its first item): planes finished
  int32_t nslots, slot_bytes, pair_cap;
  int32_t bulk;          // 1: 8 bytes per warp behind the windows hold the mbarriers of the bulk payload staging
};

// Window bytes needed to decode a stream of `neblock` bytes with its compressed bytes staged in place above the output.
// A stream that is stored compressed is shorter than neblock, so the window only has to leave the output a head start:
// blosclz literal runs expand by 1/32; an LZ4 (or zstd: no staging) stream needs a small margin only.
__host__ __device__ __forceinline__ int decode_window_bytes(int neblock, int fmt) {
  if (fmt == kFmtBlosclz) return (neblock + (neblock >> 5) + 64 + 15) & ~15;
  return (neblock + 80 + 15) & ~15;
}

// item number inside a chunk -> (block j, streams [s, s_end) of ns, block bytes, stream bytes)
struct DGeom { int j, s, ns, s_end; int32_t bsize, neblock; bool scatter; };
__device__ __forceinline__ DGeom decode_geometry_of(const DInfo& d, int it) {
  DGeom g;
  const int full = d.nblocks - (d.leftover > 0 ? 1 : 0);
  const int per_block = d.serial ? 1 : d.nstreams;
  if (it < full * per_block) { g.j = it / per_block; g.s = it - g.j * per_block; g.ns = d.nstreams; }
  else { g.j = full; g.s = 0; g.ns = 1; }
  g.s_end = d.serial ? g.ns : g.s + 1;
  g.bsize = g.j < full ? d.blocksize : d.leftover;
  g.neblock = g.bsize / g.ns;
  g.scatter = g.ns > 1 && !d.serial;
  return g;
}

// dst[0..n) = src[0..n) with dst in shared memory and dst = src (mod 16): the aligned body goes through asynchronous
// 16-byte copies that are all in flight at once (a plain load / store loop pays the global latency once per 512 bytes).
__device__ __forceinline__ void stage_payload(uint8_t* dst, const uint8_t* src, int n, int lane) {
  int head = (int)((16 - ((uintptr_t)src & 15)) & 15);
  if (head > n) head = n;
  const int body = (n - head) >> 4;
  for (int g = lane; g < body; g += 32) async_copy16(dst + head + 16 * g, src + head + 16 * g);
  if (lane < head) dst[lane] = src[lane];
  for (int i = head + (body << 4) + lane; i < n; i += 32) dst[i] = src[i];
  async_copy_wait();
}

// The same staging with the aligned body moved by ONE bulk copy of the async proxy (cp.async.bulk, the 1-D form of TMA: no
// tensor map) instead of n / 16 per-lane copies: lane 0 posts the byte count on the warp's mbarrier and issues the copy,
// the lanes bring in the unaligned head and tail, everybody waits on the barrier's phase.  `bar` lives in shared memory
// (initialised with count 1 at kernel start), `phase` is the warp's running parity.
struct StageBar {
  uint64_t* bar;
  uint32_t phase;
};

#ifndef B2G_SIM
__device__ __forceinline__ void stage_payload_bulk(uint8_t* dst, const uint8_t* src, int n, int lane, StageBar& sb) {
  int head = (int)((16 - ((uintptr_t)src & 15)) & 15);
  if (head > n) head = n;
  const int body = (n - head) & ~15;
  const uint32_t bar = (uint32_t)__cvta_generic_to_shared(sb.bar);
  if (body > 0 && lane == 0) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst + head);
    // the window was written through the generic proxy (the previous stream's output): order that before the async write
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)body) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(d), "l"(src + head), "r"((uint32_t)body), "r"(bar) : "memory");
  }
  if (lane < head) dst[lane] = src[lane];
  for (int i = head + body + lane; i < n; i += 32) dst[i] = src[i];
  if (body > 0) {
    uint32_t done = 0;
    while (!done) {
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done) : "r"(bar), "r"(sb.phase) : "memory");
    }
    sb.phase ^= 1u;
  }
}
#endif

// One stream (blosc2.c:1783-1923) decoded by one warp into win[0..neblock).  `wbytes` = usable bytes of the window:
// lz4 / blosclz streams are first copied to its top (16-byte loads, same alignment phase as the source) and decoded in
// place, so the byte-serial part of the decoder only ever touches the window.
template <bool Z, bool STAGE, bool SMEM>
__device__ __forceinline__ int decode_stream(int fmt, const uint8_t* payload, int32_t csize, int32_t avail,
                                             uint8_t* win, int32_t wbytes, int32_t neblock, ZWork* zw, StageBar* sb) {
  const int lane = lane_id();
  if (csize == 0) {
    fill_bytes(win, 0, neblock, lane, 32);
    return 0;
  }
  if (csize < 0) {
    if (avail < 1) return kErrRead;
    const uint8_t token = payload[0];
    if (!(token & 1)) return kErrRunLength;
    if (csize < -255) return kErrRunLength;
    fill_bytes(win, (uint8_t)(-csize), neblock, lane, 32);
    return 0;
  }
  if (csize > avail) return kErrRead;
  if (csize == neblock) {
    copy_bytes(win, payload, neblock, lane, 32);
    return 0;
  }
  int got;
  if (Z) {
    if (fmt != kFmtZstd) return kErrCodec;
    long r = zstd_decode_warp(zw, payload, csize, win, neblock, STAGE);   // STAGE: a window with slack, not the caller's buffer
    got = r < 0 ? 0 : (int)r;  // blosc2.c:545-548: zstd errors surface as 0 bytes
  } else {
    if (fmt != kFmtLz4 && fmt != kFmtBlosclz) return kErrCodec;
    got = -2;
    if (STAGE) {
      const int m = (int)((uintptr_t)payload & 15);
      const int a = ((wbytes - csize - m) & ~15) + m;  // a + csize <= wbytes, (win + a) = payload (mod 16)
      if (wbytes - csize - m >= 0) {
        if (SMEM) {
#ifndef B2G_SIM
          if (sb) stage_payload_bulk(win + a, payload, csize, lane, *sb);   // one bulk copy (async proxy) + head / tail
          else
#endif
            stage_payload(win + a, payload, csize, lane);  // asynchronous 16-byte copies, all in flight at once
        } else {
          copy_bytes(win + a, payload, csize, lane, 32);
        }
        __syncwarp();
        got = fmt == kFmtLz4 ? lz4_decode_warp_t<true, SMEM>(win + a, csize, win, neblock, a)
                             : blosclz_decode_batched<true, SMEM>(win + a, csize, win, neblock, a);
      }
    }
    if (got == -2) {  // not staged, or the output caught up with the staged input: decode from the chunk itself
      __syncwarp();
      got = fmt == kFmtLz4 ? lz4_decode_warp_t<false, SMEM>(payload, csize, win, neblock, 0)
                           : blosclz_decode_warp(payload, csize, win, neblock);
    }
  }
  return got == neblock ? 0 : kErrData;
}

// Streams [s, s_end) of one block, decoded by one warp into `win` and un-filtered into the raw block.
struct StreamJob {
  const uint8_t* src;   // the chunk
  uint8_t* raw;         // the raw block
  int32_t srcsize, first, s, s_end, ns, neblock, bsize;
  int32_t item;         // the work item (batch-wide numbering)
  bool scatter;
  uint8_t fmt, filter, typesize, version;
};

// raw[i * ns + s] = plane[i]: the unshuffle restricted to one plane (shuffle-generic.h:62-81), byte stores.
__device__ __forceinline__ void scatter_plane(uint8_t* raw, int ns, int s, const uint8_t* plane, int neblock, int lane) {
  uint8_t* o = raw + s;
  int i = lane;
  for (; i + 7 * 32 < neblock; i += 8 * 32) {  // 8 loads in flight, then 8 stores
    uint8_t v[8];
#pragma unroll
    for (int u = 0; u < 8; u++) v[u] = plane[i + 32 * u];
#pragma unroll
    for (int u = 0; u < 8; u++) o[(size_t)(i + 32 * u) * ns] = v[u];
  }
  for (; i < neblock; i += 32) o[(size_t)i * ns] = plane[i];
}

__device__ __forceinline__ int32_t mail_read(int32_t* p) {
#ifdef B2G_SIM
  return *(volatile int32_t*)p;
#else
  int32_t v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
#endif
}

// A decoded byte plane (win[0..neblock), plane jb.s of jb.ns) on its way into the raw block, see DecodeArgs::stage.
__device__ __forceinline__ void pair_planes(const DecodeArgs& a, const StreamJob& jb, const uint8_t* win) {
  const int lane = lane_id();
  const int ns = jb.ns, s = jb.s, neblock = jb.neblock;
  const int first = jb.item - s;  // the block's first item names the block
  int old = 0;
  __syncwarp();
  if (lane == 0) old = atomicAdd(&a.blkcnt[first], 1);
  old = __shfl_sync(kFull, old, 0);
  if (old < ns - 1) {
    // not the last plane of the block: park it
    int slot = -1;
    if (lane == 0) {
      const int t = (int)((unsigned)atomicAdd(&a.ctl->ring, 1) % (unsigned)a.nslots);
      if (atomicCAS(&a.slotflag[t], 0, 1) == 0) slot = t;
    }
    slot = __shfl_sync(kFull, slot, 0);
    if (slot >= 0) {
      uint4* d4 = (uint4*)(a.stage + (size_t)slot * (size_t)a.slot_bytes);
      const uint4* s4 = (const uint4*)win;
      for (int g = lane; g < (neblock >> 4); g += 32) d4[g] = s4[g];
    } else {
      scatter_plane(jb.raw, ns, s, win, neblock, lane);  // no free slot: the old way
    }
    __threadfence();
    __syncwarp();
    if (lane == 0) atomicExch(&a.mail[jb.item], slot >= 0 ? slot + 1 : -1);
    return;
  }
  // the last plane: collect the others (their warps have passed the counter, the parking takes them a microsecond)
  const uint8_t* pl[4];
  int32_t m[4];
  bool all = true;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    pl[k] = win;
    m[k] = 0;
    if (k < ns && k != s) {
      int32_t v = 0;
      if (lane == 0) {
        while ((v = mail_read(&a.mail[first + k])) == 0) B2G_SPIN_PAUSE();
      }
      v = __shfl_sync(kFull, v, 0);
      m[k] = v;
      if (v > 0) pl[k] = a.stage + (size_t)(v - 1) * (size_t)a.slot_bytes;
      else all = false;
    }
  }
  __threadfence();
  if (all) {
    uint4* o4 = (uint4*)jb.raw;
    if (ns == 2) {  // 8 elements per lane and step: 8 bytes of each plane -> 16 bytes; 8 steps' loads in flight
      const int steps = neblock >> 3;
      int g = lane;
      for (; g + 7 * 32 < steps; g += 8 * 32) {
        uint2 p0[8], p1[8];
#pragma unroll
        for (int u = 0; u < 8; u++) {
          p0[u] = *(const uint2*)(pl[0] + 8 * (g + 32 * u));
          p1[u] = *(const uint2*)(pl[1] + 8 * (g + 32 * u));
        }
#pragma unroll
        for (int u = 0; u < 8; u++)
          o4[g + 32 * u] = make_uint4(__byte_perm(p0[u].x, p1[u].x, 0x5140), __byte_perm(p0[u].x, p1[u].x, 0x7362),
                                      __byte_perm(p0[u].y, p1[u].y, 0x5140), __byte_perm(p0[u].y, p1[u].y, 0x7362));
      }
      for (; g < steps; g += 32) {
        const uint2 p0 = *(const uint2*)(pl[0] + 8 * g), p1 = *(const uint2*)(pl[1] + 8 * g);
        o4[g] = make_uint4(__byte_perm(p0.x, p1.x, 0x5140), __byte_perm(p0.x, p1.x, 0x7362),
                           __byte_perm(p0.y, p1.y, 0x5140), __byte_perm(p0.y, p1.y, 0x7362));
      }
    } else {        // ns == 4: 4 elements per lane and step: 4 bytes of each plane -> 16 bytes
      const int steps = neblock >> 2;
      int g = lane;
      for (; g + 3 * 32 < steps; g += 4 * 32) {
        uint32_t p[4][4];
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
          for (int k = 0; k < 4; k++) p[u][k] = *(const uint32_t*)(pl[k] + 4 * (g + 32 * u));
#pragma unroll
        for (int u = 0; u < 4; u++) {
          const uint32_t t0 = __byte_perm(p[u][0], p[u][1], 0x5140), t1 = __byte_perm(p[u][0], p[u][1], 0x7362);
          const uint32_t u0 = __byte_perm(p[u][2], p[u][3], 0x5140), u1 = __byte_perm(p[u][2], p[u][3], 0x7362);
          o4[g + 32 * u] = make_uint4(__byte_perm(t0, u0, 0x5410), __byte_perm(t0, u0, 0x7632),
                                      __byte_perm(t1, u1, 0x5410), __byte_perm(t1, u1, 0x7632));
        }
      }
      for (; g < steps; g += 32) {
        const uint32_t p0 = *(const uint32_t*)(pl[0] + 4 * g), p1 = *(const uint32_t*)(pl[1] + 4 * g);
        const uint32_t p2 = *(const uint32_t*)(pl[2] + 4 * g), p3 = *(const uint32_t*)(pl[3] + 4 * g);
        const uint32_t t0 = __byte_perm(p0, p1, 0x5140), t1 = __byte_perm(p0, p1, 0x7362);
        const uint32_t u0 = __byte_perm(p2, p3, 0x5140), u1 = __byte_perm(p2, p3, 0x7362);
        o4[g] = make_uint4(__byte_perm(t0, u0, 0x5410), __byte_perm(t0, u0, 0x7632), __byte_perm(t1, u1, 0x5410),
                           __byte_perm(t1, u1, 0x7632));
      }
    }
    if (lane == 0) atomicAdd(&a.ctl->paired, 1);
    // The parked planes have been read for the last time: tell L2 so.  Without this every staged plane is written back
    // to HBM once it ages out of the cache (measured: 53 MB of extra DRAM writes for 268 MB of output).
#ifndef B2G_SIM
#pragma unroll
    for (int k = 0; k < 4; k++) {
      if (k < ns && k != s && (((uintptr_t)pl[k]) & 127) == 0) {
        for (int o = lane * 128; o + 128 <= neblock; o += 32 * 128)
          asm volatile("discard.global.L2 [%0], 128;" ::"l"(pl[k] + o) : "memory");
      }
    }
#endif
  } else {
    scatter_plane(jb.raw, ns, s, win, neblock, lane);
#pragma unroll
    for (int k = 0; k < 4; k++)
      if (k < ns && k != s && m[k] > 0) scatter_plane(jb.raw, ns, k, pl[k], neblock, lane);
  }
  __syncwarp();
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < 4; k++)
      if (k < ns && k != s && m[k] > 0) atomicExch(&a.slotflag[m[k] - 1], 0);
  }
}

// The decoded stream(s) of a work item, from the window into the raw block: a byte plane is paired or scattered, a
// whole block is un(bit)shuffled (pipeline_backward).
__device__ __forceinline__ void unfilter_streams(const DecodeArgs& a, const StreamJob& jb, const uint8_t* win) {
  const int lane = lane_id();
  if (jb.scatter) {
    const bool pair = a.stage != nullptr && (jb.ns == 2 || jb.ns == 4) && (jb.neblock & 15) == 0 &&
                      jb.neblock <= a.slot_bytes && (((uintptr_t)jb.raw) & 15) == 0 &&
                      jb.item - jb.s + jb.ns <= a.pair_cap;  // the same answer for every plane of the block
    if (pair) pair_planes(a, jb, win);
    else scatter_plane(jb.raw, jb.ns, jb.s, win, jb.neblock, lane);
  } else if (win != jb.raw) {
    filter_backward(jb.filter, jb.typesize, jb.bsize, win, jb.raw, jb.version, lane, 32);
  }
}

template <bool Z, bool STAGE, bool SMEM>
__device__ __forceinline__ int decode_streams(const DecodeArgs& a, const StreamJob& jb, uint8_t* win, int32_t wbytes,
                                              ZWork* zw, StageBar* sb) {
  // stream directory: each csize tells where the next stream starts (warp-uniform loads)
  int32_t pos = jb.first;
  if (pos <= 0 || pos >= jb.srcsize) return kErrData;
  if (jb.neblock == 0) return kErrWrite;
  for (int k = 0; k < jb.s_end; k++) {
    if (jb.srcsize - pos < 4) return kErrRead;
    const int32_t cs = ld_le32(jb.src + pos);
    pos += 4;
    if (cs > 0 && jb.srcsize - pos < cs) return kErrRead;
    if (cs < 0 && jb.srcsize - pos < 1) return kErrRead;
    if (k >= jb.s) {
      const int32_t o = jb.scatter ? 0 : k * jb.neblock;  // a plane has the window to itself
      const int rc = decode_stream<Z, STAGE, SMEM>(jb.fmt, jb.src + pos, cs, jb.srcsize - pos, win + o, wbytes - o,
                                             jb.neblock, zw, sb);
      if (rc < 0) return rc;
    }
    pos += cs > 0 ? cs : (cs < 0 ? 1 : 0);
  }
  __syncwarp();
  unfilter_streams(a, jb, win);
  return 0;
}

template <bool Z>
__global__ void __launch_bounds__(512, 1) k_decode(DecodeArgs a) {
  const int warp = warp_id(), lane = lane_id();
  uint8_t* win_s = dyn_smem() + (size_t)warp * (size_t)a.wstride;
  ZWork* zw = Z ? (ZWork*)(win_s + a.wcap) : nullptr;
  const int gwarp = (int)blockIdx.x * (int)(blockDim.x >> 5) + warp;
  // payload staging by bulk copy: one mbarrier per warp behind the windows (a.bulk: the launch reserved the room)
  StageBar sbar{nullptr, 0};
  StageBar* sb = nullptr;
  (void)sbar;
#ifndef B2G_SIM
  if (!Z && a.bulk) {
    sbar.bar = (uint64_t*)(dyn_smem() + (size_t)(blockDim.x >> 5) * (size_t)a.wstride) + warp;
    if (lane == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(sbar.bar)) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    sb = &sbar;
  }
#endif
  int hint = 0;
  for (;;) {
    int item = 0;
    if (lane == 0) item = atomicAdd(&a.ctl->ticket, 1);
    item = __shfl_sync(kFull, item, 0);
    if (item >= a.ctl->total_items) break;
    const int c = find_chunk(a.base, a.n, item, hint);
    hint = c;
    const DInfo d = a.info[c];
    const DChunk ch = a.chunks[c];
    const int it = item - d.item_base;

    if (d.special || d.memcpyed) {  // one item per block
      const int j = it;
      const int32_t bsize = (j == d.nblocks - 1 && d.leftover > 0) ? d.leftover : d.blocksize;
      uint8_t* raw = ch.dst + (size_t)j * d.blocksize;
      if (d.memcpyed) {
        copy_bytes(raw, ch.src + d.hdr + (size_t)j * d.blocksize, bsize, lane, 32);
      } else if (d.special == 1 || d.special == 4) {  // blosc2.c:1694-1718
        fill_bytes(raw, 0, bsize, lane, 32);  // 4 = uninitialised: any content is legal, zeros are stable
      } else if (d.special == 2) {
        if (d.typesize == 4) {
          for (int i = lane; i < bsize / 4; i += 32) st_le32(raw + 4 * i, 0x7fc00000);
        } else {
          for (int i = lane; i < bsize / 8; i += 32) { st_le32(raw + 8 * i, 0); st_le32(raw + 8 * i + 4, 0x7ff80000); }
        }
      } else {
        const size_t o = (size_t)j * d.blocksize;
        for (int i = lane; i < bsize; i += 32) raw[i] = ch.src[kHdrExt + (o + i) % d.vts];
      }
      continue;
    }

    const DGeom g = decode_geometry_of(d, it);
    const int j = g.j;
    const int32_t bsize = g.bsize, neblock = g.neblock;
    uint8_t* raw = ch.dst + (size_t)j * d.blocksize;
    const int32_t srcsize = min(ch.src_bytes, d.cbytes);
    const bool direct = !g.scatter && d.filter == kFilterNone;  // no filter: the window is only a staging area
    const int32_t need = decode_window_bytes(g.scatter ? neblock : bsize, d.fmt);
    // Three window kinds, three instantiations: the shared-memory one must be its own code path so that every access
    // of the byte-serial decoder compiles to LDS / STS instead of generic loads and stores.
    StreamJob jb;
    jb.src = ch.src; jb.raw = raw; jb.srcsize = srcsize; jb.first = ld_le32(ch.src + d.hdr + 4 * j);
    jb.s = g.s; jb.s_end = g.s_end; jb.ns = g.ns; jb.neblock = neblock; jb.bsize = bsize; jb.scatter = g.scatter;
    jb.fmt = d.fmt; jb.filter = d.filter; jb.typesize = d.typesize; jb.version = d.version; jb.item = item;
    int err;
    if (need <= a.wcap) err = decode_streams<Z, true, true>(a, jb, win_s, a.wcap, zw, sb);
    else if (direct) err = decode_streams<Z, false, false>(a, jb, raw, bsize, zw, nullptr);
    else err = decode_streams<Z, true, false>(a, jb, a.slab + (size_t)gwarp * (size_t)a.slab_stride, need, zw, nullptr);
    if (err && lane == 0) set_status(a.res, c, err);
    __syncwarp();  // the window is free for the next item
  }
}

// ---- compress --------------------------------------------------------------------------------------------------------
// Result of one stream item (k_compress -> k_place -> k_finalize).
struct CItem {
  int32_t csize;   // the stream's csize field (0 zero run, < 0 run of a byte value, == neblock stored, else encoded size)
  int32_t psize;   // payload bytes that follow the csize field
  int32_t offset;  // k_place: byte offset of the csize field inside the chunk
  int32_t stored;  // payload = the filtered stream itself (not kept: k_finalize re-filters it from the source)
};

// item number inside a chunk -> (block j, stream s of ns, block bytes, stream bytes)
struct StreamGeom { int j, s, ns, bsize, neblock; };
__device__ __forceinline__ StreamGeom stream_geometry(const CInfo& d, int ts, int it) {
  StreamGeom g;
  const int full = d.nblocks - (d.leftover > 0 ? 1 : 0);
  const int nsf = d.dont_split ? 1 : ts;
  if (it < full * nsf) { g.j = it / nsf; g.s = it - g.j * nsf; g.ns = nsf; }
  else { g.j = full; g.s = 0; g.ns = 1; }
  g.bsize = g.j < full ? d.blocksize : d.leftover;
  g.neblock = g.bsize / g.ns;
  return g;
}

struct CompressArgs {
  const DChunk* chunks;          // src = raw elements, dst = chunk out
  const CInfo* info;
  CState* state;
  const int32_t* base;           // [n + 1]
  BatchCtl* ctl;
  CItem* items;                  // [total_items] what each stream item produced
  uint8_t* scratch;              // [total_items][scratch_stride] the encoded streams, in item order
  int64_t scratch_stride;
  int32_t n;
  int32_t win_bytes;             // shared-memory window of one warp: the largest stream of the batch
  int32_t htab_bytes;            // per-warp hash table bytes (a power of two); each warp's work area is
                                 // htab_bytes + enc_work_bytes(tile), see enc_work()
  int32_t tile;                  // encoder tile (positions), enc_tile_for(largest stream of the batch)
  int32_t team_prime;            // k_compress_team: positions in front of a segment that prime its hash table
  int32_t team_prime_step;       // 1 = every one of them, 4 = every fourth
  uint8_t* team_bodies;          // k_compress_team: [gridDim.x][team_body_bytes] the segment bodies of the item a CTA is on
  int64_t team_body_bytes;
  CParams p;
};

// Is in[0..n) one repeated byte?  (get_run, blosc2.c:1054-1076)
__device__ __forceinline__ bool warp_is_run(const uint8_t* in, int n, int lane) {
  const uint8_t v = in[0];
  bool same = true;
  if ((((uintptr_t)in) & 15) == 0) {
    const uint32_t w = 0x01010101u * v;
    const uint4* p = (const uint4*)in;
    const int groups = n >> 4;
    for (int g = lane; g < groups && same; g += 32) {
      uint4 x = p[g];
      same = (x.x == w) && (x.y == w) && (x.z == w) && (x.w == w);
    }
    for (int i = (groups << 4) + lane; i < n; i += 32) same = same && in[i] == v;
  } else {
    for (int i = lane; i < n && same; i += 32) same = in[i] == v;
  }
  return __all_sync(kFull, same);
}

// Byte plane s of a raw block into the window: win[i] = raw[i * ts + s] (the forward shuffle restricted to one plane,
// shuffle-generic.h:34-53).  Every plane's warp streams the whole raw block with 128-bit loads (the block is read from
// HBM once; the other planes' warps find it in L2).
__device__ __forceinline__ void gather_plane(int ts, int s, int neblock, const uint8_t* raw, uint8_t* win, int lane,
                                             int nthr) {
  const int bytes = neblock * ts;
  if ((ts == 2 || ts == 4 || ts == 8 || ts == 16) && (((uintptr_t)raw | (uintptr_t)win) & 15) == 0 && (bytes & 15) == 0) {
    const uint4* r4 = (const uint4*)raw;
    const int groups = bytes >> 4;
    if (ts == 2) {
      const uint32_t sel = s ? 0x7531u : 0x6420u;
      int g = lane;
      for (; g + 7 * nthr < groups; g += 8 * nthr) {  // eight loads in flight
        uint4 v[8];
#pragma unroll
        for (int u = 0; u < 8; u++) v[u] = r4[g + u * nthr];
#pragma unroll
        for (int u = 0; u < 8; u++)
          *(uint2*)(win + 8 * (g + u * nthr)) = make_uint2(__byte_perm(v[u].x, v[u].y, sel), __byte_perm(v[u].z, v[u].w, sel));
      }
      for (; g < groups; g += nthr) {
        const uint4 v = r4[g];
        *(uint2*)(win + 8 * g) = make_uint2(__byte_perm(v.x, v.y, sel), __byte_perm(v.z, v.w, sel));
      }
    } else if (ts == 4) {
      const uint32_t sel = (uint32_t)s | ((uint32_t)(4 + s) << 4);
      for (int g = lane; g < groups; g += nthr) {
        const uint4 v = r4[g];
        const uint32_t lo = __byte_perm(v.x, v.y, sel), hi = __byte_perm(v.z, v.w, sel);
        *(uint32_t*)(win + 4 * g) = __byte_perm(lo, hi, 0x5410);
      }
    } else if (ts == 8) {
      const uint32_t sel = (uint32_t)(s & 3) | ((uint32_t)(4 + (s & 3)) << 4);
      for (int g = lane; g < groups; g += nthr) {
        const uint4 v = r4[g];
        *(uint16_t*)(win + 2 * g) = (uint16_t)(s < 4 ? __byte_perm(v.x, v.z, sel) : __byte_perm(v.y, v.w, sel));
      }
    } else {
      for (int g = lane; g < groups; g += nthr) {
        const uint4 v = r4[g];
        const uint32_t w = s < 4 ? v.x : (s < 8 ? v.y : (s < 12 ? v.z : v.w));
        win[g] = (uint8_t)(w >> (8 * (s & 3)));
      }
    }
    return;
  }
  const uint8_t* r = raw + s;
  for (int i = lane; i < neblock; i += nthr) win[i] = r[(size_t)i * ts];
}

// Work item = one stream (a byte plane of a split block, or a whole unsplit / leftover block), one independent warp per
// item: gather + filter into the warp's window, run detection, encode into the item's scratch slot.  Where the stream
// lands in the chunk depends on the sizes of all earlier streams, so placement is a separate pass (k_place, k_finalize):
// no warp ever waits for another one.
// One instantiation per (codec, tile): the encoder is a large loop nest, and a kernel that holds all four variants
// inline loses 5 % to instruction fetch.
template <int CODEC, int TILE>
__global__ void __launch_bounds__(256, 2) k_compress(CompressArgs a) {
  const int warp = warp_id(), lane = lane_id();
  const int wstride = a.win_bytes + a.htab_bytes + enc_work_bytes(TILE);
  uint8_t* win = dyn_smem() + (size_t)warp * (size_t)wstride;   // [win_bytes] the filtered stream
  uint8_t* work = win + a.win_bytes;
  const int ts = a.p.typesize;
  int hint = 0;
  for (;;) {
    int item = 0;
    if (lane == 0) item = atomicAdd(&a.ctl->ticket, 1);
    item = __shfl_sync(kFull, item, 0);
    if (item >= a.ctl->total_items) break;
    const int c = find_chunk(a.base, a.n, item, hint);
    hint = c;
    const CInfo d = a.info[c];
    const int it = item - d.item_base;
    if (d.status < 0 || d.memcpyed || it >= d.nitems) continue;  // k_finalize stores these chunks
    const StreamGeom g = stream_geometry(d, ts, it);
    const uint8_t* raw = a.chunks[c].src + (size_t)g.j * d.blocksize;
    if (g.ns > 1) gather_plane(ts, g.s, g.neblock, raw, win, lane, 32);
    else filter_forward(a.p.filter, ts, g.bsize, raw, win, lane, 32);
    __syncwarp();
    CItem r;
    r.offset = 0;
    r.stored = 0;
    if (warp_is_run(win, g.neblock, lane)) {
      const uint8_t v = win[0];
      r.csize = -(int32_t)v;
      r.psize = v ? 1 : 0;
    } else {
      int hlog = 28 - __clz(g.neblock);              // floor(log2(neblock)) - 3: table bytes <= neblock / 4
      const int hmax = 30 - __clz(a.htab_bytes);     // log2(htab_bytes / 2)
      hlog = hlog < 6 ? 6 : (hlog > hmax ? hmax : hlog);
      const EncWork ew = enc_work(work, a.htab_bytes, hlog, TILE, CODEC == kCodecBlosclz || a.p.clevel > 5 ? 4 : (a.p.clevel > 2 ? 5 : 6));   // == plan_min_match (b2g_plan.h is host code)
      uint8_t* out = a.scratch + (size_t)item * (size_t)a.scratch_stride;
      const int cap = g.neblock - 1;  // only a strictly smaller result is worth keeping (blosc2.c:1204-1211)
      const int n = lz_encode_warp<CODEC, TILE>(win, g.neblock, out, cap, ew);
      // below level 9 a stream that the codec shrinks by next to nothing is stored (like the reference's blosclz, which
      // gives up on a block whose estimated ratio is under 1.0 .. 1.2 depending on the level, blosclz.c:463-468): the
      // bytes saved are within 0.8 % / 1.6 % / 3 % of the stream, and a stored stream decodes as a copy
      const int margin = a.p.clevel >= 9 ? 0 : (a.p.clevel >= 6 ? g.neblock >> 7 : (a.p.clevel >= 3 ? g.neblock >> 6 : g.neblock >> 5));
      if (n <= 0 || n >= g.neblock - margin) { r.stored = 1; r.psize = r.csize = g.neblock; }
      else r.psize = r.csize = n;
    }
    if (lane == 0) a.items[item] = r;
    __syncwarp();  // the window is free for the next item
  }
}

// The same work item encoded by a TEAM: the CTA's warps share one window and encode consecutive segments of the stream
// side by side (lz_encode_warp<SEG>), then join their bodies into the item's scratch slot (lz_join_team).  A 16 KiB
// plane keeps four warps busy for one window instead of one, so the SM holds as many warps as its registers allow
// instead of as many as 227 KiB / (window + table) -- the encoder is a latency-bound dependent chain, more resident
// warps are what it lacks.  Costs: a match may not cross a segment boundary, and a segment's table only knows the
// team_prime positions in front of it (ratio: see DESIGN.md); every item costs five CTA barriers.
// The bodies live in a per-CTA area (they are dead once the item is joined: a few MB for the whole grid, L2-resident);
// the item's scratch slot receives the joined stream, as in k_compress.
struct TeamCtl {
  EncSeg seg[kTeamMaxWarps];
  int32_t item;
  int32_t notrun;
};

template <int CODEC, int TILE>
__global__ void __launch_bounds__(256, 2) k_compress_team(CompressArgs a) {
  const int warp = warp_id(), lane = lane_id(), tid = (int)threadIdx.x, nthr = (int)blockDim.x, S = nthr >> 5;
  const int wbytes = a.htab_bytes + enc_work_bytes(TILE);
  uint8_t* win = dyn_smem();                                      // [win_bytes] the filtered stream, the whole team's
  uint8_t* work = win + a.win_bytes + (size_t)warp * (size_t)wbytes;
  TeamCtl* tc = (TeamCtl*)(win + a.win_bytes + (size_t)S * (size_t)wbytes);
  const int ts = a.p.typesize;
  int hint = 0;
  for (;;) {
    __syncthreads();  // the team is done with the previous item: window, segment records, flags
    if (tid == 0) { tc->item = atomicAdd(&a.ctl->ticket, 1); tc->notrun = 0; }
    __syncthreads();
    const int item = tc->item;
    if (item >= a.ctl->total_items) break;
    const int c = find_chunk(a.base, a.n, item, hint);
    hint = c;
    const CInfo d = a.info[c];
    const int it = item - d.item_base;
    if (d.status < 0 || d.memcpyed || it >= d.nitems) continue;  // k_finalize stores these chunks
    const StreamGeom g = stream_geometry(d, ts, it);
    const uint8_t* raw = a.chunks[c].src + (size_t)g.j * d.blocksize;
    if (g.ns > 1) gather_plane(ts, g.s, g.neblock, raw, win, tid, nthr);
    else filter_forward(a.p.filter, ts, g.bsize, raw, win, tid, nthr);
    __syncthreads();
    const int n = g.neblock;
    const int seglen = team_segment_len(n, S, TILE);
    const int sb = warp * seglen, se = min(n, sb + seglen);
    const bool same = sb >= n || (win[sb] == win[0] && warp_is_run(win + sb, se - sb, lane));
    if (!same && lane == 0) tc->notrun = 1;
    __syncthreads();
    CItem r;
    r.offset = 0;
    r.stored = 0;
    if (!tc->notrun) {
      const uint8_t v = win[0];
      r.csize = -(int32_t)v;
      r.psize = v ? 1 : 0;
    } else {
      const int cap = n - 1;  // only a strictly smaller result is worth keeping (blosc2.c:1204-1211)
      int total = 0;
      if (CODEC != kCodecBlosclz || (n >= 16 && cap >= 66)) {  // blosclz.c:480-482
        int hlog = 28 - __clz(n);   // as k_compress: floor(log2(n)) - 3, at most what the host's table size allows
        const int hmax = 30 - __clz(a.htab_bytes);
        hlog = hlog < 6 ? 6 : (hlog > hmax ? hmax : hlog);
        const EncWork ew = enc_work(work, a.htab_bytes, hlog, TILE, CODEC == kCodecBlosclz || a.p.clevel > 5 ? 4 : (a.p.clevel > 2 ? 5 : 6));
        uint8_t* out = a.scratch + (size_t)item * (size_t)a.scratch_stride;
        uint8_t* bodies = a.team_bodies + (size_t)blockIdx.x * (size_t)a.team_body_bytes;  // this CTA's, item after item
        const int bstride = team_body_stride(seglen);
        if (sb < n) {
          lz_encode_warp<CODEC, TILE, true>(win, n, bodies + (size_t)warp * (size_t)bstride, bstride, ew, sb, se, a.team_prime,
                                            &tc->seg[warp], a.team_prime_step);
        } else if (lane == 0) {
          tc->seg[warp].p0 = -1;
          tc->seg[warp].bytes = 0;
          tc->seg[warp].end = 0;
        }
        __syncthreads();
        total = lz_join_team<CODEC>(win, n, tc->seg, S, bodies, bstride, out, cap, warp, lane);
        if (CODEC == kCodecBlosclz && total > 0) {
          __syncthreads();
          if (tid == 0) out[0] |= 32;  // codec marker (blosclz.c:613)
        }
      }
      const int margin = a.p.clevel >= 9 ? 0 : (a.p.clevel >= 6 ? n >> 7 : (a.p.clevel >= 3 ? n >> 6 : n >> 5));
      if (total <= 0 || total >= n - margin) { r.stored = 1; r.psize = r.csize = n; }
      else r.psize = r.csize = total;
    }
    if (tid == 0) a.items[item] = r;
  }
}

// One CTA per chunk: exclusive scan of the stream sizes -> where every stream goes (blosc2.c:1956-1959 bstarts), the
// chunk total, and the two whole-chunk outcomes (does not fit -> stored chunk; nothing but zero runs -> 32-byte chunk).
struct PlaceArgs {
  const DChunk* chunks;
  const CInfo* info;
  CItem* items;
  CState* state;
  int32_t n;
};

__global__ void __launch_bounds__(256) k_place(PlaceArgs a) {
  int32_t* part = (int32_t*)dyn_smem();  // [blockDim.x] dynamic shared memory: 2 x blockDim.x ints
  int32_t* nz = part + blockDim.x;
  const int tid = (int)threadIdx.x, nt = (int)blockDim.x;
  for (int c = (int)blockIdx.x; c < a.n; c += (int)gridDim.x) {
    const CInfo d = a.info[c];
    if (d.status < 0 || d.memcpyed) {
      if (tid == 0) { CState st; st.overflow = 0; st.nonzero_streams = 0; st.done_blocks = 0; st.total = 0; a.state[c] = st; }
      continue;
    }
    CItem* it = a.items + d.item_base;
    const int per = (d.nitems + nt - 1) / nt;
    const int lo = min(d.nitems, tid * per), hi = min(d.nitems, lo + per);
    int32_t sum = 0, nonzero = 0;
    for (int i = lo; i < hi; i++) { sum += 4 + it[i].psize; nonzero += it[i].csize != 0; }
    __syncthreads();
    part[tid] = sum;
    nz[tid] = nonzero;
    __syncthreads();
    for (int o = 1; o < nt; o <<= 1) {  // inclusive scan over the thread partials
      const int32_t v = tid >= o ? part[tid - o] : 0;
      const int32_t w = tid >= o ? nz[tid - o] : 0;
      __syncthreads();
      part[tid] += v;
      nz[tid] += w;
      __syncthreads();
    }
    int32_t off = kHdrExt + 4 * d.nblocks + part[tid] - sum;
    for (int i = lo; i < hi; i++) { it[i].offset = off; off += 4 + it[i].psize; }
    if (tid == nt - 1) {
      CState st;
      st.total = kHdrExt + 4 * d.nblocks + part[tid];
      st.nonzero_streams = nz[tid];
      st.overflow = st.total > min(a.chunks[c].dst_cap, d.nbytes + kHdrExt) ? 1 : 0;
      st.done_blocks = 0;
      a.state[c] = st;
    }
  }
}

struct FinalizeArgs {
  const DChunk* chunks;
  const CInfo* info;
  const CState* state;
  const CItem* items;
  const uint8_t* scratch;
  int64_t scratch_stride;
  DResult* res;
  const int32_t* base;
  int32_t n;
  int32_t total_items;
  int32_t win_bytes;     // dynamic shared memory: a stored stream is re-filtered through it
  CParams p;
};

// One CTA per item number: the chunk header by the chunk's first item; the item's stream moved from its scratch slot to
// its place in the chunk ([csize][payload], bstarts for the first stream of a block); for stored chunks the raw copy of
// one block per item number.
__global__ void __launch_bounds__(kCtaThreads) k_finalize(FinalizeArgs a) {
  const int tid = (int)threadIdx.x;
  uint8_t* win = dyn_smem();
  const int ts = a.p.typesize;
  for (int item = (int)blockIdx.x; item < a.total_items; item += (int)gridDim.x) {
    const int c = find_chunk(a.base, a.n, item, 0);
    const CInfo d = a.info[c];
    const DChunk ch = a.chunks[c];
    const int it = item - d.item_base;
    if (d.status < 0) {
      if (it == 0 && tid == 0) { a.res[c].status = d.status; a.res[c].out_bytes = 0; }
      continue;
    }
    const CState st = a.state[c];
    const bool stored = d.memcpyed || st.overflow;
    const bool fits = !stored || d.nbytes + kHdrExt <= ch.dst_cap;
    const bool zero = !stored && st.nonzero_streams == 0;
    if (it == 0 && tid == 0) {
      uint8_t* h = ch.dst;
      int32_t cbytes = 0;
      if (fits) {
        uint8_t flags = kFlagShuffle | kFlagBitshuffle;  // both bits = "extended header" marker (blosc2.c:2413)
        uint8_t b2flags = 0;
        if (!d.memcpyed) {
          const uint8_t fmt = a.p.codec == kCodecBlosclz ? kFmtBlosclz : kFmtLz4;
          flags |= (uint8_t)(d.dont_split << 4) | (uint8_t)(fmt << 5);
        }
        if (stored) {
          flags |= kFlagMemcpyed;
          cbytes = d.nbytes + kHdrExt;
        } else if (zero) {
          b2flags = 1 << 4;  // BLOSC2_SPECIAL_ZERO (blosc2.c:2520-2529)
          cbytes = kHdrExt;
        } else {
          cbytes = st.total;
        }
        for (int i = 0; i < kHdrExt; i++) h[i] = 0;
        h[0] = 5;  // BLOSC2_VERSION_FORMAT_STABLE
        h[1] = 1;
        h[2] = flags;
        h[3] = a.p.typesize;
        st_le32(h + 4, d.nbytes);
        st_le32(h + 8, d.blocksize);
        st_le32(h + 12, cbytes);
        h[16 + 5] = a.p.filter;
        h[22] = a.p.codec;
        h[31] = b2flags;
      }
      a.res[c].status = 0;
      a.res[c].out_bytes = cbytes;  // 0 = does not fit, like blosc2_compress_ctx returning 0
    }
    if (stored) {
      if (fits && it < d.nblocks) {
        const bool is_left = (it == d.nblocks - 1) && d.leftover > 0;
        const int32_t bsize = is_left ? d.leftover : d.blocksize;
        copy_bytes(ch.dst + kHdrExt + (size_t)it * d.blocksize, ch.src + (size_t)it * d.blocksize, bsize, tid,
                   kCtaThreads);
      }
      continue;
    }
    if (zero || it >= d.nitems) continue;
    const CItem r = a.items[item];
    const StreamGeom g = stream_geometry(d, ts, it);
    uint8_t* o = ch.dst + r.offset;
    if (tid == 0) {
      st_le32(o, r.csize);
      if (g.s == 0) st_le32(ch.dst + kHdrExt + 4 * g.j, r.offset);
    }
    if (r.stored) {  // the payload is the filtered stream: filter it again, straight from the source
      const uint8_t* raw = ch.src + (size_t)g.j * d.blocksize;
      __syncthreads();  // the previous item's copy has left the window
      if (g.ns > 1) gather_plane(ts, g.s, g.neblock, raw, win, tid, kCtaThreads);
      else filter_forward(a.p.filter, ts, g.bsize, raw, win, tid, kCtaThreads);
      __syncthreads();
      copy_bytes(o + 4, win, r.psize, tid, kCtaThreads);
    } else if (r.csize < 0) {
      if (tid == 0) o[4] = 0x01;  // run token (blosc2.c:1186-1199)
    } else {
      copy_bytes(o + 4, a.scratch + (size_t)item * (size_t)a.scratch_stride, r.psize, tid, kCtaThreads);
    }
  }
}

// ---- standalone filters ------------------------------------------------------------------------------------------
struct FilterArgs {
  const uint8_t* src;
  uint8_t* dst;
  int64_t n;
  int32_t blocksize, typesize, filter, forward, version, nblocks;
};

__global__ void __launch_bounds__(kCtaThreads) k_filter(FilterArgs a) {
  uint8_t* win = dyn_smem();
  const int tid = (int)threadIdx.x;
  for (int b = (int)blockIdx.x; b < a.nblocks; b += (int)gridDim.x) {
    const int64_t o = (int64_t)b * a.blocksize;
    const int32_t bsize = (int32_t)min((int64_t)a.blocksize, a.n - o);
    __syncthreads();
    if (a.forward) {
      filter_forward(a.filter, a.typesize, bsize, a.src + o, win, tid, kCtaThreads);
      __syncthreads();
      copy_bytes(a.dst + o, win, bsize, tid, kCtaThreads);
    } else {
      copy_bytes(win, a.src + o, bsize, tid, kCtaThreads);
      __syncthreads();
      filter_backward(a.filter, a.typesize, bsize, win, a.dst + o, a.version, tid, kCtaThreads);
    }
  }
}

}  // namespace b2g
