// b2g_dev.cuh -- device-side structures and helpers shared by the b2g kernels (sm_100a).
//
// Vocabulary follows blosc2: a CHUNK is split into BLOCKS (default 32 KiB); a block is stored as 1 or
// `typesize` STREAMS (the byte planes of the shuffled block), each framed by an int32 csize
// (reference: thirdparty/c-blosc2/blosc/blosc2.c:1080-1332 blosc_c, :1568-1937 blosc_d).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2g {

constexpr int kHdrMin = 16, kHdrExt = 32;
constexpr int kMaxStreams = 16;        // stune.h:24 MAX_STREAMS
constexpr int kMinBuffer = 32;         // b2h:193 BLOSC_MIN_BUFFERSIZE
constexpr int kMaxBlocksize = 536866816;

enum : uint8_t { kFlagShuffle = 1, kFlagMemcpyed = 2, kFlagBitshuffle = 4, kFlagDelta = 8, kFlagDontSplit = 0x10 };
enum : uint8_t { kFmtBlosclz = 0, kFmtLz4 = 1, kFmtZlib = 3, kFmtZstd = 4 };
enum : uint8_t { kFilterNone = 0, kFilterShuffle = 1, kFilterBitshuffle = 2 };
enum : int { kErrData = -3, kErrRead = -5, kErrWrite = -6, kErrCodec = -7, kErrCodecParam = -8, kErrVersion = -10,
             kErrHeader = -11, kErrParam = -12, kErrRunLength = -17, kErrFilter = -18, kErrMaxBuf = -35,
             kErrBlocksize = -101, kErrCtxCodec = -102 };

// One chunk of a batch as the kernels see it (device pointers).
struct DChunk {
  const uint8_t* src;
  uint8_t* dst;
  int32_t src_bytes;
  int32_t dst_cap;
};

// Decompress plan: one per chunk, written by k_plan_decode from the chunk's own header.
struct DInfo {
  int32_t nbytes, blocksize, cbytes, nblocks, leftover;
  int32_t item_base[2];  // first work item of this chunk in class 0 (lz4/blosclz/stored/special) and 1 (zstd)
  int32_t status;        // header-level status (0 = ok)
  uint8_t version, flags, typesize, filter, fmt, dont_split, special, memcpyed, hdr, pad[3];
};

// Compress plan: one per chunk, written on the host (all geometry is known there).
struct CInfo {
  int32_t nbytes, blocksize, nblocks, leftover;
  int32_t item_base;
  int32_t memcpyed;      // clevel 0 or nbytes < 32: stored chunk
  int32_t dont_split;
  int32_t status;
};

// Per-chunk running state of a compress batch (device, zero-initialised per batch).
struct CState {
  unsigned long long chain;  // decoupled look-back chain: (blocks_published << 32) | bytes_used
  int32_t overflow;          // some block did not fit dst_capacity -> chunk becomes memcpyed
  int32_t nonzero_streams;   // streams that are not zero runs (0 -> SPECIAL_ZERO chunk)
  int32_t done_blocks;
  int32_t err;
};

__device__ __forceinline__ int32_t ld_le32(const uint8_t* p) {
  return (int32_t)((uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24));
}
__device__ __forceinline__ void st_le32(uint8_t* p, int32_t v) {
  p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

// 8 bytes from an arbitrarily aligned global address, using only aligned 8-byte loads that contain at
// least one requested byte (never touches a word that lies wholly outside [p, p+8)).
__device__ __forceinline__ uint64_t ldg_u64_unaligned(const uint8_t* p) {
  uintptr_t a = (uintptr_t)p;
  const uint64_t* w = (const uint64_t*)(a & ~(uintptr_t)7);
  unsigned sh = (unsigned)(a & 7) * 8;
  uint64_t lo = __ldg(w);
  if (sh == 0) return lo;
  uint64_t hi = __ldg(w + 1);
  return (lo >> sh) | (hi << (64 - sh));
}

// Transpose an 8x8 bit matrix held in a u64 (byte r = row r, bit c = column c)  ->  byte c = column c.
__device__ __forceinline__ uint64_t transpose8x8(uint64_t x) {
  uint64_t t;
  t = (x ^ (x >> 7)) & 0x00AA00AA00AA00AAull; x = x ^ t ^ (t << 7);
  t = (x ^ (x >> 14)) & 0x0000CCCC0000CCCCull; x = x ^ t ^ (t << 14);
  t = (x ^ (x >> 28)) & 0x00000000F0F0F0F0ull; x = x ^ t ^ (t << 28);
  return x;
}

}  // namespace b2g
