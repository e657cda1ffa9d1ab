// b2g_dev.cuh -- device-side structures and helpers shared by the b2g kernels (sm_100a).
//
// Vocabulary follows blosc2: a CHUNK is split into BLOCKS (default 32 KiB); a block is stored as 1 or
// `typesize` STREAMS (the byte planes of the shuffled block), each framed by an int32 csize
// (reference: thirdparty/c-blosc2/blosc/blosc2.c:1080-1332 blosc_c, :1568-1937 blosc_d).
//
// The same sources compile two ways: with nvcc for sm_100a (the product), and with g++ under tests/sim/cusim.h
// (B2G_SIM; a CPU fiber simulator used only by the test-suite to exercise this exact code without a GPU).
#pragma once
#ifndef B2G_SIM
#include <cuda_runtime.h>
#endif
#include <stdint.h>

namespace b2g {

constexpr int kHdrMin = 16, kHdrExt = 32;
constexpr int kMaxStreams = 16;        // stune.h:24 MAX_STREAMS
constexpr int kMinBuffer = 32;         // b2h:193 BLOSC_MIN_BUFFERSIZE
constexpr int kMaxBlocksize = 536866816;
constexpr int kCtaThreads = 128;       // every block-level kernel runs 4 warps per CTA
constexpr int kCtaWarps = kCtaThreads / 32;
// encoder teams (k_compress_team, b2g_plan.h): bytes of a stream per warp, warps per team, encoder tile, and how
// many positions in front of its segment a warp's hash table is primed with.  Measured on 32 KiB streams (cfg 1: LZ4
// on unsplit uint8 blocks; cfg 3: blosclz on bit-shuffled blocks): four warps x 8 KiB primed with the 8 KiB in front
// compress 2.07x / 1.47x faster than one warp per stream at -0.3 % / +0.3 % ratio; 4 KiB segments or less priming
// are faster still but cost 2-4 % ratio.  On 16 KiB streams four warps cost 0.55 % ratio for 8 %; two warps gain 5 % at
// -0.08 % when the second one primes its table with every fourth position only (plan_team_prime_step).
constexpr int kTeamSegment = 8192;
constexpr int kTeamMaxWarps = 8;
constexpr int kTeamTile = 256;   // 512 leaves room for 14 warps per SM on 16 KiB planes, 256 for 16: k_compress 1.50 -> 1.43 ms
constexpr int kTeamPrime = 8192;
constexpr unsigned kFull = 0xffffffffu;

enum : uint8_t { kFlagShuffle = 1, kFlagMemcpyed = 2, kFlagBitshuffle = 4, kFlagDelta = 8, kFlagDontSplit = 0x10 };
enum : uint8_t { kFmtBlosclz = 0, kFmtLz4 = 1, kFmtZlib = 3, kFmtZstd = 4 };
enum : uint8_t { kCodecBlosclz = 0, kCodecLz4 = 1, kCodecLz4hc = 2, kCodecZstd = 5 };
enum : uint8_t { kFilterNone = 0, kFilterShuffle = 1, kFilterBitshuffle = 2 };
enum : int { kErrFailure = -1, kErrData = -3, kErrRead = -5, kErrWrite = -6, kErrCodec = -7, kErrCodecParam = -8,
             kErrVersion = -10, kErrHeader = -11, kErrParam = -12, kErrRunLength = -17, kErrFilter = -18,
             kErrMaxBuf = -35, kErrBlocksize = -101, kErrCtxCodec = -102 };

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
  int32_t item_base;     // first work item (= block) of this chunk in the batch-wide item numbering
  int32_t nitems;        // work items: streams (blocks for special / memcpyed chunks); 0 for empty / failed chunks
  int32_t vts;           // special "repeated value" chunks: size of the value (cbytes - 32)
  int32_t nstreams;      // streams a full block is stored as (typesize when split, else 1)
  uint8_t version, flags, typesize, filter, fmt, dont_split, special, memcpyed, hdr, cls, serial, pad;
};

// Result of a batch, per chunk (device -> host).
struct DResult {
  int32_t out_bytes;
  int32_t status;
};

// Compress plan: one per chunk, written on the host (all geometry is known there).
struct CInfo {
  int32_t nbytes, blocksize, nblocks, leftover;
  int32_t item_base;
  int32_t memcpyed;      // clevel 0 or nbytes < 32: stored chunk
  int32_t dont_split;
  int32_t status;        // host-side validation result (0 = ok)
  int32_t nitems;        // stream work items of the chunk (0 for stored chunks); the chunk occupies
                         // max(nitems, nblocks, 1) item numbers so that k_finalize has one per block to copy
  int32_t max_stream;    // largest stream of the chunk in bytes (sizes the encoder window)
};

// Per-chunk running state of a compress batch (device, zero-initialised per batch).
struct CState {
  int32_t overflow;          // some block did not fit dst_capacity -> chunk becomes memcpyed
  int32_t nonzero_streams;   // streams that are not zero runs (0 -> SPECIAL_ZERO chunk)
  int32_t done_blocks;
  int32_t total;             // bytes used by the chunk when every block fitted
};

// Compress parameters shared by the whole batch.
struct CParams {
  uint8_t codec, clevel, typesize, filter;
};

// Batch-wide counters (device).
struct BatchCtl {
  int32_t total_items;   // written by the plan kernel (decode) or the host (compress)
  int32_t ticket;        // dynamic work distribution
  int32_t ring;          // plane staging: next slot to try (k_decode)
  int32_t paired;        // blocks written out interleaved from staged planes (statistics)
  int32_t pad2[4];
};

__device__ __forceinline__ int32_t ld_le32(const uint8_t* p) {
  return (int32_t)((uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24));
}
__device__ __forceinline__ void st_le32(uint8_t* p, int32_t v) {
  p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24);
}

// Transpose an 8x8 bit matrix held in a u64 (byte r = row r, bit c = column c)  ->  byte c = column c.
__device__ __forceinline__ uint64_t transpose8x8(uint64_t x) {
  uint64_t t;
  t = (x ^ (x >> 7)) & 0x00AA00AA00AA00AAull; x = x ^ t ^ (t << 7);
  t = (x ^ (x >> 14)) & 0x0000CCCC0000CCCCull; x = x ^ t ^ (t << 14);
  t = (x ^ (x >> 28)) & 0x00000000F0F0F0F0ull; x = x ^ t ^ (t << 28);
  return x;
}

__device__ __forceinline__ int lane_id() { return (int)(threadIdx.x & 31); }
__device__ __forceinline__ int warp_id() { return (int)(threadIdx.x >> 5); }

// Dynamic shared memory base of the running CTA.
#ifdef B2G_SIM
__device__ __forceinline__ uint8_t* dyn_smem() { return cusim::S().cur->cta->smem; }
#define B2G_SPIN_PAUSE() cusim::spin_yield()
#else
extern __shared__ __align__(1024) uint8_t b2g_dyn_smem_[];
__device__ __forceinline__ uint8_t* dyn_smem() { return b2g_dyn_smem_; }
#define B2G_SPIN_PAUSE() __nanosleep(64)
#endif

// Volatile / release-acquire accessors for cross-CTA flags.
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
#ifdef B2G_SIM
  return *(const volatile unsigned long long*)p;
#else
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
#endif
}
__device__ __forceinline__ void st_release_u64(unsigned long long* p, unsigned long long v) {
#ifdef B2G_SIM
  *(volatile unsigned long long*)p = v;
#else
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
#endif
}

// Record the first error of a chunk (0 = ok, errors are negative).
__device__ __forceinline__ void set_status(DResult* res, int chunk, int code) {
  if (code < 0) atomicCAS(&res[chunk].status, 0, code);
}

}  // namespace b2g
