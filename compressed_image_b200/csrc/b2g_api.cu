// b2g_api.cu -- host side of the C ABI declared in include/b2g.h: contexts, batch planning, stream-ordered
// staging and kernel launches.  No CPU codec lives here: every byte of (de)compression happens in the sm_100a
// kernels of b2g_kernels.cuh; without a usable CUDA device every data call fails with B2G_ERR_NO_DEVICE.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <climits>
#include <condition_variable>
#include <functional>
#include <thread>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/b2g.h"
#include "b2g_interleave.cuh"
#include "b2g_kernels.cuh"
#include "b2g_plan.h"

using namespace b2g;

namespace {

thread_local std::string g_last_error;

bool cuda_fail(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return false;
  g_last_error = std::string(what) + ": " + cudaGetErrorString(e);
  return true;
}
#define CU(call)                                  \
  do {                                            \
    if (cuda_fail((call), #call)) return B2G_ERR_NO_DEVICE; \
  } while (0)

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = std::max(bytes, (size_t)4096);
    want += want / 4;
    if (cuda_fail(cudaMalloc(&p, want), "cudaMalloc")) return B2G_ERR_NO_DEVICE;
    cap = want;
    return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
struct HostBuf {  // pinned
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    size_t want = std::max(bytes, (size_t)4096);
    want += want / 4;
    if (cuda_fail(cudaHostAlloc(&p, want, cudaHostAllocDefault), "cudaHostAlloc")) return B2G_ERR_NO_DEVICE;
    cap = want;
    return 0;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

constexpr int kMaxSmemWindow = 160 * 1024;  // larger blocks decode through the global slab

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

// One pipeline lane: a stream plus every per-batch work buffer, so that sub-batches of a host-pointer call can be in
// flight on different lanes at once (H2D of one overlapping the kernels of another and the D2H of a third).  The
// device-resident entry points use lane 0 on the caller's stream.
constexpr int kLanes = 16; // lanes a context owns; host-pointer batches rotate over the first g_lanes of them, the
                           // staged path gives worker thread t the lanes t and t + g_lanes
constexpr int kStagePieces = 8;
static const int g_lanes = [] { const char* e = getenv("B2G_LANES"); int v = e ? atoi(e) : 6; return v < 1 ? 1 : (v > kLanes / 2 ? kLanes / 2 : v); }();
struct Lane {
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};   // [0] descriptors uploaded, [1] plan done, [2] kernels done
  cudaEvent_t res_ready = nullptr;                   // per-chunk results landed in h_res
  DevBuf d_chunks, d_info, d_res, d_base, d_ctl, d_cinfo, d_state, d_items, d_scratch, d_slab, d_stage, d_pair, d_team;
  HostBuf h_chunks, h_res, h_cinfo, h_base;
  HostBuf h_in, h_out;                               // pinned bounce buffers of the staged (pageable-caller) path
  cudaEvent_t piece[kStagePieces] = {};              // D2H pieces of a staged sub-batch landed in h_out
  size_t st_out_lo = 0, st_piece = 0;                // staged sub-batch between its issue and its drain
  int st_npieces = 0, st_per = 1;
  std::vector<size_t> st_loc;
  void release() {
    for (DevBuf* b : {&d_chunks, &d_info, &d_res, &d_base, &d_ctl, &d_cinfo, &d_state, &d_items, &d_scratch, &d_slab, &d_stage, &d_pair, &d_team}) b->release();
    for (HostBuf* b : {&h_chunks, &h_res, &h_cinfo, &h_base, &h_in, &h_out}) b->release();
    for (auto& e : ev) if (e) cudaEventDestroy(e);
    for (auto& e : piece) if (e) cudaEventDestroy(e);
    if (res_ready) cudaEventDestroy(res_ready);
    if (stream) cudaStreamDestroy(stream);
  }
};

// Worker threads of a context, one per lane, for host-pointer batches whose buffers are PAGEABLE (what
// compressed::channel<T> hands over: std::span over a std::vector, wrapper.h:167-185, 236-259).  cudaMemcpyAsync on
// pageable memory is a synchronous bounce inside the driver at a fraction of the link speed; instead thread t takes the
// sub-batches t, t + T, ... of the batch and runs each one through its lane: memcpy of the inputs into the lane's pinned
// buffer (pipelined with the H2D copies of the pieces already there), kernels, D2H into the lane's pinned buffer,
// memcpy out to the caller's buffers.  T memcpy threads keep the host memory system as busy as the link.
struct HostPool {
  std::vector<std::thread> th;
  std::mutex m;
  std::condition_variable go, done;
  uint64_t gen = 0;
  int nsub = 0, running = 0, rc = 0;
  bool quit = false;
  std::function<int(int, int)> issue, drain;   // (sub-batch, lane) -> rc: everything up to the enqueued D2H copies /
                                               // the rest (wait, memcpy out to the caller)
};

struct b2g_ctx {
  b2g_params p{};
  int dev = 0, sms = 0;
  Lane lanes[kLanes];
  float last_ms[2] = {0, 0};
  std::atomic<int64_t> launches{0};
  // geometry
  int blocksize = 32768;   // context blocksize (compress) / largest block the decode path is sized for (slab, plan)
  int stream_hint = 32768; // stream size the decode windows are sized for on the device-resident path
  int win_bytes = 32768;   // shared-memory window (standalone filters)
  bool geometry_ready = false;
  bool pair_planes = true; // B2G_PAIR_PLANES=0 at creation: every plane scatters its own bytes
  // device + pinned work buffers
  DevBuf d_in, d_out;      // staging of the host-pointer calls (chunks / raw data of a whole batch)
  // in-flight batch (device-resident API, or an asynchronous host-pointer batch)
  b2g_chunk* pend = nullptr;
  int pend_n = 0;
  cudaStream_t pend_stream = nullptr;
  int pend_kind = 0;                 // 0 device-resident, 1 host decompress, 2 host compress, 3 staged (worker threads)
  HostPool* pool = nullptr;
  std::vector<size_t> st_in_off, st_out_off, st_in_len, st_out_len;   // staged batch: per chunk (caller's numbering)
  std::vector<int> pend_bounds;      // sub-batch boundaries of the host-pointer batch in flight
  std::vector<size_t> pend_out_off;  // compress: where each chunk sits in d_out
  std::vector<int> pend_order;       // batch position -> index into the caller's array (batches are processed in host
                                     // address order so that neighbouring buffers move with one copy)
  // host-pointer batches: pinned descriptor / plan / result arrays for the WHOLE batch (sub-batches use slices, so
  // nothing the copy engines still have to read is ever overwritten), one event per sub-batch, and a stream of its own
  // for the size-dependent D2H copies of compress
  HostBuf h_desc_all, h_res_all, h_cinfo_all, h_base_all;
  std::vector<cudaEvent_t> sub_done;
  cudaStream_t d2h_stream = nullptr;
};

namespace {

std::mutex g_attr_mutex;
size_t g_smem_limit[64][16];  // [device][kernel]

int raise_smem_limit(int dev, int which, const void* fn, size_t bytes) {
  std::lock_guard<std::mutex> lock(g_attr_mutex);
  size_t& cur = g_smem_limit[dev & 63][which];
  if (bytes <= cur) return 0;
  CU(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  cur = bytes;
  return 0;
}

size_t slab_stride(const b2g_ctx* c);

// Compress launch geometry for a batch whose largest stream is `max_stream` bytes: like decode, every warp is an
// independent worker (window + hash table + encoder work area), so the CTA shape only decides how many fit an SM.
constexpr int kMaxEncoderStream = 192 * 1024;  // window + 8 KiB table + work area must fit 227 KiB
struct CompGeom {
  int win, htab, warps, grid, tile, which;
  int team, prime;        // team > 1: k_compress_team, one stream per CTA of `team` warps
  size_t scratch_stride;  // per stream item
  const void* fn;
  size_t smem;
};
// Experiment switches for the encoder teams: B2G_TEAM = 0 / 1 one warp per stream for every batch, N > 1 teams of N;
// B2G_TEAM_PRIME = positions primed in front of a segment; B2G_TEAM_HLOG = log2 of the per-warp table entries.
static const int g_team_env = [] { const char* e = getenv("B2G_TEAM"); return e ? atoi(e) : -1; }();
static const int g_team_prime_env = [] { const char* e = getenv("B2G_TEAM_PRIME"); return e ? atoi(e) : -1; }();
static const int g_team_sparse_env = [] { const char* e = getenv("B2G_TEAM_SPARSE"); return e ? atoi(e) : -1; }();
static const int g_team_hlog_env = [] { const char* e = getenv("B2G_TEAM_HLOG"); return e ? atoi(e) : 0; }();

int compress_geometry(b2g_ctx* c, int max_stream, CompGeom& g) {
  g.win = std::max((int)align_up((size_t)std::max(max_stream, 1), 16), 1024);
  g.team = g_team_env < 0 ? plan_team_warps(max_stream) : std::min(std::max(g_team_env, 1), (int)kTeamMaxWarps);
  g.prime = 0;
  g.scratch_stride = (size_t)g.win;  // an encoded stream is kept only when it is smaller than the stream
  const bool lz4codec = c->p.codec == B2G_LZ4 || c->p.codec == B2G_LZ4HC;
  if (g.team > 1) {
    g.prime = g_team_prime_env < 0 ? kTeamPrime : (g_team_prime_env + 127) / 128 * 128;
    g.tile = kTeamTile;
    g.htab = g_team_hlog_env > 0 ? (2 << std::min(std::max(g_team_hlog_env, 6), 12)) : plan_htab_bytes(max_stream, c->p.clevel);
    const size_t per_warp = (size_t)g.htab + (size_t)enc_work_bytes(g.tile);
    while (g.team > 1 && (size_t)g.win + (size_t)g.team * per_warp + sizeof(TeamCtl) > 227 * 1024) g.team--;  // very large streams
  }
  if (g.team > 1) {
    g.warps = g.team;
    g.smem = (size_t)g.win + (size_t)g.team * ((size_t)g.htab + (size_t)enc_work_bytes(g.tile)) + sizeof(TeamCtl);
    g.which = lz4codec ? 4 : 5;
    g.fn = lz4codec ? (const void*)k_compress_team<kCodecLz4, kTeamTile> : (const void*)k_compress_team<kCodecBlosclz, kTeamTile>;
    if (int rc = raise_smem_limit(c->dev, 8 + g.which, g.fn, g.smem)) return rc;   // slots 12, 13
    int occ = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, g.fn, 32 * g.warps, g.smem));
    g.grid = c->sms * std::max(occ, 1);
    return 0;
  }
  g.htab = plan_htab_bytes(max_stream, c->p.clevel);
  g.tile = enc_tile_for(max_stream);
  const size_t per = (size_t)g.win + (size_t)g.htab + (size_t)enc_work_bytes(g.tile);
  int best = 0;
  g.warps = 1;
  for (int w = 1; w <= 8; w++) {  // k_compress is bounded to 256 threads
    const size_t cta = (size_t)w * per;
    if (cta > 227 * 1024) break;
    int n = (int)((228 * 1024) / (cta + 1024));
    n = std::min(n, std::min(32, 64 / w));
    if (w * n >= best) { best = w * n; g.warps = w; }
  }
  g.smem = (size_t)g.warps * per;
  if (g.smem > 227 * 1024) return B2G_ERR_BLOCKSIZE;
  // lz4hc chunks hold ordinary LZ4 block streams (format id 1, blosc2.c:2457): the one LZ4 encoder writes both
  const bool lz4 = c->p.codec == B2G_LZ4 || c->p.codec == B2G_LZ4HC;
  g.which = (lz4 ? 0 : 2) + (g.tile == 512 ? 0 : 1);
  g.fn = g.which == 0 ? (const void*)k_compress<kCodecLz4, 512> : g.which == 1 ? (const void*)k_compress<kCodecLz4, 1024>
       : g.which == 2 ? (const void*)k_compress<kCodecBlosclz, 512> : (const void*)k_compress<kCodecBlosclz, 1024>;
  if (int rc = raise_smem_limit(c->dev, 8 + g.which, g.fn, g.smem)) return rc;
  int occ = 0;
  CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, g.fn, 32 * g.warps, g.smem));
  g.grid = c->sms * std::max(occ, 1);
  const size_t max_warps = std::max((size_t)c->sms, ((size_t)1 << 30) / slab_stride(c));
  g.grid = (int)std::min<size_t>((size_t)g.grid, std::max<size_t>(max_warps / g.warps, 1));
  return 0;
}

// Largest blocksize seen so far: bounds the global-slab windows and the standalone-filter window.
int set_geometry(b2g_ctx* c, int blocksize) {
  if (c->geometry_ready && blocksize <= c->blocksize) return 0;
  c->blocksize = std::max(blocksize, 1);
  const int win = (int)align_up((size_t)std::min(c->blocksize, kMaxSmemWindow), 16);
  c->win_bytes = std::max(win, 1024);
  // The dynamic shared-memory limit is a per-function (per-device) attribute shared by every context of the process:
  // only ever raise it, or a context sized for small blocks would cut off one sized for large blocks.
  if (int rc = raise_smem_limit(c->dev, 3, (const void*)k_filter, (size_t)c->win_bytes)) return rc;
  c->geometry_ready = true;
  return 0;
}

size_t slab_stride(const b2g_ctx* c) {
  return align_up((size_t)decode_window_bytes(std::max(c->blocksize, c->p.blocksize), kFmtBlosclz) + 256, 256);
}

// Decode launch geometry for streams of `stream_bytes`: every warp is an independent worker with its own window, so
// the CTA shape only matters for how many warps fit an SM's shared memory (228 KiB, 1 KiB reserved per CTA).
struct DecGeom {
  int wcap, wstride, warps, grid, bulk;
  size_t smem;
};
// lz4 / blosclz streams are decoded in place in batches of up to 32 sequences, and a batch may only produce as many bytes
// as lie between its output and its own (staged) input.  A stream that barely compresses runs with that distance down at
// the window's slack, so the windows get kDecodeSlack bytes more than decoding needs; whatever shared memory the warp
// count then leaves over is handed out as well (16 KiB lz4 streams: 13 windows of 17872 bytes per SM).
constexpr int kDecodeSlack = 1024, kDecodeSlackMax = 4096;
int decode_geometry(b2g_ctx* c, int stream_bytes, bool zstd, DecGeom& g, bool any_blosclz = true) {
  if (stream_bytes > 96 * 1024) stream_bytes = 16 * 1024;  // such streams decode through the global slab anyway
  static const int slack = [] { const char* e = getenv("B2G_DECODE_SLACK"); return e ? std::max(atoi(e), 0) : kDecodeSlack; }();
  const int need = std::max(decode_window_bytes(std::max(stream_bytes, 1), any_blosclz ? kFmtBlosclz : kFmtLz4), 1024);
  const int extra = zstd ? (int)align_up(sizeof(ZWork), 16) : 0;
  g.wcap = need + (zstd ? 0 : slack);
  g.wstride = g.wcap + extra;
  if ((size_t)g.wstride > 227 * 1024) { g.wcap = need; g.wstride = need + extra; }
  int best = 0, ctas = 1;
  g.warps = 1;
  for (int w = 1; w <= 16; w++) {
    const size_t cta = (size_t)w * g.wstride;
    if (cta > 227 * 1024) break;
    int n = (int)((228 * 1024) / (cta + 1024));
    n = std::min(n, std::min(32, 64 / w));
    if (w * n >= best) { best = w * n; g.warps = w; ctas = n; }
  }
  if (!zstd) {  // the same warps, the largest windows
    const size_t per_cta = std::min<size_t>((228 * 1024) / ctas - 1024, 227 * 1024);
    const int room = (int)(per_cta / g.warps) & ~15;
    g.wcap = std::max(g.wcap, std::min(room, need + std::max(slack, kDecodeSlackMax)));
    g.wstride = g.wcap;
  }
  g.smem = (size_t)g.warps * g.wstride;
  static const bool bulk_on = [] { const char* e = getenv("B2G_BULK_STAGE"); return !(e && e[0] == '0'); }();
  g.bulk = (!zstd && bulk_on && g.smem + 8 * (size_t)g.warps <= 227 * 1024) ? 1 : 0;   // one mbarrier per warp behind the windows
  if (g.bulk) g.smem += 8 * (size_t)g.warps;
  const void* fn = zstd ? (const void*)k_decode<true> : (const void*)k_decode<false>;
  if (int rc = raise_smem_limit(c->dev, zstd ? 1 : 0, fn, g.smem)) return rc;
  int occ = 0;
  if (zstd) CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_decode<true>, 32 * g.warps, g.smem));
  else CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_decode<false>, 32 * g.warps, g.smem));
  g.grid = c->sms * std::max(occ, 1);
  // the slab holds one window per warp of the grid: keep it below 1 GiB for very large blocks
  const size_t max_warps = std::max((size_t)c->sms, ((size_t)1 << 30) / slab_stride(c));
  g.grid = (int)std::min<size_t>((size_t)g.grid, std::max<size_t>(max_warps / g.warps, 1));
  return 0;
}

// any_blosclz = false: the caller knows that no chunk of the batch is blosclz (the windows are then sized for lz4; a
// blosclz stream that shows up anyway decodes through the global slab)
int enqueue_decode(b2g_ctx* c, Lane& L, const DChunk* h_desc, int n, cudaStream_t s, bool want_zstd, int stream_bytes,
                   DResult* h_res = nullptr, bool any_blosclz = true) {
  DecGeom g, gz;
  if (int rc = decode_geometry(c, stream_bytes, false, g, any_blosclz)) return rc;
  gz = g;
  if (want_zstd) if (int rc = decode_geometry(c, stream_bytes, true, gz)) return rc;
  // slab: windows for streams larger than the shared-memory window, one per warp of the grid
  // (sized for THIS launch: a running maximum over earlier geometries would multiply a later large-block stride by a
  // warp count that belonged to small blocks)
  if (int rc = L.d_slab.ensure(slab_stride(c) * (size_t)std::max(g.grid * g.warps, gz.grid * gz.warps))) return rc;
  if (int rc = L.d_chunks.ensure(sizeof(DChunk) * (size_t)n)) return rc;
  if (int rc = L.d_info.ensure(sizeof(DInfo) * (size_t)n)) return rc;
  if (int rc = L.d_res.ensure(sizeof(DResult) * (size_t)n)) return rc;
  if (int rc = L.d_base.ensure(sizeof(int32_t) * 2 * ((size_t)n + 1))) return rc;
  if (int rc = L.d_ctl.ensure(sizeof(BatchCtl) * 2)) return rc;
  if (!h_res) {
    if (int rc = L.h_res.ensure(sizeof(DResult) * (size_t)n)) return rc;
    h_res = (DResult*)L.h_res.p;
  }
  // upper bound of the batch's stream items (sizes the plane-pairing mailboxes)
  int64_t item_cap = 0;
  const int hint_block = std::max(std::min(c->p.blocksize, c->blocksize), 1024);
  for (int i = 0; i < n; i++) item_cap += ((int64_t)std::max(h_desc[i].dst_cap, 0) / hint_block + 2) * kMaxStreams;
  // plane pairing (DecodeArgs::stage): a ring of staging slots, one flag per slot, a mailbox and a counter per item for
  // each of the two kernels; everything but the slots is zeroed per batch
  const int64_t pair_cap = c->pair_planes ? std::min<int64_t>(item_cap, (int64_t)1 << 24) : 0;
  const int nslots = 2048;
  const size_t slot_bytes = align_up((size_t)g.wcap, 128);   // whole 128-byte lines: the reader discards them from L2
  const size_t pair_ints = (size_t)nslots + 4 * (size_t)pair_cap;
  if (pair_cap > 0) {
    if (int rc = L.d_stage.ensure((size_t)nslots * slot_bytes)) return rc;
    if (int rc = L.d_pair.ensure(sizeof(int32_t) * pair_ints)) return rc;
    CU(cudaMemsetAsync(L.d_pair.p, 0, sizeof(int32_t) * pair_ints, s));
  }
  CU(cudaMemcpyAsync(L.d_chunks.p, h_desc, sizeof(DChunk) * (size_t)n, cudaMemcpyHostToDevice, s));
  CU(cudaEventRecord(L.ev[0], s));
  PlanArgs pa;
  pa.chunks = (const DChunk*)L.d_chunks.p;
  pa.info = (DInfo*)L.d_info.p;
  pa.res = (DResult*)L.d_res.p;
  pa.base = (int32_t*)L.d_base.p;
  pa.ctl = (BatchCtl*)L.d_ctl.p;
  pa.n = n;
  pa.max_block = std::max(c->blocksize, c->p.blocksize);
  pa.zstd_ok = want_zstd ? 1 : 0;
  k_plan_decode<<<1, kPlanThreads, kPlanSmemBytes, s>>>(pa);
  c->launches++;
  CU(cudaEventRecord(L.ev[1], s));
  DecodeArgs da;
  da.chunks = pa.chunks;
  da.info = pa.info;
  da.res = pa.res;
  da.base = pa.base;
  da.ctl = pa.ctl;
  da.n = n;
  da.wcap = g.wcap;
  da.wstride = g.wstride;
  da.slab = (uint8_t*)L.d_slab.p;
  da.slab_stride = (int64_t)slab_stride(c);
  da.stage = pair_cap > 0 ? (uint8_t*)L.d_stage.p : nullptr;
  da.slotflag = (int32_t*)L.d_pair.p;
  da.mail = da.slotflag + nslots;
  da.blkcnt = da.mail + 2 * pair_cap;
  da.nslots = nslots;
  da.slot_bytes = (int32_t)slot_bytes;
  da.pair_cap = (int32_t)pair_cap;
  da.bulk = g.bulk;
  k_decode<false><<<g.grid, 32 * g.warps, g.smem, s>>>(da);
  c->launches++;
  if (want_zstd) {
    da.base = pa.base + (n + 1);
    da.ctl = pa.ctl + 1;
    da.wcap = gz.wcap;
    da.wstride = gz.wstride;
    da.bulk = 0;
    da.mail += pair_cap;     // the zstd class numbers its items from 0 again
    da.blkcnt += pair_cap;
    k_decode<true><<<gz.grid, 32 * gz.warps, gz.smem, s>>>(da);
    c->launches++;
  }
  CU(cudaEventRecord(L.ev[2], s));
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(h_res, L.d_res.p, sizeof(DResult) * (size_t)n, cudaMemcpyDeviceToHost, s));
  CU(cudaEventRecord(L.res_ready, s));
  return 0;
}

// ci / base: pinned plan arrays of n / n + 1 entries (nullptr: the lane's own); h_res: where the results land
int enqueue_compress(b2g_ctx* c, Lane& L, const DChunk* h_desc, const int32_t* nbytes, int n, cudaStream_t s,
                     CInfo* ci = nullptr, int32_t* base = nullptr, DResult* h_res = nullptr) {
  const b2g_params& p = c->p;
  // encode: blosclz, lz4, and lz4hc as plain LZ4 streams under the lz4hc codec id (unsplit blocks like the reference's,
  // stune.c:200-209; the high-compression match search itself is not part of this build); no zstd encoder
  if (p.codec != B2G_BLOSCLZ && p.codec != B2G_LZ4 && p.codec != B2G_LZ4HC) return B2G_ERR_CODEC_SUPPORT;
  if (p.typesize == 0) return B2G_ERR_INVALID_PARAM;
  if (p.clevel > 9) return B2G_ERR_CODEC_PARAM;
  if (!ci) {
    if (int rc = L.h_cinfo.ensure(sizeof(CInfo) * (size_t)n)) return rc;
    if (int rc = L.h_base.ensure(sizeof(int32_t) * ((size_t)n + 1))) return rc;
    ci = (CInfo*)L.h_cinfo.p;
    base = (int32_t*)L.h_base.p;
  }
  int64_t items = 0;
  int max_stream = 1;
  CParams cp;
  cp.codec = p.codec; cp.clevel = p.clevel; cp.typesize = p.typesize; cp.filter = p.filter;
  for (int i = 0; i < n; i++) {
    const int nit = plan_compress_chunk(ci[i], cp, c->p.blocksize, nbytes[i], h_desc[i].dst_cap, kMaxEncoderStream);
    if (ci[i].status == 0) max_stream = std::max(max_stream, ci[i].max_stream);
    ci[i].item_base = (int32_t)items;
    base[i] = (int32_t)items;
    items += nit;
    if (items > INT32_MAX) return B2G_ERR_MAX_BUFSIZE;
  }
  base[n] = (int32_t)items;
  CompGeom g;
  if (int rc = compress_geometry(c, max_stream, g)) return rc;
  const size_t scratch_stride = g.scratch_stride;
  if (int rc = L.d_items.ensure(sizeof(CItem) * (size_t)items)) return rc;
  if (int rc = L.d_scratch.ensure(scratch_stride * (size_t)items + 256)) return rc;
  if (int rc = L.d_chunks.ensure(sizeof(DChunk) * (size_t)n)) return rc;
  if (int rc = L.d_cinfo.ensure(sizeof(CInfo) * (size_t)n)) return rc;
  if (int rc = L.d_state.ensure(sizeof(CState) * (size_t)n)) return rc;
  if (int rc = L.d_res.ensure(sizeof(DResult) * (size_t)n)) return rc;
  if (int rc = L.d_base.ensure(sizeof(int32_t) * ((size_t)n + 1))) return rc;
  if (int rc = L.d_ctl.ensure(sizeof(BatchCtl) * 2)) return rc;
  if (!h_res) {
    if (int rc = L.h_res.ensure(sizeof(DResult) * (size_t)n)) return rc;
    h_res = (DResult*)L.h_res.p;
  }
  CU(cudaMemcpyAsync(L.d_chunks.p, h_desc, sizeof(DChunk) * (size_t)n, cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(L.d_cinfo.p, ci, sizeof(CInfo) * (size_t)n, cudaMemcpyHostToDevice, s));
  CU(cudaMemcpyAsync(L.d_base.p, base, sizeof(int32_t) * ((size_t)n + 1), cudaMemcpyHostToDevice, s));
  CU(cudaEventRecord(L.ev[0], s));
  CU(cudaMemsetAsync(L.d_ctl.p, 0, sizeof(BatchCtl) * 2, s));
  CU(cudaMemcpyAsync(L.d_ctl.p, &base[n], sizeof(int32_t), cudaMemcpyHostToDevice, s));  // ctl[0].total_items
  CU(cudaEventRecord(L.ev[1], s));
  CompressArgs ca;
  ca.chunks = (const DChunk*)L.d_chunks.p;
  ca.info = (const CInfo*)L.d_cinfo.p;
  ca.state = (CState*)L.d_state.p;
  ca.base = (const int32_t*)L.d_base.p;
  ca.ctl = (BatchCtl*)L.d_ctl.p;
  ca.items = (CItem*)L.d_items.p;
  ca.scratch = (uint8_t*)L.d_scratch.p;
  ca.scratch_stride = (int64_t)scratch_stride;
  ca.n = n;
  ca.win_bytes = g.win;
  ca.htab_bytes = g.htab;
  ca.tile = g.tile;
  ca.team_prime = g.prime;
  ca.team_prime_step = g_team_sparse_env < 0 ? plan_team_prime_step(max_stream) : (g_team_sparse_env ? 4 : 1);
  ca.p = cp;
  const int cgrid = (int)std::min<int64_t>(g.grid, g.team > 1 ? items : (items + g.warps - 1) / g.warps);
  ca.team_bodies = nullptr;
  ca.team_body_bytes = 0;
  if (g.team > 1) {
    ca.team_body_bytes = (int64_t)plan_team_body_bytes(g.win);
    if (int rc = L.d_team.ensure((size_t)ca.team_body_bytes * (size_t)std::max(cgrid, 1) + 256)) return rc;
    ca.team_bodies = (uint8_t*)L.d_team.p;
  }
  {
    void* kargs[] = {(void*)&ca};
    CU(cudaLaunchKernel(g.fn, dim3((unsigned)cgrid), dim3((unsigned)(32 * g.warps)), kargs, g.smem, s));
  }
  c->launches++;
  PlaceArgs pl;
  pl.chunks = ca.chunks;
  pl.info = ca.info;
  pl.items = ca.items;
  pl.state = ca.state;
  pl.n = n;
  k_place<<<std::min(n, c->sms * 4), 256, 2 * 256 * sizeof(int32_t), s>>>(pl);
  c->launches++;
  FinalizeArgs fa;
  fa.chunks = ca.chunks;
  fa.info = ca.info;
  fa.state = ca.state;
  fa.items = ca.items;
  fa.scratch = ca.scratch;
  fa.scratch_stride = ca.scratch_stride;
  fa.res = (DResult*)L.d_res.p;
  fa.base = ca.base;
  fa.n = n;
  fa.total_items = (int32_t)items;
  fa.win_bytes = g.win;
  fa.p = ca.p;
  if (int rc = raise_smem_limit(c->dev, 4, (const void*)k_finalize, (size_t)g.win)) return rc;
  int focc = 0;
  CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&focc, k_finalize, kCtaThreads, (size_t)g.win));
  const int fgrid = (int)std::min<int64_t>((int64_t)c->sms * std::max(focc, 1), items);
  k_finalize<<<fgrid, kCtaThreads, (size_t)g.win, s>>>(fa);
  c->launches++;
  CU(cudaEventRecord(L.ev[2], s));
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(h_res, L.d_res.p, sizeof(DResult) * (size_t)n, cudaMemcpyDeviceToHost, s));
  CU(cudaEventRecord(L.res_ready, s));
  return 0;
}

int finish(b2g_ctx* c, Lane& L, b2g_chunk* chunks, int n) {
  CU(cudaEventSynchronize(L.res_ready));
  const DResult* r = (const DResult*)L.h_res.p;
  for (int i = 0; i < n; i++) {
    chunks[i].status = r[i].status;
    chunks[i].out_bytes = r[i].status < 0 ? 0 : r[i].out_bytes;
  }
  float a = 0, b = 0;
  if (cudaEventElapsedTime(&a, L.ev[0], L.ev[1]) == cudaSuccess) c->last_ms[0] += a;
  if (cudaEventElapsedTime(&b, L.ev[1], L.ev[2]) == cudaSuccess) c->last_ms[1] += b;
  static const bool timing = [] { const char* e = getenv("B2G_TIMING"); return e && e[0] == '1'; }();
  if (timing) {  // developer switch
    BatchCtl ctl[2];
    cudaMemcpy(ctl, L.d_ctl.p, sizeof(ctl), cudaMemcpyDeviceToHost);
    fprintf(stderr, "b2g timing: plan %.3f ms, kernels %.3f ms; items %d, blocks paired %d\n", a, b, ctl[0].total_items,
            ctl[0].paired);
  }
  return 0;
}

int use_device(b2g_ctx* c) {
  CU(cudaSetDevice(c->dev));
  return 0;
}

}  // namespace

// ---- interleave / deinterleave -------------------------------------------------------------------
namespace {

template <int TS, int NCH>
void launch_ilv_vec(bool to_interleaved, void* inter, const IlvPlanes4& pl, int64_t groups, int grid, cudaStream_t s) {
  if (to_interleaved) k_interleave_vec<TS, NCH><<<grid, kIlvThreads, 0, s>>>((uint4*)inter, pl, groups);
  else k_deinterleave_vec<TS, NCH><<<grid, kIlvThreads, 0, s>>>((const uint4*)inter, pl, groups);
}

template <int TS>
bool launch_ilv_vec_nch(int nch, bool to_interleaved, void* inter, const IlvPlanes4& pl, int64_t groups, int grid, cudaStream_t s) {
  switch (nch) {
    case 2: launch_ilv_vec<TS, 2>(to_interleaved, inter, pl, groups, grid, s); return true;
    case 3: launch_ilv_vec<TS, 3>(to_interleaved, inter, pl, groups, grid, s); return true;
    case 4: launch_ilv_vec<TS, 4>(to_interleaved, inter, pl, groups, grid, s); return true;
    default: return false;
  }
}

int interleave_device(b2g_ctx* c, bool to_interleaved, int typesize, int nch, void* const* planes, void* inter, int64_t n,
                      void* cuda_stream) {
  if (!c || !planes || typesize <= 0 || typesize > 255 || nch <= 0 || n < 0) return B2G_ERR_INVALID_PARAM;
  if (n == 0) return 0;
  if (!inter) return B2G_ERR_INVALID_PARAM;
  for (int k = 0; k < nch; k++)
    if (!planes[k]) return B2G_ERR_INVALID_PARAM;
  if (int rc = use_device(c)) return rc;
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->lanes[0].stream;
  const int64_t plane_bytes = n * typesize;
  if (nch == 1) {                                   // one channel: the two layouts are the same bytes
    if (to_interleaved) CU(cudaMemcpyAsync(inter, planes[0], (size_t)plane_bytes, cudaMemcpyDeviceToDevice, s));
    else CU(cudaMemcpyAsync(planes[0], inter, (size_t)plane_bytes, cudaMemcpyDeviceToDevice, s));
    return 0;
  }
  int64_t done = 0;                                 // elements handled by the vector kernel
  bool vec_ok = nch <= 4 && (typesize == 1 || typesize == 2 || typesize == 4 || typesize == 8) && ((uintptr_t)inter & 15) == 0;
  for (int k = 0; k < nch && vec_ok; k++) vec_ok = ((uintptr_t)planes[k] & 15) == 0;
  if (vec_ok && plane_bytes >= 16) {
    const int64_t groups = plane_bytes / 16;
    IlvPlanes4 pl{};
    for (int k = 0; k < nch; k++) pl.p[k] = planes[k];
    const int grid = (int)std::min<int64_t>((groups + kIlvThreads - 1) / kIlvThreads, (int64_t)c->sms * 16);
    bool ok = false;
    switch (typesize) {
      case 1: ok = launch_ilv_vec_nch<1>(nch, to_interleaved, inter, pl, groups, grid, s); break;
      case 2: ok = launch_ilv_vec_nch<2>(nch, to_interleaved, inter, pl, groups, grid, s); break;
      case 4: ok = launch_ilv_vec_nch<4>(nch, to_interleaved, inter, pl, groups, grid, s); break;
      case 8: ok = launch_ilv_vec_nch<8>(nch, to_interleaved, inter, pl, groups, grid, s); break;
    }
    if (ok) {
      c->launches++;
      CU(cudaGetLastError());
      done = groups * (16 / typesize);
    }
  }
  // any other channel count whose pixel is whole words: tiles through shared memory
  bool tile_ok = done == 0 && nch <= kIlvMaxPlanes && (typesize == 1 || typesize == 2 || typesize == 4 || typesize == 8) &&
                 (nch * typesize) % 4 == 0 && ((uintptr_t)inter & 3) == 0 && (nch * typesize / 4 | 1) * 32 <= kIlvTileWords;
  for (int k = 0; k < nch && tile_ok; k++) tile_ok = ((uintptr_t)planes[k] & 3) == 0;
  if (tile_ok && plane_bytes >= 4) {
    const int W = nch * typesize / 4, psw = W | 1;
    const int P = std::min(1024, kIlvTileWords / psw / 32 * 32);
    const int64_t nv = typesize >= 4 ? n : n / (4 / typesize) * (4 / typesize);
    IlvPlanes pl{};
    for (int k = 0; k < nch; k++) pl.p[k] = planes[k];
    const int grid = (int)std::min<int64_t>((nv + P - 1) / P, (int64_t)c->sms * 6);
#define B2G_ILV_TILE(TS)                                                                                                 \
  if (to_interleaved) k_interleave_tile<TS><<<grid, kIlvThreads, 0, s>>>((uint32_t*)inter, pl, nch, W, psw, P, nv);      \
  else k_deinterleave_tile<TS><<<grid, kIlvThreads, 0, s>>>((const uint32_t*)inter, pl, nch, W, psw, P, nv)
    switch (typesize) {
      case 1: B2G_ILV_TILE(1); break;
      case 2: B2G_ILV_TILE(2); break;
      case 4: B2G_ILV_TILE(4); break;
      default: B2G_ILV_TILE(8); break;
    }
#undef B2G_ILV_TILE
    c->launches++;
    CU(cudaGetLastError());
    done = nv;
  }
  if (done < n) {                                   // the rest (or everything), kIlvMaxPlanes channels per launch
    const int64_t rest = n - done;
    for (int c0 = 0; c0 < nch; c0 += kIlvMaxPlanes) {
      const int count = std::min(kIlvMaxPlanes, nch - c0);
      IlvPlanes pl{};
      for (int k = 0; k < count; k++) pl.p[k] = planes[c0 + k];
      if (to_interleaved) {
        const int64_t total = rest * count;
        const int grid = (int)std::min<int64_t>((total + kIlvThreads - 1) / kIlvThreads, (int64_t)c->sms * 16);
        k_interleave_elems<<<grid, kIlvThreads, 0, s>>>((uint8_t*)inter, pl, typesize, nch, c0, count, done, n);
      } else {
        const int gx = (int)std::min<int64_t>((rest + kIlvThreads - 1) / kIlvThreads, (int64_t)c->sms * 16);
        k_deinterleave_elems<<<dim3((unsigned)gx, (unsigned)count), kIlvThreads, 0, s>>>((const uint8_t*)inter, pl, typesize, nch,
                                                                                        c0, count, done, n);
      }
      c->launches++;
      CU(cudaGetLastError());
    }
  }
  return 0;
}

}  // namespace

extern "C" {

static int pool_wait(b2g_ctx* c);
static void pool_destroy(b2g_ctx* c);

b2g_ctx* b2g_create(const b2g_params* params) {
  if (!params) { g_last_error = "b2g_create: params == NULL"; return nullptr; }
  int ndev = 0;
  if (cuda_fail(cudaGetDeviceCount(&ndev), "cudaGetDeviceCount") || ndev == 0) {
    if (g_last_error.empty()) g_last_error = "no CUDA device";
    return nullptr;
  }
  b2g_ctx* c = new b2g_ctx();
  c->p = *params;
  { const char* e = getenv("B2G_PAIR_PLANES"); c->pair_planes = !(e && e[0] == '0'); }
  if (c->p.typesize == 0) c->p.typesize = 1;
  if (c->p.blocksize <= 0) c->p.blocksize = 32768;
  int dev = params->device;
  if (dev < 0 && cuda_fail(cudaGetDevice(&dev), "cudaGetDevice")) { delete c; return nullptr; }
  c->dev = dev;
  bool ok = !cuda_fail(cudaSetDevice(dev), "cudaSetDevice") &&
            !cuda_fail(cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, dev), "sm count");
  for (int l = 0; ok && l < kLanes; l++) {
    Lane& L = c->lanes[l];
    ok = !cuda_fail(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking), "cudaStreamCreate");
    for (int i = 0; ok && i < 3; i++) ok = !cuda_fail(cudaEventCreate(&L.ev[i]), "cudaEventCreate");
    if (ok) ok = !cuda_fail(cudaEventCreateWithFlags(&L.res_ready, cudaEventDisableTiming), "cudaEventCreate");
  }
  c->stream_hint = c->p.blocksize / plan_nstreams(c->p.codec, c->p.clevel, c->p.filter, c->p.typesize, c->p.blocksize);
  if (ok) ok = set_geometry(c, c->p.blocksize) == 0;
  if (!ok) { b2g_destroy(c); return nullptr; }
  return c;
}

void b2g_destroy(b2g_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->dev);
  if (c->pool) { if (c->pend && c->pend_kind == 3) pool_wait(c); pool_destroy(c); }
  for (Lane& L : c->lanes) if (L.stream) cudaStreamSynchronize(L.stream);
  for (Lane& L : c->lanes) L.release();
  c->d_in.release();
  c->d_out.release();
  for (HostBuf* b : {&c->h_desc_all, &c->h_res_all, &c->h_cinfo_all, &c->h_base_all}) b->release();
  for (cudaEvent_t e : c->sub_done) cudaEventDestroy(e);
  if (c->d2h_stream) { cudaStreamSynchronize(c->d2h_stream); cudaStreamDestroy(c->d2h_stream); }
  delete c;
}

int b2g_decompress_batch_device(b2g_ctx* c, b2g_chunk* chunks, int n, void* cuda_stream) {
  if (!c || (!chunks && n > 0) || n < 0) return B2G_ERR_INVALID_PARAM;
  if (c->pend) return B2G_ERR_FAILURE;
  if (int rc = use_device(c)) return rc;
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->lanes[0].stream;
  Lane& L = c->lanes[0];
  c->pend = chunks; c->pend_n = n; c->pend_stream = s;
  c->last_ms[0] = c->last_ms[1] = 0;
  if (n == 0) return 0;
  if (int rc = L.h_chunks.ensure(sizeof(DChunk) * (size_t)n)) { c->pend = nullptr; return rc; }
  DChunk* h = (DChunk*)L.h_chunks.p;
  for (int i = 0; i < n; i++) {
    h[i].src = (const uint8_t*)chunks[i].src;
    h[i].dst = (uint8_t*)chunks[i].dst;
    h[i].src_bytes = chunks[i].src_bytes;
    h[i].dst_cap = chunks[i].dst_capacity;
  }
  int rc = enqueue_decode(c, L, h, n, s, c->p.codec == B2G_ZSTD, c->stream_hint, nullptr,
                          c->p.codec != B2G_LZ4 && c->p.codec != B2G_LZ4HC);  // an lz4 context sizes its windows for lz4
  if (rc) c->pend = nullptr;
  return rc;
}

int b2g_compress_batch_device(b2g_ctx* c, b2g_chunk* chunks, int n, void* cuda_stream) {
  if (!c || (!chunks && n > 0) || n < 0) return B2G_ERR_INVALID_PARAM;
  if (c->pend) return B2G_ERR_FAILURE;
  if (int rc = use_device(c)) return rc;
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->lanes[0].stream;
  Lane& L = c->lanes[0];
  c->pend = chunks; c->pend_n = n; c->pend_stream = s;
  c->last_ms[0] = c->last_ms[1] = 0;
  if (n == 0) return 0;
  if (int rc = L.h_chunks.ensure(sizeof(DChunk) * (size_t)n + sizeof(int32_t) * (size_t)n)) { c->pend = nullptr; return rc; }
  DChunk* h = (DChunk*)L.h_chunks.p;
  int32_t* nb = (int32_t*)(h + n);
  for (int i = 0; i < n; i++) {
    h[i].src = (const uint8_t*)chunks[i].src;
    h[i].dst = (uint8_t*)chunks[i].dst;
    h[i].src_bytes = chunks[i].src_bytes;
    h[i].dst_cap = chunks[i].dst_capacity;
    nb[i] = chunks[i].src_bytes;
  }
  int rc = enqueue_compress(c, L, h, nb, n, s);
  if (rc) c->pend = nullptr;
  return rc;
}

static int wait_host_batch(b2g_ctx* c);

int b2g_wait(b2g_ctx* c) {
  if (!c) return B2G_ERR_INVALID_PARAM;
  if (!c->pend) return 0;
  if (int rc = use_device(c)) return rc;
  if (c->pend_kind != 0) return wait_host_batch(c);
  b2g_chunk* chunks = c->pend;
  int n = c->pend_n;
  c->pend = nullptr;
  if (n == 0) return 0;
  return finish(c, c->lanes[0], chunks, n);
}

int b2g_chunk_sizes(const void* chunk, int32_t* nbytes, int32_t* cbytes, int32_t* blocksize) {
  // blosc2_cbuffer_sizes (blosc2.c:3696-3712) + the header checks of read_chunk_header (blosc2.c:672-711)
  const uint8_t* s = (const uint8_t*)chunk;
  int32_t nb, bs, cb;
  memcpy(&nb, s + 4, 4); memcpy(&bs, s + 8, 4); memcpy(&cb, s + 12, 4);
  int rc = 0;
  if (s[0] > 5) rc = B2G_ERR_VERSION_SUPPORT;
  else if (cb < kHdrMin) rc = B2G_ERR_INVALID_HEADER;
  else if (bs <= 0 || (nb > 0 && bs > nb)) rc = B2G_ERR_INVALID_HEADER;
  else if (bs > kMaxBlocksize) rc = B2G_ERR_INVALID_HEADER;
  else if (s[3] == 0) rc = B2G_ERR_INVALID_HEADER;
  if (rc < 0) nb = cb = bs = 0;
  if (nbytes) *nbytes = nb;
  if (cbytes) *cbytes = cb;
  if (blocksize) *blocksize = bs;
  return rc;
}

// Host-pointer calls are pipelined: the batch is cut into sub-batches of about B2G_SUBBATCH_MB (default 32) MiB of
// payload which rotate over the context's lanes, so the H2D copies of one sub-batch overlap the kernels of the previous
// one and the D2H copies of the one before (PCIe is full duplex; the copy engines and the SMs run concurrently).
static std::vector<int> split_batch(const std::vector<size_t>& payload, bool staged = false) {
  static const size_t target_pinned = [] {
    const char* e = getenv("B2G_SUBBATCH_MB");
    long mb = e ? atol(e) : 32;
    return (size_t)std::max(mb, 1L) << 20;
  }();
  static const size_t target_staged = [] {   // worker threads copy through pinned buffers: smaller pieces pipeline better
    const char* e = getenv("B2G_STAGED_SUBBATCH_MB");
    long mb = e ? atol(e) : 12;
    return (size_t)std::max(mb, 1L) << 20;
  }();
  const size_t target = staged ? target_staged : target_pinned;
  std::vector<int> bounds{0};
  size_t acc = 0;
  for (int i = 0; i < (int)payload.size(); i++) {
    acc += payload[i];
    if (acc >= target || i + 1 - bounds.back() >= 65536) { bounds.push_back(i + 1); acc = 0; }
  }
  if (bounds.back() != (int)payload.size()) bounds.push_back((int)payload.size());
  return bounds;
}

// ---- host-pointer batches ------------------------------------------------------------------------------------------
// Everything that does not depend on a device result is enqueued up front: per sub-batch the H2D copies, the kernels and
// the copy of the per-chunk results, on the lane streams in rotation.  Decompress also enqueues its D2H copies (the sizes
// are in the headers); compress copies a chunk back once its size is known, from b2g_wait(), on a stream of its own so
// that those copies never queue behind the H2D copies of later sub-batches.
static void drain_lanes(b2g_ctx* c) {
  for (Lane& L : c->lanes) if (L.stream) cudaStreamSynchronize(L.stream);
  if (c->d2h_stream) cudaStreamSynchronize(c->d2h_stream);
}

// B2G_TRACE=1 (diagnosis): when each sub-batch of a pinned host-pointer batch completed on the device and when the host
// enqueued / collected it, in ms since the first traced call of the process, on stderr (DESIGN.md section 3 reads one).
static const bool g_trace = [] { const char* e = getenv("B2G_TRACE"); return e && e[0] == '1'; }();
static cudaEvent_t g_trace_origin = nullptr;
static std::chrono::steady_clock::time_point g_trace_host0;
static double trace_host_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - g_trace_host0).count(); }

static int ensure_sub_events(b2g_ctx* c, int nsub) {
  if (g_trace && !g_trace_origin) {
    CU(cudaEventCreate(&g_trace_origin));
    CU(cudaEventRecord(g_trace_origin, c->lanes[0].stream));
    CU(cudaEventSynchronize(g_trace_origin));
    g_trace_host0 = std::chrono::steady_clock::now();
  }
  while ((int)c->sub_done.size() < nsub) {
    cudaEvent_t e = nullptr;
    CU(cudaEventCreateWithFlags(&e, g_trace ? cudaEventDefault : cudaEventDisableTiming));
    c->sub_done.push_back(e);
  }
  if (!c->d2h_stream) CU(cudaStreamCreateWithFlags(&c->d2h_stream, cudaStreamNonBlocking));
  return 0;
}


// ---- staged host-pointer batches (pageable caller memory) --------------------------------------------------------------
static bool is_pinned(const void* p) {
  if (!p) return true;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// Does any buffer of the batch live in pageable memory?  (B2G_STAGED=1 / 0 forces the answer.)
static bool batch_needs_staging(const b2g_chunk* chunks, int n) {
  static const int forced = [] { const char* e = getenv("B2G_STAGED"); return e ? atoi(e) : -1; }();
  if (forced >= 0) return forced != 0;
  for (int i = 0; i < n; i++) {
    if (chunks[i].src_bytes > 0 && !is_pinned(chunks[i].src)) return true;
    if (chunks[i].dst_capacity > 0 && !is_pinned(chunks[i].dst)) return true;
  }
  return false;
}

static void pool_thread(b2g_ctx* c, int t, int nthreads) {
  cudaSetDevice(c->dev);
  HostPool& P = *c->pool;
  uint64_t seen = 0;
  for (;;) {
    std::function<int(int, int)> issue, drain;
    int nsub = 0;
    {
      std::unique_lock<std::mutex> lk(P.m);
      P.go.wait(lk, [&] { return P.quit || P.gen != seen; });
      if (P.quit) return;
      seen = P.gen;
      issue = P.issue;
      drain = P.drain;
      nsub = P.nsub;
    }
    // two lanes per thread: sub-batch k + T is on its way (upload, kernels, D2H) while sub-batch k is copied out
    int rc = 0, prev = -1, prev_lane = 0;
    for (int k = t; k < nsub; k += nthreads) {
      const int lane = t + ((k / nthreads) & 1) * nthreads;
      int r = issue(k, lane);
      if (prev >= 0) { const int r2 = drain(prev, prev_lane); if (r2 && !rc) rc = r2; }
      if (r) { if (!rc) rc = r; prev = -1; cudaStreamSynchronize(c->lanes[lane].stream); continue; }
      prev = k;
      prev_lane = lane;
    }
    if (prev >= 0) { const int r2 = drain(prev, prev_lane); if (r2 && !rc) rc = r2; }
    {
      std::lock_guard<std::mutex> lk(P.m);
      if (rc && !P.rc) P.rc = rc;
      if (--P.running == 0) P.done.notify_all();
    }
  }
}

static int pool_dispatch(b2g_ctx* c, int nsub, std::function<int(int, int)> issue, std::function<int(int, int)> drain) {
  if (!c->pool) {
    c->pool = new HostPool();
    for (int t = 0; t < g_lanes; t++) c->pool->th.emplace_back(pool_thread, c, t, g_lanes);
  }
  for (int l = 0; l < 2 * g_lanes; l++)
    for (auto& e : c->lanes[l].piece)
      if (!e) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming | cudaEventBlockingSync));
  HostPool& P = *c->pool;
  std::lock_guard<std::mutex> lk(P.m);
  P.issue = std::move(issue);
  P.drain = std::move(drain);
  P.nsub = nsub;
  P.running = (int)P.th.size();
  P.rc = 0;
  P.gen++;
  P.go.notify_all();
  return 0;
}

static int pool_wait(b2g_ctx* c) {
  HostPool& P = *c->pool;
  std::unique_lock<std::mutex> lk(P.m);
  P.done.wait(lk, [&] { return P.running == 0; });
  return P.rc;
}

static void pool_destroy(b2g_ctx* c) {
  if (!c->pool) return;
  {
    std::lock_guard<std::mutex> lk(c->pool->m);
    c->pool->quit = true;
    c->pool->go.notify_all();
  }
  for (auto& t : c->pool->th) t.join();
  delete c->pool;
  c->pool = nullptr;
}

// Caller memory -> the lane's pinned buffer -> device, the H2D copy of a piece overlapping the memcpy of the next one.
struct StagedUpload {
  Lane& L;
  uint8_t* d_base;      // device address of region offset 0
  size_t sent = 0, filled = 0;
  int rc = 0;
  static constexpr size_t kPiece = (size_t)2 << 20;
  void put(size_t off, const void* src, size_t len) {
    memcpy((uint8_t*)L.h_in.p + off, src, len);
    filled = off + len;
    if (filled - sent >= kPiece) flush();
  }
  void flush() {
    if (filled > sent && !rc &&
        cuda_fail(cudaMemcpyAsync(d_base + sent, (uint8_t*)L.h_in.p + sent, filled - sent, cudaMemcpyHostToDevice, L.stream),
                  "cudaMemcpyAsync"))
      rc = B2G_ERR_NO_DEVICE;
    sent = filled;
  }
};

int b2g_decompress_batch_async(b2g_ctx* c, b2g_chunk* chunks, int n, void* cuda_stream) {
  if (!c || (!chunks && n > 0) || n < 0) return B2G_ERR_INVALID_PARAM;
  if (c->pend) return B2G_ERR_FAILURE;
  if (n == 0) return 0;
  if (int rc = use_device(c)) return rc;
  (void)cuda_stream;  // host-pointer batches run on the context's own lane streams
  std::vector<size_t> in_off(n), out_off(n), in_len(n), out_len(n), payload(n);
  size_t in_total = 0, out_total = 0;
  int maxblock = 0, maxstream = 0;
  bool any_zstd = false, any_blosclz = false;
  for (int i = 0; i < n; i++) {
    const uint8_t* src = (const uint8_t*)chunks[i].src;
    int32_t avail = chunks[i].src_bytes;
    int32_t nb = 0, cb = 0, bs = 0;
    size_t take = 0, give = 0;
    if (src && avail >= kHdrMin) {
      memcpy(&nb, src + 4, 4); memcpy(&bs, src + 8, 4); memcpy(&cb, src + 12, 4);
      take = (size_t)std::max(kHdrMin, std::min(avail, cb < kHdrMin ? kHdrExt : cb));  // INT32_MAX: trust the header
      if (avail != INT32_MAX) take = std::min(take, (size_t)avail);
      if (nb > 0 && nb <= chunks[i].dst_capacity) give = (size_t)nb;
      // only a header that can decode at all may size the work buffers (plan_one rejects the others per chunk)
      if (bs > 0 && bs <= kMaxBlocksize && nb > 0 && bs <= nb && nb <= chunks[i].dst_capacity) {
        maxblock = std::max(maxblock, bs);
        const int ts = src[3];  // split blocks are stored as `typesize` streams (flag bit 4 clear)
        const bool split = !(src[2] & kFlagDontSplit) && !(src[2] & kFlagMemcpyed) && ts >= 1 && ts <= kMaxStreams;
        maxstream = std::max(maxstream, split ? bs / ts : bs);
      }
      if ((src[2] >> 5) == kFmtZstd) any_zstd = true;
      if ((src[2] >> 5) == kFmtBlosclz && !(src[2] & kFlagMemcpyed)) any_blosclz = true;
    } else if (src && avail > 0) {
      take = (size_t)avail;
    }
    in_off[i] = in_total; in_len[i] = take; in_total += align_up(take + 8, 256);
    out_off[i] = out_total; out_len[i] = give; out_total += align_up(give + 8, 256);
    payload[i] = take + give;
  }
  // Process the batch in the order of the destination addresses: outputs that are neighbours in host memory (the chunks
  // of a channel decoded into one array) become neighbours in d_out and go back with ONE copy per sub-batch.
  std::vector<int>& ord = c->pend_order;
  ord.resize(n);
  for (int i = 0; i < n; i++) ord[i] = i;
  std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) { return (uintptr_t)chunks[a].dst < (uintptr_t)chunks[b].dst; });
  in_total = out_total = 0;
  std::vector<size_t> pay(n);
  for (int k = 0; k < n; k++) {
    const int i = ord[k];
    in_off[i] = in_total; in_total += align_up(in_len[i] + 8, 256);
    const int p = k ? ord[k - 1] : -1;
    if (p >= 0 && out_len[p] && out_len[i] && (const uint8_t*)chunks[p].dst + out_len[p] == (const uint8_t*)chunks[i].dst)
      out_off[i] = out_off[p] + out_len[p];
    else
      out_off[i] = align_up(out_total, 256);
    out_total = out_off[i] + out_len[i] + 8;
    pay[k] = payload[i];
  }
  if (int rc = set_geometry(c, std::max(maxblock, c->blocksize))) return rc;
  if (int rc = c->d_in.ensure(in_total + 256)) return rc;
  if (int rc = c->d_out.ensure(out_total + 512)) return rc;
  if (int rc = c->h_desc_all.ensure(sizeof(DChunk) * (size_t)n)) return rc;
  if (int rc = c->h_res_all.ensure(sizeof(DResult) * (size_t)n)) return rc;
  c->last_ms[0] = c->last_ms[1] = 0;
  bool staged = batch_needs_staging(chunks, n);
  c->pend_bounds = split_batch(pay, staged);
  if (staged && c->pend_bounds.size() <= 2) {   // one sub-batch: nothing to overlap, the worker hand-off only adds latency
    staged = false;
    c->pend_bounds = split_batch(pay, false);
  }
  const int nsub = (int)c->pend_bounds.size() - 1;
  if (int rc = ensure_sub_events(c, nsub)) return rc;
  DChunk* desc = (DChunk*)c->h_desc_all.p;
  DResult* res = (DResult*)c->h_res_all.p;
  if (staged) {
    // pageable caller memory: every sub-batch runs on a worker thread through its lane's pinned buffers
    c->st_in_off = in_off; c->st_out_off = out_off; c->st_in_len = in_len; c->st_out_len = out_len;
    const int zstd = any_zstd, blz = any_blosclz, mstream = std::max(maxstream, 1);
    auto issue = [c, chunks, desc, res, zstd, blz, mstream](int k, int lane) -> int {
      Lane& L = c->lanes[lane];
      const std::vector<int>& ord = c->pend_order;
      const std::vector<size_t>&in_off = c->st_in_off, &out_off = c->st_out_off, &in_len = c->st_in_len, &out_len = c->st_out_len;
      const int lo = c->pend_bounds[k], cnt = c->pend_bounds[k + 1] - lo;
      cudaStream_t s = L.stream;
      const size_t in_lo = in_off[ord[lo]];
      const size_t in_hi = in_off[ord[lo + cnt - 1]] + in_len[ord[lo + cnt - 1]];
      if (int rc = L.h_in.ensure(in_hi - in_lo + 16)) return rc;
      StagedUpload up{L, (uint8_t*)c->d_in.p + in_lo};
      size_t out_lo = (size_t)-1, out_hi = 0;
      for (int q = lo; q < lo + cnt; q++) {
        const int i = ord[q];
        DChunk& d = desc[q];
        d.src = (const uint8_t*)c->d_in.p + in_off[i];
        d.dst = (uint8_t*)c->d_out.p + out_off[i];
        d.src_bytes = chunks[i].src_bytes == INT32_MAX ? (int32_t)in_len[i]
                                                       : std::min<int32_t>(chunks[i].src_bytes, (int32_t)in_len[i]);
        d.dst_cap = chunks[i].dst_capacity;
        if (in_len[i]) up.put(in_off[i] - in_lo, chunks[i].src, in_len[i]);
        if (out_len[i]) { out_lo = std::min(out_lo, out_off[i]); out_hi = std::max(out_hi, out_off[i] + out_len[i]); }
      }
      up.flush();
      if (up.rc) return up.rc;
      if (int rc = enqueue_decode(c, L, desc + lo, cnt, s, zstd, mstream, res + lo, blz)) return rc;
      L.st_out_lo = out_lo;
      L.st_piece = 1;
      L.st_npieces = 0;
      if (out_hi > out_lo) {
        const size_t span = out_hi - out_lo;
        if (int rc = L.h_out.ensure(span + 16)) return rc;
        L.st_piece = std::max<size_t>((size_t)4 << 20, align_up((span + kStagePieces - 1) / kStagePieces, 4096));
        for (size_t o = 0; o < span; o += L.st_piece, L.st_npieces++) {
          CU(cudaMemcpyAsync((uint8_t*)L.h_out.p + o, (uint8_t*)c->d_out.p + out_lo + o, std::min(L.st_piece, span - o),
                             cudaMemcpyDeviceToHost, s));
          CU(cudaEventRecord(L.piece[L.st_npieces], s));
        }
      }
      return 0;
    };
    auto drain = [c, chunks, res](int k, int lane) -> int {
      Lane& L = c->lanes[lane];
      const std::vector<int>& ord = c->pend_order;
      const std::vector<size_t>&out_off = c->st_out_off, &out_len = c->st_out_len;
      const int lo = c->pend_bounds[k], cnt = c->pend_bounds[k + 1] - lo;
      CU(cudaEventSynchronize(L.res_ready));
      int waited = 0;
      for (int q = lo; q < lo + cnt; q++) {
        const int i = ord[q];
        chunks[i].status = res[q].status;
        chunks[i].out_bytes = res[q].status < 0 ? 0 : res[q].out_bytes;
        if (!out_len[i]) continue;
        const size_t o = out_off[i] - L.st_out_lo;
        const int last = (int)((o + out_len[i] - 1) / L.st_piece);
        for (; waited <= last && waited < L.st_npieces; waited++) CU(cudaEventSynchronize(L.piece[waited]));
        memcpy(chunks[i].dst, (uint8_t*)L.h_out.p + o, out_len[i]);
      }
      CU(cudaStreamSynchronize(L.stream));
      return 0;
    };
    if (int rc = pool_dispatch(c, nsub, issue, drain)) return rc;
    c->pend = chunks; c->pend_n = n; c->pend_kind = 3;
    return 0;
  }
  // An error after the first enqueue must not leave copies and kernels in flight on the caller's buffers and the staging
  // areas: the whole loop runs in a lambda, and a failure drains every lane before it is reported.
  auto enqueue_all = [&]() -> int {
  for (int k = 0; k < nsub; k++) {
    Lane& L = c->lanes[k % g_lanes];
    const int lo = c->pend_bounds[k], cnt = c->pend_bounds[k + 1] - lo;
    cudaStream_t s = L.stream;
    for (int q = lo; q < lo + cnt; q++) {
      const int i = ord[q];
      DChunk& d = desc[q];
      d.src = (const uint8_t*)c->d_in.p + in_off[i];
      d.dst = (uint8_t*)c->d_out.p + out_off[i];
      d.src_bytes = chunks[i].src_bytes == INT32_MAX ? (int32_t)in_len[i]
                                                     : std::min<int32_t>(chunks[i].src_bytes, (int32_t)in_len[i]);
      d.dst_cap = chunks[i].dst_capacity;
      if (in_len[i]) CU(cudaMemcpyAsync((void*)d.src, chunks[i].src, in_len[i], cudaMemcpyHostToDevice, s));
    }
    if (int rc = enqueue_decode(c, L, desc + lo, cnt, s, any_zstd, std::max(maxstream, 1), res + lo, any_blosclz)) return rc;
    for (int q = lo; q < lo + cnt;) {  // one copy per run of neighbouring outputs
      const int i = ord[q];
      size_t len = out_len[i];
      int e = q + 1;
      while (e < lo + cnt && out_len[ord[e]] && len && out_off[ord[e]] == out_off[i] + len &&
             (uint8_t*)chunks[ord[e]].dst == (uint8_t*)chunks[i].dst + len) {
        len += out_len[ord[e]];
        e++;
      }
      if (len) CU(cudaMemcpyAsync(chunks[i].dst, desc[q].dst, len, cudaMemcpyDeviceToHost, s));
      q = e;
    }
    CU(cudaEventRecord(c->sub_done[k], s));
  }
  return 0;
  };
  const double trace_t0 = g_trace ? trace_host_ms() : 0;
  if (int rc = enqueue_all()) { drain_lanes(c); return rc; }
  if (g_trace) fprintf(stderr, "b2g-trace ctx=%p enqueue host=%.3f .. %.3f (%d sub-batches)\n", (void*)c, trace_t0, trace_host_ms(), nsub);
  c->pend = chunks; c->pend_n = n; c->pend_kind = 1;
  return 0;
}

int b2g_compress_batch_async(b2g_ctx* c, b2g_chunk* chunks, int n, void* cuda_stream) {
  if (!c || (!chunks && n > 0) || n < 0) return B2G_ERR_INVALID_PARAM;
  if (c->pend) return B2G_ERR_FAILURE;
  if (n == 0) return 0;
  if (int rc = use_device(c)) return rc;
  (void)cuda_stream;
  // Process the batch in the order of the source addresses: raw chunks that are neighbours in host memory (a channel
  // handed over as one array) become neighbours in d_in and are uploaded with ONE copy per sub-batch.
  std::vector<int>& ord = c->pend_order;
  ord.resize(n);
  for (int i = 0; i < n; i++) ord[i] = i;
  std::stable_sort(ord.begin(), ord.end(), [&](int a, int b) { return (uintptr_t)chunks[a].src < (uintptr_t)chunks[b].src; });
  std::vector<size_t> in_off(n), payload(n);
  c->pend_out_off.assign(n, 0);
  size_t in_total = 0, out_total = 0;
  for (int k = 0; k < n; k++) {
    const int i = ord[k];
    const size_t sz = (size_t)std::max(chunks[i].src_bytes, 0);
    const int32_t cap = std::max(std::min<int64_t>(chunks[i].dst_capacity, (int64_t)sz + B2G_MAX_OVERHEAD), (int64_t)0);
    const int p = k ? ord[k - 1] : -1;
    const size_t psz = p >= 0 ? (size_t)std::max(chunks[p].src_bytes, 0) : 0;
    if (p >= 0 && psz && sz && (const uint8_t*)chunks[p].src + psz == (const uint8_t*)chunks[i].src)
      in_off[i] = in_off[p] + psz;
    else
      in_off[i] = align_up(in_total, 256);
    in_total = in_off[i] + sz + 8;
    c->pend_out_off[i] = out_total; out_total += align_up((size_t)cap + 8, 256);
    payload[k] = sz;
  }
  if (int rc = c->d_in.ensure(in_total + 512)) return rc;
  if (int rc = c->d_out.ensure(out_total + 256)) return rc;
  bool staged = batch_needs_staging(chunks, n);
  c->pend_bounds = split_batch(payload, staged);
  if (staged && c->pend_bounds.size() <= 2) {
    staged = false;
    c->pend_bounds = split_batch(payload, false);
  }
  const int nsub = (int)c->pend_bounds.size() - 1;
  if (int rc = c->h_desc_all.ensure((sizeof(DChunk) + sizeof(int32_t)) * (size_t)n)) return rc;
  if (int rc = c->h_res_all.ensure(sizeof(DResult) * (size_t)n)) return rc;
  if (int rc = c->h_cinfo_all.ensure(sizeof(CInfo) * (size_t)n)) return rc;
  if (int rc = c->h_base_all.ensure(sizeof(int32_t) * ((size_t)n + (size_t)nsub))) return rc;
  if (int rc = ensure_sub_events(c, nsub)) return rc;
  c->last_ms[0] = c->last_ms[1] = 0;
  DChunk* desc = (DChunk*)c->h_desc_all.p;
  int32_t* nb = (int32_t*)(desc + n);
  DResult* res = (DResult*)c->h_res_all.p;
  CInfo* ci = (CInfo*)c->h_cinfo_all.p;
  int32_t* base = (int32_t*)c->h_base_all.p;
  if (staged) {
    c->st_in_off = in_off;
    auto issue = [c, chunks, desc, nb, res, ci, base](int k, int lane) -> int {
      Lane& L = c->lanes[lane];
      const std::vector<int>& ord = c->pend_order;
      const std::vector<size_t>& in_off = c->st_in_off;
      const int lo = c->pend_bounds[k], cnt = c->pend_bounds[k + 1] - lo;
      cudaStream_t s = L.stream;
      const size_t in_lo = in_off[ord[lo]];
      const size_t in_hi = in_off[ord[lo + cnt - 1]] + (size_t)std::max(chunks[ord[lo + cnt - 1]].src_bytes, 0);
      if (int rc = L.h_in.ensure(in_hi - in_lo + 16)) return rc;
      StagedUpload up{L, (uint8_t*)c->d_in.p + in_lo};
      size_t out_cap = 0;
      for (int q = lo; q < lo + cnt; q++) {
        const int i = ord[q];
        DChunk& d = desc[q];
        d.src = (const uint8_t*)c->d_in.p + in_off[i];
        d.dst = (uint8_t*)c->d_out.p + c->pend_out_off[i];
        d.src_bytes = chunks[i].src_bytes;
        d.dst_cap = std::min(chunks[i].dst_capacity, std::max(chunks[i].src_bytes, 0) + B2G_MAX_OVERHEAD);
        nb[q] = chunks[i].src_bytes;
        if (chunks[i].src_bytes > 0) up.put(in_off[i] - in_lo, chunks[i].src, (size_t)chunks[i].src_bytes);
        out_cap += align_up((size_t)std::max(d.dst_cap, 0) + 16, 16);
      }
      up.flush();
      if (up.rc) return up.rc;
      if (int rc = enqueue_compress(c, L, desc + lo, nb + lo, cnt, s, ci + lo, base + lo + k, res + lo)) return rc;
      return L.h_out.ensure(out_cap + 16);
    };
    auto drain = [c, chunks, res](int k, int lane) -> int {
      Lane& L = c->lanes[lane];
      const std::vector<int>& ord = c->pend_order;
      const int lo = c->pend_bounds[k], cnt = c->pend_bounds[k + 1] - lo;
      cudaStream_t s = L.stream;
      CU(cudaEventSynchronize(L.res_ready));             // the sizes are known: fetch the chunks, densely packed
      size_t off = 0;
      int np = 0;                                         // an event every ~1/8 of the chunks: copy out while the rest lands
      const int per = std::max(1, (cnt + kStagePieces - 1) / kStagePieces);
      L.st_loc.resize((size_t)cnt);
      for (int q = lo; q < lo + cnt; q++) {
        const int i = ord[q];
        chunks[i].status = res[q].status;
        chunks[i].out_bytes = res[q].status < 0 ? 0 : res[q].out_bytes;
        L.st_loc[(size_t)(q - lo)] = off;
        if (chunks[i].out_bytes > 0) {
          CU(cudaMemcpyAsync((uint8_t*)L.h_out.p + off, (uint8_t*)c->d_out.p + c->pend_out_off[i], (size_t)chunks[i].out_bytes,
                             cudaMemcpyDeviceToHost, s));
          off += align_up((size_t)chunks[i].out_bytes, 16);
        }
        if ((q - lo + 1) % per == 0 || q + 1 == lo + cnt) CU(cudaEventRecord(L.piece[np++], s));
      }
      int waited = 0;
      for (int q = lo; q < lo + cnt; q++) {
        const int i = ord[q];
        const int need = std::min((q - lo) / per, np - 1);
        for (; waited <= need; waited++) CU(cudaEventSynchronize(L.piece[waited]));
        if (chunks[i].out_bytes > 0)
          memcpy(chunks[i].dst, (uint8_t*)L.h_out.p + L.st_loc[(size_t)(q - lo)], (size_t)chunks[i].out_bytes);
      }
      CU(cudaStreamSynchronize(s));
      return 0;
    };
    if (int rc = pool_dispatch(c, nsub, issue, drain)) return rc;
    c->pend = chunks; c->pend_n = n; c->pend_kind = 3;
    return 0;
  }
  // An error after the first enqueue must not leave copies and kernels in flight on the caller's buffers and the staging
  // areas: the whole loop runs in a lambda, and a failure drains every lane before it is reported.
  auto enqueue_all = [&]() -> int {
  for (int k = 0; k < nsub; k++) {
    Lane& L = c->lanes[k % g_lanes];
    const int lo = c->pend_bounds[k], cnt = c->pend_bounds[k + 1] - lo;
    cudaStream_t s = L.stream;
    for (int q = lo; q < lo + cnt; q++) {
      const int i = ord[q];
      DChunk& d = desc[q];
      d.src = (const uint8_t*)c->d_in.p + in_off[i];
      d.dst = (uint8_t*)c->d_out.p + c->pend_out_off[i];
      d.src_bytes = chunks[i].src_bytes;
      d.dst_cap = std::min(chunks[i].dst_capacity, std::max(chunks[i].src_bytes, 0) + B2G_MAX_OVERHEAD);
      nb[q] = chunks[i].src_bytes;
    }
    for (int q = lo; q < lo + cnt;) {  // one copy per run of neighbouring inputs
      const int i = ord[q];
      size_t len = (size_t)std::max(chunks[i].src_bytes, 0);
      int e = q + 1;
      while (e < lo + cnt && len && chunks[ord[e]].src_bytes > 0 && in_off[ord[e]] == in_off[i] + len &&
             (const uint8_t*)chunks[ord[e]].src == (const uint8_t*)chunks[i].src + len) {
        len += (size_t)chunks[ord[e]].src_bytes;
        e++;
      }
      if (len) CU(cudaMemcpyAsync((void*)desc[q].src, chunks[i].src, len, cudaMemcpyHostToDevice, s));
      q = e;
    }
    if (int rc = enqueue_compress(c, L, desc + lo, nb + lo, cnt, s, ci + lo, base + lo + k, res + lo)) return rc;
    CU(cudaEventRecord(c->sub_done[k], s));
  }
  return 0;
  };
  const double trace_t0 = g_trace ? trace_host_ms() : 0;
  if (int rc = enqueue_all()) { drain_lanes(c); return rc; }
  if (g_trace) fprintf(stderr, "b2g-trace ctx=%p enqueue host=%.3f .. %.3f (%d sub-batches)\n", (void*)c, trace_t0, trace_host_ms(), nsub);
  c->pend = chunks; c->pend_n = n; c->pend_kind = 2;
  return 0;
}

// Completes a host-pointer batch: per sub-batch, in order, the results; for compress the chunks are copied back as soon
// as their sizes are known.
static int wait_host_batch(b2g_ctx* c) {
  b2g_chunk* chunks = c->pend;
  const int kind = c->pend_kind;
  c->pend = nullptr;
  c->pend_kind = 0;
  if (kind == 3) return pool_wait(c);   // staged: the worker threads have filled status / out_bytes
  const int nsub = (int)c->pend_bounds.size() - 1;
  const DResult* r = (const DResult*)c->h_res_all.p;
  int rc_all = 0;
  if (g_trace) fprintf(stderr, "b2g-trace ctx=%p kind=%d wait begins host=%.3f\n", (void*)c, kind, trace_host_ms());
  // Sub-batches finish in whatever order the copy engines served their lanes (traced: 0, 4, 2, 6, 1, 5, 3, 7).  A compress
  // batch is collected in COMPLETION order, so that a sub-batch's chunk copies are issued when its sizes arrive and not
  // when every earlier-numbered one has arrived too; a decompress batch has nothing to issue and is read in order.
  std::vector<char> seen((size_t)nsub, 0);
  for (int done = 0; done < nsub; done++) {
    int k = -1;
    if (kind == 2) {
      for (;;) {
        bool failed = false;
        for (int j = 0; j < nsub && k < 0; j++) {
          if (seen[(size_t)j]) continue;
          const cudaError_t q = cudaEventQuery(c->sub_done[j]);
          if (q == cudaSuccess) k = j;
          else if (q != cudaErrorNotReady) { failed = cuda_fail(q, "cudaEventQuery"); k = j; }
        }
        if (failed) rc_all = B2G_ERR_NO_DEVICE;
        if (k >= 0) break;
        std::this_thread::yield();
      }
      if (rc_all) break;
    } else {
      k = done;
      if (cuda_fail(cudaEventSynchronize(c->sub_done[k]), "cudaEventSynchronize")) { rc_all = B2G_ERR_NO_DEVICE; break; }
    }
    seen[(size_t)k] = 1;
    if (g_trace) {
      float ms = 0;
      cudaEventElapsedTime(&ms, g_trace_origin, c->sub_done[k]);
      fprintf(stderr, "b2g-trace ctx=%p kind=%d sub=%d done_dev=%.3f seen_host=%.3f\n", (void*)c, kind, k, ms, trace_host_ms());
    }
    for (int q = c->pend_bounds[k]; q < c->pend_bounds[k + 1]; q++) {
      const int i = c->pend_order[q];
      chunks[i].status = r[q].status;
      chunks[i].out_bytes = r[q].status < 0 ? 0 : r[q].out_bytes;
      if (kind == 2 && chunks[i].out_bytes > 0 &&
          cuda_fail(cudaMemcpyAsync(chunks[i].dst, (uint8_t*)c->d_out.p + c->pend_out_off[i], (size_t)chunks[i].out_bytes,
                                    cudaMemcpyDeviceToHost, c->d2h_stream), "cudaMemcpyAsync"))
        rc_all = B2G_ERR_NO_DEVICE;
    }
  }
  for (Lane& L : c->lanes) if (cuda_fail(cudaStreamSynchronize(L.stream), "cudaStreamSynchronize")) rc_all = B2G_ERR_NO_DEVICE;
  if (c->d2h_stream && cuda_fail(cudaStreamSynchronize(c->d2h_stream), "cudaStreamSynchronize")) rc_all = B2G_ERR_NO_DEVICE;
  if (g_trace) fprintf(stderr, "b2g-trace ctx=%p kind=%d wait ends host=%.3f\n", (void*)c, kind, trace_host_ms());
  return rc_all;
}

int b2g_decompress_batch(b2g_ctx* c, b2g_chunk* chunks, int n, void* cuda_stream) {
  if (int rc = b2g_decompress_batch_async(c, chunks, n, cuda_stream)) return rc;
  return b2g_wait(c);
}

int b2g_compress_batch(b2g_ctx* c, b2g_chunk* chunks, int n, void* cuda_stream) {
  if (int rc = b2g_compress_batch_async(c, chunks, n, cuda_stream)) return rc;
  return b2g_wait(c);
}

int b2g_filter_device(b2g_ctx* c, int filter, int forward, int typesize, int format_version, const void* src,
                      void* dst, int64_t n, int32_t blocksize, void* cuda_stream) {
  if (!c || n < 0 || blocksize <= 0 || typesize <= 0 || typesize > 255) return B2G_ERR_INVALID_PARAM;
  if (filter < 0 || filter > 2) return B2G_ERR_FILTER_PIPELINE;
  if (n == 0) return 0;
  if (int rc = use_device(c)) return rc;
  if (int rc = set_geometry(c, blocksize)) return rc;
  if (blocksize > c->win_bytes) return B2G_ERR_BLOCKSIZE;
  cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : c->lanes[0].stream;
  FilterArgs fa;
  fa.src = (const uint8_t*)src;
  fa.dst = (uint8_t*)dst;
  fa.n = n;
  fa.blocksize = blocksize;
  fa.typesize = typesize;
  fa.filter = filter;
  fa.forward = forward;
  fa.version = format_version;
  const int64_t nblocks = (n + blocksize - 1) / blocksize;
  if (nblocks > INT32_MAX) return B2G_ERR_MAX_BUFSIZE;
  fa.nblocks = (int32_t)nblocks;
  const int grid = (int)std::min<int64_t>(nblocks, (int64_t)c->sms * 8);
  k_filter<<<grid, kCtaThreads, (size_t)c->win_bytes, s>>>(fa);
  c->launches++;
  CU(cudaGetLastError());
  return 0;
}

int b2g_interleave_device(b2g_ctx* c, int typesize, int nchannels, const void* const* planes, void* interleaved,
                          int64_t n_elems, void* cuda_stream) {
  return interleave_device(c, true, typesize, nchannels, (void* const*)planes, interleaved, n_elems, cuda_stream);
}

int b2g_deinterleave_device(b2g_ctx* c, int typesize, int nchannels, const void* interleaved, void* const* planes,
                            int64_t n_elems, void* cuda_stream) {
  return interleave_device(c, false, typesize, nchannels, planes, (void*)interleaved, n_elems, cuda_stream);
}

// Host-pointer form: the planes and the interleaved buffer are staged through the context's device buffers.
static int interleave_host(b2g_ctx* c, bool to_interleaved, int typesize, int nch, void* const* planes, void* inter, int64_t n) {
  if (!c || !planes || typesize <= 0 || typesize > 255 || nch <= 0 || n < 0) return B2G_ERR_INVALID_PARAM;
  if (c->pend) return B2G_ERR_FAILURE;              // d_in / d_out belong to the batch in flight
  if (n == 0) return 0;
  if (!inter) return B2G_ERR_INVALID_PARAM;
  for (int k = 0; k < nch; k++)
    if (!planes[k]) return B2G_ERR_INVALID_PARAM;
  if (int rc = use_device(c)) return rc;
  const size_t plane_bytes = (size_t)n * typesize, pstride = align_up(plane_bytes, 256), total = plane_bytes * nch;
  DevBuf& d_pl = to_interleaved ? c->d_in : c->d_out;
  DevBuf& d_il = to_interleaved ? c->d_out : c->d_in;
  if (int rc = d_pl.ensure(pstride * nch)) return rc;
  if (int rc = d_il.ensure(total)) return rc;
  cudaStream_t s = c->lanes[0].stream;
  std::vector<void*> dp((size_t)nch);
  for (int k = 0; k < nch; k++) dp[k] = (uint8_t*)d_pl.p + (size_t)k * pstride;
  if (to_interleaved) {
    for (int k = 0; k < nch; k++) CU(cudaMemcpyAsync(dp[k], planes[k], plane_bytes, cudaMemcpyHostToDevice, s));
  } else {
    CU(cudaMemcpyAsync(d_il.p, inter, total, cudaMemcpyHostToDevice, s));
  }
  if (int rc = interleave_device(c, to_interleaved, typesize, nch, dp.data(), d_il.p, n, s)) return rc;
  if (to_interleaved) {
    CU(cudaMemcpyAsync(inter, d_il.p, total, cudaMemcpyDeviceToHost, s));
  } else {
    for (int k = 0; k < nch; k++) CU(cudaMemcpyAsync(planes[k], dp[k], plane_bytes, cudaMemcpyDeviceToHost, s));
  }
  CU(cudaStreamSynchronize(s));
  return 0;
}

int b2g_interleave(b2g_ctx* c, int typesize, int nchannels, const void* const* planes, void* interleaved, int64_t n_elems) {
  return interleave_host(c, true, typesize, nchannels, (void* const*)planes, interleaved, n_elems);
}

int b2g_deinterleave(b2g_ctx* c, int typesize, int nchannels, const void* interleaved, void* const* planes, int64_t n_elems) {
  return interleave_host(c, false, typesize, nchannels, planes, (void*)interleaved, n_elems);
}

void* b2g_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cuda_fail(cudaHostAlloc(&p, std::max(bytes, (size_t)1), cudaHostAllocDefault), "cudaHostAlloc")) return nullptr;
  return p;
}

void b2g_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

b2g_ctx* b2g_clone(const b2g_ctx* c) {
  if (!c) return nullptr;
  b2g_params p = c->p;
  p.device = c->dev;
  return b2g_create(&p);
}

int b2g_get_params(const b2g_ctx* c, b2g_params* out) {
  if (!c || !out) return B2G_ERR_INVALID_PARAM;
  *out = c->p;
  out->device = c->dev;
  return 0;
}

int64_t b2g_launch_count(const b2g_ctx* c) { return c ? c->launches.load() : (int64_t)0; }

int b2g_last_kernel_ms(const b2g_ctx* c, float out_ms[2]) {
  if (!c || !out_ms) return B2G_ERR_INVALID_PARAM;
  out_ms[0] = c->last_ms[0];
  out_ms[1] = c->last_ms[1];
  return 0;
}

const char* b2g_strerror(int code) {
  switch (code) {
    case B2G_OK: return "ok";
    case B2G_ERR_FAILURE: return "generic failure";
    case B2G_ERR_DATA: return "compressed data is corrupt (a stream did not decode to its block size)";
    case B2G_ERR_READ_BUFFER: return "chunk is truncated";
    case B2G_ERR_WRITE_BUFFER: return "destination buffer too small";
    case B2G_ERR_CODEC_SUPPORT: return "codec not supported";
    case B2G_ERR_CODEC_PARAM: return "invalid codec parameter";
    case B2G_ERR_VERSION_SUPPORT: return "chunk format version not supported";
    case B2G_ERR_INVALID_HEADER: return "invalid chunk header";
    case B2G_ERR_INVALID_PARAM: return "invalid parameter";
    case B2G_ERR_RUN_LENGTH: return "invalid run-length stream";
    case B2G_ERR_FILTER_PIPELINE: return "filter pipeline not supported";
    case B2G_ERR_MAX_BUFSIZE: return "buffer larger than 2 GiB - 32";
    case B2G_ERR_NO_DEVICE: return "no usable CUDA device / CUDA runtime error";
    case B2G_ERR_BLOCKSIZE: return "blocksize exceeds what this context was sized for";
    case B2G_ERR_CTX_CODEC: return "zstd chunk on a context not created for zstd";
    default: return "unknown error";
  }
}

const char* b2g_last_error(void) { return g_last_error.c_str(); }

int b2g_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

const char* b2g_version(void) { return "b2g 0.1.0 (sm_100a; blosc2 chunk format 5)"; }

}  // extern "C"

// ---- allocation-free decompress (b2g_decompress_batch_device_noalloc) ---------------------------------------------------
// For callers that cannot allocate (CUDA graph capture around them, serving loops, harnesses that forbid cudaMalloc): the
// same plan + decode kernels as b2g_decompress_batch_device, with EVERY work buffer -- chunk descriptors, per-chunk plans,
// results, item offsets, batch counters, the plane-pairing mailboxes and the ring of staging slots -- carved out of one
// file-scope device array.  Every GPU that loads the library holds these kNoAllocScratchBytes until the process exits.
// The array's name is only a host-side stand-in here: the device address comes from cudaGetSymbolAddress, once per device.
constexpr size_t kNoAllocScratchBytes = (size_t)8 << 20;
static __device__ __align__(256) uint8_t g_noalloc_scratch[kNoAllocScratchBytes];

namespace {

constexpr int kNoAllocSlots = 64;          // staging slots for parked planes (a full ring falls back to the byte scatter)
constexpr int kNoAllocPairCap = 131072;    // stream items that can take part in plane pairing; later ones scatter

struct NoAllocDev {
  bool ready = false;
  uint8_t* scratch = nullptr;              // device address of g_noalloc_scratch on this device
  DecGeom g{};
  size_t off_info = 0, off_res = 0, off_base = 0, off_ctl = 0, off_pair = 0, off_stage = 0, slot_bytes = 0;
};
NoAllocDev g_noalloc[64];
std::mutex g_noalloc_mutex;

// Once per device, with that device current.
int noalloc_setup(int dev, NoAllocDev& D) {
  if (D.ready) return 0;
  void* ptr = nullptr;
  if (cuda_fail(cudaGetSymbolAddress(&ptr, g_noalloc_scratch), "cudaGetSymbolAddress(g_noalloc_scratch)")) return B2G_ERR_NO_DEVICE;
  if (!ptr) { g_last_error = "cudaGetSymbolAddress returned a null address"; return B2G_ERR_NO_DEVICE; }
  b2g_ctx tmp;                             // decode_geometry only reads dev, sms and the block sizes
  tmp.dev = dev;
  tmp.blocksize = tmp.p.blocksize = B2G_NOALLOC_MAX_BLOCKSIZE;
  CU(cudaDeviceGetAttribute(&tmp.sms, cudaDevAttrMultiProcessorCount, dev));
  if (int rc = decode_geometry(&tmp, B2G_NOALLOC_MAX_BLOCKSIZE, false, D.g, true)) return rc;
  size_t o = 0;
  o += align_up(sizeof(DChunk) * B2G_NOALLOC_MAX_CHUNKS, 256); D.off_info = o;
  o += align_up(sizeof(DInfo) * B2G_NOALLOC_MAX_CHUNKS, 256);  D.off_res = o;
  o += align_up(sizeof(DResult) * B2G_NOALLOC_MAX_CHUNKS, 256); D.off_base = o;
  o += align_up(sizeof(int32_t) * 2 * (B2G_NOALLOC_MAX_CHUNKS + 1), 256); D.off_ctl = o;
  o += align_up(sizeof(BatchCtl) * 2, 256); D.off_pair = o;
  o += align_up(sizeof(int32_t) * ((size_t)kNoAllocSlots + 4 * (size_t)kNoAllocPairCap), 256); D.off_stage = o;
  D.slot_bytes = align_up((size_t)D.g.wcap, 128);
  o += D.slot_bytes * kNoAllocSlots;
  if (o > kNoAllocScratchBytes) { g_last_error = "b2g no-alloc scratch layout exceeds the static array"; return B2G_ERR_FAILURE; }
  D.scratch = (uint8_t*)ptr;
  D.ready = true;
  return 0;
}

}  // namespace

extern "C" int b2g_decompress_batch_device_noalloc(int device, b2g_chunk* chunks, int n, void* cuda_stream) {
  if ((!chunks && n > 0) || n < 0) return B2G_ERR_INVALID_PARAM;
  if (n > B2G_NOALLOC_MAX_CHUNKS) {
    g_last_error = "b2g_decompress_batch_device_noalloc: " + std::to_string(n) + " chunks in one call, the static scratch holds " +
                   std::to_string(B2G_NOALLOC_MAX_CHUNKS) + " (B2G_NOALLOC_MAX_CHUNKS); split the batch";
    return B2G_ERR_MAX_BUFSIZE;
  }
  if (n == 0) return 0;
  if (device < 0) CU(cudaGetDevice(&device));
  if (device >= 64) return B2G_ERR_INVALID_PARAM;
  CU(cudaSetDevice(device));
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);   // one call at a time: every call uses the same device array
  NoAllocDev& D = g_noalloc[device];
  if (int rc = noalloc_setup(device, D)) return rc;
  cudaStream_t s = (cudaStream_t)cuda_stream;          // nullptr = the default stream; no stream is created here
  DChunk h[B2G_NOALLOC_MAX_CHUNKS];                    // host stack, pageable: the copies below stage it before returning
  DResult r[B2G_NOALLOC_MAX_CHUNKS];
  for (int i = 0; i < n; i++) {
    h[i].src = (const uint8_t*)chunks[i].src;
    h[i].dst = (uint8_t*)chunks[i].dst;
    h[i].src_bytes = chunks[i].src_bytes;
    h[i].dst_cap = chunks[i].dst_capacity;
  }
  uint8_t* base = D.scratch;
  const size_t pair_ints = (size_t)kNoAllocSlots + 4 * (size_t)kNoAllocPairCap;
  // the array keeps the last call's contents: slot flags, mailboxes and block counters must start at zero (the batch
  // counters, plans, offsets and results are written by k_plan_decode before anything reads them)
  CU(cudaMemsetAsync(base + D.off_pair, 0, sizeof(int32_t) * pair_ints, s));
  CU(cudaMemcpyAsync(base, h, sizeof(DChunk) * (size_t)n, cudaMemcpyHostToDevice, s));
  PlanArgs pa;
  pa.chunks = (const DChunk*)base;
  pa.info = (DInfo*)(base + D.off_info);
  pa.res = (DResult*)(base + D.off_res);
  pa.base = (int32_t*)(base + D.off_base);
  pa.ctl = (BatchCtl*)(base + D.off_ctl);
  pa.n = n;
  pa.max_block = B2G_NOALLOC_MAX_BLOCKSIZE;            // larger blocks: that chunk fails with B2G_ERR_BLOCKSIZE
  pa.zstd_ok = 0;                                      // zstd chunks: B2G_ERR_CTX_CODEC (their work area is not in the array)
  k_plan_decode<<<1, kPlanThreads, kPlanSmemBytes, s>>>(pa);
  DecodeArgs da;
  da.chunks = pa.chunks;
  da.info = pa.info;
  da.res = pa.res;
  da.base = pa.base;
  da.ctl = pa.ctl;
  da.n = n;
  da.wcap = D.g.wcap;
  da.wstride = D.g.wstride;
  da.slab = nullptr;                                   // every stream of a block <= the maximum fits its shared-memory window
  da.slab_stride = 0;
  da.stage = base + D.off_stage;
  da.slotflag = (int32_t*)(base + D.off_pair);
  da.mail = da.slotflag + kNoAllocSlots;
  da.blkcnt = da.mail + 2 * kNoAllocPairCap;
  da.nslots = kNoAllocSlots;
  da.slot_bytes = (int32_t)D.slot_bytes;
  da.pair_cap = kNoAllocPairCap;
  da.bulk = D.g.bulk;
  k_decode<false><<<D.g.grid, 32 * D.g.warps, D.g.smem, s>>>(da);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(r, base + D.off_res, sizeof(DResult) * (size_t)n, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  for (int i = 0; i < n; i++) {
    chunks[i].status = r[i].status;
    chunks[i].out_bytes = r[i].status < 0 ? 0 : r[i].out_bytes;
  }
  return 0;
}
