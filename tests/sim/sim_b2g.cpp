// sim_b2g.cpp -- TEST INFRASTRUCTURE ONLY.  Runs the product's device code (compressed_image_b200/csrc/*.cuh,
// unmodified) on the CPU under tests/sim/cusim.h and exposes flat entry points for the ctypes tests.  The host
// orchestration mirrors b2g_api.cu (plan -> decode kernels; host plan -> compress -> finalize) on plain host arrays.
#include "cusim.h"

#include "../../compressed_image_b200/csrc/b2g_kernels.cuh"
#include "../../compressed_image_b200/csrc/b2g_plan.h"

using namespace b2g;

static size_t au(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int g_last_paired = 0;
// encoder teams: -1 = as the host library decides (plan_team_warps), 0 / 1 = k_compress (one warp per stream), S > 1 =
// k_compress_team with S warps; prime < 0 = kTeamPrime
static int g_team = -1, g_team_prime = -1, g_team_hlog = 0, g_team_sparse = -1;

extern "C" {

// Decompress a batch.  grid = number of simulated persistent CTAs; win_bytes = shared window of one warp (streams
// that need more use the slab path).  out[i] = nbytes or 0, status[i] = 0 or error.
int sim_decompress_batch(int n, const void* const* srcs, const int32_t* srcsizes, void* const* dsts,
                         const int32_t* caps, int32_t* out, int32_t* status, int win_bytes, int max_block, int grid,
                         int zstd_ok, int pair) {
  std::vector<DChunk> ch(n);
  std::vector<DInfo> info(n);
  std::vector<DResult> res(n);
  std::vector<int32_t> base(2 * (n + 1));
  BatchCtl ctl[2];
  memset(ctl, 0, sizeof(ctl));
  for (int i = 0; i < n; i++) {
    ch[i].src = (const uint8_t*)srcs[i];
    ch[i].dst = (uint8_t*)dsts[i];
    ch[i].src_bytes = srcsizes[i];
    ch[i].dst_cap = caps[i];
  }
  win_bytes = (int)au(win_bytes, 16);
  PlanArgs pa;
  pa.chunks = ch.data(); pa.info = info.data(); pa.res = res.data(); pa.base = base.data(); pa.ctl = ctl;
  pa.n = n; pa.max_block = max_block; pa.zstd_ok = zstd_ok;
  cusim::launch(dim3(1), dim3(kPlanThreads), kPlanSmemBytes, [&] { k_plan_decode(pa); });
  const int warps = 2;  // independent workers per simulated CTA
  const size_t stride = au((size_t)decode_window_bytes(max_block, kFmtBlosclz) + 256, 256);
  std::vector<uint8_t> slab(stride * grid * warps + 256);
  DecodeArgs da;
  da.chunks = ch.data(); da.info = info.data(); da.res = res.data(); da.base = base.data(); da.ctl = &ctl[0];
  da.n = n; da.wcap = win_bytes; da.wstride = win_bytes; da.slab = slab.data(); da.slab_stride = (int64_t)stride;
  // plane pairing: few slots on purpose (the no-free-slot path must run too); pair == 0 switches it off
  long long pair_cap = 0;
  for (int i = 0; i < n; i++) pair_cap += ((long long)std::max(caps[i], 0) / 64 + 2) * kMaxStreams;
  const int nslots = pair == 2 ? 1 : 3;
  std::vector<uint8_t> stage((size_t)nslots * win_bytes + 16);
  std::vector<int32_t> pairmeta(nslots + 4 * pair_cap + 4, 0);
  da.stage = pair ? stage.data() : nullptr;
  da.slotflag = pairmeta.data(); da.mail = da.slotflag + nslots; da.blkcnt = da.mail + 2 * pair_cap;
  da.nslots = nslots; da.slot_bytes = win_bytes; da.pair_cap = (int32_t)pair_cap; da.bulk = 0;
  cusim::launch(dim3(grid), dim3(32 * warps), (size_t)da.wstride * warps, [&] { k_decode<false>(da); });
  if (zstd_ok) {
    da.base = base.data() + (n + 1);
    da.ctl = &ctl[1];
    da.wstride = win_bytes + (int)au(sizeof(ZWork), 16);
    da.mail += pair_cap; da.blkcnt += pair_cap;
    cusim::launch(dim3(grid), dim3(32 * warps), (size_t)da.wstride * warps, [&] { k_decode<true>(da); });
  }
  for (int i = 0; i < n; i++) {
    status[i] = res[i].status;
    out[i] = res[i].status < 0 ? 0 : res[i].out_bytes;
  }
  g_last_paired = ctl[0].paired + ctl[1].paired;
  for (int k = 0; k < nslots; k++) if (pairmeta[k] != 0) return -1;  // a staging slot was never given back
  return 0;
}

int sim_compress_batch(int codec, int clevel, int typesize, int filter, int blocksize, int n,
                       const void* const* srcs, const int32_t* nbytes, void* const* dsts, const int32_t* caps,
                       int32_t* out, int32_t* status, int win_bytes, int grid, int hlog) {
  std::vector<DChunk> ch(n);
  std::vector<CInfo> info(n);
  std::vector<CState> state(n);
  std::vector<DResult> res(n);
  std::vector<int32_t> base(n + 1);
  memset(state.data(), 0, sizeof(CState) * n);
  CParams p;
  p.codec = (uint8_t)codec; p.clevel = (uint8_t)clevel; p.typesize = (uint8_t)typesize; p.filter = (uint8_t)filter;
  int64_t items = 0;
  int max_stream = 1;
  for (int i = 0; i < n; i++) {
    ch[i].src = (const uint8_t*)srcs[i];
    ch[i].dst = (uint8_t*)dsts[i];
    ch[i].src_bytes = nbytes[i];
    ch[i].dst_cap = caps[i];
    int nit = plan_compress_chunk(info[i], p, blocksize, nbytes[i], caps[i], win_bytes);
    if (info[i].status == 0) max_stream = std::max(max_stream, info[i].max_stream);
    info[i].item_base = (int32_t)items;
    base[i] = (int32_t)items;
    items += nit;
  }
  base[n] = (int32_t)items;
  BatchCtl ctl[2];
  memset(ctl, 0, sizeof(ctl));
  ctl[0].total_items = (int32_t)items;
  const int warps = 2;  // independent workers per simulated CTA
  const int wwin = std::max((int)au(max_stream, 16), 1024);
  const int team = g_team < 0 ? plan_team_warps(max_stream) : std::min(g_team, (int)kTeamMaxWarps);
  const size_t sstride = (size_t)wwin;
  std::vector<uint8_t> team_bodies(team > 1 ? (size_t)plan_team_body_bytes(wwin) * (size_t)grid + 256 : 1);
  std::vector<CItem> citems(items + 1);
  std::vector<uint8_t> scratch(sstride * items + 256);
  CompressArgs ca;
  ca.chunks = ch.data(); ca.info = info.data(); ca.state = state.data(); ca.base = base.data(); ca.ctl = ctl;
  ca.items = citems.data(); ca.scratch = scratch.data(); ca.scratch_stride = (int64_t)sstride; ca.n = n;
  const bool lz4 = codec == kCodecLz4 || codec == kCodecLz4hc;
  ca.win_bytes = wwin; ca.p = p;
  if (team > 1) {
    ca.team_bodies = team_bodies.data(); ca.team_body_bytes = plan_team_body_bytes(wwin);
    ca.team_prime = g_team_prime < 0 ? kTeamPrime : g_team_prime;
    ca.team_prime_step = g_team_sparse < 0 ? plan_team_prime_step(max_stream) : (g_team_sparse ? 4 : 1);
    ca.htab_bytes = g_team_hlog > 0 ? (2 << g_team_hlog) : hlog > 0 ? (2 << hlog) : plan_htab_bytes(max_stream, clevel);  // as b2g_api.cu
    ca.tile = kTeamTile;
    const size_t smem = (size_t)wwin + (size_t)team * (ca.htab_bytes + enc_work_bytes(kTeamTile)) + sizeof(TeamCtl);
    cusim::launch(dim3(grid), dim3(32 * team), smem, [&] {
      if (lz4) k_compress_team<kCodecLz4, kTeamTile>(ca); else k_compress_team<kCodecBlosclz, kTeamTile>(ca);
    });
  } else {
    const int htab_bytes = hlog > 0 ? (2 << hlog) : plan_htab_bytes(max_stream, clevel);
    ca.htab_bytes = htab_bytes; ca.team_prime = 0; ca.team_prime_step = 1; ca.team_bodies = nullptr; ca.team_body_bytes = 0;
    ca.tile = enc_tile_for(max_stream);
    const size_t smem = (size_t)warps * ((size_t)wwin + htab_bytes + enc_work_bytes(ca.tile));
    cusim::launch(dim3(grid), dim3(32 * warps), smem, [&] {
      if (ca.tile == 512) { if (lz4) k_compress<kCodecLz4, 512>(ca); else k_compress<kCodecBlosclz, 512>(ca); }
      else { if (lz4) k_compress<kCodecLz4, 1024>(ca); else k_compress<kCodecBlosclz, 1024>(ca); }
    });
  }
  PlaceArgs pl;
  pl.chunks = ch.data(); pl.info = info.data(); pl.items = citems.data(); pl.state = state.data(); pl.n = n;
  cusim::launch(dim3(std::min(grid, n)), dim3(256), 2 * 256 * sizeof(int32_t), [&] { k_place(pl); });
  FinalizeArgs fa;
  fa.chunks = ch.data(); fa.info = info.data(); fa.state = state.data(); fa.items = citems.data();
  fa.scratch = scratch.data(); fa.scratch_stride = (int64_t)sstride; fa.res = res.data(); fa.base = base.data();
  fa.n = n; fa.total_items = (int32_t)items; fa.win_bytes = wwin; fa.p = p;
  cusim::launch(dim3(std::min<int>(grid, (int)items)), dim3(kCtaThreads), (size_t)wwin, [&] { k_finalize(fa); });
  for (int i = 0; i < n; i++) {
    status[i] = res[i].status;
    out[i] = res[i].status < 0 ? 0 : res[i].out_bytes;
  }
  return 0;
}

int sim_filter(int filter, int forward, int typesize, int version, const void* src, void* dst, int64_t n,
               int blocksize, int grid) {
  FilterArgs fa;
  fa.src = (const uint8_t*)src; fa.dst = (uint8_t*)dst; fa.n = n; fa.blocksize = blocksize; fa.typesize = typesize;
  fa.filter = filter; fa.forward = forward; fa.version = version;
  fa.nblocks = (int32_t)((n + blocksize - 1) / blocksize);
  if (fa.nblocks == 0) return 0;
  cusim::launch(dim3(std::min(grid, fa.nblocks)), dim3(kCtaThreads), au(blocksize, 16) + 64, [&] { k_filter(fa); });
  return 0;
}

// Raw stream codecs (one warp), for focused tests.
int sim_stream_decode(int fmt, const void* src, int n, void* dst, int cap) {
  int result = 0;
  cusim::launch(dim3(1), dim3(32), sizeof(ZWork) + 64, [&] {
    int r;
    if (fmt == kFmtLz4) r = lz4_decode_warp((const uint8_t*)src, n, (uint8_t*)dst, cap);
    else if (fmt == kFmtBlosclz) r = blosclz_decode_warp((const uint8_t*)src, n, (uint8_t*)dst, cap);
    else r = (int)zstd_decode_warp((ZWork*)dyn_smem(), (const uint8_t*)src, n, (uint8_t*)dst, cap);
    if (threadIdx.x == 0) result = r;
  });
  return result;
}

int sim_stream_encode(int codec, const void* src, int n, void* dst, int cap, int hlog) {
  int result = 0;
  const size_t in_bytes = au((size_t)n + 16, 16);
  const int tile = enc_tile_for(n);
  cusim::launch(dim3(1), dim3(32), in_bytes + (2u << hlog) + enc_work_bytes(tile), [&] {
    uint8_t* in = dyn_smem();
    for (int i = (int)threadIdx.x; i < n; i += 32) in[i] = ((const uint8_t*)src)[i];
    __syncwarp();
    const EncWork ew = enc_work(in + in_bytes, 2 << hlog, hlog, tile);
    int r;
    if (tile == 512) {
      r = codec == kCodecLz4 ? lz_encode_warp<kCodecLz4, 512>(in, n, (uint8_t*)dst, cap, ew)
                             : lz_encode_warp<kCodecBlosclz, 512>(in, n, (uint8_t*)dst, cap, ew);
    } else {
      r = codec == kCodecLz4 ? lz_encode_warp<kCodecLz4, 1024>(in, n, (uint8_t*)dst, cap, ew)
                             : lz_encode_warp<kCodecBlosclz, 1024>(in, n, (uint8_t*)dst, cap, ew);
    }
    if (threadIdx.x == 0) result = r;
  });
  return result;
}

void sim_set_team(int warps, int prime, int hlog) { g_team = warps; g_team_prime = prime; g_team_hlog = hlog; }
void sim_set_team_sparse(int sparse) { g_team_sparse = sparse; }
unsigned long long sim_collectives() { return cusim::S().collectives; }
int sim_last_paired() { return g_last_paired; }
// lz4 / blosclz batched decode counters: [batches, sequences executed in batches, (unused), matches copied in stream
// order by the whole warp, sequences taken by the general path]
void sim_decode_stats(unsigned long long* out5, int reset) {
  SimDecodeStats& s = sim_decode_stats();
  out5[0] = s.batches; out5[1] = s.batch_seqs; out5[2] = s.lane_matches; out5[3] = s.ordered_matches; out5[4] = s.general_seqs;
  if (reset) s = SimDecodeStats{0, 0, 0, 0, 0};
}

}  // extern "C"
