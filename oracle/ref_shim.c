/* oracle/ref_shim.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Flat, ctypes-friendly entry points over the UNMODIFIED reference library (vendored c-blosc2 compiled by
 * oracle/Makefile into oracle/_ref/libblosc2_ref.so).  This file is ours; it only #includes the reference's
 * public header where it lies under /root/reference and calls the same functions the reference wrapper calls
 * (compressed_image/include/compressed/blosc2/wrapper.h:134-277, 338-413), with the same cparams the
 * wrapper builds (wrapper.h:338-359: BLOSC2_CPARAMS_DEFAULTS + blocksize, typesize, BLOSC_AUTO_SPLIT,
 * clevel, nthreads, compcode), plus an explicit filter knob (filters[5]) for the bit-shuffle config.
 *
 * The *_batch entry points are the "R2" CPU baseline of SURVEY.md section 8(d): one nthreads=1 context per
 * host thread, chunks dealt round-robin over all threads.
 */
#include <blosc2.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static blosc2_context* make_cctx(int codec, int clevel, int typesize, int filter, int blocksize, int nthreads) {
  blosc2_cparams cp = BLOSC2_CPARAMS_DEFAULTS;
  cp.blocksize = blocksize;
  cp.typesize = typesize;
  cp.splitmode = BLOSC_AUTO_SPLIT;
  cp.clevel = (uint8_t)clevel;
  cp.nthreads = (int16_t)(nthreads < 1 ? 1 : nthreads);
  cp.compcode = (uint8_t)codec;
  cp.filters[BLOSC2_MAX_FILTERS - 1] = (uint8_t)filter;
  return blosc2_create_cctx(cp);
}

static blosc2_context* make_dctx(int nthreads) {
  blosc2_dparams dp = BLOSC2_DPARAMS_DEFAULTS;
  dp.nthreads = (int16_t)(nthreads < 1 ? 1 : nthreads);
  return blosc2_create_dctx(dp);
}

int ref_compress(int codec, int clevel, int typesize, int filter, int blocksize, int nthreads,
                 const void* src, int32_t nbytes, void* dst, int32_t cap) {
  blosc2_context* c = make_cctx(codec, clevel, typesize, filter, blocksize, nthreads);
  if (!c) return -1000;
  int r = blosc2_compress_ctx(c, src, nbytes, dst, cap);
  blosc2_free_ctx(c);
  return r;
}

int ref_decompress(const void* src, int32_t srcsize, void* dst, int32_t cap, int nthreads) {
  blosc2_context* c = make_dctx(nthreads);
  if (!c) return -1000;
  int r = blosc2_decompress_ctx(c, src, srcsize, dst, cap);
  blosc2_free_ctx(c);
  return r;
}

int ref_cbuffer_sizes(const void* chunk, int32_t* nbytes, int32_t* cbytes, int32_t* blocksize) {
  return blosc2_cbuffer_sizes(chunk, nbytes, cbytes, blocksize);
}

typedef struct {
  int tid, nworkers, n, compress;
  int codec, clevel, typesize, filter, blocksize;
  const void* const* srcs;
  const int32_t* sizes;
  void* const* dsts;
  const int32_t* caps;
  int32_t* out;
} job_t;

static void* worker(void* p) {
  job_t* j = (job_t*)p;
  blosc2_context* c = j->compress ? make_cctx(j->codec, j->clevel, j->typesize, j->filter, j->blocksize, 1)
                                  : make_dctx(1);
  for (int i = j->tid; i < j->n; i += j->nworkers) {
    if (j->compress)
      j->out[i] = blosc2_compress_ctx(c, j->srcs[i], j->sizes[i], j->dsts[i], j->caps[i]);
    else
      j->out[i] = blosc2_decompress_ctx(c, j->srcs[i], j->sizes[i], j->dsts[i], j->caps[i]);
  }
  blosc2_free_ctx(c);
  return NULL;
}

static int run_batch(job_t proto, int nworkers) {
  if (nworkers < 1) nworkers = 1;
  if (nworkers > proto.n) nworkers = proto.n > 0 ? proto.n : 1;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nworkers);
  job_t* jobs = (job_t*)malloc(sizeof(job_t) * nworkers);
  for (int t = 0; t < nworkers; t++) {
    jobs[t] = proto;
    jobs[t].tid = t;
    jobs[t].nworkers = nworkers;
    pthread_create(&th[t], NULL, worker, &jobs[t]);
  }
  for (int t = 0; t < nworkers; t++) pthread_join(th[t], NULL);
  free(th);
  free(jobs);
  return 0;
}

int ref_compress_batch(int codec, int clevel, int typesize, int filter, int blocksize, int n,
                       const void* const* srcs, const int32_t* nbytes, void* const* dsts, const int32_t* caps,
                       int32_t* out_cbytes, int nworkers) {
  job_t j = {0, 0, n, 1, codec, clevel, typesize, filter, blocksize, srcs, nbytes, dsts, caps, out_cbytes};
  return run_batch(j, nworkers);
}

int ref_decompress_batch(int n, const void* const* srcs, const int32_t* srcsizes, void* const* dsts,
                         const int32_t* caps, int32_t* out_nbytes, int nworkers) {
  job_t j = {0, 0, n, 0, 0, 0, 0, 0, 0, srcs, srcsizes, dsts, caps, out_nbytes};
  return run_batch(j, nworkers);
}

/* blosc2_chunk_zeros / blosc2_chunk_repeatval with the cparams the wrapper builds (blosc2.c:4304-4347, 4444-4489);
 * repeatval == NULL -> zeros.  -> chunk bytes written, or negative. */
int ref_special_chunk(int typesize, int blocksize, int32_t nbytes, const void* repeatval, void* dst, int32_t cap) {
  blosc2_cparams cp = BLOSC2_CPARAMS_DEFAULTS;
  cp.blocksize = blocksize;
  cp.typesize = typesize;
  cp.splitmode = BLOSC_AUTO_SPLIT;
  cp.clevel = 9;
  cp.nthreads = 1;
  cp.compcode = BLOSC_LZ4;
  return repeatval ? blosc2_chunk_repeatval(cp, nbytes, dst, cap, repeatval) : blosc2_chunk_zeros(cp, nbytes, dst, cap);
}

/* Raw stream codecs + filters of the reference, for pinning the restatement one layer at a time. */
extern int LZ4_compress_fast(const char*, char*, int, int, int);
extern int LZ4_decompress_safe(const char*, char*, int, int);
extern int LZ4_compress_HC(const char*, char*, int, int, int);
extern int blosclz_compress(int, const void*, int, void*, int, blosc2_context*);
extern int blosclz_decompress(const void*, int, void*, int);
extern size_t ZSTD_compress(void*, size_t, const void*, size_t, int);
extern size_t ZSTD_decompress(void*, size_t, const void*, size_t);
extern unsigned ZSTD_isError(size_t);
extern void shuffle(int32_t, int32_t, const uint8_t*, uint8_t*);
extern void unshuffle(int32_t, int32_t, const uint8_t*, uint8_t*);
extern int32_t bitshuffle(int32_t, int32_t, const uint8_t*, uint8_t*);
extern int32_t bitunshuffle(int32_t, int32_t, const uint8_t*, uint8_t*, uint8_t);

int ref_lz4_compress(const void* s, int n, void* d, int cap) { return LZ4_compress_fast(s, d, n, cap, 1); }
int ref_lz4hc_compress(const void* s, int n, void* d, int cap, int lvl) { return LZ4_compress_HC(s, d, n, cap, lvl); }
int ref_lz4_decompress(const void* s, int n, void* d, int cap) { return LZ4_decompress_safe(s, d, n, cap); }
int ref_blosclz_compress(int clevel, int typesize, int filter, const void* s, int n, void* d, int cap) {
  /* blosclz reads ctx->src (NULL check only), ctx->typesize, ctx->filter_flags... give it a real context. */
  blosc2_context* c = make_cctx(BLOSC_BLOSCLZ, clevel, typesize, filter, 0, 1);
  int r = blosclz_compress(clevel, s, n, d, cap, c);
  blosc2_free_ctx(c);
  return r;
}
int ref_blosclz_decompress(const void* s, int n, void* d, int cap) { return blosclz_decompress(s, n, d, cap); }
long ref_zstd_compress(const void* s, long n, void* d, long cap, int lvl) {
  size_t r = ZSTD_compress(d, cap, s, n, lvl);
  return ZSTD_isError(r) ? -1 : (long)r;
}
long ref_zstd_decompress(const void* s, long n, void* d, long cap) {
  size_t r = ZSTD_decompress(d, cap, s, n);
  return ZSTD_isError(r) ? -1 : (long)r;
}
int ref_shuffle(int ts, int n, const void* s, void* d) { shuffle(ts, n, s, d); return n; }
int ref_unshuffle(int ts, int n, const void* s, void* d) { unshuffle(ts, n, s, d); return n; }
int ref_bitshuffle(int ts, int n, const void* s, void* d) { return bitshuffle(ts, n, s, d); }
int ref_bitunshuffle(int ts, int n, const void* s, void* d, int ver) { return bitunshuffle(ts, n, s, d, (uint8_t)ver); }

/* ---- contiguous frames -------------------------------------------------------------------
 * What compressed-image does to export a channel: schunk::to_schunk() (compressed/blosc2/schunk.h:107-121: a default
 * schunk from wrapper.h:301-310, every chunk appended with copy) followed by blosc2_schunk_to_buffer (schunk.c:357). */
int64_t ref_chunks_to_cframe(int n, const void* const* chunks, void* dst, int64_t cap) {
  blosc2_init();
  blosc2_cparams cp = BLOSC2_CPARAMS_DEFAULTS;
  blosc2_dparams dp = BLOSC2_DPARAMS_DEFAULTS;
  blosc2_storage st = BLOSC2_STORAGE_DEFAULTS;
  st.cparams = &cp;
  st.dparams = &dp;
  blosc2_schunk* sc = blosc2_schunk_new(&st);
  if (!sc) return -1;
  for (int i = 0; i < n; i++) {
    if (blosc2_schunk_append_chunk(sc, (uint8_t*)chunks[i], true) < 0) { blosc2_schunk_free(sc); return -2; }
  }
  uint8_t* cframe = NULL;
  bool needs_free = false;
  int64_t len = blosc2_schunk_to_buffer(sc, &cframe, &needs_free);
  if (len > 0 && len <= cap) memcpy(dst, cframe, (size_t)len);
  else if (len > cap) len = -3;
  if (needs_free) free(cframe);
  blosc2_schunk_free(sc);
  return len;
}

/* blosc2_schunk_from_buffer (schunk.c:676): out[0] nchunks, out[1] nbytes, out[2] cbytes, out[3] typesize,
 * out[4] chunksize, out[5] blocksize. */
int ref_cframe_info(const void* cframe, int64_t len, int64_t* out) {
  blosc2_init();
  blosc2_schunk* sc = blosc2_schunk_from_buffer((uint8_t*)cframe, len, false);
  if (!sc) return -1;
  out[0] = sc->nchunks; out[1] = sc->nbytes; out[2] = sc->cbytes; out[3] = sc->typesize; out[4] = sc->chunksize;
  out[5] = sc->blocksize;
  blosc2_schunk_free(sc);
  return 0;
}

/* chunk i of a frame, still compressed (blosc2_schunk_get_chunk) -> cbytes, or negative */
int ref_cframe_get_chunk(const void* cframe, int64_t len, int64_t i, void* dst, int32_t cap) {
  blosc2_init();
  blosc2_schunk* sc = blosc2_schunk_from_buffer((uint8_t*)cframe, len, false);
  if (!sc) return -1;
  uint8_t* chunk = NULL;
  bool needs_free = false;
  int cb = blosc2_schunk_get_chunk(sc, i, &chunk, &needs_free);
  if (cb > 0 && cb <= cap) memcpy(dst, chunk, (size_t)cb);
  else if (cb > cap) cb = -3;
  if (needs_free) free(chunk);
  blosc2_schunk_free(sc);
  return cb;
}

/* chunk i of a frame, decompressed by the reference (blosc2_schunk_decompress_chunk) -> nbytes, or negative */
int ref_cframe_decompress_chunk(const void* cframe, int64_t len, int64_t i, void* dst, int32_t cap) {
  blosc2_init();
  blosc2_schunk* sc = blosc2_schunk_from_buffer((uint8_t*)cframe, len, false);
  if (!sc) return -1;
  int nb = blosc2_schunk_decompress_chunk(sc, i, dst, cap);
  blosc2_schunk_free(sc);
  return nb;
}
