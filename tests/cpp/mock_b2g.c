/* tests/cpp/mock_b2g.c -- TEST INFRASTRUCTURE ONLY: a stand-in for libb2g.so so that the C++ host layer
 * (include/compressed/...) can be exercised on a machine without a GPU.  It is never built into or loaded by the
 * product.  "Compression" stores the data as a memcpyed blosc2 chunk (32-byte extended header, flag bit 1), which is a
 * valid chunk of the format.  The asynchronous calls are DEFERRED: they only record the batch, b2g_wait() runs it --
 * so a caller that reuses a buffer, drops a descriptor or reads a result before waiting is caught, and a second batch
 * on a context that still has one in flight fails like the real library does. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/b2g.h"

struct b2g_ctx {
  b2g_params p;
  b2g_chunk* pend;
  int pend_n, pend_kind;
  int64_t launches;
};

static int do_compress(b2g_ctx* c, b2g_chunk* k) {
  uint8_t* d = (uint8_t*)k->dst;
  const int32_t nb = k->src_bytes, cb = nb + 32;
  if (nb < 0) return k->status = B2G_ERR_INVALID_PARAM;
  if (k->dst_capacity < cb) return k->status = B2G_ERR_WRITE_BUFFER;
  memset(d, 0, 32);
  d[0] = 5; d[1] = 1; d[2] = 0x07; d[3] = c->p.typesize ? c->p.typesize : 1;
  int32_t bs = c->p.blocksize > 0 ? c->p.blocksize : 32768;
  if (bs > nb) bs = nb;
  memcpy(d + 4, &nb, 4); memcpy(d + 8, &bs, 4); memcpy(d + 12, &cb, 4);
  d[21] = c->p.filter; d[22] = c->p.codec;
  memcpy(d + 32, k->src, (size_t)nb);
  k->out_bytes = cb;
  return k->status = 0;
}

static int do_decompress(b2g_chunk* k) {
  const uint8_t* s = (const uint8_t*)k->src;
  int32_t nb, cb;
  if (k->src_bytes < 32) return k->status = B2G_ERR_READ_BUFFER;
  memcpy(&nb, s + 4, 4); memcpy(&cb, s + 12, 4);
  if (s[0] != 5 || !(s[2] & 0x02) || cb != nb + 32) return k->status = B2G_ERR_INVALID_HEADER;
  if (cb > k->src_bytes) return k->status = B2G_ERR_READ_BUFFER;
  if (nb > k->dst_capacity) return k->status = B2G_ERR_WRITE_BUFFER;
  memcpy(k->dst, s + 32, (size_t)nb);
  k->out_bytes = nb;
  return k->status = 0;
}

b2g_ctx* b2g_create(const b2g_params* p) {
  if (!p) return NULL;
  b2g_ctx* c = (b2g_ctx*)calloc(1, sizeof(b2g_ctx));
  c->p = *p;
  return c;
}
b2g_ctx* b2g_clone(const b2g_ctx* c) { return c ? b2g_create(&c->p) : NULL; }
void b2g_destroy(b2g_ctx* c) { free(c); }

static int enqueue(b2g_ctx* c, b2g_chunk* chunks, int n, int kind) {
  if (!c || (!chunks && n) || n < 0) return B2G_ERR_INVALID_PARAM;
  if (c->pend) return B2G_ERR_FAILURE;                 /* one batch in flight per context */
  for (int i = 0; i < n; i++) { chunks[i].status = -99; chunks[i].out_bytes = -99; }   /* not valid before b2g_wait */
  c->pend = n ? chunks : NULL; c->pend_n = n; c->pend_kind = kind;
  return 0;
}
int b2g_wait(b2g_ctx* c) {
  if (!c) return B2G_ERR_INVALID_PARAM;
  b2g_chunk* k = c->pend;
  c->pend = NULL;
  for (int i = 0; k && i < c->pend_n; i++) { if (c->pend_kind == 2) do_compress(c, &k[i]); else do_decompress(&k[i]); }
  c->launches++;
  return 0;
}
int b2g_compress_batch_async(b2g_ctx* c, b2g_chunk* k, int n, void* s) { (void)s; return enqueue(c, k, n, 2); }
int b2g_decompress_batch_async(b2g_ctx* c, b2g_chunk* k, int n, void* s) { (void)s; return enqueue(c, k, n, 1); }
int b2g_compress_batch(b2g_ctx* c, b2g_chunk* k, int n, void* s) { int rc = b2g_compress_batch_async(c, k, n, s); return rc ? rc : b2g_wait(c); }
int b2g_decompress_batch(b2g_ctx* c, b2g_chunk* k, int n, void* s) { int rc = b2g_decompress_batch_async(c, k, n, s); return rc ? rc : b2g_wait(c); }
int b2g_compress_batch_device(b2g_ctx* c, b2g_chunk* k, int n, void* s) { (void)c; (void)k; (void)n; (void)s; return B2G_ERR_NO_DEVICE; }
int b2g_decompress_batch_device(b2g_ctx* c, b2g_chunk* k, int n, void* s) { (void)c; (void)k; (void)n; (void)s; return B2G_ERR_NO_DEVICE; }

int b2g_chunk_sizes(const void* chunk, int32_t* nbytes, int32_t* cbytes, int32_t* blocksize) {
  const uint8_t* s = (const uint8_t*)chunk;
  int32_t nb, bs, cb;
  memcpy(&nb, s + 4, 4); memcpy(&bs, s + 8, 4); memcpy(&cb, s + 12, 4);
  int rc = 0;
  if (s[0] > 5) rc = B2G_ERR_VERSION_SUPPORT;
  else if (cb < 16 || s[3] == 0 || bs < 0 || (nb > 0 && bs > nb)) rc = B2G_ERR_INVALID_HEADER;
  if (rc < 0) nb = cb = bs = 0;
  if (nbytes) *nbytes = nb;
  if (cbytes) *cbytes = cb;
  if (blocksize) *blocksize = bs;
  return rc;
}

int b2g_interleave(b2g_ctx* c, int ts, int nch, const void* const* planes, void* inter, int64_t n) {
  if (!c || ts <= 0 || nch <= 0 || n < 0) return B2G_ERR_INVALID_PARAM;
  for (int64_t i = 0; i < n; i++)
    for (int k = 0; k < nch; k++) memcpy((uint8_t*)inter + (i * nch + k) * ts, (const uint8_t*)planes[k] + i * ts, (size_t)ts);
  return 0;
}
int b2g_deinterleave(b2g_ctx* c, int ts, int nch, const void* inter, void* const* planes, int64_t n) {
  if (!c || ts <= 0 || nch <= 0 || n < 0) return B2G_ERR_INVALID_PARAM;
  for (int64_t i = 0; i < n; i++)
    for (int k = 0; k < nch; k++) memcpy((uint8_t*)planes[k] + i * ts, (const uint8_t*)inter + (i * nch + k) * ts, (size_t)ts);
  return 0;
}
int b2g_interleave_device(b2g_ctx* c, int ts, int nch, const void* const* p, void* i, int64_t n, void* s) { (void)c; (void)ts; (void)nch; (void)p; (void)i; (void)n; (void)s; return B2G_ERR_NO_DEVICE; }
int b2g_deinterleave_device(b2g_ctx* c, int ts, int nch, const void* i, void* const* p, int64_t n, void* s) { (void)c; (void)ts; (void)nch; (void)p; (void)i; (void)n; (void)s; return B2G_ERR_NO_DEVICE; }
int b2g_filter_device(b2g_ctx* c, int f, int fw, int ts, int v, const void* s, void* d, int64_t n, int32_t bs, void* st) { (void)c; (void)f; (void)fw; (void)ts; (void)v; (void)s; (void)d; (void)n; (void)bs; (void)st; return B2G_ERR_NO_DEVICE; }
int b2g_get_params(const b2g_ctx* c, b2g_params* out) { if (!c || !out) return -12; *out = c->p; return 0; }
void* b2g_host_alloc(size_t bytes) { return malloc(bytes ? bytes : 1); }
void b2g_host_free(void* p) { free(p); }
int64_t b2g_launch_count(const b2g_ctx* c) { return c ? c->launches : 0; }
int b2g_last_kernel_ms(const b2g_ctx* c, float out[2]) { (void)c; out[0] = out[1] = 0; return 0; }
const char* b2g_strerror(int code) { (void)code; return "mock"; }
const char* b2g_last_error(void) { return "mock"; }
int b2g_device_count(void) { return 1; }
const char* b2g_version(void) { return "mock"; }
