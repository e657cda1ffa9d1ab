/* b2g.h -- C ABI of the B200-native blosc2 chunk engine ("b2g" = blosc2 on GPU).
 *
 * This is the drop-in boundary for the per-chunk compress / decompress path of EmilDohne/compressed-image.
 * The reference reaches that path through the `compressed::blosc2` wrapper
 * (compressed_image/include/compressed/blosc2/wrapper.h), which calls c-blosc2 ONE chunk at a time; every
 * entry point below names the reference interface it replaces.  Plain pointers and sizes only: no C++, CUDA
 * or torch types cross this boundary (the CUDA stream is passed as an opaque `void*`, NULL = the context's own
 * stream).
 *
 * Chunks produced here are ordinary blosc2 chunks (32-byte extended header, bstarts, per-stream int32 csize
 * framing: thirdparty/c-blosc2/README_CHUNK_FORMAT.rst) and are decodable by the reference; chunks produced by
 * the reference (blosclz, lz4, lz4hc, zstd; shuffle, bitshuffle or no filter) are decoded here bit-exactly.
 *
 * There is no CPU fallback: every function that touches data runs sm_100a kernels and fails with
 * B2G_ERR_NO_DEVICE if no CUDA device is usable.
 */
#ifndef B2G_H
#define B2G_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define B2G_API __declspec(dllexport)
#else
#define B2G_API __attribute__((visibility("default")))
#endif

/* codec ids: identical to c-blosc2's BLOSC_BLOSCLZ/LZ4/LZ4HC/ZSTD (include/blosc2.h:318-322), which is what
 * compressed::blosc2::codec_to_blosc2 (wrapper.h:74-96) hands to cparams.compcode. */
enum { B2G_BLOSCLZ = 0, B2G_LZ4 = 1, B2G_LZ4HC = 2, B2G_ZSTD = 5 };

/* filter ids: c-blosc2's BLOSC_NOFILTER / BLOSC_SHUFFLE / BLOSC_BITSHUFFLE (cparams.filters[5]). The reference
 * wrapper never sets it, so its chunks always carry SHUFFLE (include/blosc2.h:1246-1253). */
enum { B2G_NOFILTER = 0, B2G_SHUFFLE = 1, B2G_BITSHUFFLE = 2 };

/* per-chunk status / return codes: c-blosc2's numbering (include/blosc2.h:444-482), so the C++ wrapper can
 * throw the same std::runtime_error text as wrapper.h:146-149 / 254-257. Codes <= -100 are b2g's own. */
enum {
  B2G_OK = 0,
  B2G_ERR_FAILURE = -1,
  B2G_ERR_DATA = -3,            /* stream did not decode to exactly the expected size */
  B2G_ERR_READ_BUFFER = -5,     /* chunk truncated */
  B2G_ERR_WRITE_BUFFER = -6,    /* destination too small */
  B2G_ERR_CODEC_SUPPORT = -7,   /* zlib / user codecs / dictionaries; encode with zstd */
  B2G_ERR_CODEC_PARAM = -8,
  B2G_ERR_VERSION_SUPPORT = -10,
  B2G_ERR_INVALID_HEADER = -11,
  B2G_ERR_INVALID_PARAM = -12,
  B2G_ERR_RUN_LENGTH = -17,
  B2G_ERR_FILTER_PIPELINE = -18, /* delta / trunc-prec / user filters */
  B2G_ERR_MAX_BUFSIZE = -35,
  B2G_ERR_NO_DEVICE = -100,     /* no usable CUDA device / driver error (see b2g_last_error) */
  B2G_ERR_BLOCKSIZE = -101,     /* blocksize exceeds what this context was sized for (b2g_params.blocksize) */
  B2G_ERR_CTX_CODEC = -102      /* device-resident decode of a zstd chunk on a context not created for zstd */
};

#define B2G_MAX_OVERHEAD 32 /* = BLOSC2_MAX_OVERHEAD: compress dst_capacity >= nbytes + 32 always succeeds */

/* Replaces blosc2_cparams / blosc2_dparams as filled in by compressed::blosc2::create_blosc2_cparams<T>
 * (wrapper.h:338-359) and create_decompression_context (wrapper.h:381-413).  splitmode is always
 * BLOSC_AUTO_SPLIT, as in the reference.  `nthreads` has no equivalent: the grid is the thread pool. */
typedef struct b2g_params {
  uint8_t codec;     /* B2G_*: codec for compress. For decompress contexts: B2G_ZSTD sizes the context for zstd
                        chunks on the device-resident path (host-pointer calls size themselves from the headers). */
  uint8_t clevel;    /* 0..9 (0 = stored); reference clamps with util::ensure_compression_level (util.h:86-94) */
  uint8_t typesize;  /* sizeof(T) */
  uint8_t filter;    /* B2G_SHUFFLE is what the reference always uses */
  int32_t blocksize; /* bytes; 0 = 32768 (compressed::s_default_blocksize, constants.h:11). Also the largest
                        blocksize the device-resident decode path is sized for. */
  int32_t device;    /* CUDA device ordinal; -1 = current device */
} b2g_params;

/* One chunk of a batch.  The reference moves exactly this information through
 * blosc2_compress_ctx(ctx, src, srcsize, dest, destsize) / blosc2_decompress_ctx(ctx, src, srcsize, dest,
 * destsize) (b2c:2577, b2c:2955). */
typedef struct b2g_chunk {
  const void* src;      /* compress: raw elements.   decompress: blosc2 chunk */
  void* dst;            /* compress: chunk out.      decompress: raw elements out */
  int32_t src_bytes;    /* compress: nbytes.         decompress: bytes available at src (INT32_MAX = trust header,
                           as compressed::blosc2::decompress does, wrapper.h:247) */
  int32_t dst_capacity; /* compress: >= nbytes + 32 to be sure to succeed */
  int32_t out_bytes;    /* OUT: cbytes (compress) / nbytes (decompress); 0 when status < 0 */
  int32_t status;       /* OUT: B2G_OK or a negative B2G_ERR_* */
} b2g_chunk;

typedef struct b2g_ctx b2g_ctx;

/* Replaces blosc2_create_cctx / blosc2_create_dctx / blosc2_free_ctx (wrapper.h:41-67, 362-413).  One context
 * serves both directions.  A context is not thread-safe (neither is a blosc2_context); distinct contexts may
 * be used concurrently.  Returns NULL on failure (see b2g_last_error). */
B2G_API b2g_ctx* b2g_create(const b2g_params* params);
B2G_API void b2g_destroy(b2g_ctx* ctx);

/* Replaces the per-chunk loop around blosc2_compress_ctx in compressed::blosc2::compress<T>
 * (wrapper.h:134-185; callers schunk.h:85-104, 196-208, iterator.h:289).  src/dst are HOST pointers (pageable or
 * pinned).  All n chunks are staged, compressed by one launch and copied back; the call returns when every
 * out_bytes/status is final.  Returns B2G_OK, or a negative code if the batch as a whole could not run
 * (per-chunk failures are reported in status and do not fail the call). */
B2G_API int b2g_compress_batch(b2g_ctx* ctx, b2g_chunk* chunks, int n, void* cuda_stream);

/* Replaces the per-chunk loop around blosc2_decompress_ctx in compressed::blosc2::decompress<T>
 * (wrapper.h:236-277; callers schunk.h:123-141, 164-180, lazyschunk.h:186,237).  Codec, filter, typesize and
 * blocksize come from each chunk's own header.  HOST pointers. */
B2G_API int b2g_decompress_batch(b2g_ctx* ctx, b2g_chunk* chunks, int n, void* cuda_stream);

/* Asynchronous forms of the two host-pointer calls (no reference equivalent: blosc2 contexts are synchronous).  The
 * call returns as soon as the H2D copies, the kernels and -- for decompress -- the D2H copies are enqueued on the
 * context's own streams; b2g_wait() completes the batch (compress: copies each chunk back once its size is known) and
 * fills out_bytes/status.  `chunks` and every src/dst buffer must stay valid until then; pinned host memory keeps the
 * copies asynchronous.  One batch may be in flight per context: a decompress batch on one context and a compress batch
 * on another use the PCIe link in both directions at once.  b2g_X_batch(...) == b2g_X_batch_async(...) + b2g_wait(). */
B2G_API int b2g_compress_batch_async(b2g_ctx* ctx, b2g_chunk* chunks, int n, void* cuda_stream);
B2G_API int b2g_decompress_batch_async(b2g_ctx* ctx, b2g_chunk* chunks, int n, void* cuda_stream);

/* Device-resident variants: src/dst are DEVICE pointers on the context's device.  The work is enqueued on the
 * stream and the call returns without waiting; out_bytes/status of `chunks` are filled by b2g_wait(), until
 * which `chunks` must stay alive.  One batch may be in flight per context. */
B2G_API int b2g_compress_batch_device(b2g_ctx* ctx, b2g_chunk* chunks, int n, void* cuda_stream);
B2G_API int b2g_decompress_batch_device(b2g_ctx* ctx, b2g_chunk* chunks, int n, void* cuda_stream);
B2G_API int b2g_wait(b2g_ctx* ctx);

/* Replaces blosc2_cbuffer_sizes (b2c:3696-3712) behind compressed::blosc2::chunk_num_elements<T>
 * (wrapper.h:447-468).  Host-only; reads the first 16 bytes of a chunk in host memory. */
B2G_API int b2g_chunk_sizes(const void* chunk, int32_t* nbytes, int32_t* cbytes, int32_t* blocksize);

/* Standalone filters on device-resident blocks (K1-K4 of SURVEY.md 8a: blosc/shuffle.c:421-514).  n bytes each;
 * enqueued on the stream, not waited for. `format_version` as in the chunk header byte 0 (2 = Blosc1 rule). */
B2G_API int b2g_filter_device(b2g_ctx* ctx, int filter, int forward, int typesize, int format_version,
                              const void* src, void* dst, int64_t n, int32_t blocksize, void* cuda_stream);

/* Replace compressed::image_algo::interleave<T> / deinterleave<T> (compressed_image/include/compressed/
 * image_algo.h:34-65, :84-111; call site image.h:1880) for device-resident buffers:
 *     interleaved[i * nchannels + c] = planes[c][i],   elements of `typesize` bytes, i < n_elems.
 * `planes` is a HOST array of nchannels DEVICE pointers (n_elems * typesize bytes each); `interleaved` holds
 * n_elems * nchannels * typesize bytes.  Enqueued on the stream, not waited for.  nchannels <= 0 is
 * B2G_ERR_INVALID_PARAM (the reference throws std::invalid_argument for an empty span list). */
B2G_API int b2g_interleave_device(b2g_ctx* ctx, int typesize, int nchannels, const void* const* planes,
                                  void* interleaved, int64_t n_elems, void* cuda_stream);
B2G_API int b2g_deinterleave_device(b2g_ctx* ctx, int typesize, int nchannels, const void* interleaved,
                                    void* const* planes, int64_t n_elems, void* cuda_stream);
/* The same two operations on HOST buffers (what the reference's call site holds, image.h:1880): copied to the device,
 * run by the kernels above, copied back; returns when the result is in place.  Not while a batch is in flight. */
B2G_API int b2g_interleave(b2g_ctx* ctx, int typesize, int nchannels, const void* const* planes, void* interleaved,
                           int64_t n_elems);
B2G_API int b2g_deinterleave(b2g_ctx* ctx, int typesize, int nchannels, const void* interleaved, void* const* planes,
                             int64_t n_elems);

/* Page-locked (pinned) host memory for buffers that the asynchronous host-pointer calls read or write: only pinned
 * buffers keep b2g_*_batch_async asynchronous (pageable ones are staged or bounced by the driver).  No reference
 * counterpart (blosc2 contexts are synchronous); compressed::channel_iterator<T> takes its chunk buffers from here.
 * b2g_host_alloc returns NULL on failure (b2g_last_error says why). */
B2G_API void* b2g_host_alloc(size_t bytes);
B2G_API void b2g_host_free(void* p);

/* A new context with the parameters of `ctx` (its own streams and work buffers): what an iterator takes so that its
 * batches in flight never collide with calls on the channel's own contexts (iterators/iterator.h:209-210 shares them,
 * which is safe there because blosc2 contexts never have work in flight between calls). */
B2G_API b2g_ctx* b2g_clone(const b2g_ctx* ctx);

/* The parameters `ctx` was created with (device resolved to the ordinal it runs on).  Lets a caller that pools
 * contexts match a pooled one against the context it would otherwise clone. */
B2G_API int b2g_get_params(const b2g_ctx* ctx, b2g_params* out);

/* Number of kernels this library has launched through `ctx` since creation (bench.py's gpu_launches). */
B2G_API int64_t b2g_launch_count(const b2g_ctx* ctx);
/* Milliseconds the last batch's kernels took on the device (CUDA events on the launching stream), split by
 * kernel: [0] plan/finalize, [1] main codec kernel.  Valid after b2g_wait / a host-pointer batch call. */
B2G_API int b2g_last_kernel_ms(const b2g_ctx* ctx, float out_ms[2]);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  Decompresses n blosc2 chunks
 * like b2g_decompress_batch_device -- src / dst are DEVICE pointers on `device` (-1 = the current device), codec, filter,
 * typesize and blocksize come from each chunk's header, out_bytes / status are filled per chunk -- but needs no context and
 * allocates NOTHING: no cudaMalloc / cudaFree / cudaHostAlloc, no stream or event creation.  All work buffers live in one
 * static device array of 8 MiB inside the library, which every GPU that loads the library holds until the process exits.
 * Limits: n <= B2G_NOALLOC_MAX_CHUNKS per call (more: B2G_ERR_MAX_BUFSIZE, b2g_last_error() says why); chunks whose
 * blocksize exceeds B2G_NOALLOC_MAX_BLOCKSIZE fail individually with B2G_ERR_BLOCKSIZE; lz4 / lz4hc / blosclz / stored /
 * special-value chunks only (a zstd chunk fails with B2G_ERR_CTX_CODEC).  Synchronous: the work runs on `cuda_stream`
 * (NULL = the default stream) and the call returns when status / out_bytes are final.  Every call shares the one array,
 * which keeps the previous call's contents (each call clears what must start at zero).  Calls that overlapped on the GPU
 * would race on it: the entry point is synchronous and takes a process-wide lock, so calls made through it never do, and
 * nothing else in the library touches the array. */
#define B2G_NOALLOC_MAX_CHUNKS 1024
#define B2G_NOALLOC_MAX_BLOCKSIZE 65536
B2G_API int b2g_decompress_batch_device_noalloc(int device, b2g_chunk* chunks, int n, void* cuda_stream);

B2G_API const char* b2g_strerror(int code);
B2G_API const char* b2g_last_error(void); /* thread-local text of the last B2G_ERR_NO_DEVICE-class failure */
B2G_API int b2g_device_count(void);
B2G_API const char* b2g_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B2G_H */
