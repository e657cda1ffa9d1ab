/* oracle/b2o.h -- TEST INFRASTRUCTURE ONLY (never imported, linked or executed by the product path).
 *
 * Plain-C CPU restatement of the reference's per-chunk compress / decompress path
 * (compressed-image -> vendored c-blosc2 2.17.1.dev + LZ4 1.10.0 + zstd 1.5.7).  Each function cites the
 * reference file:line it follows.  Paths are relative to /root/reference/thirdparty/c-blosc2 unless noted:
 *   b2c  = blosc/blosc2.c            L4 = internal-complibs/lz4-1.10.0/lz4.c
 *   Z/   = internal-complibs/zstd-1.5.7/
 *
 * Parity status: PINNED.  tests/test_oracle_*.py check this restatement against (a) the reference itself
 * compiled from its own sources into oracle/_ref/libblosc2_ref.so (oracle/Makefile), (b) the c-blosc2
 * compat/ golden .cdata chunks re-framed as committed fixtures under tests/golden/.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use this.
 */
#ifndef B2O_H
#define B2O_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* codec ids = b2h:318-322 ; filter ids = b2h BLOSC_NOFILTER/SHUFFLE/BITSHUFFLE */
enum { B2O_BLOSCLZ = 0, B2O_LZ4 = 1, B2O_LZ4HC = 2, B2O_ZSTD = 5 };
enum { B2O_NOFILTER = 0, B2O_SHUFFLE = 1, B2O_BITSHUFFLE = 2 };

/* error numbering = include/blosc2.h:444-482 */
enum {
  B2O_ERR_FAILURE = -1, B2O_ERR_DATA = -3, B2O_ERR_READ_BUFFER = -5, B2O_ERR_WRITE_BUFFER = -6,
  B2O_ERR_CODEC_SUPPORT = -7, B2O_ERR_CODEC_PARAM = -8, B2O_ERR_VERSION_SUPPORT = -10,
  B2O_ERR_INVALID_HEADER = -11, B2O_ERR_INVALID_PARAM = -12, B2O_ERR_RUN_LENGTH = -17,
  B2O_ERR_FILTER_PIPELINE = -18, B2O_ERR_MAX_BUFSIZE = -35
};

/* whole-chunk path (b2c:2577 blosc2_compress_ctx, b2c:2955 blosc2_decompress_ctx, b2c:3696 cbuffer_sizes) */
int b2o_compress(int codec, int clevel, int typesize, int filter, int32_t blocksize, const void* src,
                 int32_t nbytes, void* dst, int32_t dstcap);
int b2o_decompress(const void* src, int32_t srcsize, void* dst, int32_t dstcap);
int b2o_chunk_sizes(const void* chunk, int32_t* nbytes, int32_t* cbytes, int32_t* blocksize);

/* filters (blosc/shuffle.c:421-514) */
void b2o_shuffle(int ts, int32_t n, const uint8_t* src, uint8_t* dst);
void b2o_unshuffle(int ts, int32_t n, const uint8_t* src, uint8_t* dst);
int32_t b2o_bitshuffle(int ts, int32_t n, const uint8_t* src, uint8_t* dst);
int32_t b2o_bitunshuffle(int ts, int32_t n, const uint8_t* src, uint8_t* dst, int format_version);

/* stream codecs */
int b2o_lz4_compress(const uint8_t* src, int n, uint8_t* dst, int cap);   /* LZ4_compress_fast accel=1 */
int b2o_lz4_decompress(const uint8_t* src, int n, uint8_t* dst, int cap); /* LZ4_decompress_safe */
int b2o_blosclz_compress(int clevel, const uint8_t* src, int n, uint8_t* dst, int cap);
int b2o_blosclz_decompress(const uint8_t* src, int n, uint8_t* dst, int cap);
long b2o_zstd_decompress(const uint8_t* src, long n, uint8_t* dst, long cap); /* one frame; <0 on error */

#ifdef __cplusplus
}
#endif
#endif
