// compressed/blosc2/wrapper.h -- the `compressed::blosc2` seam of compressed-image, re-pointed at the b2g C ABI.
//
// The reference's wrapper (compressed_image/include/compressed/blosc2/wrapper.h) is the only place that calls into
// c-blosc2; every caller above it (schunk, channel, iterator, image) uses the signatures below.  This header keeps
// those signatures -- context_ptr, create_compression_context<T>, create_decompression_context, compress<T>,
// decompress<T>, min_compressed_size, chunk_num_elements<T>, codec mapping -- and adds what the GPU needs to be worth
// calling: compress_batch / decompress_batch, which hand MANY chunks to one launch (include/b2g.h).
//
// Error behaviour follows the reference: a negative engine status throws std::runtime_error (wrapper.h:146-149,
// 254-257), an output buffer beyond INT32_MAX throws std::out_of_range (wrapper.h:241-246).
#pragma once

#include <algorithm>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <memory>
#include <span>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../b2g.h"
#include "../constants.h"
#include "../enums.h"
#include "../macros.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	namespace blosc2
	{
		/// A b2g context serves both directions; the reference keeps one blosc2_context per direction
		/// (wrapper.h:41-67), so both of its pointer types map onto the same handle here.
		struct context_deleter
		{
			void operator()(b2g_ctx* ctx) const noexcept { b2g_destroy(ctx); }
		};
		using context_ptr = std::unique_ptr<b2g_ctx, context_deleter>;
		using context_raw_ptr = b2g_ctx*;

		/// enums::codec -> blosc2 compcode (wrapper.h:74-96).
		inline std::uint8_t codec_to_blosc2(enums::codec compcode)
		{
			switch (compcode)
			{
			case enums::codec::blosclz: return B2G_BLOSCLZ;
			case enums::codec::lz4: return B2G_LZ4;
			case enums::codec::lz4hc: return B2G_LZ4HC;
			case enums::codec::zstd: return B2G_ZSTD;
			}
			return B2G_BLOSCLZ;
		}

		/// blosc2 compcode -> enums::codec (wrapper.h:100-119).
		inline enums::codec blosc2_to_codec(std::uint8_t compcode)
		{
			switch (compcode)
			{
			case B2G_LZ4: return enums::codec::lz4;
			case B2G_LZ4HC: return enums::codec::lz4hc;
			case B2G_ZSTD: return enums::codec::zstd;
			default: return enums::codec::blosclz;
			}
		}

		namespace detail
		{
			inline context_ptr make_context(std::uint8_t codec, std::uint8_t level, std::size_t typesize, std::size_t block_size)
			{
				b2g_params p{};
				p.codec = codec;
				p.clevel = level;
				p.typesize = static_cast<std::uint8_t>(typesize);
				p.filter = B2G_SHUFFLE;  // the reference never overrides blosc2's default filter pipeline
				p.blocksize = static_cast<std::int32_t>(block_size);
				p.device = -1;
				context_ptr ctx(b2g_create(&p));
				if (!ctx)
				{
					throw std::runtime_error(std::string("Unable to create a b2g context: ") + b2g_last_error());
				}
				return ctx;
			}

			[[noreturn]] inline void throw_status(const char* what, int code)
			{
				throw std::runtime_error(std::string(what) + ": " + b2g_strerror(code) + " (code " + std::to_string(code) + ")");
			}
		}

		/// wrapper.h:362-378.  `nthreads` is accepted for source compatibility: the CUDA grid is the thread pool.
		template <typename T>
		context_ptr create_compression_context(std::size_t /*nthreads*/, enums::codec codec, std::uint8_t compression_level,
			std::size_t block_size = s_default_blocksize)
		{
			return detail::make_context(codec_to_blosc2(codec), compression_level, sizeof(T), block_size);
		}

		/// wrapper.h:381-413.  Decoding takes codec, filter, typesize and blocksize from each chunk's own header; the
		/// context is created zstd-capable so any chunk the reference can produce decodes.
		inline context_ptr create_decompression_context(std::size_t /*nthreads*/)
		{
			return detail::make_context(B2G_ZSTD, 9, 1, s_default_blocksize);
		}

		/// wrapper.h:416-439
		inline constexpr std::size_t min_compressed_size(std::size_t chunk_size) { return chunk_size + B2G_MAX_OVERHEAD; }
		inline constexpr std::size_t min_decompressed_size(std::size_t chunk_size) { return chunk_size; }

		/// One chunk of a batch: raw elements in, blosc2 chunk out (compress) or the other way round (decompress).
		template <typename T>
		struct compress_job
		{
			std::span<const T> data;
			std::span<std::byte> chunk;
			std::size_t cbytes = 0;  // out
		};
		template <typename T>
		struct decompress_job
		{
			std::span<T> buffer;
			std::span<const std::byte> chunk;
			std::size_t nbytes = 0;  // out
		};

		/// All jobs in ONE launch (replaces the per-chunk loops of schunk.h:85-104).  Throws like compress<T>.
		template <typename T>
		void compress_batch(context_raw_ptr context, std::span<compress_job<T>> jobs)
		{
			std::vector<b2g_chunk> batch(jobs.size());
			for (std::size_t i = 0; i < jobs.size(); ++i)
			{
				if (jobs[i].data.size() * sizeof(T) > static_cast<std::size_t>(std::numeric_limits<std::int32_t>::max()))
				{
					throw std::out_of_range("Chunk too large: a blosc2 chunk holds at most INT32_MAX bytes");
				}
				batch[i].src = jobs[i].data.data();
				batch[i].dst = jobs[i].chunk.data();
				batch[i].src_bytes = static_cast<std::int32_t>(jobs[i].data.size() * sizeof(T));
				batch[i].dst_capacity = static_cast<std::int32_t>(
					std::min<std::size_t>(jobs[i].chunk.size(), std::numeric_limits<std::int32_t>::max()));
			}
			const int rc = b2g_compress_batch(context, batch.data(), static_cast<int>(batch.size()), nullptr);
			if (rc < 0) detail::throw_status("Error while compressing blosc2 chunk", rc);
			for (std::size_t i = 0; i < jobs.size(); ++i)
			{
				if (batch[i].status < 0) detail::throw_status("Error while compressing blosc2 chunk", batch[i].status);
				if (batch[i].out_bytes == 0 && !jobs[i].data.empty())
				{
					throw std::runtime_error("Error while compressing blosc2 chunk: destination buffer too small");
				}
				jobs[i].cbytes = static_cast<std::size_t>(batch[i].out_bytes);
			}
		}

		/// All jobs in ONE launch (replaces the per-chunk loops of schunk.h:123-141).  Throws like decompress<T>.
		template <typename T>
		void decompress_batch(context_raw_ptr context, std::span<decompress_job<T>> jobs)
		{
			std::vector<b2g_chunk> batch(jobs.size());
			for (std::size_t i = 0; i < jobs.size(); ++i)
			{
				if (jobs[i].buffer.size() * sizeof(T) > static_cast<std::size_t>(std::numeric_limits<std::int32_t>::max()))
				{
					throw std::out_of_range("Buffer too large: a blosc2 chunk decompresses to at most INT32_MAX bytes");
				}
				batch[i].src = jobs[i].chunk.data();
				batch[i].dst = jobs[i].buffer.data();
				// the reference passes INT32_MAX and trusts the header's cbytes (wrapper.h:247); a span knows its size
				batch[i].src_bytes = static_cast<std::int32_t>(
					std::min<std::size_t>(jobs[i].chunk.size(), std::numeric_limits<std::int32_t>::max()));
				batch[i].dst_capacity = static_cast<std::int32_t>(jobs[i].buffer.size() * sizeof(T));
			}
			const int rc = b2g_decompress_batch(context, batch.data(), static_cast<int>(batch.size()), nullptr);
			if (rc < 0) detail::throw_status("Error while decompressing blosc2 chunk", rc);
			for (std::size_t i = 0; i < jobs.size(); ++i)
			{
				if (batch[i].status < 0) detail::throw_status("Error while decompressing blosc2 chunk", batch[i].status);
				jobs[i].nbytes = static_cast<std::size_t>(batch[i].out_bytes);
			}
		}

		/// wrapper.h:134-223: one chunk.  Returns cbytes.
		template <typename T>
		std::size_t compress(context_raw_ptr context, std::span<const T> data, std::span<std::byte> chunk)
		{
			compress_job<T> job{ data, chunk, 0 };
			compress_batch<T>(context, std::span<compress_job<T>>(&job, 1));
			return job.cbytes;
		}
		template <typename T>
		std::size_t compress(context_raw_ptr context, std::span<T> data, std::span<std::byte> chunk)
		{
			return compress<T>(context, std::span<const T>(data.data(), data.size()), chunk);
		}
		template <typename T>
		std::size_t compress(context_ptr& context, std::span<const T> data, std::span<std::byte> chunk)
		{
			return compress<T>(context.get(), data, chunk);
		}
		template <typename T>
		std::size_t compress(context_ptr& context, std::span<T> data, std::span<std::byte> chunk)
		{
			return compress<T>(context.get(), std::span<const T>(data.data(), data.size()), chunk);
		}

		/// wrapper.h:236-277: one chunk.  Returns the decompressed size in bytes.
		template <typename T>
		std::size_t decompress(context_raw_ptr context, std::span<T> buffer, std::span<const std::byte> chunk)
		{
			decompress_job<T> job{ buffer, chunk, 0 };
			decompress_batch<T>(context, std::span<decompress_job<T>>(&job, 1));
			return job.nbytes;
		}
		template <typename T>
		std::size_t decompress(context_ptr& context, std::span<T> buffer, std::span<const std::byte> chunk)
		{
			return decompress<T>(context.get(), buffer, chunk);
		}

		/// wrapper.h:447-468: number of T elements a chunk decompresses to, from its header.
		template <typename T>
		std::size_t chunk_num_elements(std::span<const std::byte> chunk)
		{
			if (chunk.size() < 16)
			{
				throw std::runtime_error("Unable to read blosc2 chunk header: chunk is shorter than 16 bytes");
			}
			std::int32_t nbytes = 0, cbytes = 0, blocksize = 0;
			const int rc = b2g_chunk_sizes(chunk.data(), &nbytes, &cbytes, &blocksize);
			if (rc < 0) detail::throw_status("Unable to read blosc2 chunk header", rc);
			return static_cast<std::size_t>(nbytes) / sizeof(T);
		}
		template <typename T>
		std::size_t chunk_num_elements(const std::vector<std::byte>& chunk)
		{
			return chunk_num_elements<T>(std::span<const std::byte>(chunk.data(), chunk.size()));
		}
	}
}
