// compressed/blosc2/schunk.h -- the super-chunk of a channel: a vector of independently compressed blosc2 chunks.
//
// Mirrors the interface of the reference's schunk<T> / lazy_schunk<T> / schunk_mixin<T>
// (compressed_image/include/compressed/blosc2/schunk.h:22-276, lazyschunk.h:47-370, schunk_mixin.h:24-262) in ONE class:
// a chunk is either compressed bytes or a lazy fill value (lazyschunk.h:82-105) that costs sizeof(T) until it is
// written.  Where the reference loops over chunks calling c-blosc2 once per chunk (schunk.h:85-104, 123-141) this class
// fills job lists and makes ONE batched call (blosc2::compress_batch / decompress_batch).
//
// Exceptions are the reference's: std::out_of_range for a bad chunk index (schunk_mixin.h:204-214),
// std::invalid_argument for a buffer / chunk of the wrong size (schunk.h:168-176, schunk_mixin.h:216-246),
// std::runtime_error for engine failures.
#pragma once

#include <algorithm>
#include <cstddef>
#include <cstring>
#include <span>
#include <stdexcept>
#include <string>
#include <variant>
#include <vector>

#include "wrapper.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	namespace blosc2
	{
		template <typename T>
		struct schunk
		{
			/// A lazy chunk: `nbytes` of one repeated element (lazyschunk.h:60-80).
			struct lazy_chunk
			{
				T value{};
				std::size_t nbytes = 0;
			};
			using chunk_t = std::variant<std::vector<std::byte>, lazy_chunk>;

			schunk() = default;

			/// An empty super-chunk (schunk.h:50-56).
			schunk(std::size_t block_size, std::size_t chunk_size) : m_BlockSize(block_size), m_ChunkSize(chunk_size)
			{
				validate_chunk_size(chunk_size);
			}

			/// Compress `data` into chunks of chunk_size bytes (the last one may be shorter): schunk.h:65-105, with every
			/// chunk of the channel in one launch.
			schunk(std::span<const T> data, std::size_t block_size, std::size_t chunk_size, blosc2::context_ptr& compression_ctx)
				: m_BlockSize(block_size), m_ChunkSize(chunk_size)
			{
				validate_chunk_size(chunk_size);
				const std::size_t per_chunk = chunk_size / sizeof(T);
				std::vector<compress_job<T>> jobs;
				std::vector<std::vector<std::byte>> buffers;
				for (std::size_t offset = 0; offset < data.size(); offset += per_chunk)
				{
					const std::size_t count = std::min(per_chunk, data.size() - offset);
					buffers.emplace_back(blosc2::min_compressed_size(count * sizeof(T)));
					jobs.push_back({ data.subspan(offset, count), std::span<std::byte>(buffers.back()), 0 });
				}
				blosc2::compress_batch<T>(compression_ctx.get(), std::span<compress_job<T>>(jobs));
				for (std::size_t i = 0; i < jobs.size(); ++i)
				{
					buffers[i].resize(jobs[i].cbytes);
					buffers[i].shrink_to_fit();
					m_Chunks.emplace_back(std::move(buffers[i]));
				}
			}

			/// The two halves of the constructor above, for callers that put the chunks of SEVERAL super-chunks into one
			/// launch (compressed::image<T>: every channel of the image): plan() sizes the output buffers and appends one
			/// job per chunk to `jobs`; after blosc2::compress_batch() over the combined list, adopt() takes this
			/// super-chunk's `count` results starting at jobs[first].
			struct compress_plan
			{
				std::vector<std::vector<std::byte>> buffers;
				std::size_t first = 0, count = 0;
			};
			compress_plan plan(std::span<const T> data, std::vector<compress_job<T>>& jobs) const
			{
				compress_plan pl;
				pl.first = jobs.size();
				const std::size_t per_chunk = m_ChunkSize / sizeof(T);
				for (std::size_t offset = 0; offset < data.size(); offset += per_chunk)
				{
					const std::size_t count = std::min(per_chunk, data.size() - offset);
					pl.buffers.emplace_back(blosc2::min_compressed_size(count * sizeof(T)));
					jobs.push_back({ data.subspan(offset, count), std::span<std::byte>(pl.buffers.back()), 0 });
				}
				pl.count = jobs.size() - pl.first;
				return pl;
			}
			void adopt(compress_plan&& pl, std::span<const compress_job<T>> jobs)
			{
				for (std::size_t i = 0; i < pl.count; ++i)
				{
					pl.buffers[i].resize(jobs[pl.first + i].cbytes);
					pl.buffers[i].shrink_to_fit();
					m_Chunks.emplace_back(std::move(pl.buffers[i]));
				}
			}
			/// One decompress job per stored chunk of this super-chunk into consecutive pieces of `out` (lazy chunks are
			/// filled on the spot); the caller runs blosc2::decompress_batch() over the combined list.
			void collect_decompress(std::span<T> out, std::vector<decompress_job<T>>& jobs) const
			{
				std::size_t offset = 0;
				for (std::size_t i = 0; i < m_Chunks.size(); ++i)
				{
					const std::size_t count = this->chunk_elements(i);
					auto piece = out.subspan(offset, count);
					if (const auto* lazy = std::get_if<lazy_chunk>(&m_Chunks[i])) std::fill(piece.begin(), piece.end(), lazy->value);
					else jobs.push_back({ piece, std::span<const std::byte>(std::get<0>(m_Chunks[i])), 0 });
					offset += count;
				}
			}

			/// `num_elements` copies of `value`, never compressed until written (lazyschunk.h:82-105).
			schunk(T value, std::size_t num_elements, std::size_t block_size, std::size_t chunk_size)
				: m_BlockSize(block_size), m_ChunkSize(chunk_size)
			{
				validate_chunk_size(chunk_size);
				const std::size_t total = num_elements * sizeof(T);
				for (std::size_t offset = 0; offset < total; offset += chunk_size)
				{
					m_Chunks.emplace_back(lazy_chunk{ value, std::min(chunk_size, total - offset) });
				}
			}

			// ---- whole super-chunk -----------------------------------------------------------------------------------
			/// schunk.h:123-141 / lazyschunk.h:176-194: every chunk decompressed into one vector, one launch.
			std::vector<T> to_uncompressed(blosc2::context_ptr& decompression_ctx) const
			{
				std::vector<T> data(this->size());
				std::vector<decompress_job<T>> jobs;
				std::size_t offset = 0;
				for (std::size_t i = 0; i < m_Chunks.size(); ++i)
				{
					const std::size_t count = this->chunk_elements(i);
					auto out = std::span<T>(data.data() + offset, count);
					if (const auto* lazy = std::get_if<lazy_chunk>(&m_Chunks[i])) std::fill(out.begin(), out.end(), lazy->value);
					else jobs.push_back({ out, std::span<const std::byte>(std::get<0>(m_Chunks[i])), 0 });
					offset += count;
				}
				if (!jobs.empty()) blosc2::decompress_batch<T>(decompression_ctx.get(), std::span<decompress_job<T>>(jobs));
				return data;
			}

			// ---- one chunk ---------------------------------------------------------------------------------------------
			/// schunk.h:143-180: decompress chunk `index` into a new vector / into `buffer` (which must have its size).
			std::vector<T> chunk(blosc2::context_raw_ptr decompression_ctx, std::size_t index) const
			{
				validate_chunk_index(index);
				std::vector<T> out(this->chunk_elements(index));
				this->chunk(decompression_ctx, std::span<T>(out), index);
				return out;
			}
			std::vector<T> chunk(blosc2::context_ptr& decompression_ctx, std::size_t index) const
			{
				return this->chunk(decompression_ctx.get(), index);
			}
			void chunk(blosc2::context_raw_ptr decompression_ctx, std::span<T> buffer, std::size_t index) const
			{
				validate_chunk_index(index);
				if (buffer.size() != this->chunk_elements(index))
				{
					throw std::invalid_argument("Invalid buffer size passed to chunk(): expected " +
						std::to_string(this->chunk_elements(index)) + " elements, got " + std::to_string(buffer.size()));
				}
				if (const auto* lazy = std::get_if<lazy_chunk>(&m_Chunks[index]))
				{
					std::fill(buffer.begin(), buffer.end(), lazy->value);
					return;
				}
				blosc2::decompress<T>(decompression_ctx, buffer, std::span<const std::byte>(std::get<0>(m_Chunks[index])));
			}
			void chunk(blosc2::context_ptr& decompression_ctx, std::span<T> buffer, std::size_t index) const
			{
				this->chunk(decompression_ctx.get(), buffer, index);
			}

			/// Decompress several chunks in one launch (no reference equivalent: the point of the batched ABI).
			void chunks(blosc2::context_raw_ptr decompression_ctx, std::span<const std::size_t> indices,
				std::span<const std::span<T>> buffers) const
			{
				std::vector<decompress_job<T>> jobs;
				for (std::size_t k = 0; k < indices.size(); ++k)
				{
					validate_chunk_index(indices[k]);
					if (buffers[k].size() != this->chunk_elements(indices[k]))
					{
						throw std::invalid_argument("Invalid buffer size passed to chunks()");
					}
					if (const auto* lazy = std::get_if<lazy_chunk>(&m_Chunks[indices[k]]))
						std::fill(buffers[k].begin(), buffers[k].end(), lazy->value);
					else jobs.push_back({ buffers[k], std::span<const std::byte>(std::get<0>(m_Chunks[indices[k]])), 0 });
				}
				if (!jobs.empty()) blosc2::decompress_batch<T>(decompression_ctx, std::span<decompress_job<T>>(jobs));
			}

			/// schunk.h:182-208: replace chunk `index` by already compressed bytes / by compressing `uncompressed`.
			void set_chunk(std::vector<std::byte> compressed, std::size_t index)
			{
				validate_chunk_index(index);
				if (blosc2::chunk_num_elements<std::byte>(compressed) != this->chunk_bytes(index))
				{
					throw std::invalid_argument("Invalid chunk passed to set_chunk(): it decompresses to a different size");
				}
				m_Chunks[index] = std::move(compressed);
			}
			void set_chunk(std::span<const std::byte> compressed, std::size_t index)
			{
				this->set_chunk(std::vector<std::byte>(compressed.begin(), compressed.end()), index);
			}
			void set_chunk(blosc2::context_ptr& compression_ctx, std::span<T> uncompressed, std::size_t index)
			{
				this->set_chunk(compression_ctx.get(), std::span<const T>(uncompressed.data(), uncompressed.size()), index);
			}
			void set_chunk(blosc2::context_raw_ptr compression_ctx, std::span<const T> uncompressed, std::size_t index)
			{
				validate_chunk_index(index);
				if (uncompressed.size() != this->chunk_elements(index))
				{
					throw std::invalid_argument("Invalid chunk size passed to set_chunk(): expected " +
						std::to_string(this->chunk_elements(index)) + " elements, got " + std::to_string(uncompressed.size()));
				}
				std::vector<std::byte> buffer(blosc2::min_compressed_size(uncompressed.size() * sizeof(T)));
				buffer.resize(blosc2::compress<T>(compression_ctx, uncompressed, std::span<std::byte>(buffer)));
				buffer.shrink_to_fit();
				m_Chunks[index] = std::move(buffer);
			}

			/// Replace several chunks in one launch (no reference equivalent).
			void set_chunks(blosc2::context_raw_ptr compression_ctx, std::span<const std::size_t> indices,
				std::span<const std::span<const T>> data)
			{
				std::vector<compress_job<T>> jobs;
				std::vector<std::vector<std::byte>> buffers;
				for (std::size_t k = 0; k < indices.size(); ++k)
				{
					validate_chunk_index(indices[k]);
					if (data[k].size() != this->chunk_elements(indices[k]))
					{
						throw std::invalid_argument("Invalid chunk size passed to set_chunks()");
					}
					buffers.emplace_back(blosc2::min_compressed_size(data[k].size() * sizeof(T)));
				}
				for (std::size_t k = 0; k < indices.size(); ++k) jobs.push_back({ data[k], std::span<std::byte>(buffers[k]), 0 });
				blosc2::compress_batch<T>(compression_ctx, std::span<compress_job<T>>(jobs));
				for (std::size_t k = 0; k < indices.size(); ++k)
				{
					buffers[k].resize(jobs[k].cbytes);
					m_Chunks[indices[k]] = std::move(buffers[k]);
				}
			}

			/// schunk.h:213-248: only the last chunk may be shorter than chunk_bytes().
			void append_chunk(std::vector<std::byte> compressed)
			{
				if (!m_Chunks.empty() && this->chunk_bytes(m_Chunks.size() - 1) != m_ChunkSize)
				{
					throw std::invalid_argument("Cannot append a chunk after a chunk that is smaller than the chunk size");
				}
				if (blosc2::chunk_num_elements<std::byte>(compressed) > m_ChunkSize)
				{
					throw std::invalid_argument("Cannot append a chunk that is larger than the chunk size");
				}
				m_Chunks.emplace_back(std::move(compressed));
			}
			void append_chunk(blosc2::context_ptr& compression_ctx, std::span<T> uncompressed)
			{
				std::vector<std::byte> buffer(blosc2::min_compressed_size(uncompressed.size() * sizeof(T)));
				buffer.resize(blosc2::compress<T>(compression_ctx, uncompressed, std::span<std::byte>(buffer)));
				this->append_chunk(std::move(buffer));
			}

			// ---- sizes (schunk_mixin.h:118-194) ------------------------------------------------------------------------
			std::size_t num_chunks() const noexcept { return m_Chunks.size(); }
			std::size_t chunk_bytes() const noexcept { return m_ChunkSize; }
			std::size_t chunk_bytes(std::size_t index) const
			{
				validate_chunk_index(index);
				if (const auto* lazy = std::get_if<lazy_chunk>(&m_Chunks[index])) return lazy->nbytes;
				return blosc2::chunk_num_elements<std::byte>(std::get<0>(m_Chunks[index]));
			}
			std::size_t chunk_elements() const noexcept { return m_ChunkSize / sizeof(T); }
			std::size_t chunk_elements(std::size_t index) const { return this->chunk_bytes(index) / sizeof(T); }
			/// compressed bytes held (a lazy chunk counts sizeof(T): lazyschunk.h:312-327)
			std::size_t csize() const noexcept
			{
				std::size_t total = 0;
				for (const auto& c : m_Chunks) total += std::holds_alternative<lazy_chunk>(c) ? sizeof(T) : std::get<0>(c).size();
				return total;
			}
			/// uncompressed ELEMENTS
			std::size_t size() const
			{
				std::size_t total = 0;
				for (std::size_t i = 0; i < m_Chunks.size(); ++i) total += this->chunk_bytes(i);
				return total / sizeof(T);
			}
			std::size_t byte_size() const { return this->size() * sizeof(T); }
			std::size_t max_chunk_size() const noexcept { return m_ChunkSize; }
			std::size_t max_block_size() const noexcept { return m_BlockSize; }
			bool is_lazy(std::size_t index) const { validate_chunk_index(index); return std::holds_alternative<lazy_chunk>(m_Chunks[index]); }
			/// the compressed bytes of chunk `index` (a lazy chunk has none: std::invalid_argument)
			std::span<const std::byte> chunk_data(std::size_t index) const
			{
				validate_chunk_index(index);
				if (std::holds_alternative<lazy_chunk>(m_Chunks[index])) throw std::invalid_argument("chunk is lazy");
				return std::span<const std::byte>(std::get<0>(m_Chunks[index]));
			}

			void validate_chunk_index(std::size_t index) const
			{
				if (index >= m_Chunks.size())
				{
					throw std::out_of_range("Cannot access index " + std::to_string(index) +
						" in schunk. Total amount of chunks is " + std::to_string(m_Chunks.size()));
				}
			}

		private:
			static void validate_chunk_size(std::size_t chunk_size)
			{
				if (chunk_size == 0 || chunk_size % sizeof(T) != 0)
				{
					throw std::invalid_argument("chunk_size must be a non-zero multiple of sizeof(T)");
				}
			}

			std::vector<chunk_t> m_Chunks;
			std::size_t m_BlockSize = s_default_blocksize;
			std::size_t m_ChunkSize = s_default_chunksize;
		};

		/// The reference's second container type: same class here.
		template <typename T>
		using lazy_schunk = schunk<T>;
	}
}
