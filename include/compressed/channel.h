// compressed/channel.h -- compressed::channel<T>: one image channel held as independently compressed blosc2 chunks.
//
// Public interface of the reference's channel<T> (compressed_image/include/compressed/channel.h:29-585), chunk_span<T>
// (containers/chunk_span.h:20-110) and channel_iterator<T> (iterators/iterator.h:28-330) on top of the b2g engine:
// same constructors, static factories, getters, get_chunk / set_chunk / get_decompressed, and the same
// decompress -> user modifies -> recompress iteration.  All chunks of a constructor or of get_decompressed() go to the
// GPU as ONE batch.
#pragma once

#include <cstddef>
#include <cstdint>
#include <iterator>
#include <memory>
#include <span>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "blosc2/schunk.h"
#include "constants.h"
#include "enums.h"
#include "util.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	namespace container
	{
		/// A decompressed chunk handed out by the iterator: a span plus where it sits in the image
		/// (containers/chunk_span.h:38-62).
		template <typename T>
		struct chunk_span : public std::span<T>
		{
			chunk_span() = default;
			chunk_span(std::span<T> data, std::size_t width, std::size_t height, std::size_t chunk_index, std::size_t chunk_size)
				: std::span<T>(data), m_ChunkSize(chunk_size), m_Width(width), m_Height(height), m_ChunkIndex(chunk_index) {}

			/// image x / y coordinate of element `index` of this chunk
			std::size_t x(std::size_t index) const noexcept { return global_index(index) % m_Width; }
			std::size_t y(std::size_t index) const noexcept { return global_index(index) / m_Width; }
			std::size_t chunk_index() const noexcept { return m_ChunkIndex; }

		private:
			std::size_t global_index(std::size_t index) const noexcept { return m_ChunkIndex * (m_ChunkSize / sizeof(T)) + index; }
			std::size_t m_ChunkSize = 1, m_Width = 1, m_Height = 1, m_ChunkIndex = 0;
		};
	}

	/// Forward iterator over the chunks of a channel (iterators/iterator.h:97-141): dereferencing decompresses the
	/// current chunk into a buffer owned by the iterator and returns a chunk_span over it; advancing (and destroying
	/// the iterator) compresses the buffer back into the channel.
	///
	/// Double-buffered: while the caller works on chunk i, chunk i + 1 is already being decompressed
	/// (b2g_decompress_batch_async) and chunk i - 1 is being compressed (b2g_compress_batch_async), on two contexts
	/// cloned from the channel's (b2g_clone) so that calls on the channel from inside the loop never meet a context with
	/// a batch in flight.  Three buffers rotate; a chunk_span stays valid until the iterator has advanced twice.
	template <typename T>
	struct channel_iterator
	{
		using iterator_category = std::forward_iterator_tag;
		using difference_type = std::ptrdiff_t;
		using value_type = container::chunk_span<T>;

		channel_iterator() = default;
		channel_iterator(blosc2::schunk<T>* schunk, blosc2::context_raw_ptr compression_ctx,
			blosc2::context_raw_ptr decompression_ctx, std::size_t chunk_index, std::size_t width, std::size_t height)
			: m_Schunk(schunk), m_Cctx(compression_ctx), m_Dctx(decompression_ctx), m_ChunkIndex(chunk_index), m_Width(width),
			  m_Height(height)
		{
			if (width == 0 || height == 0)
			{
				throw std::runtime_error("passed zero width or height to iterator which is not valid");
			}
		}
		/// A copy refers to the same position but starts with nothing loaded and nothing in flight.
		channel_iterator(const channel_iterator& other)
			: m_Schunk(other.m_Schunk), m_Cctx(other.m_Cctx), m_Dctx(other.m_Dctx), m_ChunkIndex(other.m_ChunkIndex),
			  m_Width(other.m_Width), m_Height(other.m_Height) {}
		channel_iterator(channel_iterator&& other) noexcept
			: m_Schunk(other.m_Schunk), m_Cctx(other.m_Cctx), m_Dctx(other.m_Dctx), m_ChunkIndex(other.m_ChunkIndex),
			  m_Width(other.m_Width), m_Height(other.m_Height), m_Pipe(std::move(other.m_Pipe)), m_Loaded(other.m_Loaded),
			  m_Cur(other.m_Cur)
		{
			other.m_Loaded = npos;
		}
		channel_iterator& operator=(const channel_iterator& other)
		{
			if (this != &other)
			{
				flush();
				m_Schunk = other.m_Schunk; m_Cctx = other.m_Cctx; m_Dctx = other.m_Dctx;
				m_ChunkIndex = other.m_ChunkIndex; m_Width = other.m_Width; m_Height = other.m_Height;
			}
			return *this;
		}
		~channel_iterator()
		{
			try { flush(); } catch (...) {}  // iterator.h:72-93: the last chunk is written back on scope exit
			if (m_Pipe && s_Pool.size() < 8) s_Pool.push_back(std::move(m_Pipe));   // contexts + pinned buffers for the next loop
		}

		value_type operator*()
		{
			if (!m_Schunk || m_ChunkIndex >= m_Schunk->num_chunks())
			{
				throw std::runtime_error("Invalid Iterator struct encountered, cannot dereference item");
			}
			if (m_Loaded != m_ChunkIndex)
			{
				pipeline& p = pipe();
				start_compress();                                    // chunk m_Loaded leaves while the next one is fetched
				if (p.fetch_index == m_ChunkIndex)
				{
					wait_fetch();
					m_Cur = p.fetch_buf;
				}
				else                                                 // first dereference, or a jump
				{
					wait_fetch();
					m_Cur = free_buffer(npos);
					p.buf[m_Cur].resize(m_Schunk->chunk_elements(m_ChunkIndex));
					m_Schunk->chunk(p.dctx.get(), p.buf[m_Cur].span(), m_ChunkIndex);
				}
				m_Loaded = m_ChunkIndex;
				start_fetch(m_ChunkIndex + 1);
			}
			return value_type(m_Pipe->buf[m_Cur].span(), m_Width, m_Height, m_ChunkIndex, m_Schunk->chunk_bytes());
		}

		channel_iterator& operator++()
		{
			if (!m_Schunk || m_ChunkIndex >= m_Schunk->num_chunks())
			{
				throw std::out_of_range("Iterator: unable to increment past the end of the channel");
			}
			++m_ChunkIndex;
			return *this;
		}

		bool operator==(const channel_iterator& other) const noexcept
		{
			return m_Schunk == other.m_Schunk && m_ChunkIndex == other.m_ChunkIndex;
		}
		bool operator!=(const channel_iterator& other) const noexcept { return !(*this == other); }
		std::size_t chunk_index() const noexcept { return m_ChunkIndex; }

	private:
		static constexpr std::size_t npos = static_cast<std::size_t>(-1);

		/// A growable array of page-locked host memory (b2g_host_alloc): what the iterator's batches read and write, so
		/// that b2g_*_batch_async really returns while the copies run (on pageable memory they are a blocking bounce).
		template <typename U>
		struct pinned_array
		{
			pinned_array() = default;
			pinned_array(const pinned_array&) = delete;
			pinned_array& operator=(const pinned_array&) = delete;
			~pinned_array() { b2g_host_free(m_Ptr); }
			void resize(std::size_t count)
			{
				if (count > m_Capacity)
				{
					b2g_host_free(m_Ptr);
					m_Ptr = static_cast<U*>(b2g_host_alloc(count * sizeof(U)));
					m_Capacity = m_Ptr ? count : 0;
					if (!m_Ptr) throw std::runtime_error(std::string("Unable to allocate pinned host memory: ") + b2g_last_error());
				}
				m_Size = count;
			}
			U* data() noexcept { return m_Ptr; }
			std::size_t size() const noexcept { return m_Size; }
			std::span<U> span() noexcept { return std::span<U>(m_Ptr, m_Size); }
		private:
			U* m_Ptr = nullptr;
			std::size_t m_Size = 0, m_Capacity = 0;
		};

		/// Everything a batch in flight points at lives here, on the heap, so that moving the iterator moves nothing.
		struct pipeline
		{
			blosc2::context_ptr dctx, cctx;
			b2g_params key{};                 // parameters of the compress context the pipeline was cloned for
			pinned_array<T> buf[3];
			pinned_array<std::byte> cbuf;
			b2g_chunk fetch_job{}, comp_job{};
			std::size_t fetch_index = npos, comp_index = npos;
			std::size_t fetch_buf = 0, comp_buf = 0;
		};

		/// Pipelines that finished iterators left behind, per thread: creating two contexts (streams, events, device work
		/// buffers on first use) and four pinned buffers costs tens of milliseconds, a loop over a 64 MiB channel a few.
		static inline thread_local std::vector<std::unique_ptr<pipeline>> s_Pool;

		static bool same_params(const b2g_params& a, const b2g_params& b) noexcept
		{
			return a.codec == b.codec && a.clevel == b.clevel && a.typesize == b.typesize && a.filter == b.filter &&
				a.blocksize == b.blocksize && a.device == b.device;
		}

		pipeline& pipe()
		{
			if (!m_Pipe)
			{
				b2g_params want{};
				if (b2g_get_params(m_Cctx, &want) == 0)
				{
					for (std::size_t k = 0; k < s_Pool.size(); ++k)
					{
						if (!same_params(s_Pool[k]->key, want)) continue;
						m_Pipe = std::move(s_Pool[k]);
						s_Pool.erase(s_Pool.begin() + static_cast<std::ptrdiff_t>(k));
						m_Pipe->fetch_index = m_Pipe->comp_index = npos;
						return *m_Pipe;
					}
				}
				m_Pipe = std::make_unique<pipeline>();
				m_Pipe->key = want;
				m_Pipe->dctx = blosc2::context_ptr(b2g_clone(m_Dctx));
				m_Pipe->cctx = blosc2::context_ptr(b2g_clone(m_Cctx));
				if (!m_Pipe->dctx || !m_Pipe->cctx)
				{
					m_Pipe.reset();
					throw std::runtime_error(std::string("Unable to create the iterator's b2g contexts: ") + b2g_last_error());
				}
			}
			return *m_Pipe;
		}

		/// A buffer that is neither the loaded one, nor being compressed, nor `other`.
		std::size_t free_buffer(std::size_t other) const
		{
			for (std::size_t k = 0; k < 3; ++k)
			{
				if (k == other) continue;
				if (m_Loaded != npos && k == m_Cur) continue;
				if (m_Pipe->comp_index != npos && k == m_Pipe->comp_buf) continue;
				return k;
			}
			return 0;  // unreachable: at most two of the three are taken
		}

		void start_fetch(std::size_t index)
		{
			pipeline& p = *m_Pipe;
			if (index >= m_Schunk->num_chunks() || m_Schunk->is_lazy(index)) return;   // lazy chunks are fills: nothing to overlap
			const std::size_t k = free_buffer(npos);
			p.buf[k].resize(m_Schunk->chunk_elements(index));
			const auto chunk = m_Schunk->chunk_data(index);
			p.fetch_job = b2g_chunk{ chunk.data(), p.buf[k].data(), static_cast<int32_t>(chunk.size()),
				static_cast<int32_t>(p.buf[k].size() * sizeof(T)), 0, 0 };
			const int rc = b2g_decompress_batch_async(p.dctx.get(), &p.fetch_job, 1, nullptr);
			if (rc < 0) blosc2::detail::throw_status("Error while decompressing blosc2 chunk", rc);
			p.fetch_index = index;
			p.fetch_buf = k;
		}

		void wait_fetch()
		{
			pipeline& p = *m_Pipe;
			if (p.fetch_index == npos) return;
			p.fetch_index = npos;
			const int rc = b2g_wait(p.dctx.get());
			if (rc < 0 || p.fetch_job.status < 0)
				blosc2::detail::throw_status("Error while decompressing blosc2 chunk", rc < 0 ? rc : p.fetch_job.status);
		}

		void start_compress()
		{
			if (!m_Pipe || m_Loaded == npos) return;
			pipeline& p = *m_Pipe;
			wait_compress();
			const std::size_t bytes = p.buf[m_Cur].size() * sizeof(T);
			p.cbuf.resize(blosc2::min_compressed_size(bytes));
			p.comp_job = b2g_chunk{ p.buf[m_Cur].data(), p.cbuf.data(), static_cast<int32_t>(bytes),
				static_cast<int32_t>(p.cbuf.size()), 0, 0 };
			const int rc = b2g_compress_batch_async(p.cctx.get(), &p.comp_job, 1, nullptr);
			if (rc < 0) blosc2::detail::throw_status("Error while compressing blosc2 chunk", rc);
			p.comp_index = m_Loaded;
			p.comp_buf = m_Cur;
			m_Loaded = npos;
		}

		void wait_compress()
		{
			pipeline& p = *m_Pipe;
			if (p.comp_index == npos) return;
			const std::size_t index = p.comp_index;
			p.comp_index = npos;
			const int rc = b2g_wait(p.cctx.get());
			if (rc < 0 || p.comp_job.status < 0)
				blosc2::detail::throw_status("Error while compressing blosc2 chunk", rc < 0 ? rc : p.comp_job.status);
			m_Schunk->set_chunk(std::vector<std::byte>(p.cbuf.data(), p.cbuf.data() + p.comp_job.out_bytes), index);
		}

		void flush()
		{
			if (!m_Pipe) return;
			try { wait_fetch(); } catch (...) {}     // a prefetch nobody will look at
			start_compress();
			wait_compress();
		}

		blosc2::schunk<T>* m_Schunk = nullptr;
		blosc2::context_raw_ptr m_Cctx = nullptr, m_Dctx = nullptr;
		std::size_t m_ChunkIndex = 0, m_Width = 1, m_Height = 1;
		std::unique_ptr<pipeline> m_Pipe;
		std::size_t m_Loaded = npos, m_Cur = 0;
	};

	template <typename T>
	struct channel
	{
		using value_type = T;
		using iterator = channel_iterator<T>;

		channel(channel&&) = default;
		channel& operator=(channel&&) = default;
		channel(const channel&) = delete;
		channel& operator=(const channel&) = delete;

		/// An empty channel (channel.h:71-81).
		channel()
		{
			create_contexts();
			m_Schunk = std::make_unique<blosc2::schunk<T>>(s_default_blocksize, s_default_chunksize);
		}

		/// Compress `data` (width * height elements) into scanline-aligned chunks (channel.h:97-133).
		channel(const std::span<const T> data, std::size_t width, std::size_t height,
			enums::codec compression_codec = enums::codec::lz4, std::uint8_t compression_level = 9,
			std::size_t block_size = s_default_blocksize, std::size_t chunk_size = s_default_chunksize)
			: m_Codec(compression_codec), m_CompressionLevel(util::ensure_compression_level(compression_level)),
			  m_Width(width), m_Height(height), m_BlockSize(block_size)
		{
			if (data.size() != width * height)
			{
				throw std::runtime_error("Invalid channel data passed. Expected its size to match up to width * height (" +
					std::to_string(width) + " * " + std::to_string(height) + ") which would be " +
					std::to_string(width * height) + ". Instead got " + std::to_string(data.size()));
			}
			create_contexts();
			const std::size_t chunk_size_aligned = util::align_chunk_to_scanlines_bytes<T>(width, chunk_size);
			m_Schunk = std::make_unique<blosc2::schunk<T>>(data, block_size, chunk_size_aligned, m_CompressionContext);
		}

		/// The constructor above in two steps, for compressed::image<T>, which compresses the chunks of ALL its channels
		/// with one launch: deferred() validates and sets the channel up without compressing, plan_compress() appends
		/// the channel's chunk jobs to the image-wide list, adopt_compressed() takes the results.
		struct deferred_t {};
		channel(deferred_t, std::size_t data_size, std::size_t width, std::size_t height, enums::codec compression_codec,
			std::uint8_t compression_level, std::size_t block_size, std::size_t chunk_size)
			: m_Codec(compression_codec), m_CompressionLevel(util::ensure_compression_level(compression_level)),
			  m_Width(width), m_Height(height), m_BlockSize(block_size)
		{
			if (data_size != width * height)
			{
				throw std::runtime_error("Invalid channel data passed. Expected its size to match up to width * height (" +
					std::to_string(width) + " * " + std::to_string(height) + ") which would be " +
					std::to_string(width * height) + ". Instead got " + std::to_string(data_size));
			}
			create_contexts();
			m_Schunk = std::make_unique<blosc2::schunk<T>>(block_size, util::align_chunk_to_scanlines_bytes<T>(width, chunk_size));
		}
		typename blosc2::schunk<T>::compress_plan plan_compress(std::span<const T> data,
			std::vector<blosc2::compress_job<T>>& jobs) const { return m_Schunk->plan(data, jobs); }
		void adopt_compressed(typename blosc2::schunk<T>::compress_plan&& pl, std::span<const blosc2::compress_job<T>> jobs)
		{
			m_Schunk->adopt(std::move(pl), jobs);
		}
		blosc2::context_raw_ptr compression_context() const noexcept { return m_CompressionContext.get(); }
		blosc2::context_raw_ptr decompression_context() const noexcept { return m_DecompressionContext.get(); }

		/// Adopt an existing super-chunk (channel.h:150-201).
		channel(blosc2::schunk<T> schunk, std::size_t width, std::size_t height,
			enums::codec compression_codec = enums::codec::lz4, std::uint8_t compression_level = 9)
			: m_Codec(compression_codec), m_CompressionLevel(util::ensure_compression_level(compression_level)),
			  m_Width(width), m_Height(height), m_BlockSize(schunk.max_block_size())
		{
			if (schunk.size() != width * height)
			{
				throw std::runtime_error("Invalid schunk passed to compressed::channel constructor. Expected a size of " +
					std::to_string(width * height) + " but instead got " + std::to_string(schunk.size()));
			}
			create_contexts();
			m_Schunk = std::make_unique<blosc2::schunk<T>>(std::move(schunk));
		}

		/// Lazy channels: nothing is compressed until a chunk is written (channel.h:219-304).
		static channel full(std::size_t width, std::size_t height, T fill_value,
			enums::codec compression_codec = enums::codec::lz4, std::uint8_t compression_level = 9,
			std::size_t block_size = s_default_blocksize, std::size_t chunk_size = s_default_chunksize)
		{
			const std::size_t chunk_size_aligned = util::align_chunk_to_scanlines_bytes<T>(width, chunk_size);
			blosc2::schunk<T> lazy(fill_value, width * height, block_size, chunk_size_aligned);
			channel ch(std::move(lazy), width, height, compression_codec, compression_level);
			ch.m_BlockSize = block_size;
			return ch;
		}
		static channel zeros(std::size_t width, std::size_t height, enums::codec compression_codec = enums::codec::lz4,
			std::uint8_t compression_level = 9, std::size_t block_size = s_default_blocksize,
			std::size_t chunk_size = s_default_chunksize)
		{
			return full(width, height, static_cast<T>(0), compression_codec, compression_level, block_size, chunk_size);
		}
		static channel full_like(const channel& other, T fill_value)
		{
			return full(other.width(), other.height(), fill_value, other.compression(), other.compression_level(),
				other.block_size(), other.chunk_size());
		}
		static channel zeros_like(const channel& other) { return full_like(other, static_cast<T>(0)); }

		/// Chunk iteration (channel.h:309-327).
		iterator begin() { return iterator(m_Schunk.get(), m_CompressionContext.get(), m_DecompressionContext.get(), 0, m_Width, m_Height); }
		iterator end()
		{
			return iterator(m_Schunk.get(), m_CompressionContext.get(), m_DecompressionContext.get(), m_Schunk->num_chunks(),
				m_Width, m_Height);
		}

		blosc2::context_raw_ptr compression_context() { return m_CompressionContext.get(); }
		blosc2::context_raw_ptr decompression_context() { return m_DecompressionContext.get(); }

		/// channel.h:343-348.  The grid is the thread pool; only the block size has an effect.
		void update_nthreads(std::size_t nthreads, std::size_t block_size = s_default_blocksize)
		{
			m_Nthreads = nthreads;
			m_BlockSize = block_size;
			create_contexts();
		}

		std::size_t width() const noexcept { return m_Width; }
		std::size_t height() const noexcept { return m_Height; }
		enums::codec compression() const noexcept { return m_Codec; }
		std::uint8_t compression_level() const noexcept { return m_CompressionLevel; }
		std::size_t compressed_bytes() const { return m_Schunk->csize(); }
		/// uncompressed size in ELEMENTS (channel.h:397-416)
		std::size_t uncompressed_size() const { return m_Schunk->size(); }
		std::size_t num_chunks() const { return m_Schunk->num_chunks(); }
		std::size_t block_size() const { return m_Schunk->max_block_size(); }
		std::size_t chunk_size() const noexcept { return m_Schunk->chunk_bytes(); }
		std::size_t chunk_elems() const { return m_Schunk->chunk_elements(); }
		std::size_t chunk_size(std::size_t chunk_index) const { return m_Schunk->chunk_bytes(chunk_index); }
		std::size_t chunk_elems(std::size_t chunk_index) const { return m_Schunk->chunk_elements(chunk_index); }

		/// channel.h:502-516
		void get_chunk(std::span<T> buffer, std::size_t chunk_idx) const
		{
			m_Schunk->chunk(m_DecompressionContext.get(), buffer, chunk_idx);
		}
		/// channel.h:527-538
		void set_chunk(std::span<T> buffer, std::size_t chunk_idx)
		{
			m_Schunk->set_chunk(m_CompressionContext.get(), std::span<const T>(buffer.data(), buffer.size()), chunk_idx);
		}
		/// Several chunks per launch (b2g extension of get_chunk / set_chunk).
		void get_chunks(std::span<const std::size_t> chunk_indices, std::span<const std::span<T>> buffers) const
		{
			m_Schunk->chunks(m_DecompressionContext.get(), chunk_indices, buffers);
		}
		void set_chunks(std::span<const std::size_t> chunk_indices, std::span<const std::span<const T>> buffers)
		{
			m_Schunk->set_chunks(m_CompressionContext.get(), chunk_indices, buffers);
		}
		/// channel.h:545-558: the whole channel, every chunk in one launch.
		std::vector<T> get_decompressed() const
		{
			return m_Schunk->to_uncompressed(m_DecompressionContext);
		}

		/// Direct access to the chunk container (the reference exposes it through its iterators and to_schunk()).
		blosc2::schunk<T>& schunk() noexcept { return *m_Schunk; }
		const blosc2::schunk<T>& schunk() const noexcept { return *m_Schunk; }

		bool operator==(const channel<T>& other) const noexcept { return this == &other; }

	private:
		void create_contexts()
		{
			m_CompressionContext = blosc2::create_compression_context<T>(m_Nthreads, m_Codec, m_CompressionLevel, m_BlockSize);
			m_DecompressionContext = blosc2::create_decompression_context(m_Nthreads);
		}

		std::unique_ptr<blosc2::schunk<T>> m_Schunk;
		enums::codec m_Codec = enums::codec::lz4;
		std::uint8_t m_CompressionLevel = 9;
		std::size_t m_Nthreads = std::thread::hardware_concurrency() / 2;
		std::size_t m_Width = 1, m_Height = 1;
		std::size_t m_BlockSize = s_default_blocksize;
		mutable blosc2::context_ptr m_CompressionContext;
		mutable blosc2::context_ptr m_DecompressionContext;
	};
}
