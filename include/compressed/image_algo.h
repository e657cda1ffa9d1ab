// compressed/image_algo.h -- interleave / deinterleave with the reference's signatures and checks
// (compressed_image/include/compressed/image_algo.h:33-65 interleave, :83-111 deinterleave into spans, :130-160 into
// vectors), run on the GPU by libb2g.so (b2g_interleave / b2g_deinterleave, csrc/b2g_interleave.cuh).  No CPU path.
#pragma once
#include <algorithm>
#include <cstddef>
#include <mutex>
#include <span>
#include <stdexcept>
#include <string>
#include <vector>

#include <b2g.h>

#include "macros.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	namespace image_algo
	{
		namespace detail
		{
			/// One context for the transposes of this process (the reference functions take none); calls are serialised.
			inline b2g_ctx* context(std::unique_lock<std::mutex>& lock)
			{
				static std::mutex mutex;
				static b2g_ctx* ctx = nullptr;
				lock = std::unique_lock<std::mutex>(mutex);
				if (!ctx)
				{
					b2g_params p{};
					p.codec = 1; p.clevel = 9; p.typesize = 1; p.filter = B2G_SHUFFLE; p.blocksize = 32768; p.device = -1;
					ctx = b2g_create(&p);
					if (!ctx) throw std::runtime_error(std::string("image_algo: no usable CUDA device: ") + b2g_last_error());
				}
				return ctx;
			}
		}

		template <typename T>
		void interleave(std::span<T> buffer, const std::vector<std::span<const T>>& spans)
		{
			if (spans.empty())
				throw std::invalid_argument("Interleave: No spans provided for interleaving.");
			for (const auto& span : spans)
				if (span.size() != spans.front().size())
					throw std::invalid_argument("Interleave: All input spans must have the same size.");
			if (buffer.size() != spans.front().size() * spans.size())
				throw std::invalid_argument("Interleave: Provided buffer is not large enough to hold all the elements to interleave.");
			if (spans.front().empty()) return;
			std::vector<const void*> planes;
			for (const auto& span : spans) planes.push_back(span.data());
			std::unique_lock<std::mutex> lock;
			b2g_ctx* ctx = detail::context(lock);
			const int rc = b2g_interleave(ctx, static_cast<int>(sizeof(T)), static_cast<int>(planes.size()), planes.data(),
				buffer.data(), static_cast<int64_t>(spans.front().size()));
			if (rc < 0) throw std::runtime_error(std::string("Interleave: ") + b2g_strerror(rc));
		}

		template <typename T>
		void deinterleave(std::span<const T> interleaved, std::vector<std::span<T>>& channel_spans)
		{
			if (channel_spans.empty())
				throw std::invalid_argument("Deinterleave: No spans provided for deinterleaving.");
			const std::size_t size = channel_spans.front().size();
			if (!std::all_of(channel_spans.begin(), channel_spans.end(), [size](const std::span<T>& s) { return s.size() == size; }))
				throw std::invalid_argument("Deinterleave: All output spans must have the same size.");
			if (interleaved.size() != size * channel_spans.size())
				throw std::invalid_argument("Deinterleave: Input buffer size does not match the expected size for deinterleaving.");
			if (size == 0) return;
			std::vector<void*> planes;
			for (auto& span : channel_spans) planes.push_back(span.data());
			std::unique_lock<std::mutex> lock;
			b2g_ctx* ctx = detail::context(lock);
			const int rc = b2g_deinterleave(ctx, static_cast<int>(sizeof(T)), static_cast<int>(planes.size()), interleaved.data(),
				planes.data(), static_cast<int64_t>(size));
			if (rc < 0) throw std::runtime_error(std::string("Deinterleave: ") + b2g_strerror(rc));
		}

		template <typename T>
		void deinterleave(std::span<const T> interleaved, std::vector<std::vector<T>>& channel_vecs)
		{
			std::vector<std::span<T>> spans;
			for (auto& v : channel_vecs) spans.push_back(std::span<T>(v.data(), v.size()));
			deinterleave<T>(interleaved, spans);
		}
	}
}
