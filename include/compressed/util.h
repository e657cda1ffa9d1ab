// compressed/util.h -- the two helpers of the reference's util.h that the chunk path depends on
// (compressed_image/include/compressed/util.h:86-94 ensure_compression_level, :115-131 align_chunk_to_scanlines_bytes).
#pragma once
#include <cstddef>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "macros.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	namespace util
	{
		namespace detail
		{
			/// Allocator whose construct() default-initialises: a vector of T resized to n does not zero its n elements.
			template <typename T, typename A = std::allocator<T>>
			struct default_init_allocator : public A
			{
				using A::A;
				template <typename U>
				struct rebind
				{
					using other = default_init_allocator<U, typename std::allocator_traits<A>::template rebind_alloc<U>>;
				};
				template <typename U>
				void construct(U* ptr) noexcept(std::is_nothrow_default_constructible_v<U>)
				{
					::new (static_cast<void*>(ptr)) U;
				}
				template <typename U, typename... Args>
				void construct(U* ptr, Args&&... args)
				{
					std::allocator_traits<A>::construct(static_cast<A&>(*this), ptr, std::forward<Args>(args)...);
				}
			};
		}

		/// util.h:56-57 -- a std::vector that skips value-initialisation on resize (scratch / swap buffers that are about to
		/// be overwritten by a decompress batch anyway).
		template <typename T>
		using default_init_vector = std::vector<T, detail::default_init_allocator<T>>;

		/// Clamp a compression level into blosc2's 0..9.
		inline std::uint8_t ensure_compression_level(std::size_t level)
		{
			return static_cast<std::uint8_t>(level > 9 ? 9 : level);
		}

		/// Largest multiple of one scanline (width * sizeof(T)) that fits into chunk_size; throws when not even one fits.
		template <typename T>
		std::size_t align_chunk_to_scanlines_bytes(std::size_t width, std::size_t chunk_size)
		{
			const std::size_t scanline = width * sizeof(T);
			const std::size_t lines = scanline ? chunk_size / scanline : 0;
			if (lines == 0)
			{
				throw std::runtime_error(
					"Unable to align chunk size to scanlines as the size of a scanline exceeds the chunk size. Got a scanline "
					"size of " + std::to_string(scanline) + " while the max size of the chunks is " + std::to_string(chunk_size));
			}
			return lines * scanline;
		}
	}
}
