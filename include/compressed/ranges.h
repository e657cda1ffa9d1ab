// compressed/ranges.h -- ranges::zip over several channels / images (ranges.h:20-140): lock-step iteration, stops with
// the shortest range.  std::views::zip is C++23; this is the C++20 subset the reference's call sites use
// (for (auto [a, b] : compressed::ranges::zip(channel_a, channel_b))).
#pragma once

#include <cstddef>
#include <iterator>
#include <tuple>
#include <utility>

#include "macros.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	namespace ranges
	{
		template <typename... Ranges>
		struct zip_view
		{
			std::tuple<Ranges&...> m_Ranges;

			struct iterator
			{
				using iterator_category = std::forward_iterator_tag;
				using difference_type = std::ptrdiff_t;
				using value_type = std::tuple<decltype(*std::begin(std::declval<Ranges&>()))...>;
				using reference = value_type;
				using pointer = void;

				std::tuple<decltype(std::begin(std::declval<Ranges&>()))...> m_Iters;

				value_type operator*() { return std::apply([](auto&... it) { return value_type(*it...); }, m_Iters); }
				iterator& operator++() { std::apply([](auto&... it) { (++it, ...); }, m_Iters); return *this; }
				iterator operator++(int) { iterator tmp = *this; ++*this; return tmp; }
				/// equal as soon as ANY component has reached its end: the shortest range ends the zip
				bool operator==(const iterator& other) const { return any_equal(other, std::index_sequence_for<Ranges...>{}); }
				bool operator!=(const iterator& other) const { return !(*this == other); }

			private:
				template <std::size_t... I>
				bool any_equal(const iterator& other, std::index_sequence<I...>) const
				{
					return ((std::get<I>(m_Iters) == std::get<I>(other.m_Iters)) || ...);
				}
			};

			iterator begin() { return iterator{ std::apply([](auto&... r) { return std::make_tuple(std::begin(r)...); }, m_Ranges) }; }
			iterator end() { return iterator{ std::apply([](auto&... r) { return std::make_tuple(std::end(r)...); }, m_Ranges) }; }
		};

		template <typename... Ranges>
		zip_view<Ranges...> zip(Ranges&... ranges) { return zip_view<Ranges...>{ std::tie(ranges...) }; }
	}
}
