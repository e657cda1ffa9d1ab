// compressed/image.h -- compressed::image<T>: named channels of equal size, each a compressed::channel<T>.
//
// The part of the reference's image<T> (compressed_image/include/compressed/image.h) that sits on the chunk path:
// constructors from raw channel data (:87-258) and from channels (:276-325), channel() (:1192-1210), channels()
// (:1289-1300), get_decompressed() (:1307-1315), add / remove / extract_channel (:940-1098), the size / ratio getters
// (:1160-1175, 1339-1445) and update_nthreads (:1399-1405).  Not provided: read() (OpenImageIO file decode, :360-896) and
// the json metadata accessors -- neither is on the chunk path.
//
// Every channel's chunks are compressed / decompressed as one batch (one launch per channel).
#pragma once

#include <cstddef>
#include <optional>
#include <span>
#include <stdexcept>
#include <string>
#include <string_view>
#include <vector>

#include "channel.h"

namespace NAMESPACE_COMPRESSED_IMAGE
{
	template <typename T>
	struct image
	{
		image() = default;
		image(image&&) = default;
		image& operator=(image&&) = default;
		image(const image&) = delete;
		image& operator=(const image&) = delete;

		/// image.h:87-159
		image(std::vector<std::span<const T>> channels, std::size_t width, std::size_t height,
			std::vector<std::string> channel_names = {}, enums::codec compression_codec = enums::codec::lz4,
			std::size_t compression_level = 9, std::size_t block_size = s_default_blocksize,
			std::size_t chunk_size = s_default_chunksize)
			: m_Width(width), m_Height(height)
		{
			init_names(channels.size(), std::move(channel_names));
			const auto level = util::ensure_compression_level(compression_level);
			for (std::size_t i = 0; i < channels.size(); ++i)
			{
				try
				{
					m_Channels.emplace_back(channels[i], width, height, compression_codec, level, block_size, chunk_size);
				}
				catch (const std::exception& e)
				{
					throw std::runtime_error("Unable to construct channel '" + m_ChannelNames[i] + "' (index " +
						std::to_string(i) + "): " + e.what());
				}
			}
		}

		/// image.h:189-258
		image(std::vector<std::vector<T>> channels, std::size_t width, std::size_t height,
			std::vector<std::string> channel_names = {}, enums::codec compression_codec = enums::codec::lz4,
			std::size_t compression_level = 9, std::size_t block_size = s_default_blocksize,
			std::size_t chunk_size = s_default_chunksize)
			: image(to_spans(channels), width, height, std::move(channel_names), compression_codec, compression_level,
				block_size, chunk_size) {}

		/// image.h:276-325
		image(std::vector<compressed::channel<T>> channels, std::size_t width, std::size_t height,
			std::vector<std::string> channel_names = {})
			: m_Width(width), m_Height(height)
		{
			init_names(channels.size(), std::move(channel_names));
			for (std::size_t i = 0; i < channels.size(); ++i)
			{
				if (channels[i].width() != width || channels[i].height() != height)
				{
					throw std::invalid_argument("Invalid channel at index " + std::to_string(i) +
						": its dimensions do not match the image's");
				}
			}
			m_Channels = std::move(channels);
		}

		/// image.h:940-975
		void add_channel(compressed::channel<T> _channel, std::optional<std::string> name = std::nullopt)
		{
			if (m_Channels.empty()) { m_Width = _channel.width(); m_Height = _channel.height(); }
			if (_channel.width() != m_Width || _channel.height() != m_Height)
			{
				throw std::invalid_argument("Unable to add channel '" + name.value_or("") +
					"': its dimensions do not match the image's");
			}
			m_ChannelNames.push_back(name.value_or(""));
			m_Channels.push_back(std::move(_channel));
		}

		/// image.h:993-1043
		void add_channel(std::span<const T> data, std::size_t width, std::size_t height,
			std::optional<std::string> name = std::nullopt, enums::codec compression_codec = enums::codec::lz4,
			std::uint8_t compression_level = 9, std::size_t block_size = s_default_blocksize,
			std::size_t chunk_size = s_default_chunksize)
		{
			if (!m_Channels.empty() && (width != m_Width || height != m_Height))
			{
				throw std::invalid_argument("Unable to add channel '" + name.value_or("") +
					"': its dimensions do not match the image's");
			}
			add_channel(compressed::channel<T>(data, width, height, compression_codec, compression_level, block_size,
				chunk_size), std::move(name));
		}

		/// image.h:1048-1098
		compressed::channel<T> extract_channel(std::size_t index)
		{
			if (index >= m_Channels.size())
			{
				throw std::out_of_range("Cannot access channel index " + std::to_string(index) + ", the image has " +
					std::to_string(m_Channels.size()) + " channels");
			}
			auto ret = std::move(m_Channels[index]);
			m_Channels.erase(m_Channels.begin() + static_cast<std::ptrdiff_t>(index));
			m_ChannelNames.erase(m_ChannelNames.begin() + static_cast<std::ptrdiff_t>(index));
			return ret;
		}
		compressed::channel<T> extract_channel(const std::string_view name) { return extract_channel(get_channel_offset(name)); }
		void remove_channel(std::size_t index) { (void)extract_channel(index); }
		void remove_channel(const std::string_view name) { (void)extract_channel(name); }

		/// image.h:1192-1210
		compressed::channel<T>& channel(std::size_t index)
		{
			if (index >= m_Channels.size())
			{
				throw std::out_of_range("Cannot access channel index " + std::to_string(index) + ", the image has " +
					std::to_string(m_Channels.size()) + " channels");
			}
			return m_Channels[index];
		}
		compressed::channel<T>& channel(const std::string_view name) { return m_Channels[get_channel_offset(name)]; }

		std::vector<compressed::channel<T>>& channels() { return m_Channels; }
		const std::vector<compressed::channel<T>>& channels() const { return m_Channels; }

		/// image.h:1307-1315
		std::vector<std::vector<T>> get_decompressed() const
		{
			std::vector<std::vector<T>> result;
			result.reserve(m_Channels.size());
			for (const auto& ch : m_Channels) result.push_back(ch.get_decompressed());
			return result;
		}

		/// image.h:1326-1337
		std::size_t get_channel_offset(const std::string_view channelname) const
		{
			for (std::size_t i = 0; i < m_ChannelNames.size(); ++i)
			{
				if (m_ChannelNames[i] == channelname) return i;
			}
			throw std::invalid_argument("Unknown channelname '" + std::string(channelname) + "' encountered");
		}

		/// image.h:1160-1175
		double compression_ratio() const noexcept
		{
			std::size_t compressed = 0, uncompressed = 0;
			for (const auto& ch : m_Channels)
			{
				compressed += ch.compressed_bytes();
				uncompressed += ch.uncompressed_size() * sizeof(T);
			}
			return compressed ? static_cast<double>(uncompressed) / static_cast<double>(compressed) : 0.0;
		}

		std::size_t width() const noexcept { return m_Width; }
		std::size_t height() const noexcept { return m_Height; }
		std::size_t num_channels() const noexcept { return m_Channels.size(); }
		std::vector<std::string> channelnames() const noexcept { return m_ChannelNames; }
		void channelnames(std::vector<std::string> _channelnames)
		{
			if (_channelnames.size() != m_Channels.size())
			{
				throw std::invalid_argument("Invalid number of channelnames: expected " + std::to_string(m_Channels.size()) +
					" but got " + std::to_string(_channelnames.size()));
			}
			m_ChannelNames = std::move(_channelnames);
		}
		void update_nthreads(std::size_t nthreads)
		{
			for (auto& ch : m_Channels) ch.update_nthreads(nthreads, ch.block_size());
		}
		/// image.h:1413-1445: the chunk / block size shared by the channels (of the first one when they differ)
		std::size_t chunk_size() const { return m_Channels.empty() ? s_default_chunksize : m_Channels.front().chunk_size(); }
		std::size_t block_size() const { return m_Channels.empty() ? s_default_blocksize : m_Channels.front().block_size(); }

	private:
		static std::vector<std::span<const T>> to_spans(const std::vector<std::vector<T>>& channels)
		{
			std::vector<std::span<const T>> spans;
			spans.reserve(channels.size());
			for (const auto& c : channels) spans.emplace_back(c.data(), c.size());
			return spans;
		}
		void init_names(std::size_t count, std::vector<std::string> names)
		{
			if (!names.empty() && names.size() != count)
			{
				throw std::invalid_argument("Invalid number of channel names: expected " + std::to_string(count) +
					" but got " + std::to_string(names.size()));
			}
			if (names.empty())
			{
				for (std::size_t i = 0; i < count; ++i) names.push_back(std::to_string(i));
			}
			m_ChannelNames = std::move(names);
		}

		std::vector<compressed::channel<T>> m_Channels;
		std::vector<std::string> m_ChannelNames;
		std::size_t m_Width = 1, m_Height = 1;
	};
}
