// compressed/image.h -- compressed::image<T>: named channels of equal size, each a compressed::channel<T>.
//
// The part of the reference's image<T> (compressed_image/include/compressed/image.h) that sits on the chunk path:
// constructors from raw channel data (:87-258) and from channels (:276-325), channel() (:1192-1210), channels()
// (:1289-1300), get_decompressed() (:1307-1315), add / remove / extract_channel (:940-1098), the size / ratio getters
// (:1160-1175, 1339-1445), print_statistics (:1122-1156), the metadata accessors (:1379-1394, when nlohmann/json.hpp is on
// the include path) and update_nthreads (:1399-1405).  Not provided: read() (OpenImageIO file decode, :360-896), which is
// not on the chunk path.
//
// The chunks of ALL channels of an image are compressed (constructor) / decompressed (get_decompressed) as ONE batch: one
// launch per image, where the reference loops channel by channel, chunk by chunk (:119-159, :1307-1315).
#pragma once

#include <cstddef>
#include <iostream>
#include <optional>
#include <span>
#include <stdexcept>
#include <string>
#include <string_view>
#include <vector>

#include "channel.h"
#include "fwd.h"
#include "json_alias.h"

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
			// every chunk of every channel in one compress batch
			std::vector<blosc2::compress_job<T>> jobs;
			std::vector<typename blosc2::schunk<T>::compress_plan> plans;
			for (std::size_t i = 0; i < channels.size(); ++i)
			{
				try
				{
					m_Channels.emplace_back(typename compressed::channel<T>::deferred_t{}, channels[i].size(), width, height,
						compression_codec, level, block_size, chunk_size);
					plans.push_back(m_Channels.back().plan_compress(channels[i], jobs));
				}
				catch (const std::exception& e)
				{
					throw std::runtime_error("Unable to construct channel '" + m_ChannelNames[i] + "' (index " +
						std::to_string(i) + "): " + e.what());
				}
			}
			if (!jobs.empty())
			{
				blosc2::compress_batch<T>(m_Channels.front().compression_context(), std::span<blosc2::compress_job<T>>(jobs));
			}
			for (std::size_t i = 0; i < m_Channels.size(); ++i)
			{
				m_Channels[i].adopt_compressed(std::move(plans[i]), std::span<const blosc2::compress_job<T>>(jobs));
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

		/// image.h:1307-1315: every chunk of every channel in one decompress batch.
		std::vector<std::vector<T>> get_decompressed() const
		{
			std::vector<std::vector<T>> result;
			result.reserve(m_Channels.size());
			std::vector<blosc2::decompress_job<T>> jobs;
			for (const auto& ch : m_Channels)
			{
				result.emplace_back(ch.uncompressed_size());
				ch.schunk().collect_decompress(std::span<T>(result.back()), jobs);
			}
			if (!jobs.empty())
			{
				blosc2::decompress_batch<T>(m_Channels.front().decompression_context(), std::span<blosc2::decompress_job<T>>(jobs));
			}
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

		/// image.h:1160-1170, reproduced as it is: both sums start at 1 and the uncompressed one counts ELEMENTS, so for
		/// sizeof(T) > 1 the figure is the byte ratio divided by sizeof(T) (the Python tests of the reference pin it).
		double compression_ratio() const noexcept
		{
			std::size_t total_uncompressed = 1, total_compressed = 1;
			for (const auto& ch : m_Channels)
			{
				total_compressed += ch.compressed_bytes();
				total_uncompressed += ch.uncompressed_size();
			}
			return static_cast<double>(total_uncompressed) / static_cast<double>(total_compressed);
		}

		/// image.h:1122-1156: a summary on stdout (sizes as the reference prints them: compressed bytes, uncompressed
		/// elements).
		void print_statistics()
		{
			std::size_t compressed_size = 0, uncompressed_size = 0, num_chunks = 0;
			for (const auto& ch : m_Channels)
			{
				compressed_size += ch.compressed_bytes();
				uncompressed_size += ch.uncompressed_size();
				num_chunks += ch.num_chunks();
			}
			std::cout << "Statistics for image buffer:\n";
			std::cout << " Width:             " << m_Width << "\n";
			std::cout << " Height:            " << m_Height << "\n";
			std::cout << " Channels:          " << m_Channels.size() << "\n";
			std::cout << " Channelnames:      [";
			for (std::size_t i = 0; i < m_ChannelNames.size(); ++i) std::cout << (i ? ", " : "") << m_ChannelNames[i];
			std::cout << "]\n";
			std::cout << " --------------     \n";
			std::cout << " Compressed Size:   " << compressed_size << "\n";
			std::cout << " Uncompressed Size: " << uncompressed_size << "\n";
			std::cout << " Compression ratio: " << static_cast<double>(uncompressed_size) / static_cast<double>(compressed_size) << "x\n";
			std::cout << " Num Chunks:        " << num_chunks << "\n";
#ifdef COMPRESSED_IMAGE_HAVE_JSON
			std::cout << " Metadata:          \n " << m_Metadata.dump(4) << std::endl;
#else
			std::cout << " Metadata:          \n (nlohmann/json.hpp not on the include path)" << std::endl;
#endif
		}

#ifdef COMPRESSED_IMAGE_HAVE_JSON
		/// image.h:1379-1394: arbitrary user metadata, neither authored nor interpreted by the image.
		void metadata(const json_ordered& _metadata) noexcept { m_Metadata = _metadata; }
		json_ordered& metadata() noexcept { return m_Metadata; }
		const json_ordered& metadata() const noexcept { return m_Metadata; }
#endif

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
#ifdef COMPRESSED_IMAGE_HAVE_JSON
		json_ordered m_Metadata;
#endif
		std::size_t m_Width = 1, m_Height = 1;
	};
}
