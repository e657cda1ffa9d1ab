// tests/cpp/test_compressed.cpp -- the reference's OpenImageIO-free C++ test cases, re-hosted on include/compressed/*.h
// (which runs every chunk through libb2g.so on the GPU).  Mirrors:
//   test/src/test_schunk.cpp:17-74        empty schunk; 4096 iota elements, chunk 256 B, block 64
//   test/src/test_channel.cpp:20-125      size mismatch throws; 50- and 8192-element round trips; attribute getters;
//                                         iterate-read and iterate-modify-reread
//   test/src/test_chunk_span.cpp:16-57    x() / y() from the chunk index
//   test/src/test_image.cpp:552-858       construct from vectors, chunk == / < / > data size, modify then re-read
//   python/test/test_channel.py:173-237   get_chunk / set_chunk round trip and the exception types
//   test/src/util.h:40-56                 image_algo::deinterleave as the reference's test helper uses it (+ interleave back)
// Needs a CUDA device (there is no CPU path); run by tests/test_gpu_cpp.py under -m gpu.
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <numeric>
#include <string>
#include <vector>

#include "compressed/image.h"
#include "compressed/image_algo.h"
// the reference's include paths for the same things (forwarding headers)
#include "compressed/blosc2/lazyschunk.h"
#include "compressed/blosc2/schunk_mixin.h"
#include "compressed/blosc2/typedefs.h"
#include "compressed/blosc2/util.h"
#include "compressed/containers/chunk_span.h"
#include "compressed/fwd.h"
#include "compressed/iterators/iterator.h"
#include "compressed/json_alias.h"
#include "compressed/ranges.h"

static int g_failed = 0, g_checks = 0;
#define CHECK(cond)                                                                      \
	do {                                                                                 \
		++g_checks;                                                                      \
		if (!(cond)) { ++g_failed; std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); } \
	} while (0)
#define CHECK_THROWS_AS(expr, type)                                                      \
	do {                                                                                 \
		++g_checks;                                                                      \
		bool ok_ = false;                                                                \
		try { expr; } catch (const type&) { ok_ = true; } catch (...) {}                 \
		if (!ok_) { ++g_failed; std::printf("FAILED %s:%d: %s should throw %s\n", __FILE__, __LINE__, #expr, #type); } \
	} while (0)

namespace ci = compressed;

template <typename T>
static void test_schunk()
{
	{
		ci::blosc2::schunk<T> super_chunk(128, 4096);
		auto ctx = ci::blosc2::create_decompression_context(4);
		CHECK(super_chunk.to_uncompressed(ctx).size() == 0);
		CHECK(super_chunk.num_chunks() == 0);
	}
	std::vector<T> data(4096);
	std::iota(data.begin(), data.end(), T{ 0 });
	auto ctx = ci::blosc2::create_compression_context<T>(4, ci::enums::codec::lz4, 9, 128);
	ci::blosc2::schunk<T> super_chunk(std::span<const T>(data), 64, 256, ctx);
	auto dctx = ci::blosc2::create_decompression_context(4);
	auto decompressed = super_chunk.to_uncompressed(dctx);
	CHECK(decompressed.size() == 4096);
	CHECK(decompressed == data);
	CHECK(super_chunk.num_chunks() == 4096 * sizeof(T) / 256);
	CHECK(super_chunk.byte_size() / sizeof(T) == 4096);
	CHECK(super_chunk.chunk(dctx, 0).size() == 256 / sizeof(T));
	CHECK_THROWS_AS(super_chunk.chunk(dctx, super_chunk.num_chunks()), std::out_of_range);
}

template <typename T>
static void test_channel_roundtrip(ci::enums::codec codec)
{
	std::vector<T> small(50, static_cast<T>(25));
	ci::channel<T> ch_small(std::span<const T>(small), 10, 5, codec);
	CHECK(ch_small.get_decompressed() == small);

	std::vector<T> data(8192);
	for (size_t i = 0; i < data.size(); ++i) data[i] = static_cast<T>((i * 7) % 251);
	ci::channel<T> ch(std::span<const T>(data), 128, 64, codec, 5, 128, 4096);
	CHECK(ch.get_decompressed() == data);
	CHECK(ch.width() == 128);
	CHECK(ch.height() == 64);
	CHECK(ch.compression() == codec);
	CHECK(ch.compression_level() == 5);
	CHECK(ch.uncompressed_size() == 8192);
	CHECK(ch.chunk_size() == 4096 / (128 * sizeof(T)) * (128 * sizeof(T)));
	CHECK(ch.num_chunks() == (8192 * sizeof(T) + ch.chunk_size() - 1) / ch.chunk_size());
	CHECK(ch.compressed_bytes() > 0);

	// get_chunk / set_chunk over all chunks, buffers of the wrong size and bad indices
	for (size_t i = 0; i < ch.num_chunks(); ++i)
	{
		std::vector<T> buf(ch.chunk_elems(i));
		ch.get_chunk(std::span<T>(buf), i);
		for (auto& v : buf) v = static_cast<T>(v + 1);
		ch.set_chunk(std::span<T>(buf), i);
	}
	auto after = ch.get_decompressed();
	bool all_plus_one = true;
	for (size_t i = 0; i < data.size(); ++i) all_plus_one = all_plus_one && after[i] == static_cast<T>(data[i] + 1);
	CHECK(all_plus_one);
	std::vector<T> wrong(ch.chunk_elems(0) + 1);
	CHECK_THROWS_AS(ch.get_chunk(std::span<T>(wrong), 0), std::invalid_argument);
	CHECK_THROWS_AS(ch.set_chunk(std::span<T>(wrong), 0), std::invalid_argument);
	std::vector<T> right(ch.chunk_elems(0));
	CHECK_THROWS_AS(ch.get_chunk(std::span<T>(right), ch.num_chunks()), std::out_of_range);
	CHECK_THROWS_AS(ch.set_chunk(std::span<T>(right), ch.num_chunks()), std::out_of_range);

	// batched chunk access: all chunks in one launch each way
	std::vector<std::vector<T>> bufs;
	std::vector<std::span<T>> spans;
	std::vector<std::span<const T>> cspans;
	std::vector<size_t> idx;
	for (size_t i = 0; i < ch.num_chunks(); ++i) { bufs.emplace_back(ch.chunk_elems(i)); idx.push_back(i); }
	for (auto& b : bufs) { spans.emplace_back(b); }
	ch.get_chunks(idx, spans);
	for (auto& b : bufs) { for (auto& v : b) v = static_cast<T>(v - 1); cspans.emplace_back(b.data(), b.size()); }
	ch.set_chunks(idx, cspans);
	CHECK(ch.get_decompressed() == data);
}

static void test_channel_misc()
{
	std::vector<uint16_t> data(100);
	CHECK_THROWS_AS(ci::channel<uint16_t>(std::span<const uint16_t>(data), 20, 20), std::runtime_error);
	CHECK_THROWS_AS(ci::channel<uint16_t>(std::span<const uint16_t>(data), 10, 10, ci::enums::codec::lz4, 9, 64, 8),
		std::runtime_error);  // a scanline does not fit the chunk size

	// iterate and read, then iterate, modify and re-read (test_channel.cpp:92-125)
	std::vector<uint16_t> vec(128);
	std::iota(vec.begin(), vec.end(), uint16_t{ 0 });
	ci::channel<uint16_t> ch(std::span<const uint16_t>(vec), 16, 8, ci::enums::codec::lz4, 9, 64, 64);
	CHECK(ch.num_chunks() == 4);
	size_t seen = 0;
	for (auto chunk : ch)
	{
		for (size_t i = 0; i < chunk.size(); ++i)
		{
			CHECK(chunk[i] == vec[seen]);
			CHECK(chunk.x(i) == seen % 16);
			CHECK(chunk.y(i) == seen / 16);
			++seen;
		}
	}
	CHECK(seen == 128);
	for (auto chunk : ch) for (auto& v : chunk) v = static_cast<uint16_t>(v * 2);
	auto doubled = ch.get_decompressed();
	bool ok = true;
	for (size_t i = 0; i < vec.size(); ++i) ok = ok && doubled[i] == vec[i] * 2;
	CHECK(ok);
	auto it = ch.end();
	CHECK_THROWS_AS(++it, std::out_of_range);

	// lazy channels: nothing compressed until written (channel.h:219-304, python/test/test_channel.py:92-170)
	auto z = ci::channel<float>::zeros(64, 64, ci::enums::codec::blosclz, 2, 256, 1024);
	CHECK(z.compressed_bytes() == z.num_chunks() * sizeof(float));
	CHECK(z.get_decompressed() == std::vector<float>(64 * 64, 0.0f));
	auto f = ci::channel<float>::full_like(z, 7.5f);
	CHECK(f.get_decompressed() == std::vector<float>(64 * 64, 7.5f));
	std::vector<float> buf(f.chunk_elems(1), 1.0f);
	f.set_chunk(std::span<float>(buf), 1);
	auto mixed = f.get_decompressed();
	CHECK(mixed[0] == 7.5f);
	CHECK(mixed[f.chunk_elems(0)] == 1.0f);
	CHECK(f.compressed_bytes() > f.num_chunks() * sizeof(float));
}

template <typename T>
static void test_image(size_t chunk_size)
{
	const size_t w = 32, h = 16;
	std::vector<std::vector<T>> chans(3, std::vector<T>(w * h));
	for (size_t c = 0; c < 3; ++c)
		for (size_t i = 0; i < w * h; ++i) chans[c][i] = static_cast<T>(i * (c + 1));
	ci::image<T> img(chans, w, h, { "R", "G", "B" }, ci::enums::codec::lz4, 9, 256, chunk_size);
	CHECK(img.num_channels() == 3);
	CHECK(img.width() == w);
	CHECK(img.height() == h);
	CHECK(img.get_decompressed() == chans);
	CHECK(img.channel("G").get_decompressed() == chans[1]);
	CHECK(img.channel(2).get_decompressed() == chans[2]);
	CHECK(img.get_channel_offset("B") == 2);
	CHECK_THROWS_AS(img.channel(3), std::out_of_range);
	CHECK_THROWS_AS(img.channel("A"), std::invalid_argument);
	CHECK(img.compression_ratio() > 0.0);
	// modify every channel through chunk iteration, re-read (test_image.cpp:780-858 does this through zip)
	for (auto& ch : img.channels())
		for (auto chunk : ch) for (auto& v : chunk) v = static_cast<T>(v + 3);
	auto out = img.get_decompressed();
	bool ok = true;
	for (size_t c = 0; c < 3; ++c)
		for (size_t i = 0; i < w * h; ++i) ok = ok && out[c][i] == static_cast<T>(chans[c][i] + 3);
	CHECK(ok);
	// add / extract / remove
	std::vector<T> alpha(w * h, static_cast<T>(1));
	img.add_channel(std::span<const T>(alpha), w, h, "A");
	CHECK(img.num_channels() == 4);
	CHECK(img.channel("A").get_decompressed() == alpha);
	CHECK_THROWS_AS(img.add_channel(std::span<const T>(alpha), w / 2, h * 2, "X"), std::invalid_argument);
	auto r = img.extract_channel("R");
	CHECK(r.width() == w);
	img.remove_channel(0);
	CHECK(img.num_channels() == 2);
	CHECK(img.channelnames() == (std::vector<std::string>{ "B", "A" }));
}

// The double-buffered iterator over 23 chunks (the last one short): contents, every third chunk modified, calls on the
// channel's own contexts from inside the loop, an abandoned iteration flushing its last chunk.
static void test_iterator_pipeline()
{
	const size_t w = 256, h = 23 * 16 - 5, per = 16 * w;
	std::vector<uint16_t> flat(w * h);
	for (size_t i = 0; i < flat.size(); ++i) flat[i] = static_cast<uint16_t>((i * 13 + (i >> 9)) & 0x3ff);
	ci::channel<uint16_t> ch(std::span<const uint16_t>(flat), w, h, ci::enums::codec::lz4, 9, 32768, per * sizeof(uint16_t));
	CHECK(ch.num_chunks() == 23);
	size_t i = 0;
	bool ok = true;
	for (auto chunk : ch)
	{
		for (size_t k = 0; k < chunk.size() && ok; ++k) ok = chunk[k] == flat[i * per + k];
		if (i % 3 == 0)
			for (size_t k = 0; k < chunk.size(); ++k) { chunk[k] = static_cast<uint16_t>(chunk[k] + 7); flat[i * per + k] = static_cast<uint16_t>(flat[i * per + k] + 7); }
		if (i == 5)
		{
			std::vector<uint16_t> tmp(per);
			ch.get_chunk(std::span<uint16_t>(tmp), 1);
			for (size_t k = 0; k < per && ok; ++k) ok = tmp[k] == flat[per + k];
			std::fill(tmp.begin(), tmp.end(), uint16_t{ 9 });
			ch.set_chunk(std::span<uint16_t>(tmp), 20);
			std::fill(flat.begin() + 20 * per, flat.begin() + 21 * per, uint16_t{ 9 });
		}
		++i;
	}
	CHECK(ok);
	CHECK(i == 23);
	CHECK(ch.get_decompressed() == flat);
	i = 0;
	for (auto chunk : ch)
	{
		for (size_t k = 0; k < chunk.size(); ++k) { chunk[k] = 1; flat[i * per + k] = 1; }
		if (++i == 3) break;
	}
	CHECK(ch.get_decompressed() == flat);
}

template <typename T>
static void test_image_algo(size_t nch, size_t n)
{
	std::vector<T> pixels(nch * n);
	for (size_t i = 0; i < pixels.size(); ++i) pixels[i] = static_cast<T>((i * 2654435761u) >> 7);
	std::vector<std::vector<T>> channels(nch, std::vector<T>(n));
	ci::image_algo::deinterleave(std::span<const T>(pixels), channels);
	bool ok = true;
	for (size_t c = 0; c < nch && ok; ++c)
		for (size_t i = 0; i < n; ++i)
			if (channels[c][i] != pixels[i * nch + c]) { ok = false; break; }
	CHECK(ok);
	std::vector<T> back(nch * n, T{});
	std::vector<std::span<const T>> spans;
	for (auto& c : channels) spans.push_back(std::span<const T>(c));
	ci::image_algo::interleave(std::span<T>(back), spans);
	CHECK(back == pixels);
	std::vector<T> small(nch * n - 1);
	CHECK_THROWS_AS(ci::image_algo::interleave(std::span<T>(small), spans), std::invalid_argument);
	std::vector<std::span<const T>> none;
	CHECK_THROWS_AS(ci::image_algo::interleave(std::span<T>(back), none), std::invalid_argument);
	if (nch > 1)
	{
		channels[1].resize(n + 1);
		CHECK_THROWS_AS(ci::image_algo::deinterleave(std::span<const T>(pixels), channels), std::invalid_argument);
	}
}

// image<T>: one launch for the whole image (round 2), the statistics helpers of image.h:1122-1170, metadata, zip.
static void test_image_batched()
{
	const size_t w = 256, h = 128, nch = 5;
	std::vector<std::vector<uint16_t>> data(nch, std::vector<uint16_t>(w * h));
	for (size_t c = 0; c < nch; ++c)
		for (size_t i = 0; i < w * h; ++i) data[c][i] = static_cast<uint16_t>((i / 7 + c * 31) & 0x3ff);
	ci::image<uint16_t> img(data, w, h, { "R", "G", "B", "A", "Z" }, ci::enums::codec::lz4, 9, 32768, 16 * w * sizeof(uint16_t));
	CHECK(img.num_channels() == nch);
	CHECK(img.channel(0).num_chunks() == h / 16);
	// all 5 x 8 chunks went through the FIRST channel's context, as one batch: 3 kernels (compress, place, finalize) in
	// the engine, 1 batch in the CPU mock the host-logic test links against
	const bool mock = std::string(b2g_version()) == "mock";
	CHECK(b2g_launch_count(img.channel(0).compression_context()) == (mock ? 1 : 3));
	CHECK(b2g_launch_count(img.channel(1).compression_context()) == 0);
	const auto back = img.get_decompressed();
	CHECK(back.size() == nch);
	for (size_t c = 0; c < nch; ++c) CHECK(back[c] == data[c]);
	CHECK(b2g_launch_count(img.channel(0).decompression_context()) == (mock ? 1 : 2));   // plan + decode, once for the image
	// compression_ratio(): the reference's arithmetic (both sums seeded with 1, uncompressed counted in elements)
	size_t comp = 1, unc = 1;
	for (const auto& ch : img.channels()) { comp += ch.compressed_bytes(); unc += ch.uncompressed_size(); }
	CHECK(img.compression_ratio() == static_cast<double>(unc) / static_cast<double>(comp));
	img.print_statistics();
#ifdef COMPRESSED_IMAGE_HAVE_JSON
	img.metadata({ { "author", "test" }, { "frame", 12 } });
	CHECK(img.metadata()["author"] == "test");
	CHECK(img.metadata().dump() == "{\"author\":\"test\",\"frame\":12}");      // ordered_json keeps insertion order
#endif
	// a lazy channel next to compressed ones
	img.add_channel(ci::channel<uint16_t>::full(w, h, 7), "fill");
	const auto back2 = img.get_decompressed();
	CHECK(back2.size() == nch + 1 && back2[nch] == std::vector<uint16_t>(w * h, 7) && back2[2] == data[2]);
	// ranges::zip: lock-step chunk iteration over two channels, modifications written back
	auto& a = img.channel("R");
	auto& b = img.channel("G");
	size_t steps = 0;
	for (auto [ca, cb] : ci::ranges::zip(a, b))
	{
		CHECK(ca.size() == cb.size());
		for (size_t i = 0; i < ca.size(); ++i) ca[i] = cb[i];
		++steps;
	}
	CHECK(steps == a.num_chunks());
	CHECK(a.get_decompressed() == data[1]);
	ci::blosc2::lazy_schunk<float> lazy(1.5f, 1000, 4096, 1024);
	CHECK(lazy.num_chunks() == 4 && lazy.is_lazy(3));
}

// The reference's "modify" usage (benchmark/main.cpp:139-157) on a 64 MiB uint16 channel, timed: iterate, change every
// value, let the iterator recompress.  Prints the time; checks the data.  Not run against the CPU mock.
static void bench_iterator_modify()
{
	if (std::string(b2g_version()) == "mock") return;
	const size_t w = 8192, h = 4096;
	std::vector<uint16_t> data(w * h);
	uint32_t x = 12345;
	for (size_t i = 0; i < data.size(); ++i) { x = x * 1664525u + 1013904223u; data[i] = static_cast<uint16_t>(((i % w) >> 2) + ((x >> 28) & 3)); }
	ci::channel<uint16_t> ch(std::span<const uint16_t>(data), w, h);
	double best = 1e30;
	for (int rep = 0; rep < 4; ++rep)
	{
		const auto t0 = std::chrono::steady_clock::now();
		for (auto chunk : ch)
			for (auto& v : chunk) v ^= 1;
		const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
		if (rep > 0 && ms < best) best = ms;
	}
	std::printf("iterator modify loop, 64 MiB uint16 channel, %zu chunks: %.2f ms (best of 3)\n", ch.num_chunks(), best);
	const auto back = ch.get_decompressed();       // four passes of ^= 1 cancel out
	CHECK(back == data);
}

int main()
{
	if (b2g_device_count() < 1)
	{
		std::printf("no CUDA device: compressed::* has no CPU path\n");
		try { auto ctx = compressed::blosc2::create_decompression_context(1); } catch (const std::runtime_error& e)
		{
			std::printf("create_decompression_context failed loudly as expected: %s\n", e.what());
			return 3;
		}
		return 2;
	}
	test_schunk<uint8_t>(); test_schunk<uint16_t>(); test_schunk<uint32_t>(); test_schunk<float>();
	for (auto codec : { compressed::enums::codec::lz4, compressed::enums::codec::blosclz })
	{
		test_channel_roundtrip<uint8_t>(codec); test_channel_roundtrip<int8_t>(codec);
		test_channel_roundtrip<uint16_t>(codec); test_channel_roundtrip<int16_t>(codec);
		test_channel_roundtrip<uint32_t>(codec); test_channel_roundtrip<int32_t>(codec);
		test_channel_roundtrip<float>(codec);
	}
	test_channel_misc();
	for (size_t chunk : { size_t(32 * 16 * 4), size_t(1024), size_t(768), size_t(4194304) })
	{
		test_image<uint8_t>(chunk); test_image<uint16_t>(chunk); test_image<float>(chunk);
	}
	test_iterator_pipeline();
	bench_iterator_modify();
	test_image_batched();
	for (size_t nch : { size_t(1), size_t(3), size_t(4), size_t(7), size_t(16) })
	{
		test_image_algo<uint8_t>(nch, 1000003); test_image_algo<uint16_t>(nch, 64 * 64);
		test_image_algo<float>(nch, 12345); test_image_algo<double>(nch, 777);
	}
	std::printf("%d checks, %d failed\n", g_checks, g_failed);
	return g_failed ? 1 : 0;
}
