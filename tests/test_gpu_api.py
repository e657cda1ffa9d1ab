"""The reference's own OIIO-free API tests, re-hosted on compressed_image_b200 (-m gpu):
python/test/test_channel.py:21-237, python/test/test_image.py:167-223, test/src/test_schunk.cpp:39-74,
test/src/test_channel.cpp:20-125, test/src/test_image.py:552-858 (construct / iterate / modify semantics)."""
import numpy as np
import pytest

import _libs

pytestmark = pytest.mark.gpu

DTYPES = [np.uint8, np.int8, np.uint16, np.int16, np.uint32, np.int32, np.float16, np.float32]
SHAPES = [(64, 64), (123, 456), (2048, 16), (1920, 1080)]


def _c():
    import compressed_image_b200 as compressed
    return compressed


def _rand(dtype, shape, seed=0):
    r = np.random.default_rng(seed)
    a = (r.integers(0, 40, size=shape) + np.arange(shape[1])[None, :] // 7).astype(dtype)
    return a


def test_invalid_dtype():
    c = _c()
    with pytest.raises(ValueError):
        c.Channel.zeros(np.bool_, 1, 1)
    with pytest.raises(ValueError):
        c.Channel.full(np.bool_, 100, 1, 1)
    with pytest.raises(ValueError):
        c.Channel(np.array((1, 1), np.bool_), 1, 1)


def test_channel_size_mismatch_throws():
    """test/src/test_channel.cpp:20-41."""
    c = _c()
    with pytest.raises(RuntimeError):
        c.Channel(np.zeros(50, np.uint16), 10, 10)
    with pytest.raises(RuntimeError):    # util.h:121-131: a scanline larger than the chunk size
        c.Channel(np.zeros(4096, np.uint16), 4096, 1, chunk_size=4096)


@pytest.mark.parametrize("width,height", SHAPES)
@pytest.mark.parametrize("dtype", DTYPES)
class TestChannelParametrized:
    def test_roundtrip(self, width, height, dtype):
        c = _c()
        arr = _rand(dtype, (height, width))
        ch = c.Channel(arr, width, height)
        assert ch.shape == (height, width) and ch.dtype == np.dtype(dtype)
        assert np.array_equal(ch.get_decompressed(), arr)
        assert ch.uncompressed_size() == width * height

    def test_initialization_with_small_chunksize(self, width, height, dtype):
        c = _c()
        arr = _rand(dtype, (height, width), seed=1)
        isz = np.dtype(dtype).itemsize
        ch = c.Channel(arr, width, height, chunk_size=width * isz, compression_codec=c.Codec.blosclz,
                       compression_level=2)
        assert ch.num_chunks() == height
        assert ch.chunk_size() == width * isz
        assert ch.height == height and ch.width == width and ch.shape == (height, width)
        assert ch.uncompressed_size() == width * height
        assert ch.compression() == c.Codec.blosclz
        assert ch.compression_level() == 2
        assert np.array_equal(ch.get_decompressed(), arr)

    def test_full_and_zeros(self, width, height, dtype):
        c = _c()
        ch = c.Channel.full(dtype, 25, width, height, compression_codec=c.Codec.blosclz, compression_level=2)
        assert ch.shape == (height, width) and ch.uncompressed_size() == width * height
        assert ch.compression() == c.Codec.blosclz and ch.compression_level() == 2
        assert np.array_equal(ch.get_decompressed(), np.full((height, width), 25, dtype))
        z = c.Channel.zeros(dtype, width, height)
        assert np.array_equal(z.get_decompressed(), np.zeros((height, width), dtype))

    def test_full_like_zeros_like(self, width, height, dtype):
        c = _c()
        base = c.Channel(np.zeros((height, width), dtype), width, height)
        ch = c.Channel.full_like(base, 25)
        assert ch.shape == base.shape and ch.uncompressed_size() == base.uncompressed_size()
        assert ch.compression() == base.compression() and ch.compression_level() == base.compression_level()
        assert np.array_equal(ch.get_decompressed(), np.full((height, width), 25, dtype))
        assert np.array_equal(c.Channel.zeros_like(base).get_decompressed(), np.zeros((height, width), dtype))

    def test_modify_chunk(self, width, height, dtype):
        c = _c()
        isz = np.dtype(dtype).itemsize
        ch = c.Channel(np.zeros((height, width), dtype), width, height, chunk_size=width * isz)
        chunk0 = ch.get_chunk(0)
        assert ch.chunk_size() == width * isz
        assert chunk0.dtype == dtype and chunk0.shape == (width,)
        assert np.array_equal(chunk0, np.zeros_like(chunk0))
        ch.set_chunk(0, np.full_like(chunk0, 100))
        assert np.array_equal(ch.get_chunk(0), np.full_like(chunk0, 100))

    def test_set_invalid_chunk_size(self, width, height, dtype):
        c = _c()
        isz = np.dtype(dtype).itemsize
        ch = c.Channel(np.zeros((height, width), dtype), width, height, chunk_size=width * isz)
        with pytest.raises(ValueError):
            ch.set_chunk(0, np.zeros((width, 1), dtype))
        with pytest.raises(ValueError):
            ch.set_chunk(0, np.zeros((width + 20), dtype))
        with pytest.raises(IndexError):
            ch.set_chunk(height + 100, np.zeros((width), dtype))
        with pytest.raises(IndexError):
            ch.get_chunk(height + 100)

    def test_iter_over_channel(self, width, height, dtype):
        c = _c()
        ch = c.Channel(np.zeros((height, width), dtype), width, height, chunk_size=max(width * 16, 4096))
        buf = np.ndarray((ch.chunk_elems(),), dtype=ch.dtype)
        for i in range(ch.num_chunks() - 1):
            ch.get_chunk(i, buf)
            buf[:] = i % 100
            ch.set_chunk(i, buf)
        last = ch.get_chunk(ch.num_chunks() - 1)
        last[:] = (ch.num_chunks() - 1) % 100
        ch.set_chunk(ch.num_chunks() - 1, last)
        for i in range(ch.num_chunks()):
            assert np.all(ch.get_chunk(i) == i % 100)


def test_channel_4096_default_chunks():
    """The largest shape of the reference's parametrisation (4096 x 4096), default 4 MiB chunks."""
    c = _c()
    for dtype in (np.uint8, np.float16, np.float32):
        arr = (_rand(np.uint8, (4096, 4096), seed=3) >> 3).astype(dtype)
        ch = c.Channel(arr, 4096, 4096)
        assert ch.num_chunks() == 4 * np.dtype(dtype).itemsize
        assert np.array_equal(ch.get_decompressed(), arr)
        assert ch.compressed_bytes() < arr.nbytes


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.uint32, np.float32])
def test_schunk_iota_roundtrip(dtype):
    """test/src/test_schunk.cpp:39-74: 4096 iota elements, chunk 256 B, block 64."""
    c = _c()
    isz = np.dtype(dtype).itemsize
    data = np.arange(4096).astype(dtype)
    w = 64 // isz
    ch = c.Channel(data, w, 4096 // w, block_size=64, chunk_size=256)
    assert ch.num_chunks() == 4096 * isz // 256
    assert ch.uncompressed_size() == 4096
    assert ch.chunk_elems(0) == 256 // isz
    assert np.array_equal(ch.get_decompressed().reshape(-1), data)
    for chunk in ch.chunks():                       # every stored chunk is a blosc2 chunk the reference decodes
        rc, _ = (_libs.ref_decompress if _libs.ref() else _libs.oracle_decompress)(chunk)
        assert rc == 256


def test_channel_cpp_roundtrips():
    """test/src/test_channel.cpp:46-69 (50 and 8192 elements, chunk 4096 / block 128), :74-87 (blosclz getters)."""
    c = _c()
    d = np.arange(50, dtype=np.uint16)
    assert np.array_equal(c.Channel(d, 10, 5).get_decompressed().reshape(-1), d)
    d = (np.arange(8192) % 251).astype(np.uint16)
    ch = c.Channel(d, 128, 64, block_size=128, chunk_size=4096)
    assert np.array_equal(ch.get_decompressed().reshape(-1), d)
    ch = c.Channel(d, 128, 64, compression_codec=c.Codec.blosclz, compression_level=5, block_size=128, chunk_size=4096)
    assert ch.compression() == c.Codec.blosclz and ch.compression_level() == 5
    assert ch.width == 128 and ch.height == 64 and ch.num_chunks() == 4 and ch.block_size() == 128


def test_channel_iterate_modify_reread():
    """test/src/test_channel.cpp:92-125: iterate, modify every chunk, re-read."""
    c = _c()
    d = np.arange(128, dtype=np.uint16)
    ch = c.Channel(d, 16, 8, block_size=32, chunk_size=64)
    seen = []
    for chunk in ch:
        seen.append(chunk.copy())
        chunk[:] = 7
    assert np.array_equal(np.concatenate(seen), d)
    assert np.all(ch.get_decompressed() == 7)


@pytest.mark.parametrize("dtype", DTYPES)
class TestImage:
    @pytest.mark.parametrize("width,height", SHAPES)
    def test_add_channel(self, dtype, width, height):
        c = _c()
        image = c.Image(dtype, [], width, height)
        for _ in range(4):
            image.add_channel(np.full((height, width), 90, dtype), width, height)
        assert image.num_channels == 4
        for i in range(4):
            assert np.all(image[i].get_decompressed() == 90)

    def test_add_channel_named(self, dtype):
        c = _c()
        image = c.Image(dtype, [], 64, 64)
        for n in "RGBA":
            image.add_channel(np.full((64, 64), 90, dtype), 64, 64, n)
        assert image.num_channels == 4 and image.get_channel_names() == ["R", "G", "B", "A"]
        assert image.get_channel_index("B") == 2
        assert np.all(image["A"].get_decompressed() == 90)

    def test_add_channel_invalid(self, dtype):
        c = _c()
        image = c.Image(dtype, [], 64, 64)
        with pytest.raises(ValueError):
            image.add_channel(np.full((64, 64), 90, np.bool_), 64, 64)
        with pytest.raises(ValueError):
            image.add_channel(np.full((64, 64), 90, dtype), 32, 64)
        with pytest.raises(ValueError):
            image.add_channel(np.full((64, 32), 90, dtype), 64, 64)

    def test_construct_and_decompress(self, dtype):
        """test/src/test_image.cpp:552-700: construct from vectors, chunk == / < / > data size."""
        c = _c()
        for chunk_size in (1024, 768, 1 << 20):
            chans = [_rand(dtype, (24, 32), seed=s) for s in range(3)]
            img = c.Image(dtype, chans, 32, 24, channel_names=["R", "G", "B"], block_size=256, chunk_size=chunk_size)
            assert img.num_channels == 3 and img.shape == (3, 24, 32)
            for a, b in zip(img.get_decompressed(), chans):
                assert np.array_equal(a, b)
            assert np.array_equal(img.channel("G").get_decompressed(), chans[1])
            assert img.compression_ratio() > 0
            img.remove_channel("R")
            assert img.get_channel_names() == ["G", "B"] and img.num_channels == 2


def test_image_zip_modify_reread():
    """test/src/test_image.cpp:760-858: modify all channels chunk by chunk in lock-step (ranges::zip), re-read."""
    c = _c()
    chans = [_rand(np.float32, (64, 64), seed=s) for s in range(3)]
    img = c.Image(np.float32, chans, 64, 64, block_size=256, chunk_size=1024)
    for r, g, b in zip(img[0], img[1], img[2]):
        r[:] = 1.0
        g[:] = r + 1.0
        b[:] = g + 1.0
    out = img.get_decompressed()
    assert np.all(out[0] == 1.0) and np.all(out[1] == 2.0) and np.all(out[2] == 3.0)


def test_iterator_double_buffered_many_chunks():
    """The double-buffered iterator over 23 chunks (last one short): every third chunk modified, get_chunk() and
    set_chunk() of OTHER chunks from inside the loop, chunk views rotate over three buffers; then an abandoned
    iteration flushes its last modification, and lazy channels iterate as fills."""
    import compressed_image_b200 as c
    rng = np.random.default_rng(3)
    w, h = 256, 23 * 16 - 5
    data = (np.cumsum(rng.integers(0, 3, size=(h, w)), axis=1) & 0x3ff).astype(np.uint16)
    ch = c.Channel(data, w, h, chunk_size=16 * w * 2)
    assert ch.num_chunks() == 23
    flat = data.reshape(-1).copy()
    per = 16 * w
    seen = 0
    for i, chunk in enumerate(ch):
        assert np.array_equal(chunk, flat[i * per:i * per + chunk.size]), i
        if i % 3 == 0:
            chunk += 7
            flat[i * per:i * per + chunk.size] += 7
        if i == 5:                                            # other chunks through the channel's own context
            assert np.array_equal(ch.get_chunk(1), flat[per:2 * per])
            ch.set_chunk(20, np.full(per, 9, np.uint16))
            flat[20 * per:21 * per] = 9
        seen += 1
    assert seen == 23
    assert np.array_equal(np.asarray(ch.get_decompressed()).reshape(-1), flat)
    for i, chunk in enumerate(ch):                            # abandoned after a modification
        chunk[:] = 1
        flat[i * per:i * per + chunk.size] = 1
        if i == 2:
            break
    del chunk
    import gc
    gc.collect()
    assert np.array_equal(np.asarray(ch.get_decompressed()).reshape(-1), flat)
    lazy = c.Channel.full(np.uint16, 5, w, h, chunk_size=16 * w * 2)
    total = 0
    for i, chunk in enumerate(lazy):
        assert (chunk == 5).all()
        if i == 4:
            chunk[0] = 6
        total += chunk.size
    assert total == w * h
    out = np.asarray(lazy.get_decompressed()).reshape(-1)
    assert out[4 * per] == 6 and (np.delete(out, 4 * per) == 5).all()
    a, b = c.Channel(data, w, h, chunk_size=16 * w * 2), c.Channel(data + 1, w, h, chunk_size=16 * w * 2)
    for x, y in zip(a, b):                                    # two iterators of identical parameters at once
        assert np.array_equal(x + 1, y)
        x[:] = y
    assert np.array_equal(np.asarray(a.get_decompressed()), np.asarray(b.get_decompressed()))


# ---- round 2: codec coverage, threads, the headline config through the API ------------------------------------------
def test_codec_lz4hc_roundtrip_and_zstd_refused_early():
    """Channel(..., Codec.lz4hc) compresses (LZ4 streams under the lz4hc codec id, decodable by the reference);
    Channel(..., Codec.zstd) is refused where the channel is built, with a message that says why (no zstd encoder)."""
    c = _c()
    data = _rand(np.uint16, (300, 512), seed=3)
    ch = c.Channel(data, 512, 300, c.Codec.lz4hc, 5, chunk_size=64 * 512 * 2)
    assert ch.compression() == c.Codec.lz4hc
    assert np.array_equal(np.asarray(ch.get_decompressed()), data)
    if _libs.ref() is not None:
        for i, chunk in enumerate(ch.chunks()):
            rc, back = _libs.ref_decompress(chunk)
            assert rc == ch.chunk_size(i)
    with pytest.raises(RuntimeError, match="zstd encoder"):
        c.Channel(data, 512, 300, c.Codec.zstd)
    lazy = c.Channel.zeros(np.uint16, 512, 300, c.Codec.zstd)      # lazy zstd channels exist, writing into them is refused
    with pytest.raises(RuntimeError, match="zstd encoder"):
        lazy.set_chunk(0, np.zeros(lazy.chunk_elems(0), np.uint16))
    with pytest.raises(RuntimeError, match="encoder window"):
        c.Channel(np.zeros((1024, 1024), np.uint8), 1024, 1024, block_size=512 * 1024)


def test_channels_on_two_python_threads():
    """ctypes releases the GIL during calls: two threads working on their own channels must not meet in one b2g_ctx
    (the context pool is per thread)."""
    import threading
    c = _c()
    errs = []

    def work(seed):
        try:
            for k in range(6):
                data = _rand(np.uint16, (600, 1024), seed=seed * 100 + k)
                ch = c.Channel(data, 1024, 600, chunk_size=64 * 1024 * 2)
                assert np.array_equal(np.asarray(ch.get_decompressed()), data)
                ch.set_chunk(1, np.full(ch.chunk_elems(1), seed, np.uint16))
                assert (ch.get_chunk(1) == seed).all()
        except Exception as e:   # noqa: BLE001
            errs.append(repr(e))
    ts = [threading.Thread(target=work, args=(s,)) for s in range(4)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs


def test_cfg2_random_access_through_the_api():
    """BASELINE.json configs[1] through Channel itself: one 8192x4096 half channel (16 chunks of 4 MiB) and its alpha
    channel, get_chunk over all chunks in a seeded random order, set_chunk of modified chunks, get_decompressed."""
    c = _c()
    from compressed_image_b200 import synth
    w, h = 8192, 4096
    for channel in (0, 3):
        a = synth.half_channel(w, h, channel)
        ch = c.Channel(a, w, h)
        assert ch.num_chunks() == 16 and ch.chunk_size() == 4 << 20
        flat = a.reshape(-1).copy()
        per = ch.chunk_elems()
        order = np.random.default_rng(1234).permutation(16)
        for i in order:
            got = ch.get_chunk(int(i))
            assert np.array_equal(got.view(np.uint16), flat[i * per:(i + 1) * per].view(np.uint16))
        for i in order[:8]:
            new = (flat[i * per:(i + 1) * per].astype(np.float32) * 0.5).astype(np.float16)
            ch.set_chunk(int(i), new)
            flat[i * per:(i + 1) * per] = new
        out = np.asarray(ch.get_decompressed()).reshape(-1)
        assert np.array_equal(out.view(np.uint16), flat.view(np.uint16))


def test_iterator_groups_and_calls_on_the_channel_from_inside_the_loop():
    """The group-pipelined iterator: a chunk modified in the loop is what get_chunk() returns one step later (the
    compress in flight is collected first); set_chunk() of a chunk that is already prefetched but not yet handed out is
    what the iterator then yields; read-only iteration leaves the data alone; 4 MiB chunks (groups of 4)."""
    import compressed_image_b200 as c
    w, h = 2048, 10 * 1024 + 100
    rng = np.random.default_rng(11)
    data = (np.cumsum(rng.integers(0, 3, size=(h, w)), axis=1) & 0xfff).astype(np.uint16)
    ch = c.Channel(data, w, h)
    assert ch.num_chunks() == 11 and ch.chunk_size() == 4 << 20
    flat = data.reshape(-1).copy()
    per = ch.chunk_elems()
    for i, chunk in enumerate(ch):
        assert np.array_equal(chunk, flat[i * per:i * per + chunk.size]), i
        chunk ^= 1
        flat[i * per:i * per + chunk.size] ^= 1
        if i >= 1:
            assert np.array_equal(ch.get_chunk(i - 1), flat[(i - 1) * per:i * per]), i
        if i in (1, 2, 6) and i + 1 < ch.num_chunks():        # next chunk: same group (1, 2, 6 -> 2, 3, 7)
            new = np.full(ch.chunk_elems(i + 1), 1000 + i, np.uint16)
            ch.set_chunk(i + 1, new)
            flat[(i + 1) * per:(i + 1) * per + new.size] = new
    assert np.array_equal(np.asarray(ch.get_decompressed()).reshape(-1), flat)
    total = 0
    for chunk in ch:                                          # read-only
        total += int(chunk[0])
    assert np.array_equal(np.asarray(ch.get_decompressed()).reshape(-1), flat)
    lazy = c.Channel.zeros(np.float32, 1024, 4096)
    for chunk in lazy:
        assert not chunk.any()
    assert lazy.compressed_bytes() == 4 * lazy.num_chunks()   # still lazy: sizeof(T) per chunk (lazyschunk.h:312-327)
