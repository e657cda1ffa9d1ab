"""BASELINE.json configs at (or near) their full sizes on the GPU (-m gpu), through size-independent properties:
round trips on the device with a checksum instead of a byte-for-byte host compare, the reference library decoding a
sample of the GPU-produced chunks and producing the zstd chunks the GPU decodes, chunk counts and ratios in the expected
range.  cfg 2 (8192 x 4096 half) is bench.py's own workload and is gated there.  Sizes are cut to what finishes in
seconds (cfg 3: 4 of 16 channels, cfg 4: 12 of 256 images); the per-chunk path is the same at any count."""
import ctypes as C

import numpy as np
import pytest

import _data
import _libs

pytestmark = pytest.mark.gpu

HAVE_REF = _libs.ref() is not None
CHUNK = 4 * 1024 * 1024
BLOCK = 32768


def _device_round_trip(pieces, codec, clevel, typesize, filt, ref_sample=4):
    """compress (device-resident) -> decompress (device-resident): bit-exact on the device, every status 0; the
    reference decodes `ref_sample` of the chunks.  -> (ratio, chunks as numpy arrays of the sample)."""
    import torch
    from compressed_image_b200 import _b2g
    n = len(pieces)
    total = sum(int(p.size) for p in pieces)
    raw = torch.empty(total, dtype=torch.uint8)
    offs, o = [], 0
    for p in pieces:
        raw[o:o + p.size] = torch.from_numpy(np.ascontiguousarray(p))
        offs.append(o)
        o += int(p.size)
    d_raw = raw.cuda()
    slot = max(int(p.size) for p in pieces) + 256
    d_chunks = torch.zeros(n * slot, dtype=torch.uint8, device="cuda")
    d_out = torch.zeros(total, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    with _b2g.Context(codec, clevel, typesize, filt, BLOCK) as ctx:
        comp = ctx.make_batch([d_raw.data_ptr() + offs[i] for i in range(n)], [d_chunks.data_ptr() + i * slot for i in range(n)],
                              [int(p.size) for p in pieces], [int(p.size) + 32 for p in pieces])
        ctx.compress_batch_device(comp)
        ctx.wait()
        assert all(comp[i].status == 0 and comp[i].out_bytes > 0 for i in range(n))
        dec = ctx.make_batch([d_chunks.data_ptr() + i * slot for i in range(n)], [d_out.data_ptr() + offs[i] for i in range(n)],
                             [comp[i].out_bytes for i in range(n)], [int(p.size) for p in pieces])
        ctx.decompress_batch_device(dec)
        ctx.wait()
        assert all(dec[i].status == 0 and dec[i].out_bytes == pieces[i].size for i in range(n))
    torch.cuda.synchronize()
    assert torch.equal(d_out, d_raw), "device round trip mismatch"
    cbytes = sum(comp[i].out_bytes for i in range(n))
    sample = []
    for i in np.linspace(0, n - 1, num=min(ref_sample, n), dtype=int):
        c = d_chunks[i * slot:i * slot + comp[i].out_bytes].cpu().numpy()
        rc, back = (_libs.ref_decompress if HAVE_REF else _libs.oracle_decompress)(c)
        assert rc == pieces[i].size and np.array_equal(back, np.ascontiguousarray(pieces[i]).view(np.uint8).reshape(-1)), i
        sample.append(c)
    return total / cbytes, sample


def test_cfg1_rgba_u8_4096_image_api():
    """cfg 1: compressed::image<uint8_t> 4096 x 4096 RGBA, lz4 + byte-shuffle, default chunks: construct (full compress)
    then get_decompressed(); 4 chunks per channel; the opaque alpha plane collapses to run streams."""
    import compressed_image_b200 as c
    from compressed_image_b200 import synth
    chans = [synth.u8_rgba_channel(4096, 4096, k) for k in range(4)]
    img = c.Image(np.uint8, chans, 4096, 4096, channel_names=["R", "G", "B", "A"])
    assert [img.channel(k).num_chunks() for k in range(4)] == [4, 4, 4, 4]
    out = img.get_decompressed()
    for k in range(4):
        assert np.array_equal(out[k], chans[k])
    assert img.channel("A").compressed_bytes() < 4 * 4096   # 255 everywhere: run streams only
    for ch in img.channel(0).chunks():
        rc, back = (_libs.ref_decompress if HAVE_REF else _libs.oracle_decompress)(ch)
        assert rc == CHUNK
    assert 1.05 < img.compression_ratio() < 50


@pytest.mark.parametrize("filt", [_data.BITSHUFFLE, _data.SHUFFLE])
def test_cfg3_f32_stack_blosclz(filt):
    """cfg 3: compressed::image<float> 8192 x 8192 compositing stack (one channel of each kind: 1 GiB of the 4 GiB),
    blosclz + bit-shuffle (unsplit 32 KiB streams) and + byte-shuffle (four 8 KiB planes per block)."""
    from compressed_image_b200 import synth
    pieces = []
    for k, kind in enumerate(synth.F32_KINDS):
        pieces += synth.split_chunks(synth.f32_channel(8192, 8192, kind, channel=k), CHUNK)
    assert len(pieces) == 4 * 64
    ratio, _ = _device_round_trip(pieces, _data.BLOSCLZ, 9, 4, filt)
    assert ratio > 1.5


@pytest.mark.skipif(not HAVE_REF, reason="zstd chunks come from the reference encoder")
@pytest.mark.parametrize("clevel", [1, 5, 9])
def test_cfg4_zstd_decode_then_lz4_reencode(clevel):
    """cfg 4: 2048 x 2048 uint16 RGB images, zstd chunks from the reference (clevel <= 5: split planes, clevel 9: one
    stream per block at zstd level 22) decoded on the GPU, bit-exact, then re-encoded with lz4 on the GPU."""
    from compressed_image_b200 import _b2g, synth
    nimg = 12 if clevel < 9 else 2
    pieces = []
    for im in range(nimg):
        for ch in range(3):
            pieces += synth.split_chunks(synth.u16_sensor_channel(2048, 2048, ch, image=im), CHUNK)
    assert len(pieces) == nimg * 3 * 2
    n = len(pieces)
    R = _libs.ref()
    vp = C.c_void_p
    dsts = [np.empty(p.size + 32, np.uint8) for p in pieces]
    out_c = (C.c_int32 * n)()
    R.ref_compress_batch(_data.ZSTD, clevel, 2, _data.SHUFFLE, BLOCK, n, (vp * n)(*[p.ctypes.data for p in pieces]),
                         (C.c_int32 * n)(*[int(p.size) for p in pieces]), (vp * n)(*[d.ctypes.data for d in dsts]),
                         (C.c_int32 * n)(*[int(d.size) for d in dsts]), out_c, 16)
    zchunks = [d[:out_c[i]] for i, d in enumerate(dsts)]
    assert all(c.size > 32 and (c[2] >> 5) == 4 for c in zchunks)
    with _b2g.Context(_data.ZSTD, 9, 2, _data.SHUFFLE, BLOCK) as ctx:
        outs = ctx.decompress(zchunks)
    for p, o in zip(pieces, outs):
        assert np.array_equal(o, p)
    with _b2g.Context(_data.LZ4, 9, 2, _data.SHUFFLE, BLOCK) as ctx:
        lz = ctx.compress(outs)
        back = ctx.decompress(lz)
    for p, o in zip(pieces, back):
        assert np.array_equal(o, p)
    for i in (0, n // 2, n - 1):
        rc, b = _libs.ref_decompress(lz[i])
        assert rc == pieces[i].size and np.array_equal(b, pieces[i])


@pytest.mark.parametrize("chunk_kib", [64, 256, 1024, 4096, 16384])
@pytest.mark.parametrize("codec,filt", [(_data.LZ4, _data.SHUFFLE), (_data.BLOSCLZ, _data.SHUFFLE)])
def test_cfg5_chunk_size_sweep(chunk_kib, codec, filt):
    """cfg 5: chunk sizes 64 KiB ... 16 MiB x {lz4, blosclz} x {uint8, half, float}, 2048 px wide, 32 MiB per case."""
    from compressed_image_b200 import synth
    for ts, gen in ((1, lambda h: synth.u8_rgba_channel(2048, h, 0)),
                    (2, lambda h: synth.half_channel(2048, h, 1).view(np.uint16)),
                    (4, lambda h: synth.f32_channel(2048, h, "depth"))):
        height = (32 << 20) // (2048 * ts)
        pieces = synth.split_chunks(gen(height), chunk_kib * 1024)
        assert len(pieces) == (32 << 20) // (chunk_kib * 1024)
        ratio, _ = _device_round_trip(pieces, codec, 9, ts, filt, ref_sample=2)
        assert ratio > 1.0


# ---- full configured sizes through the HOST path (split_batch, address-ordered merging, staging threads) ----------------
def _host_round_trip_full(pieces, codec, typesize, filt, views=None):
    """One host-pointer compress batch over ALL chunks of the configuration, one host-pointer decompress batch of the
    results into ONE output array (neighbouring outputs: merged copies), pageable numpy memory (the staged path).  Checks:
    every status 0, the output equals the inputs chunk by chunk, the reference decodes a sample of the GPU chunks."""
    from compressed_image_b200 import _b2g
    n = len(pieces)
    with _b2g.Context(codec, 9, typesize, filt, BLOCK) as ctx:
        launches0 = ctx.launches
        chunks = ctx.compress(pieces)
        assert len(chunks) == n
        sizes = [int(p.size) for p in pieces]
        out = np.empty(sum(sizes), np.uint8)
        outs, o = [], 0
        for s in sizes:
            outs.append(out[o:o + s])
            o += s
        order = np.random.default_rng(7).permutation(n)             # the batch in scrambled order: sorted by address inside
        res = ctx.decompress([chunks[i] for i in order], [outs[i] for i in order])
        assert len(res) == n
        # a few hundred chunks per launch, not one launch per chunk: 1536 chunks of 4 MiB are ~200 sub-batches of 32 MiB
        assert ctx.launches - launches0 <= 12 * (sum(sizes) // (8 << 20) + 2)
    o = 0
    for i, p in enumerate(pieces):
        assert np.array_equal(out[o:o + sizes[i]], np.ascontiguousarray(p).reshape(-1).view(np.uint8)), i
        o += sizes[i]
    for i in np.linspace(0, n - 1, num=6, dtype=int):
        rc, back = (_libs.ref_decompress if HAVE_REF else _libs.oracle_decompress)(chunks[i])
        assert rc == sizes[i] and np.array_equal(back, np.ascontiguousarray(pieces[i]).reshape(-1).view(np.uint8)), i
    return sum(sizes) / sum(int(c.size) for c in chunks)


def test_cfg3_full_size_host_path():
    """cfg 3 at its configured size: 16 float channels of 8192 x 8192 = 1024 chunks, 4 GiB, blosclz + bitshuffle, one batch
    (four synthetic channels, one per kind, each standing for four channels of the stack)."""
    from compressed_image_b200 import _b2g, synth
    pieces = []
    for k, kind in enumerate(synth.F32_KINDS):
        chan = synth.split_chunks(synth.f32_channel(8192, 8192, kind, channel=k), CHUNK)
        pieces += chan * 4
    assert len(pieces) == 1024
    ratio = _host_round_trip_full(pieces, _b2g.BLOSCLZ, 4, _b2g.BITSHUFFLE)
    assert 2.0 < ratio < 5.0


def test_cfg4_full_size_host_path():
    """cfg 4 at its configured size: 256 uint16 RGB images of 2048 x 2048 = 1536 chunks, 6 GiB, lz4 + shuffle, one batch
    (32 distinct synthetic images, each standing for eight images of the batch)."""
    from compressed_image_b200 import _b2g, synth
    pieces = []
    for im in range(32):
        for ch in range(3):
            pieces += synth.split_chunks(synth.u16_sensor_channel(2048, 2048, ch, image=im), CHUNK)
    pieces = pieces * 8
    assert len(pieces) == 1536
    ratio = _host_round_trip_full(pieces, _b2g.LZ4, 2, _b2g.SHUFFLE)
    assert 1.5 < ratio < 2.5
