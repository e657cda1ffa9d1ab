"""Contiguous-frame export / import against frames written by the unmodified reference library
(tests/golden/ref_cframe.npz, made by tests/golden/make_cframe_golden.py) and, where oracle/_ref is present, against the
reference reader itself.  The frame code has no codec of its own; here the offsets chunk goes through the oracle (test
infrastructure), on the GPU box through the b2g context (test_gpu_b2frame below, -m gpu)."""
import ctypes as C
import os

import numpy as np
import pytest

import _libs
from compressed_image_b200 import b2frame

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_cframe.npz"))
CASES = ["u16_lz4_5chunks", "f32_blosclz_1chunk", "u8_lz4_12chunks"]
HAVE_REF = _libs.ref() is not None


def _oracle_engine():
    def enc(a):
        rc, c = _libs.oracle_compress(0, 5, 8, 1, 32768, np.ascontiguousarray(a).view(np.uint8))
        assert rc > 0
        return c

    def dec(chunk, out):
        rc, b = _libs.oracle_decompress(chunk)
        assert rc == out.size
        out[:] = b
    return enc, dec


def _golden_chunks(name):
    cb = G[name + "/cbytes"]
    flat = G[name + "/chunks"]
    offs = np.concatenate(([0], np.cumsum(cb)))
    return [flat[offs[i]:offs[i + 1]] for i in range(cb.size)]


def _ref_reader():
    R = _libs.ref()
    R.ref_cframe_info.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
    R.ref_cframe_decompress_chunk.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int32]
    R.ref_cframe_get_chunk.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int32]
    return R


def check_reference_reads(frame, raws_flat, nbytes):
    """The reference opens `frame` (blosc2_schunk_from_buffer) and decompresses every chunk to the original bytes."""
    R = _ref_reader()
    m = np.frombuffer(frame, np.uint8).copy()
    info = (C.c_int64 * 6)()
    assert R.ref_cframe_info(m.ctypes.data, m.size, info) == 0
    assert info[0] == len(nbytes) and info[1] == int(np.sum(nbytes))
    o = 0
    for i, nb in enumerate(nbytes):
        out = np.empty(int(nb), np.uint8)
        assert R.ref_cframe_decompress_chunk(m.ctypes.data, m.size, i, out.ctypes.data, out.size) == nb
        assert np.array_equal(out, raws_flat[o:o + int(nb)]), i
        o += int(nb)


@pytest.mark.parametrize("name", CASES)
def test_reads_reference_frames(name):
    enc, dec = _oracle_engine()
    hd, chunks = b2frame.from_cframe(G[name + "/frame"].tobytes(), dec)
    want = _golden_chunks(name)
    assert hd["nchunks"] == len(want) and hd["header_len"] == 97 and hd["trailer_len"] == 35
    assert hd["nbytes"] == int(G[name + "/nbytes"].sum()) and hd["cbytes"] == int(G[name + "/cbytes"].sum())
    assert hd["chunksize"] == int(G[name + "/nbytes"][0]) and hd["typesize"] == 8 and hd["special"] == {}
    o = 0
    for c, w, nb in zip(chunks, want, G[name + "/nbytes"]):
        assert np.array_equal(c, w)
        rc, raw = _libs.oracle_decompress(c)
        assert rc == nb and np.array_equal(raw, G[name + "/raw"][o:o + int(nb)])
        o += int(nb)


@pytest.mark.parametrize("name", CASES)
def test_writes_what_the_reference_writes(name):
    """Same chunks in -> the reference's frame byte for byte (header, chunks, trailer; and, the oracle's blosclz being
    bit-exact with the reference encoder, the offsets chunk too)."""
    enc, dec = _oracle_engine()
    mine = b2frame.to_cframe(_golden_chunks(name), enc)
    ref = G[name + "/frame"].tobytes()
    assert mine[:97] == ref[:97]
    assert mine[-35:] == ref[-35:]
    cb = int(G[name + "/cbytes"].sum())
    assert mine[97:97 + cb] == ref[97:97 + cb]
    assert mine == ref


def test_empty_frame_matches_the_reference():
    enc, dec = _oracle_engine()
    ref = G["empty/frame"].tobytes()
    hd, chunks = b2frame.from_cframe(ref, dec)
    assert chunks == [] and hd["nchunks"] == 0 and hd["nbytes"] == 0
    mine = b2frame.to_cframe([], enc)
    assert len(mine) == len(ref) == 132
    assert mine == ref


@pytest.mark.skipif(not HAVE_REF, reason="needs oracle/_ref")
@pytest.mark.parametrize("name", CASES)
def test_reference_reads_our_frames(name):
    enc, dec = _oracle_engine()
    mine = b2frame.to_cframe(_golden_chunks(name), enc, typesize=2, codec=1, clevel=9, blocksize=32768)
    check_reference_reads(mine, G[name + "/raw"], G[name + "/nbytes"])
    hd, chunks = b2frame.from_cframe(mine, dec)
    assert (hd["codec"], hd["clevel"], hd["typesize"], hd["blocksize"]) == (1, 9, 2, 32768)
    assert all(np.array_equal(a, b) for a, b in zip(chunks, _golden_chunks(name)))


def test_malformed_frames_are_rejected():
    enc, dec = _oracle_engine()
    good = bytearray(G["u16_lz4_5chunks/frame"].tobytes())
    with pytest.raises(b2frame.FrameError, match="shorter"):
        b2frame.from_cframe(bytes(good[:50]), dec)
    bad = bytearray(good); bad[5] ^= 0xff
    with pytest.raises(b2frame.FrameError, match="not a b2frame"):
        b2frame.from_cframe(bytes(bad), dec)
    with pytest.raises(b2frame.FrameError):
        b2frame.from_cframe(bytes(good[:-7]), dec)                       # frame_len beyond the buffer
    bad = bytearray(good); bad[-22:-18] = (1 << 20).to_bytes(4, "big")
    with pytest.raises(b2frame.FrameError, match="trailer"):
        b2frame.from_cframe(bytes(bad), dec)
    bad = bytearray(good); bad[26] = 1
    with pytest.raises(b2frame.FrameError, match="contiguous"):
        b2frame.from_cframe(bytes(bad), dec)
    bad = bytearray(good); bad[39:47] = (int.from_bytes(good[39:47], "big") - 8).to_bytes(8, "big")
    with pytest.raises(b2frame.FrameError):
        b2frame.from_cframe(bytes(bad), dec)                             # cbytes no longer lines up with the offsets chunk
    chunks = _golden_chunks("u16_lz4_5chunks")
    with pytest.raises(b2frame.FrameError, match="only the last chunk"):
        b2frame.to_cframe([chunks[4], chunks[0]], enc)
    with pytest.raises(b2frame.FrameError, match="header says"):
        b2frame.to_cframe([chunks[0][:-1]], enc)


def test_offsets_pointing_outside_are_rejected():
    enc, dec = _oracle_engine()
    chunks = _golden_chunks("u8_lz4_12chunks")

    def lying(a):
        a = a.copy()
        a[3] = 1 << 40
        return enc(a)
    frame = b2frame.to_cframe(chunks, lying)
    with pytest.raises(b2frame.FrameError, match="outside"):
        b2frame.from_cframe(frame, dec)


@pytest.mark.gpu
@pytest.mark.parametrize("codec,ts", [(1, 2), (0, 4)])
def test_gpu_b2frame(codec, ts):
    """Chunks compressed on the GPU, offsets chunk compressed on the GPU -> the reference opens the frame and decodes
    every chunk; a reference-written frame is opened here with the GPU decoding the offsets chunk and the chunks."""
    from compressed_image_b200 import _b2g
    rng = np.random.default_rng(ts)
    dt = {2: np.uint16, 4: np.uint32}[ts]
    raws = [(np.cumsum(rng.integers(0, 3, size=65536)) & 0x3ff).astype(dt) for _ in range(9)] + \
           [(np.cumsum(rng.integers(0, 3, size=1001)) & 0x3ff).astype(dt)]
    with _b2g.Context(codec, 9, ts, 1, 32768) as data_ctx, _b2g.Context(0, 5, 8, 1, 32768) as idx_ctx:
        enc, dec = b2frame.engine_from_context(idx_ctx)
        chunks = data_ctx.compress(raws)
        frame = b2frame.to_cframe(chunks, enc, typesize=ts, codec=codec, clevel=9, blocksize=32768)
        hd, back = b2frame.from_cframe(frame, dec)
        assert hd["nchunks"] == len(raws)
        outs = data_ctx.decompress(back)
        assert all(np.array_equal(o.view(dt), r) for o, r in zip(outs, raws))
        if HAVE_REF:
            check_reference_reads(frame, np.concatenate([r.view(np.uint8) for r in raws]), [r.nbytes for r in raws])
        for name in CASES:
            hd, cs = b2frame.from_cframe(G[name + "/frame"].tobytes(), dec)
            outs = data_ctx.decompress(cs)
            assert np.array_equal(np.concatenate(outs), G[name + "/raw"])


@pytest.mark.gpu
def test_gpu_channel_cframe_round_trip():
    """Channel.to_cframe(): the reference opens it and decodes the image; set_from_cframe() restores a channel."""
    import compressed_image_b200 as c
    rng = np.random.default_rng(9)
    data = (np.cumsum(rng.integers(0, 3, size=(300, 512)), axis=1) & 0xfff).astype(np.uint16)
    ch = c.Channel(data, 512, 300, chunk_size=64 * 1024)
    frame = ch.to_cframe()
    nb = [ch._nbytes_of(i) for i in range(ch.num_chunks())]
    assert len(nb) > 2 and nb[-1] < nb[0]
    if HAVE_REF:
        check_reference_reads(frame, data.reshape(-1).view(np.uint8), nb)
    other = c.Channel.zeros(np.uint16, 512, 300, chunk_size=64 * 1024)
    other.set_from_cframe(frame)
    assert np.array_equal(np.asarray(other.get_decompressed()).reshape(300, 512), data)
    with pytest.raises(ValueError):
        c.Channel.zeros(np.uint16, 512, 200, chunk_size=64 * 1024).set_from_cframe(frame)
