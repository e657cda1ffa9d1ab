"""Pins the CPU oracle (oracle/*.c, our restatement) against (a) the committed golden vectors and (b) the
unmodified reference library when it is available (dev container: oracle/_ref built from /root/reference)."""
import hashlib

import numpy as np
import pytest

import _data
import _golden
import _libs

HAVE_REF = _libs.ref() is not None
needs_ref = pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not built")


@pytest.mark.parametrize("name", _golden.COMPAT_NAMES)
def test_oracle_decodes_compat_golden(name):
    chunk = _golden.compat_chunk(name)
    meta = _golden.MANIFEST["compat"][name]
    rc, out = _libs.oracle_decompress(chunk)
    assert rc == meta["nbytes"], rc
    assert hashlib.sha256(out.tobytes()).hexdigest() == meta["sha256"]
    exp = _golden.compat_expected(name)
    if exp is not None:
        assert np.array_equal(out, exp)


@pytest.mark.parametrize("key", _golden.REF_KEYS)
def test_oracle_decodes_reference_chunks(key):
    case = _golden.MANIFEST["ref_chunks"][key]["case"]
    src = _data.make_case(case)
    assert hashlib.sha256(src.tobytes()).hexdigest() == _golden.MANIFEST["ref_chunks"][key]["sha256"]
    rc, out = _libs.oracle_decompress(_golden.ref_chunk(key))
    assert rc == src.size, rc
    assert np.array_equal(out, src)


@needs_ref
@pytest.mark.parametrize("key", [k for k in _golden.REF_KEYS if "_c0_" in k or "_c1_" in k])
def test_oracle_compress_matches_reference_bytes(key):
    """nthreads=1 reference output is deterministic: the restated encoders must reproduce it byte for byte."""
    case = _golden.MANIFEST["ref_chunks"][key]["case"]
    src = _data.make_case(case)
    rc, chunk = _libs.oracle_compress(case["codec"], case["clevel"], case["typesize"], case["filter"],
                                      case["blocksize"], src)
    ref_chunk = _golden.ref_chunk(key)
    assert rc == ref_chunk.size
    assert np.array_equal(chunk, ref_chunk)
    r2, back = _libs.ref_decompress(chunk)
    assert r2 == src.size and np.array_equal(back, src)


@needs_ref
@pytest.mark.parametrize("ts", [1, 2, 3, 4, 7, 8, 16, 24])
@pytest.mark.parametrize("n", [0, 1, 7, 64, 1000, 32768, 32771])
def test_filters_match_reference(ts, n):
    O, R = _libs.oracle(), _libs.ref()
    src = _data.make_data("random", max(n, 1), 1, seed=ts)[:n]
    a, b = np.zeros(n + 1, np.uint8), np.zeros(n + 1, np.uint8)
    O.b2o_shuffle(ts, n, _libs.ptr(src), _libs.ptr(a)); R.ref_shuffle(ts, n, _libs.ptr(src), _libs.ptr(b))
    assert np.array_equal(a, b)
    O.b2o_unshuffle(ts, n, _libs.ptr(src), _libs.ptr(a)); R.ref_unshuffle(ts, n, _libs.ptr(src), _libs.ptr(b))
    assert np.array_equal(a, b)
    O.b2o_bitshuffle(ts, n, _libs.ptr(src), _libs.ptr(a)); R.ref_bitshuffle(ts, n, _libs.ptr(src), _libs.ptr(b))
    assert np.array_equal(a, b)
    for ver in (2, 5):
        O.b2o_bitunshuffle(ts, n, _libs.ptr(src), _libs.ptr(a), ver)
        R.ref_bitunshuffle(ts, n, _libs.ptr(src), _libs.ptr(b), ver)
        assert np.array_equal(a, b), ver


@needs_ref
@pytest.mark.parametrize("kind", _data.KINDS)
@pytest.mark.parametrize("n", [1, 12, 13, 64, 4096, 16384, 70000])
def test_stream_codecs_match_reference(kind, n):
    O, R = _libs.oracle(), _libs.ref()
    src = _data.make_data(kind, n, 1)
    cap = n + n // 255 + 64
    a, b = np.zeros(cap, np.uint8), np.zeros(cap, np.uint8)
    for limit in (cap, max(n - 1, 1)):
        ra = O.b2o_lz4_compress(_libs.ptr(src), n, _libs.ptr(a), limit)
        rb = R.ref_lz4_compress(_libs.ptr(src), n, _libs.ptr(b), limit)
        assert ra == rb and np.array_equal(a[:ra], b[:rb])
        if ra > 0:
            out = np.zeros(n, np.uint8)
            assert O.b2o_lz4_decompress(_libs.ptr(a), ra, _libs.ptr(out), n) == n and np.array_equal(out, src)
    for lvl in (1, 5, 9):
        ra = O.b2o_blosclz_compress(lvl, _libs.ptr(src), n, _libs.ptr(a), cap)
        rb = R.ref_blosclz_compress(lvl, 1, 1, _libs.ptr(src), n, _libs.ptr(b), cap)
        assert ra == rb and np.array_equal(a[:ra], b[:rb]), (lvl, ra, rb)
        if ra > 0:
            out = np.zeros(n, np.uint8)
            assert O.b2o_blosclz_decompress(_libs.ptr(a), ra, _libs.ptr(out), n) == n
            assert np.array_equal(out, src)
    for lvl in (1, 9, 22):
        rb = R.ref_zstd_compress(_libs.ptr(src), n, _libs.ptr(b), cap, lvl)
        assert rb > 0
        out = np.zeros(n, np.uint8)
        assert O.b2o_zstd_decompress(_libs.ptr(b), rb, _libs.ptr(out), n) == n, (lvl,)
        assert np.array_equal(out, src)


@needs_ref
def test_malformed_chunks_agree_with_reference():
    """Error behaviour on truncated / corrupted chunks: oracle and reference must both reject (b2c:672-766, 1783-1923)."""
    src = _data.make_data("text", 70000, 1)
    for codec in (_data.LZ4, _data.BLOSCLZ, _data.ZSTD):
        chunk = _libs.ref_compress(codec, 5, 1, _data.SHUFFLE, 32768, src)
        for cut in (8, 20, 40, chunk.size // 2, chunk.size - 1):
            bad = chunk[:cut].copy()
            r_ref, _ = _libs.ref_decompress(bad, cap=src.size)
            r_or, _ = _libs.oracle_decompress(bad, cap=src.size)
            assert r_ref < 0 and r_or < 0, (codec, cut, r_ref, r_or)
        bad = chunk.copy(); bad[0] = 9
        assert _libs.ref_decompress(bad)[0] < 0 and _libs.oracle_decompress(bad)[0] < 0
        bad = chunk.copy(); bad[3] = 0
        assert _libs.ref_decompress(bad)[0] < 0 and _libs.oracle_decompress(bad)[0] < 0
        rc, _ = _libs.oracle_decompress(chunk, cap=src.size - 1)
        assert rc < 0 and _libs.ref_decompress(chunk, cap=src.size - 1)[0] < 0


def _lz4_offset0_stream(ml, tail=12):
    """8 literals, a match with OFFSET 0 of length ml, then `tail` last literals."""
    out = bytearray()
    mlc = ml - 4
    out.append((8 << 4) | min(mlc, 15))
    out += b"abcdefgh" + bytes([0, 0])
    if mlc >= 15:
        r = mlc - 15
        while r >= 255:
            out.append(255)
            r -= 255
        out.append(r)
    out.append(tail << 4)
    out += bytes(range(65, 65 + tail))
    return np.frombuffer(bytes(out), dtype=np.uint8).copy(), 8 + ml + tail


@pytest.mark.skipif(_libs.ref() is None, reason="reference library not built")
@pytest.mark.parametrize("ml", [4, 5, 8, 9, 16, 18, 19, 32, 33, 64, 300, 1000])
def test_lz4_offset_zero_is_zeros_like_the_reference(ml):
    """LZ4_decompress_safe accepts a match offset of 0 (lz4.c:2356 only rejects offsets beyond the prefix) and the match
    comes out as zeros whatever the destination held before (lz4.c:500, :2407).  Pins the oracle's reading of it."""
    s, cap = _lz4_offset0_stream(ml)
    for fill in (0xAB, 0x00):
        a = np.full(cap + 8, fill, np.uint8)
        b = np.full(cap + 8, fill, np.uint8)
        ra = _libs.ref().ref_lz4_decompress(s.ctypes.data, int(s.size), a.ctypes.data, cap)
        rb = _libs.oracle().b2o_lz4_decompress(s.ctypes.data, int(s.size), b.ctypes.data, cap)
        assert ra == rb == cap
        assert np.array_equal(a, b)
        assert not a[8:8 + ml].any()
