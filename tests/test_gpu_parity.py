"""Parity tests proper (-m gpu): the sm_100a kernels, called through the C ABI (include/b2g.h), against the checker
(oracle/liboracle.so, oracle/_ref/libblosc2_ref.so when it travelled with the snapshot) and the committed golden
fixtures.  Bit-exact throughout: this path is integer / byte work.

Mirrors tests/test_sim_kernels.py (which runs the same device code on the CPU simulator) case for case."""
import hashlib

import numpy as np
import pytest

import _data
import _golden
import _libs

pytestmark = pytest.mark.gpu

HAVE_REF = _libs.ref() is not None


def _gpu():
    import _gpu as g
    return g


def _checker_decompress(chunk):
    if HAVE_REF:
        return _libs.ref_decompress(chunk)
    return _libs.oracle_decompress(chunk)


def test_library_loaded_and_device_present():
    from compressed_image_b200 import _b2g
    assert _b2g.lib().b2g_device_count() >= 1
    assert b"sm_100a" in _b2g.lib().b2g_version()


# ---- filters ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("ts", [1, 2, 3, 4, 7, 8, 12, 16, 24])
@pytest.mark.parametrize("n,bs", [(32768, 32768), (70000, 32768), (1000, 256), (4099, 4096), (17, 64),
                                  (1 << 20, 32768)])
def test_filters_match_oracle(ts, n, bs):
    g, O = _gpu(), _libs.oracle()
    bs = max(bs // ts * ts, ts)
    src = _data.make_data("random", n, 1, seed=ts)
    for filt in (_data.SHUFFLE, _data.BITSHUFFLE, _data.NOFILTER):
        exp_f = np.zeros(n, np.uint8)
        exp_b = np.zeros(n, np.uint8)
        for o in range(0, n, bs):
            m = min(bs, n - o)
            s, a, b = src[o:o + m].copy(), np.zeros(m, np.uint8), np.zeros(m, np.uint8)
            if filt == _data.SHUFFLE:
                O.b2o_shuffle(ts, m, _libs.ptr(s), _libs.ptr(a)); O.b2o_unshuffle(ts, m, _libs.ptr(s), _libs.ptr(b))
            elif filt == _data.BITSHUFFLE:
                O.b2o_bitshuffle(ts, m, _libs.ptr(s), _libs.ptr(a)); O.b2o_bitunshuffle(ts, m, _libs.ptr(s), _libs.ptr(b), 5)
            else:
                a[:] = s; b[:] = s
            exp_f[o:o + m] = a; exp_b[o:o + m] = b
        assert np.array_equal(g.run_filter(filt, True, ts, src, bs), exp_f), ("fwd", filt)
        assert np.array_equal(g.run_filter(filt, False, ts, src, bs), exp_b), ("bwd", filt)
    exp = np.zeros(n, np.uint8)
    for o in range(0, n, bs):   # Blosc1 (format version 2) bit-unshuffle rule
        m = min(bs, n - o)
        s, b = src[o:o + m].copy(), np.zeros(m, np.uint8)
        O.b2o_bitunshuffle(ts, m, _libs.ptr(s), _libs.ptr(b), 2)
        exp[o:o + m] = b
    assert np.array_equal(g.run_filter(_data.BITSHUFFLE, False, ts, src, bs, version=2), exp)


# ---- whole chunks: decode ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("device", [False, True], ids=["host", "device"])
def test_decode_reference_chunks(device):
    """Every reference-produced chunk of the golden grid (tests/golden/ref_chunks.npz), in ONE batch."""
    g = _gpu()
    chunks = [_golden.ref_chunk(k) for k in _golden.REF_KEYS]
    res = g.decompress_batch(chunks, device=device)
    for k, (st, out) in zip(_golden.REF_KEYS, res):
        assert st == 0, (k, st)
        assert np.array_equal(out, _data.make_case(_golden.MANIFEST["ref_chunks"][k]["case"])), k


@pytest.mark.parametrize("key", _golden.REF_KEYS[::5])
def test_decode_reference_chunk_single(key):
    g = _gpu()
    (st, out), = g.decompress_batch([_golden.ref_chunk(key)])
    assert st == 0
    assert np.array_equal(out, _data.make_case(_golden.MANIFEST["ref_chunks"][key]["case"]))


@pytest.mark.parametrize("name", _golden.COMPAT_NAMES)
def test_decode_compat_golden(name):
    """c-blosc2's own known-answer chunks (compat/*.cdata): Blosc1 16-byte headers, 256 KiB - 2 MiB blocks."""
    g = _gpu()
    meta = _golden.MANIFEST["compat"][name]
    (st, out), = g.decompress_batch([_golden.compat_chunk(name)])
    assert st == 0, st
    assert out.size == meta["nbytes"]
    assert hashlib.sha256(out.tobytes()).hexdigest() == meta["sha256"]
    exp = _golden.compat_expected(name)
    if exp is not None:
        assert np.array_equal(out, exp)


def test_decode_malformed_chunks():
    g = _gpu()
    for codec in (_data.LZ4, _data.BLOSCLZ, _data.ZSTD):
        key = [k for k in _golden.REF_KEYS if k.startswith("text_n300000") and "_c%d_" % codec in k][0]
        chunk = _golden.ref_chunk(key)
        nbytes = 300000
        bad = []
        for cut in (8, 20, 40, chunk.size // 2, chunk.size - 1):
            bad.append(chunk[:cut].copy())
        b = chunk.copy(); b[0] = 9; bad.append(b)
        b = chunk.copy(); b[3] = 0; bad.append(b)
        b = chunk.copy(); b[32:36] = 0xFF; bad.append(b)
        b = chunk.copy(); b[chunk.size // 2:chunk.size // 2 + 64] ^= 0x5A; bad.append(b)
        for device in (False, True):
            res = g.decompress_batch(bad, caps=[nbytes] * len(bad), device=device)
            for i, (st, out) in enumerate(res):
                rc_ref, exp = _libs.oracle_decompress(bad[i], cap=nbytes)
                if rc_ref < 0:
                    assert st < 0, (codec, i, st, rc_ref)
                else:
                    assert st == 0 and np.array_equal(out, exp)
        (st, _), = g.decompress_batch([chunk], caps=[nbytes - 1])
        assert st == -6


def test_decode_special_chunks():
    g = _gpu()

    def hdr(nbytes, ts, special, cbytes=32, blocksize=None):
        h = np.zeros(32, np.uint8)
        h[0], h[1], h[2], h[3] = 5, 1, 0x05, ts
        h[4:8] = np.frombuffer(int(nbytes).to_bytes(4, "little"), np.uint8)
        h[8:12] = np.frombuffer(int(blocksize or min(nbytes, 32768)).to_bytes(4, "little"), np.uint8)
        h[12:16] = np.frombuffer(int(cbytes).to_bytes(4, "little"), np.uint8)
        h[31] = special << 4
        return h
    n = 100000
    chunks = [hdr(n, 4, 1), hdr(n, 4, 2), hdr(n, 8, 2),
              np.concatenate([hdr(n, 4, 3, cbytes=36), np.array([1, 2, 3, 4], np.uint8)])]
    for ch, (st, out) in zip(chunks, g.decompress_batch(chunks)):
        rc, exp = _libs.oracle_decompress(ch)
        assert rc == n and st == 0
        assert np.array_equal(out, exp)


# ---- whole chunks: compress -----------------------------------------------------------------------------------------
def _compress_cases():
    seen, out = set(), []
    for c in _data.REF_CHUNK_GRID:
        if c["codec"] in (_data.LZ4, _data.BLOSCLZ) and _data.case_key(c) not in seen:
            seen.add(_data.case_key(c))
            out.append(c)
    return out


@pytest.mark.parametrize("case", _compress_cases(), ids=_data.case_key)
def test_compress_chunks_decodable_by_checker(case):
    g = _gpu()
    src = _data.make_case(case)
    (st, chunk), = g.compress_batch([src], case["codec"], case["clevel"], case["typesize"], case["filter"],
                                    case["blocksize"])
    assert st == 0 and chunk.size > 0
    assert chunk.size <= src.size + 32
    nb, cb, bs = (int.from_bytes(chunk[o:o + 4].tobytes(), "little") for o in (4, 8, 12))
    assert nb == src.size and bs == chunk.size
    rc, back = _checker_decompress(chunk)          # the reference decodes every GPU-produced chunk
    assert rc == src.size, rc
    assert np.array_equal(back, src)
    rc2, back2 = _libs.oracle_decompress(chunk)
    assert rc2 == src.size and np.array_equal(back2, src)
    ref_chunk = _golden.ref_chunk(_data.case_key(case))
    assert chunk[0] == ref_chunk[0] and chunk[3] == ref_chunk[3]
    assert np.array_equal(chunk[4:12], ref_chunk[4:12])
    assert np.array_equal(chunk[16:24], ref_chunk[16:24])
    (st3, back3), = g.decompress_batch([chunk])
    assert st3 == 0 and np.array_equal(back3, src)
    # the device-resident entry points give byte-identical chunks for a single-chunk batch of <= 1 block ... and
    # always decodable ones
    (st4, chunk4), = g.compress_batch([src], case["codec"], case["clevel"], case["typesize"], case["filter"],
                                      case["blocksize"], device=True)
    assert st4 == 0
    assert np.array_equal(_checker_decompress(chunk4)[1], src)


def test_compress_ratio_next_to_reference():
    """Match choices may differ from the CPU encoder; the ratio must stay in the same league (reported, and bounded
    loosely here: within 15 % + 64 B of the reference's chunk for the same parameters)."""
    g = _gpu()
    for case in _compress_cases():
        if case["nbytes"] < 50000 or case["kind"] in ("random",):
            continue
        src = _data.make_case(case)
        (st, chunk), = g.compress_batch([src], case["codec"], case["clevel"], case["typesize"], case["filter"],
                                        case["blocksize"])
        ref = _golden.MANIFEST["ref_chunks"][_data.case_key(case)]["cbytes"]
        assert st == 0 and chunk.size <= ref * 1.15 + 64, (_data.case_key(case), chunk.size, ref)


def test_compress_batch_lookback_many_ctas():
    g = _gpu()
    kinds = ("half", "sensor", "zeros", "random", "text") * 8
    srcs = [_data.make_data(k, 2000000 + 1000 * i, 2, seed=i) for i, k in enumerate(kinds)]
    res = g.compress_batch(srcs, _data.LZ4, 9, 2, _data.SHUFFLE, 32768)
    for s, (st, chunk) in zip(srcs, res):
        assert st == 0
        rc, back = _checker_decompress(chunk)
        assert rc == s.size and np.array_equal(back, s)
    assert res[2][1].size == 32 and res[2][1][31] == 0x10
    assert res[3][1].size == srcs[3].size + 32 and (res[3][1][2] & 2)
    back = g.decompress_batch([c for _, c in res], device=True)
    for s, (st, out) in zip(srcs, back):
        assert st == 0 and np.array_equal(out, s)


@pytest.mark.parametrize("codec", [_data.LZ4, _data.BLOSCLZ])
@pytest.mark.parametrize("ts,bs", [(1, 65536), (1, 131072), (1, 163840), (1, 196608), (2, 65536), (4, 262144), (2, 24576)])
def test_compress_stream_sizes_across_the_team_geometries(codec, ts, bs):
    """Streams from 12 KiB to 192 KiB (the encoder's limit): one warp per stream, teams of 2 ... 8 warps, and the sizes
    where the shared window leaves room for fewer warps than a segment every 8 KiB asks for.  Every chunk is decoded by
    the reference / the oracle and by the kernel."""
    g = _gpu()
    srcs = [_data.make_data(k, 3 * bs + 1000 * i + 77, ts, seed=5 + i) for i, k in enumerate(("half", "text", "sparse", "ramp"))]
    res = g.compress_batch(srcs, codec, 9, ts, _data.SHUFFLE, bs)
    for s, (st, chunk) in zip(srcs, res):
        assert st == 0 and 0 < chunk.size <= s.size + 32
        rc, back = _checker_decompress(chunk)
        assert rc == s.size and np.array_equal(back, s), (ts, bs)
    for s, (st, out) in zip(srcs, g.decompress_batch([c for _, c in res])):
        assert st == 0 and np.array_equal(out, s)


def test_compress_capacity_rules():
    g = _gpu()
    src = _data.make_data("random", 50000, 1)
    (st, chunk), = g.compress_batch([src], _data.LZ4, 5, 1, _data.SHUFFLE, 32768, caps=[50000 + 31], device=True)
    assert st == 0 and chunk.size == 0
    src = _data.make_data("text", 50000, 1)
    (st, chunk), = g.compress_batch([src], _data.LZ4, 5, 1, _data.SHUFFLE, 32768, caps=[40000], device=True)
    assert st == 0 and 0 < chunk.size <= 40000
    assert np.array_equal(_checker_decompress(chunk)[1], src)
    (st, chunk), = g.compress_batch([np.zeros(0, np.uint8)], _data.LZ4, 5, 1, _data.SHUFFLE, 32768)
    assert st == 0 and chunk.size == 32
    assert _checker_decompress(chunk)[0] == 0


def test_unsupported_encoders_fail_loudly():
    """zstd ENCODE is out of scope: the call fails, it does not fall back to a CPU."""
    from compressed_image_b200 import _b2g
    g = _gpu()
    src = _data.make_data("text", 50000, 1)
    with pytest.raises(_b2g.B2GError) as e:
        g.compress_batch([src], _data.ZSTD, 5, 1, _data.SHUFFLE, 32768)
    assert e.value.code == -7


@pytest.mark.parametrize("ts", [1, 2, 4])
def test_lz4hc_codec_id_encodes_as_lz4_streams(ts):
    """Codec lz4hc: the chunks carry compcode 2 / format 1 and unsplit blocks, like the reference's lz4hc chunks
    (stune.c:200-209, blosc2.c:2457); the streams are plain LZ4, which is what lz4hc streams are.  The reference and the
    oracle decode them."""
    g = _gpu()
    srcs = [_data.make_data(k, 300000 + 8 * i, ts, seed=i) for i, k in enumerate(("half", "text", "sparse", "zeros", "random"))]
    for (st, chunk), src in zip(g.compress_batch(srcs, _data.LZ4HC, 9, ts, _data.SHUFFLE, 32768), srcs):
        assert st == 0
        raw = src.reshape(-1).view(np.uint8)
        if chunk[2] & 0x02 or chunk.size == 32:      # memcpyed / special chunk
            pass
        else:
            assert chunk[22] == 2 and (chunk[2] >> 5) == 1 and (chunk[2] & 0x10), "lz4hc header fields"
        rc, back = _checker_decompress(chunk)
        assert rc == raw.size and np.array_equal(back, raw)
    outs = g.decompress_batch([c for _, c in g.compress_batch(srcs, _data.LZ4HC, 9, ts, _data.SHUFFLE, 32768)])
    for (st, out), src in zip(outs, srcs):
        assert st == 0 and np.array_equal(out, src.reshape(-1).view(np.uint8))


def test_async_host_batches_on_two_contexts():
    """b2g_decompress_batch_async on one context and b2g_compress_batch_async on another, in flight together, give what
    the synchronous calls give; many sub-batches (B2G_SUBBATCH_MB is 32 by default, the batch is ~100 MB); a second
    batch on a busy context is refused."""
    from compressed_image_b200 import _b2g
    g = _gpu()
    kinds = ("half", "sensor", "text", "zeros", "random", "half") * 4
    srcs = [_data.make_data(k, 4000000 + 4096 * i, 2, seed=i) for i, k in enumerate(kinds)]
    ref_chunks = [c for _, c in g.compress_batch(srcs, _data.LZ4, 9, 2, _data.SHUFFLE, 32768)]
    n = len(srcs)
    with _b2g.Context(_data.LZ4, 9, 2, _data.SHUFFLE, 32768) as cd, _b2g.Context(_data.LZ4, 9, 2, _data.SHUFFLE, 32768) as cc:
        outs = [np.full(s.size + 64, 0xAB, np.uint8) for s in srcs]
        dsts = [np.full(s.size + 32 + 64, 0xAB, np.uint8) for s in srcs]
        bd = cd.make_batch([c.ctypes.data for c in ref_chunks], [o.ctypes.data for o in outs],
                           [int(c.size) for c in ref_chunks], [int(s.size) for s in srcs])
        bc = cc.make_batch([s.ctypes.data for s in srcs], [d.ctypes.data for d in dsts], [int(s.size) for s in srcs],
                           [int(s.size) + 32 for s in srcs])
        for _ in range(2):
            cd.decompress_batch_async(bd)
            cc.compress_batch_async(bc)
            with pytest.raises(_b2g.B2GError):
                cd.decompress_batch_async(bd)
            cc.wait()
            cd.wait()
            for i in range(n):
                assert bd[i].status == 0 and bd[i].out_bytes == srcs[i].size
                assert np.array_equal(outs[i][:srcs[i].size], srcs[i]) and np.all(outs[i][srcs[i].size:] == 0xAB)
                assert bc[i].status == 0 and bc[i].out_bytes == ref_chunks[i].size
                assert np.array_equal(dsts[i][:bc[i].out_bytes], ref_chunks[i])
                assert np.all(dsts[i][srcs[i].size + 32:] == 0xAB)
        cd.wait()   # nothing in flight: a no-op


def test_damaged_payloads_same_decisions_as_reference():
    """Byte flips in lz4 / blosclz payloads: the GPU decoder accepts / rejects what the reference library does, and
    equals it byte for byte where the damaged chunk still decodes (device-resident and host-pointer paths)."""
    g = _gpu()
    keys = [k for k in _golden.REF_KEYS if "_c%d_" % _data.LZ4 in k or "_c%d_" % _data.BLOSCLZ in k]
    rng = np.random.default_rng(11)
    bad = []
    for key in [k for k in keys if k.startswith("text_n300000")]:
        chunk = _golden.ref_chunk(key)
        for _ in range(24):
            b = chunk.copy()
            for _ in range(int(rng.integers(1, 4))):
                b[int(rng.integers(32, chunk.size))] ^= np.uint8(rng.integers(1, 256))
            bad.append(b)
    for device in (False, True):
        res = g.decompress_batch(bad, caps=[300000] * len(bad), device=device)
        for b, (st, out) in zip(bad, res):
            rc, exp = _libs.ref_decompress(b, cap=300000) if _libs.ref() else _libs.oracle_decompress(b, cap=300000)
            assert (st < 0) == (rc < 0), (st, rc)
            if rc >= 0:
                assert np.array_equal(out, exp)


@pytest.mark.parametrize("codec", [_data.LZ4, _data.BLOSCLZ])
@pytest.mark.parametrize("ts", [1, 2, 4])
def test_every_compression_level_is_reference_decodable(codec, ts):
    """The encoder's effort follows clevel (b2g_plan.h: table size, shortest match worth a sequence, the margin under which
    a stream is stored): whatever the level, the reference decodes every chunk bit-exactly, the GPU does too, and level 9
    compresses at least as well as the faster levels (to within a few bytes of framing)."""
    g = _gpu()
    srcs = [_data.make_data(k, 400000 + ts * i, ts, seed=i) for i, k in enumerate(("half", "sensor", "text", "ramp", "f32", "random"))]
    sizes = {}
    for clevel in (1, 2, 3, 5, 6, 8, 9):
        res = g.compress_batch(srcs, codec, clevel, ts, _data.SHUFFLE, 32768)
        total = 0
        for (st, chunk), src in zip(res, srcs):
            assert st == 0
            raw = src.reshape(-1).view(np.uint8)
            rc, back = _checker_decompress(chunk)
            assert rc == raw.size and np.array_equal(back, raw), (clevel, ts)
            total += chunk.size
        outs = g.decompress_batch([c for _, c in res])
        for (st, out), src in zip(outs, srcs):
            assert st == 0 and np.array_equal(out, src.reshape(-1).view(np.uint8))
        sizes[clevel] = total
    assert sizes[9] <= min(sizes.values()) * 1.002
    assert sizes[1] <= sizes[9] * 1.12          # the fast levels give up a few per cent, not the compression


def test_staged_host_path_reports_per_chunk_errors_like_the_direct_path():
    """A batch in pageable memory large enough for the staging threads (several sub-batches) with damaged chunks in it: every
    chunk gets the status and the bytes it gets when it is decoded alone (one sub-batch: the direct path), good chunks are
    untouched by their neighbours' failures, and a compress batch through the same path round-trips."""
    from compressed_image_b200 import _b2g
    n, nb = 10, 4 << 20
    srcs = [_data.make_data(k, nb, 2, seed=i) for i, k in enumerate(("half", "sensor", "text", "ramp", "half") * 2)]
    with _b2g.Context(_b2g.LZ4, 9, 2, _b2g.SHUFFLE, 32768, 0) as cx:
        chunks = cx.compress(srcs)                       # 40 MiB of pageable input: staged
        for c, s in zip(chunks, srcs):
            rc, back = _checker_decompress(c)
            assert rc == nb and np.array_equal(back, s)
        bad = [c.copy() for c in chunks]
        bad[2][4:8] = np.frombuffer(np.int32(-4).tobytes(), np.uint8)        # negative nbytes
        bad[5] = bad[5][:bad[5].size // 2].copy()                            # truncated
        bad[7][40:44] = 0xFF                                                 # a bstarts entry out of range
        outs = [np.full(nb, 0xAB, np.uint8) for _ in range(n)]
        b = cx.make_batch([c.ctypes.data for c in bad], [o.ctypes.data for o in outs], [int(c.size) for c in bad], [nb] * n)
        cx.decompress_batch(b)
        alone = []
        for c in bad:
            o = np.full(nb, 0xAB, np.uint8)
            b1 = cx.make_batch([c.ctypes.data], [o.ctypes.data], [int(c.size)], [nb])
            cx.decompress_batch(b1)
            alone.append((b1[0].status, b1[0].out_bytes))
        for i in range(n):
            assert (b[i].status, b[i].out_bytes) == alone[i], i
            if i in (2, 5, 7):
                assert b[i].status < 0 and b[i].out_bytes == 0
            else:
                assert b[i].status == 0 and np.array_equal(outs[i], srcs[i]), i
