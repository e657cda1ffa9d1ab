"""Functional tests of the PRODUCT's device code (compressed_image_b200/csrc/*.cuh) executed on the CPU by the fiber
simulator tests/sim/cusim.h, checked against the oracle / the reference library / the golden fixtures.

These run in the GPU-less dev container; the same checks run against the real kernels in test_gpu_*.py (-m gpu).
The checker is oracle/ (+ oracle/_ref when built): it is never on the path under test."""
import hashlib

import numpy as np
import pytest

import _data
import _golden
import _libs
import _sim

HAVE_REF = _libs.ref() is not None
FMT = {_data.LZ4: 1, _data.BLOSCLZ: 0, _data.ZSTD: 4}


def _checker_decompress(chunk):
    """Decode with the reference when available, else with the oracle."""
    if HAVE_REF:
        return _libs.ref_decompress(chunk)
    return _libs.oracle_decompress(chunk)


# ---- stream codecs -------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind", _data.KINDS)
@pytest.mark.parametrize("n", [1, 12, 13, 31, 64, 300, 4096, 16384, 40000])
def test_stream_decoders_on_oracle_streams(kind, n):
    O = _libs.oracle()
    src = _data.make_data(kind, n, 1)
    cap = n + n // 255 + 64
    buf = np.zeros(cap, np.uint8)
    r = O.b2o_lz4_compress(_libs.ptr(src), n, _libs.ptr(buf), cap)
    assert r > 0
    got, out = _sim.stream_decode(1, buf[:r], n)
    assert got == n and np.array_equal(out, src)
    for lvl in (5, 9):
        r = O.b2o_blosclz_compress(lvl, _libs.ptr(src), n, _libs.ptr(buf), cap)
        if r > 0:
            got, out = _sim.stream_decode(0, buf[:r], n)
            assert got == n and np.array_equal(out, src), lvl


@pytest.mark.skipif(not HAVE_REF, reason="zstd streams need the reference encoder")
@pytest.mark.parametrize("kind", _data.KINDS)
@pytest.mark.parametrize("n,lvl", [(1, 3), (300, 1), (4096, 3), (16384, 9), (32768, 22), (40000, 5), (200000, 3)])
def test_zstd_stream_decoder(kind, n, lvl):
    R = _libs.ref()
    src = _data.make_data(kind, n, 1)
    cap = n + n // 255 + 128
    buf = np.zeros(cap, np.uint8)
    r = R.ref_zstd_compress(_libs.ptr(src), n, _libs.ptr(buf), cap, lvl)
    assert r > 0
    got, out = _sim.stream_decode(4, buf[:r], n)
    assert got == n and np.array_equal(out, src)


@pytest.mark.parametrize("kind", _data.KINDS)
@pytest.mark.parametrize("n", [13, 16, 64, 100, 300, 4096, 16384, 40000, 70000])
@pytest.mark.parametrize("codec", [_data.LZ4, _data.BLOSCLZ])
def test_stream_encoders_decodable_by_checker(kind, n, codec):
    O = _libs.oracle()
    src = _data.make_data(kind, n, 1)
    r, enc = _sim.stream_encode(codec, src, n - 1)
    assert 0 <= r < n
    if r == 0:
        assert kind in ("random", "iota", "f32", "sensor", "half", "ramp", "text", "sparse") or n < 66, "should compress"
        return
    out = np.zeros(n, np.uint8)
    if codec == _data.LZ4:
        got = O.b2o_lz4_decompress(_libs.ptr(enc), r, _libs.ptr(out), n)
        if HAVE_REF:
            out2 = np.zeros(n, np.uint8)
            assert _libs.ref().ref_lz4_decompress(_libs.ptr(enc), r, _libs.ptr(out2), n) == n
            assert np.array_equal(out2, src)
    else:
        got = O.b2o_blosclz_decompress(_libs.ptr(enc), r, _libs.ptr(out), n)
        if HAVE_REF:
            out2 = np.zeros(n, np.uint8)
            assert _libs.ref().ref_blosclz_decompress(_libs.ptr(enc), r, _libs.ptr(out2), n) == n
            assert np.array_equal(out2, src)
    assert got == n and np.array_equal(out, src)
    # and our own decoder agrees
    got2, out3 = _sim.stream_decode(FMT[codec], enc, n)
    assert got2 == n and np.array_equal(out3, src)


def test_stream_encoders_compress_well():
    """Sanity on ratio: the warp encoders must stay in the same league as the CPU encoders."""
    O = _libs.oracle()
    for kind in ("text", "longrun", "sparse", "zeros"):
        src = _data.make_data(kind, 32768, 1)
        buf = np.zeros(40000, np.uint8)
        ref = O.b2o_lz4_compress(_libs.ptr(src), 32768, _libs.ptr(buf), 40000)
        r, _ = _sim.stream_encode(_data.LZ4, src, 32767, hlog=13)
        assert 0 < r <= ref * 1.1 + 64, (kind, r, ref)
        ref = O.b2o_blosclz_compress(9, _libs.ptr(src), 32768, _libs.ptr(buf), 40000)
        r, _ = _sim.stream_encode(_data.BLOSCLZ, src, 32767, hlog=13)
        assert 0 < r <= ref * 1.1 + 64, (kind, r, ref)


# ---- filters ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("ts", [1, 2, 3, 4, 7, 8, 12, 16, 24])
@pytest.mark.parametrize("n,bs", [(32768, 32768), (70000, 32768), (1000, 256), (4099, 4096), (17, 64)])
def test_filters_match_oracle(ts, n, bs):
    O = _libs.oracle()
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
        assert np.array_equal(_sim.run_filter(filt, True, ts, src, bs), exp_f), ("fwd", filt)
        assert np.array_equal(_sim.run_filter(filt, False, ts, src, bs), exp_b), ("bwd", filt)
    # Blosc1 (format version 2) bit-unshuffle rule
    exp = np.zeros(n, np.uint8)
    for o in range(0, n, bs):
        m = min(bs, n - o)
        s, b = src[o:o + m].copy(), np.zeros(m, np.uint8)
        O.b2o_bitunshuffle(ts, m, _libs.ptr(s), _libs.ptr(b), 2)
        exp[o:o + m] = b
    assert np.array_equal(_sim.run_filter(_data.BITSHUFFLE, False, ts, src, bs, version=2), exp)


def test_filters_unaligned_buffers():
    O = _libs.oracle()
    base = _data.make_data("random", 5000, 1)
    for ts in (2, 4, 8):
        for shift in (1, 3, 8):
            src = base[shift:shift + 4096]
            exp = np.zeros(4096, np.uint8)
            O.b2o_unshuffle(ts, 4096, _libs.ptr(np.ascontiguousarray(src)), _libs.ptr(exp))
            assert np.array_equal(_sim.run_filter(_data.SHUFFLE, False, ts, np.ascontiguousarray(src), 4096), exp)


# ---- whole chunks: decode ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("key", _golden.REF_KEYS)
def test_decode_reference_chunks(key):
    case = _golden.MANIFEST["ref_chunks"][key]["case"]
    src = _data.make_case(case)
    (st, out), = _sim.decompress_batch([_golden.ref_chunk(key)])
    assert st == 0, st
    assert np.array_equal(out, src)


@pytest.mark.parametrize("name", _golden.COMPAT_NAMES)
def test_decode_compat_golden(name):
    """Blosc1-era chunks with blocks of 256 KiB - 2 MiB: exercises the 16-byte header and the global-slab window."""
    meta = _golden.MANIFEST["compat"][name]
    (st, out), = _sim.decompress_batch([_golden.compat_chunk(name)], win_bytes=32768, grid=1)
    assert st == 0, st
    assert out.size == meta["nbytes"]
    assert hashlib.sha256(out.tobytes()).hexdigest() == meta["sha256"]


def test_decode_batch_mixed_and_order():
    keys = _golden.REF_KEYS[::7]
    chunks = [_golden.ref_chunk(k) for k in keys]
    res = _sim.decompress_batch(chunks, grid=3)
    for k, (st, out) in zip(keys, res):
        assert st == 0
        assert np.array_equal(out, _data.make_case(_golden.MANIFEST["ref_chunks"][k]["case"])), k


def test_decode_malformed_chunks():
    """Same accept / reject decisions as the reference on damaged chunks; never writes out of bounds."""
    src = _data.make_data("text", 70000, 1)
    for codec in (_data.LZ4, _data.BLOSCLZ, _data.ZSTD):
        key = [k for k in _golden.REF_KEYS if k.startswith("text_n300000") and "_c%d_" % codec in k][0]
        chunk = _golden.ref_chunk(key)
        nbytes = 300000
        bad = []
        for cut in (8, 20, 40, chunk.size // 2, chunk.size - 1):
            bad.append(chunk[:cut].copy())
        b = chunk.copy(); b[0] = 9; bad.append(b)
        b = chunk.copy(); b[3] = 0; bad.append(b)
        b = chunk.copy(); b[32:36] = 0xFF; bad.append(b)              # bstarts[0] negative
        b = chunk.copy(); b[chunk.size // 2:chunk.size // 2 + 64] ^= 0x5A; bad.append(b)  # payload damage
        res = _sim.decompress_batch(bad, caps=[nbytes] * len(bad))
        for i, (st, out) in enumerate(res):
            rc_ref = _libs.oracle_decompress(bad[i], cap=nbytes)[0]
            if rc_ref < 0:
                assert st < 0, (codec, i, st, rc_ref)
            else:  # damage that still decodes must decode identically
                assert st == 0 and np.array_equal(out, _libs.oracle_decompress(bad[i], cap=nbytes)[1])
        (st, _), = _sim.decompress_batch([chunk], caps=[nbytes - 1])
        assert st == -6


def test_decode_special_chunks():
    """SPECIAL_ZERO / NaN / repeated-value / uninit chunks (blosc2.c:1694-1718) built by hand."""
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
    res = _sim.decompress_batch(chunks)
    for ch, (st, out) in zip(chunks, res):
        rc, exp = _libs.oracle_decompress(ch)
        assert rc == n and st == 0
        assert np.array_equal(out, exp)
        if HAVE_REF:
            assert np.array_equal(_libs.ref_decompress(ch)[1], exp)


@pytest.mark.parametrize("ml", [4, 9, 19, 33, 300])
def test_lz4_offset_zero_stream(ml):
    """A match offset of 0 decodes to zeros (the reference accepts it: tests/test_oracle.py pins that)."""
    from test_oracle import _lz4_offset0_stream
    s, cap = _lz4_offset0_stream(ml)
    exp = np.zeros(cap, np.uint8)
    assert _libs.oracle().b2o_lz4_decompress(s.ctypes.data, int(s.size), exp.ctypes.data, cap) == cap
    r, out = _sim.stream_decode(_data.FMT_LZ4 if hasattr(_data, "FMT_LZ4") else 1, s, cap)
    assert r == cap and np.array_equal(out, exp)


# ---- batched sequence decode (parse ahead, lane-per-sequence copies) ---------------------------------------------------
@pytest.mark.parametrize("codec", [_data.LZ4, _data.BLOSCLZ])
def test_same_decisions_on_damaged_payloads(codec):
    """Byte flips all over the payload: the decoder agrees with the oracle on accept / reject whatever the window slack
    (= how many sequences a batch may hold), and on every byte when the damaged chunk still decodes."""
    rng = np.random.default_rng(7 + codec)
    key = [k for k in _golden.REF_KEYS if k.startswith("text_n300000") and "_c%d_" % codec in k][0]
    chunk = _golden.ref_chunk(key)
    nbytes = 300000
    bad = []
    for _ in range(40):
        b = chunk.copy()
        for _ in range(int(rng.integers(1, 4))):
            b[int(rng.integers(32, chunk.size))] ^= np.uint8(rng.integers(1, 256))
        bad.append(b)
    r1 = _sim.decompress_batch(bad, caps=[nbytes] * len(bad))
    r0 = _sim.decompress_batch(bad, caps=[nbytes] * len(bad), win_bytes=32768 + 1040 + 96)
    for i in range(len(bad)):
        rc, exp = _libs.oracle_decompress(bad[i], cap=nbytes)
        assert (r1[i][0] < 0) == (rc < 0) and (r0[i][0] < 0) == (rc < 0), (i, r1[i][0], r0[i][0], rc)
        assert r1[i][0] == r0[i][0]
        if rc >= 0:
            assert np.array_equal(r1[i][1], exp) and np.array_equal(r0[i][1], exp)


def test_gpu_compressed_chunks_decode():
    """Chunks from the simulated GPU encoder (long runs, overlapping matches, tiny tails)."""
    srcs = [_data.make_data(k, n, ts, seed=n) for k in ("iota", "sparse", "text", "f32", "zeros")
            for n, ts in ((70001, 1), (131072, 2), (65536 + 24, 4))]
    for codec, filt in ((_data.LZ4, _data.SHUFFLE), (_data.BLOSCLZ, _data.BITSHUFFLE), (_data.BLOSCLZ, _data.SHUFFLE)):
        for ts in (1, 2, 4):
            mine = [s for s in srcs if s.size % ts == 0][:6]
            enc = _sim.compress_batch(mine, codec, 9, ts, filt, 32768)
            dec = _sim.decompress_batch([c for _, c in enc], win_bytes=34816)
            for s, (st, out) in zip(mine, dec):
                assert st == 0 and np.array_equal(out, s.view(np.uint8).reshape(-1))


def _noisy_with_repeats(n, gap, mlen, seed):
    """Random bytes with a short repeat of earlier bytes every `gap` bytes: long literal runs, barely compressible, so
    an in-place decode runs with its output right behind the unread input."""
    r = np.random.default_rng(seed)
    a = r.integers(0, 256, size=n, dtype=np.uint8)
    for p in range(gap, n - mlen - 16, gap):
        back = int(r.integers(1, min(p, 60000)))
        a[p:p + mlen] = a[p - back:p - back + mlen]
    return a


@pytest.mark.parametrize("codec", [_data.LZ4, _data.BLOSCLZ])
@pytest.mark.parametrize("gap,mlen", [(40, 12), (70, 8), (200, 24), (33, 5), (1000, 40)])
def test_long_literal_runs_in_place(codec, gap, mlen):
    srcs = [_noisy_with_repeats(65536, gap, mlen, seed) for seed in (1, 2)]
    for ts, bs in ((1, 16384), (2, 32768), (1, 32768)):
        enc = _sim.compress_batch(srcs, codec, 9, ts, _data.SHUFFLE, bs)
        chunks = [c for _, c in enc]
        exp = [_libs.oracle_decompress(c)[1] for c in chunks]
        for wb in (16384 + 96, 17040, 18000, 34000):
            dec = _sim.decompress_batch(chunks, win_bytes=wb)
            for s, e, (st, out) in zip(srcs, exp, dec):
                assert st == 0 and np.array_equal(out, s) and np.array_equal(e, s), (ts, bs, wb)


# ---- plane pairing (parked planes, the last one writes the block interleaved) -----------------------------------------
@pytest.mark.parametrize("ts", [2, 4])
@pytest.mark.parametrize("codec", [_data.LZ4, _data.BLOSCLZ, _data.ZSTD])
def test_plane_pairing_matches_plain_scatter(ts, codec):
    """Split shuffled blocks decode identically with pairing on (3 staging slots), starved (1 slot: planes fall back to
    scattering their own bytes, blocks end up mixed) and off; full blocks are actually paired."""
    if codec == _data.ZSTD and not HAVE_REF:
        pytest.skip("zstd chunks need the reference encoder")
    srcs = [_data.make_data(k, 4 * 32768 + extra, ts, seed=3) for k, extra in
            (("ramp", 0), ("f32", 4 * 17), ("sparse", 0), ("random", 32768), ("zeros", 0))]
    if codec == _data.ZSTD:
        chunks = [_libs.ref_compress(codec, 3, ts, _data.SHUFFLE, 32768, s) for s in srcs]
    else:
        chunks = [c for _, c in _sim.compress_batch(srcs, codec, 9, ts, _data.SHUFFLE, 32768)]
    got = {}
    for pair in (1, 2, 0):
        res = _sim.decompress_batch(chunks, grid=3, pair=pair)
        paired = _sim.lib().sim_last_paired()
        for s, (st, out) in zip(srcs, res):
            assert st == 0 and np.array_equal(out, s.view(np.uint8).reshape(-1)), pair
        got[pair] = paired
    assert got[1] >= 1      # how many depends on the order in which the warps happen to finish
    assert got[0] == 0


def test_plane_pairing_unaligned_destination_falls_back():
    src = _data.make_data("ramp", 3 * 32768, 2, seed=5)
    (st, chunk), = _sim.compress_batch([src], _data.LZ4, 9, 2, _data.SHUFFLE, 32768)
    assert st == 0
    (st, out), = _sim.decompress_batch([chunk])
    assert st == 0 and np.array_equal(out, src)


# ---- whole chunks: compress -----------------------------------------------------------------------------------------
def _compress_cases():
    seen = set()
    for c in _data.REF_CHUNK_GRID:
        if c["codec"] not in (_data.LZ4, _data.BLOSCLZ):
            continue
        k = _data.case_key(c)
        if k not in seen:
            seen.add(k)
            yield pytest.param(c, id=k)


@pytest.mark.parametrize("case", list(_compress_cases()))
def test_compress_chunks_decodable_by_checker(case):
    src = _data.make_case(case)
    (st, chunk), = _sim.compress_batch([src], case["codec"], case["clevel"], case["typesize"], case["filter"],
                                       case["blocksize"], win_bytes=max(32768, case["blocksize"]))
    assert st == 0 and chunk.size > 0
    assert chunk.size <= src.size + 32
    nb, cb, bs = (int.from_bytes(chunk[o:o + 4].tobytes(), "little") for o in (4, 8, 12))
    assert nb == src.size and bs == chunk.size
    rc, back = _checker_decompress(chunk)
    assert rc == src.size, rc
    assert np.array_equal(back, src)
    rc2, back2 = _libs.oracle_decompress(chunk)
    assert rc2 == src.size and np.array_equal(back2, src)
    # header fields and size are what the reference would report for the same parameters
    key = _data.case_key(case)
    ref_chunk = _golden.ref_chunk(key)
    assert chunk[0] == ref_chunk[0] and chunk[3] == ref_chunk[3]
    assert np.array_equal(chunk[4:12], ref_chunk[4:12])          # nbytes, blocksize
    assert np.array_equal(chunk[16:24], ref_chunk[16:24])        # filters, compcode
    assert (chunk[2] & 0xF5) == (ref_chunk[2] & 0xF5) or (chunk[2] & 2) != (ref_chunk[2] & 2)
    # and the GPU decoder round-trips its own output
    (st3, back3), = _sim.decompress_batch([chunk])
    assert st3 == 0 and np.array_equal(back3, src)


def test_compress_batch_lookback_many_ctas():
    """Several chunks, more CTAs than blocks per chunk: exercises ticket order + decoupled look-back."""
    srcs = [_data.make_data(k, 200000 + 1000 * i, 2, seed=i) for i, k in enumerate(("half", "sensor", "zeros", "random", "text"))]
    res = _sim.compress_batch(srcs, _data.LZ4, 9, 2, _data.SHUFFLE, 32768, grid=5)
    for s, (st, chunk) in zip(srcs, res):
        assert st == 0
        rc, back = _checker_decompress(chunk)
        assert rc == s.size and np.array_equal(back, s)
    assert res[2][1].size == 32 and res[2][1][31] == 0x10      # all zeros -> SPECIAL_ZERO chunk (blosc2.c:2520-2529)
    assert res[3][1].size == srcs[3].size + 32 and (res[3][1][2] & 2)  # noise -> memcpyed chunk


@pytest.mark.parametrize("codec", [_data.LZ4, _data.BLOSCLZ])
@pytest.mark.parametrize("clevel", [1, 3, 5, 6, 8, 9])
def test_compress_levels_decodable_by_checker(codec, clevel):
    """The encoder's effort follows the level (b2g_plan.h: table size, shortest match worth a sequence, the margin under
    which a stream is stored): whatever the level, the reference / the oracle decode the chunk and so does the kernel."""
    sizes = []
    for ts, kinds in ((2, ("half", "sensor", "text")), (1, ("ramp", "longrun")), (4, ("f32",))):
        srcs = [_data.make_data(k, 150000 + 4 * i, ts, seed=7 + i) for i, k in enumerate(kinds)]
        res = _sim.compress_batch(srcs, codec, clevel, ts, _data.SHUFFLE, 32768)
        for s, (st, chunk) in zip(srcs, res):
            assert st == 0
            rc, back = _checker_decompress(chunk)
            assert rc == s.size and np.array_equal(back, s), (clevel, ts)
            sizes.append(chunk.size)
        for s, (st, out) in zip(srcs, _sim.decompress_batch([c for _, c in res])):
            assert st == 0 and np.array_equal(out, s)
    assert sum(sizes) < 6 * 150100 * 0.9          # every level still compresses this data


@pytest.mark.parametrize("codec", [_data.LZ4, _data.BLOSCLZ])
@pytest.mark.parametrize("team,prime,sparse", [(2, 0, 0), (2, 8192, 1), (3, 2048, 0), (4, 0, 0), (4, 2048, 0), (4, 8192, 1),
                                               (8, 128, 0), (8, 65536, 0)])
def test_compress_teams_decodable_by_checker(codec, team, prime, sparse):
    """k_compress_team: the warps of a CTA encode consecutive segments of one stream and join their bodies
    (lz_encode_warp<SEG>, lz_join_team).  Whatever the team size and the priming, the chunk is an ordinary one -- the
    reference / the oracle and the kernel decode it -- and it stays close to what one warp per stream makes of the same data (image planes with a primed table: within 3 %; text with cold 4 KiB tables: 25 % worse)
    makes of the same data; streams shorter than a segment, leftover blocks and run streams go through the team too."""
    cases = [(2, "half", 150000), (2, "sensor", 98304 + 10), (1, "text", 70001), (1, "ramp", 65536), (1, "longrun", 40000),
             (4, "f32", 131072 + 52), (1, "random", 33000), (2, "zeros", 66000), (1, "sparse", 50000), (1, "iota", 600)]
    try:
        for ts, kind, n in cases:
            src = _data.make_data(kind, n, ts, seed=3)
            for filt in (_data.SHUFFLE, _data.BITSHUFFLE) if ts == 4 else (_data.SHUFFLE,):
                _sim.set_team(1)
                (st0, solo), = _sim.compress_batch([src], codec, 9, ts, filt, 32768)
                _sim.set_team(team, prime, sparse=sparse)
                (st, chunk), = _sim.compress_batch([src], codec, 9, ts, filt, 32768)
                assert st == 0 and st0 == 0
                rc, back = _checker_decompress(chunk)
                assert rc == src.size and np.array_equal(back, src), (ts, kind, filt)
                (st2, out), = _sim.decompress_batch([chunk])
                assert st2 == 0 and np.array_equal(out, src)
                if n >= 8192:      # (a 600-byte stream forced through a team is cut every 256 bytes: nothing to compare)
                    assert chunk.size <= solo.size * 1.35 + 128, (ts, kind, chunk.size, solo.size)
                if kind in ("half", "sensor") and prime >= 2048:
                    assert chunk.size <= solo.size * 1.03, (ts, kind, chunk.size, solo.size)
    finally:
        _sim.set_team(-1, -1)


def test_compress_capacity_rules():
    src = _data.make_data("random", 50000, 1)
    (st, chunk), = _sim.compress_batch([src], _data.LZ4, 5, 1, _data.SHUFFLE, 32768, caps=[50000 + 31])
    assert st == 0 and chunk.size == 0                       # does not fit: 0, like blosc2_compress_ctx
    src = _data.make_data("text", 50000, 1)
    (st, chunk), = _sim.compress_batch([src], _data.LZ4, 5, 1, _data.SHUFFLE, 32768, caps=[40000])
    assert st == 0 and 0 < chunk.size <= 40000
    assert np.array_equal(_checker_decompress(chunk)[1], src)
    (st, chunk), = _sim.compress_batch([np.zeros(0, np.uint8)], _data.LZ4, 5, 1, _data.SHUFFLE, 32768)
    assert st == 0 and chunk.size == 32
    assert _checker_decompress(chunk)[0] == 0
