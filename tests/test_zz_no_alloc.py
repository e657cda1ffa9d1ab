"""b2g_decompress_batch_device_noalloc (include/b2g.h): the allocation-free decompress entry point
(no counterpart in the reference).  Called through the C ABI and checked against the existing
device-resident entry point, b2g_decompress_batch_device, bit for bit (the bar for byte work), on two calls in a row --
the second one with other chunks, so that anything the shared static scratch kept from the first call would show.

In a file of its own that pytest collects LAST; passes alone in a fresh process and after
the other tests."""
import ctypes as C

import numpy as np
import pytest

import _data

pytestmark = pytest.mark.gpu


def _setup(srcs_by_params):
    """-> list of (chunk bytes, raw bytes): GPU-compressed chunks of several parameter sets."""
    from compressed_image_b200 import _b2g
    out = []
    for (codec, ts, filt, bs), srcs in srcs_by_params:
        with _b2g.Context(codec, 9, ts, filt, bs, 0) as cx:
            for s, c in zip(srcs, cx.compress(srcs)):
                out.append((c, np.ascontiguousarray(s).reshape(-1).view(np.uint8)))
    return out


def _run_both(pairs):
    """Decode the chunks with the existing device-resident entry point and with the no-alloc one; -> two lists of
    (status, out_bytes, bytes)."""
    import torch
    from compressed_image_b200 import _b2g
    L = _b2g.lib()
    n = len(pairs)
    d_in = [torch.from_numpy(np.concatenate([c, np.zeros(16, np.uint8)])).cuda() for c, _ in pairs]
    caps = [int(r.size) for _, r in pairs]
    results = []
    for which in ("ctx", "noalloc"):
        d_out = [torch.full((max(cap, 1) + 64,), 0xAB, dtype=torch.uint8, device="cuda") for cap in caps]
        torch.cuda.synchronize()
        b = _b2g.Context.make_batch([t.data_ptr() for t in d_in], [t.data_ptr() for t in d_out],
                                    [int(c.size) for c, _ in pairs], caps)
        if which == "ctx":
            with _b2g.Context(_b2g.BLOSCLZ, 9, 1, _b2g.SHUFFLE, 65536, 0) as cx:
                cx.decompress_batch_device(b)
                cx.wait()
        else:
            rc = L.b2g_decompress_batch_device_noalloc(0, b, n, None)
            assert rc == 0, (rc, L.b2g_last_error())
        torch.cuda.synchronize()
        res = []
        for i in range(n):
            o = d_out[i].cpu().numpy()
            assert np.all(o[max(caps[i], 1):] == 0xAB), "decoder wrote past the destination capacity"
            res.append((b[i].status, b[i].out_bytes, o[:caps[i]].copy()))
        results.append(res)
    return results


def test_noalloc_decompress_matches_existing_entry_point_twice_in_a_row():
    from compressed_image_b200 import _b2g
    LZ4, BLZ, LZ4HC, SH, BSH, NOF = _b2g.LZ4, _b2g.BLOSCLZ, _b2g.LZ4HC, _b2g.SHUFFLE, _b2g.BITSHUFFLE, _b2g.NOFILTER
    first = _setup([
        ((LZ4, 2, SH, 32768), [_data.make_data(k, 600000 + 2 * i, 2, seed=i) for i, k in enumerate(("half", "sensor", "zeros", "random"))]),
        ((BLZ, 4, BSH, 32768), [_data.make_data(k, 400000 + 4 * i, 4, seed=10 + i) for i, k in enumerate(("f32", "sparse", "iota"))]),
        ((LZ4, 1, SH, 65536), [_data.make_data("text", 300001, 1, seed=3)]),
    ])
    second = _setup([
        ((LZ4, 4, SH, 32768), [_data.make_data(k, 500000 + 4 * i, 4, seed=20 + i) for i, k in enumerate(("f32", "ramp", "sparse"))]),
        ((LZ4HC, 2, SH, 16384), [_data.make_data("half", 200000, 2, seed=30)]),
        ((BLZ, 2, SH, 32768), [_data.make_data(k, 700000 + 2 * i, 2, seed=40 + i) for i, k in enumerate(("sensor", "half"))]),
        ((LZ4, 8, NOF, 32768), [_data.make_data("iota", 100000, 8, seed=50)]),
    ])
    for pairs in (first, second, first):        # calls in a row on the shared static scratch; the first set once more
        ref, got = _run_both(pairs)
        for (c, raw), (st0, nb0, out0), (st1, nb1, out1) in zip(pairs, ref, got):
            assert st0 == 0 and nb0 == raw.size and np.array_equal(out0, raw), "existing entry point"
            assert (st1, nb1) == (st0, nb0)
            assert np.array_equal(out1, out0), "no-alloc output differs from b2g_decompress_batch_device"


def test_noalloc_limits_and_per_chunk_errors():
    import torch
    from compressed_image_b200 import _b2g
    L = _b2g.lib()
    # more chunks than the static scratch holds: refused as a whole, with a text that says why
    n = _b2g.NOALLOC_MAX_CHUNKS + 1
    b = (_b2g.Chunk * n)()
    rc = L.b2g_decompress_batch_device_noalloc(0, b, n, None)
    assert rc == -35 and b"B2G_NOALLOC_MAX_CHUNKS" in L.b2g_last_error()
    assert L.b2g_decompress_batch_device_noalloc(0, None, 0, None) == 0
    assert L.b2g_decompress_batch_device_noalloc(0, None, 3, None) == -12
    # a block larger than the maximum fails that chunk only; a zstd-format header is refused per chunk; the good one decodes
    big = _data.make_data("half", 300000, 2, seed=1)
    good = _data.make_data("sensor", 300000, 2, seed=2)
    with _b2g.Context(_b2g.LZ4, 9, 2, _b2g.SHUFFLE, 262144, 0) as cx:
        c_big = cx.compress([big])[0]
    with _b2g.Context(_b2g.LZ4, 9, 2, _b2g.SHUFFLE, 32768, 0) as cx:
        c_good = cx.compress([good])[0]
    c_z = c_good.copy()
    c_z[2] = (c_z[2] & 0x1f) | (4 << 5)          # codec format bits -> zstd
    chunks = [c_big, c_good, c_z]
    d_in = [torch.from_numpy(np.concatenate([c, np.zeros(16, np.uint8)])).cuda() for c in chunks]
    d_out = [torch.zeros(300000 + 64, dtype=torch.uint8, device="cuda") for _ in chunks]
    torch.cuda.synchronize()
    bb = _b2g.Context.make_batch([t.data_ptr() for t in d_in], [t.data_ptr() for t in d_out], [int(c.size) for c in chunks],
                                 [300000] * 3)
    assert L.b2g_decompress_batch_device_noalloc(-1, bb, 3, None) == 0
    assert bb[0].status == -101 and bb[0].out_bytes == 0
    assert bb[1].status == 0 and bb[1].out_bytes == 300000
    assert np.array_equal(d_out[1].cpu().numpy()[:300000], good.reshape(-1).view(np.uint8))
    assert bb[2].status == -102 and bb[2].out_bytes == 0
