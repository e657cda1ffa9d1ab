"""Lazy chunks as blosc2 special-value chunks (second half): compressed_image_b200.channel
.special_value_chunk builds byte for byte what blosc2_chunk_zeros / blosc2_chunk_repeatval build
(thirdparty/c-blosc2/blosc/blosc2.c:4304-4347, 4444-4489); the reference, the oracle and the GPU decode them to the fill."""
import ctypes as C

import numpy as np
import pytest

import _libs

CASES = [(np.uint8, 7, 100000, 32768), (np.uint8, 0, 4096, 32768), (np.uint16, 513, 2 * 70001, 32768), (np.int16, -2, 64, 32768),
         (np.uint32, 0xDEADBEEF, 4 * 12345, 4096), (np.float16, 1.5, 2 * 9999, 32768), (np.float32, 0.0, 4 << 20, 32768),
         (np.float32, -3.25, 4 << 20, 32768), (np.int32, 0, 8, 32768), (np.float32, 2.0, 4 * 5, 32768)]


def _special(dtype, value, nbytes, bs):
    import importlib
    ch = importlib.import_module("compressed_image_b200.channel") if _have_lib() else None
    if ch is None:
        pytest.skip("libb2g.so not built")
    return ch.special_value_chunk(np.array(value).astype(dtype)[()], nbytes, bs)


def _have_lib():
    import os
    from compressed_image_b200 import _b2g
    return os.path.exists(_b2g.LIB_PATH)


@pytest.mark.parametrize("dtype,value,nbytes,bs", CASES)
def test_special_value_chunk_is_what_the_reference_builds(dtype, value, nbytes, bs):
    chunk = _special(dtype, value, nbytes, bs)
    v = np.array(value).astype(dtype).reshape(1)
    expect = np.full(nbytes // v.dtype.itemsize, v[0], dtype=dtype).view(np.uint8)
    rc, back = _libs.oracle_decompress(chunk)
    assert rc == nbytes and np.array_equal(back, expect)
    R = _libs.ref()
    if R is None:
        return
    R.ref_special_chunk.argtypes = [C.c_int, C.c_int, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32]
    dst = np.zeros(64, np.uint8)
    zero = not v.view(np.uint8).any()
    n = R.ref_special_chunk(v.dtype.itemsize, bs, nbytes, None if zero else v.ctypes.data, dst.ctypes.data, 64)
    assert n == chunk.size and np.array_equal(dst[:n], chunk), "bytes differ from blosc2_chunk_zeros / _repeatval"
    rc, back = _libs.ref_decompress(chunk)
    assert rc == nbytes and np.array_equal(back, expect)


@pytest.mark.gpu
def test_special_value_chunks_decode_on_the_device():
    """The device-resident entry point expands special-value chunks in HBM (an O(1)-input fill), mixed with ordinary
    chunks in one batch."""
    import torch
    from compressed_image_b200 import Channel, _b2g
    from compressed_image_b200.channel import special_value_chunk
    chunks, expect = [], []
    for dtype, value, nbytes, bs in CASES:
        v = np.array(value).astype(dtype)[()]
        chunks.append(special_value_chunk(v, nbytes, bs))
        expect.append(np.full(nbytes // np.dtype(dtype).itemsize, v, dtype=dtype).view(np.uint8))
    data = (np.arange(300 * 512) // 5).astype(np.uint16)
    with _b2g.Context(_b2g.LZ4, 9, 2, _b2g.SHUFFLE, 32768, 0) as cx:
        chunks.append(cx.compress([data])[0])
        expect.append(data.view(np.uint8))
        d_in = [torch.from_numpy(np.concatenate([c, np.zeros(16, np.uint8)])).cuda() for c in chunks]
        d_out = [torch.full((e.size + 64,), 0xAB, dtype=torch.uint8, device="cuda") for e in expect]
        torch.cuda.synchronize()
        b = cx.make_batch([t.data_ptr() for t in d_in], [t.data_ptr() for t in d_out], [int(c.size) for c in chunks],
                          [int(e.size) for e in expect])
        cx.decompress_batch_device(b)
        cx.wait()
    for i, e in enumerate(expect):
        o = d_out[i].cpu().numpy()
        assert b[i].status == 0 and b[i].out_bytes == e.size
        assert np.array_equal(o[:e.size], e) and np.all(o[e.size:] == 0xAB)
    # a lazy channel hands its chunks over as special chunks: 32 / 36 bytes each instead of a compressed fill buffer
    lazy = Channel.full(np.float32, 2.5, 1024, 512, chunk_size=64 * 1024 * 4)
    sp = lazy.chunks(special=True)
    assert len(sp) == lazy.num_chunks() and all(c.size == 36 for c in sp)
    lazy.set_chunk(1, np.arange(lazy.chunk_elems(1), dtype=np.float32))
    sp = lazy.chunks(special=True)
    assert sp[0].size == 36 and sp[1].size > 36
    with _b2g.Context(_b2g.LZ4, 9, 4, _b2g.SHUFFLE, 32768, 0) as cx:
        outs = cx.decompress(sp)
    full = np.concatenate(outs).view(np.float32)
    assert np.array_equal(full, np.asarray(lazy.get_decompressed()).reshape(-1))
