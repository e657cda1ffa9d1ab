"""Helpers that drive the PRODUCT through its C ABI (compressed_image_b200/libb2g.so via ctypes) for the -m gpu tests.
Same call shapes as tests/_sim.py so the GPU tests mirror the simulator tests one to one."""
import ctypes as C

import numpy as np

from compressed_image_b200 import _b2g

_ctxs = {}


def ctx(codec=1, clevel=9, typesize=1, filt=1, blocksize=32768):
    key = (codec, clevel, typesize, filt, blocksize)
    if key not in _ctxs:
        _ctxs[key] = _b2g.Context(*key)
    return _ctxs[key]


def _nbytes_of(c):
    return max(int.from_bytes(c[4:8].tobytes(), "little", signed=True), 0) if c.size >= 16 else 0


def decompress_batch(chunks, caps=None, srcsizes=None, context=None, device=False):
    """-> list of (status, bytes).  device=True runs the device-resident entry points on torch buffers."""
    n = len(chunks)
    chunks = [np.ascontiguousarray(c) for c in chunks]
    if caps is None:
        caps = [_nbytes_of(c) for c in chunks]
    if srcsizes is None:
        srcsizes = [int(c.size) for c in chunks]
    cx = context or ctx(codec=5)   # a zstd-capable context decodes everything
    if not device:
        outs = [np.full(max(cap, 1) + 64, 0xAB, dtype=np.uint8) for cap in caps]
        b = cx.make_batch([c.ctypes.data for c in chunks], [o.ctypes.data for o in outs], srcsizes, caps)
        cx.decompress_batch(b)
        res = []
        for i in range(n):
            assert np.all(outs[i][max(caps[i], 1):] == 0xAB), "decoder wrote past the destination capacity"
            res.append((b[i].status, outs[i][:b[i].out_bytes].copy()))
        return res
    import torch
    d_in = [torch.from_numpy(np.concatenate([c, np.zeros(16, np.uint8)])).cuda() for c in chunks]
    d_out = [torch.full((max(cap, 1) + 64,), 0xAB, dtype=torch.uint8, device="cuda") for cap in caps]
    torch.cuda.synchronize()
    b = cx.make_batch([t.data_ptr() for t in d_in], [t.data_ptr() for t in d_out], srcsizes, caps)
    cx.decompress_batch_device(b)
    cx.wait()
    res = []
    for i in range(n):
        o = d_out[i].cpu().numpy()
        assert np.all(o[max(caps[i], 1):] == 0xAB), "decoder wrote past the destination capacity"
        res.append((b[i].status, o[:b[i].out_bytes].copy()))
    return res


def compress_batch(srcs, codec, clevel, typesize, filt, blocksize, caps=None, device=False):
    """-> list of (status, chunk bytes)."""
    n = len(srcs)
    srcs = [np.ascontiguousarray(s).view(np.uint8).reshape(-1) for s in srcs]
    if caps is None:
        caps = [int(s.size) + 32 for s in srcs]
    cx = ctx(codec, clevel, typesize, filt, blocksize)
    nbytes = [int(s.size) for s in srcs]
    if not device:
        dsts = [np.full(max(cap, 1) + 64, 0xAB, dtype=np.uint8) for cap in caps]
        b = cx.make_batch([s.ctypes.data for s in srcs], [d.ctypes.data for d in dsts], nbytes, caps)
        cx.compress_batch(b)
        res = []
        for i in range(n):
            assert np.all(dsts[i][max(caps[i], 1):] == 0xAB), "encoder wrote past the destination capacity"
            res.append((b[i].status, dsts[i][:b[i].out_bytes].copy()))
        return res
    import torch
    d_in = [torch.from_numpy(np.concatenate([s, np.zeros(16, np.uint8)])).cuda() for s in srcs]
    d_out = [torch.full((max(cap, 1) + 64,), 0xAB, dtype=torch.uint8, device="cuda") for cap in caps]
    torch.cuda.synchronize()
    b = cx.make_batch([t.data_ptr() for t in d_in], [t.data_ptr() for t in d_out], nbytes, caps)
    cx.compress_batch_device(b)
    cx.wait()
    res = []
    for i in range(n):
        o = d_out[i].cpu().numpy()
        assert np.all(o[max(caps[i], 1):] == 0xAB), "encoder wrote past the destination capacity"
        res.append((b[i].status, o[:b[i].out_bytes].copy()))
    return res


def run_filter(filt, forward, typesize, src, blocksize, version=5):
    import torch
    src = np.ascontiguousarray(src)
    d_src = torch.from_numpy(src).cuda()
    d_dst = torch.full((src.size + 16,), 0xAB, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    cx = ctx()
    cx.filter_device(filt, forward, typesize, d_src.data_ptr(), d_dst.data_ptr(), int(src.size), blocksize, version)
    cx.wait()
    torch.cuda.synchronize()
    import ctypes
    # filter_device enqueues on the context's own stream: synchronise the whole device before reading back
    o = d_dst.cpu().numpy()
    assert np.all(o[src.size:] == 0xAB)
    return o[:src.size]
