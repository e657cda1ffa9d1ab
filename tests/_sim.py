"""ctypes bindings for tests/sim/libb2gsim.so: the product's device code compiled for the CPU fiber simulator
(tests/sim/cusim.h).  Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SIM_DIR = os.path.join(HERE, "sim")
vp, i32, ci, i64 = C.c_void_p, C.c_int32, C.c_int, C.c_int64
pi32 = C.POINTER(C.c_int32)
pvp = C.POINTER(C.c_void_p)
_lib = None


def lib():
    global _lib
    if _lib is None:
        subprocess.run(["make", "-s", "-C", SIM_DIR], check=True, capture_output=True)
        L = C.CDLL(os.path.join(SIM_DIR, "libb2gsim.so"))
        L.sim_decompress_batch.argtypes = [ci, pvp, pi32, pvp, pi32, pi32, pi32, ci, ci, ci, ci, ci]
        L.sim_compress_batch.argtypes = [ci, ci, ci, ci, ci, ci, pvp, pi32, pvp, pi32, pi32, pi32, ci, ci, ci]
        L.sim_filter.argtypes = [ci, ci, ci, ci, vp, vp, i64, ci, ci]
        L.sim_stream_decode.argtypes = [ci, vp, ci, vp, ci]
        L.sim_stream_encode.argtypes = [ci, vp, ci, vp, ci, ci]
        L.sim_collectives.restype = C.c_ulonglong
        L.sim_last_paired.restype = C.c_int
        _lib = L
    return _lib


def decode_stats(reset=True):
    """-> dict of the batched-decode counters since the last reset."""
    out = (C.c_ulonglong * 5)()
    lib().sim_decode_stats(out, 1 if reset else 0)
    return dict(zip(("batches", "batch_seqs", "lane_matches", "ordered_matches", "general_seqs"), [int(v) for v in out]))


def _ptrs(arrs):
    return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


def decompress_batch(chunks, caps=None, win_bytes=33280, max_block=None, grid=2, zstd_ok=True, srcsizes=None,
                     pair=1):
    """-> list of (status, bytes)."""
    n = len(chunks)
    chunks = [np.ascontiguousarray(c) for c in chunks]
    if caps is None:
        caps = [max(int.from_bytes(c[4:8].tobytes(), "little", signed=True), 0) if c.size >= 16 else 0 for c in chunks]
    outs = [np.full(max(cap, 1) + 64, 0xAB, dtype=np.uint8) for cap in caps]
    if srcsizes is None:
        srcsizes = [int(c.size) for c in chunks]
    if max_block is None:
        max_block = 1
        for c in chunks:
            if c.size >= 16:
                max_block = max(max_block, int.from_bytes(c[8:12].tobytes(), "little", signed=True))
        max_block = max(min(max_block, 1 << 26), win_bytes)
    out = (C.c_int32 * n)()
    st = (C.c_int32 * n)()
    rc = lib().sim_decompress_batch(n, _ptrs(chunks), (C.c_int32 * n)(*srcsizes), _ptrs(outs), (C.c_int32 * n)(*caps),
                                    out, st, win_bytes, max_block, grid, 1 if zstd_ok else 0, pair)
    assert rc == 0 or any(s < 0 for s in st), "a plane staging slot was leaked"

    res = []
    for i in range(n):
        assert np.all(outs[i][max(caps[i], 1):] == 0xAB), "decoder wrote past the destination capacity"
        res.append((st[i], outs[i][:out[i]].copy()))
    return res


def compress_batch(srcs, codec, clevel, typesize, filt, blocksize, caps=None, win_bytes=32768, grid=2, hlog=0):
    """-> list of (status, chunk bytes)."""
    n = len(srcs)
    srcs = [np.ascontiguousarray(s).view(np.uint8).reshape(-1) for s in srcs]
    if caps is None:
        caps = [int(s.size) + 32 for s in srcs]
    dsts = [np.full(max(cap, 1) + 64, 0xAB, dtype=np.uint8) for cap in caps]
    nbytes = [int(s.size) for s in srcs]
    out = (C.c_int32 * n)()
    st = (C.c_int32 * n)()
    lib().sim_compress_batch(codec, clevel, typesize, filt, blocksize, n, _ptrs(srcs), (C.c_int32 * n)(*nbytes),
                             _ptrs(dsts), (C.c_int32 * n)(*caps), out, st, win_bytes, grid, hlog)
    res = []
    for i in range(n):
        assert np.all(dsts[i][max(caps[i], 1):] == 0xAB), "encoder wrote past the destination capacity"
        res.append((st[i], dsts[i][:out[i]].copy()))
    return res


def set_team(warps=-1, prime=-1, hlog=0, sparse=-1):
    """Encoder teams in sim_compress_batch: warps -1 = as the library decides (plan_team_warps), 0 / 1 = one warp per
    stream (k_compress), N > 1 = k_compress_team with N warps; prime -1 = the library's default; sparse 1 / 0 = prime with
    every fourth position / every position, -1 = as the library decides (plan_team_prime_step)."""
    L = lib()
    L.sim_set_team.argtypes = [C.c_int] * 3
    L.sim_set_team.restype = None
    L.sim_set_team(warps, prime, hlog)
    L.sim_set_team_sparse.argtypes = [C.c_int]
    L.sim_set_team_sparse.restype = None
    L.sim_set_team_sparse(sparse)


def run_filter(filt, forward, typesize, src, blocksize, version=5, grid=2):
    src = np.ascontiguousarray(src)
    dst = np.zeros(src.size + 16, dtype=np.uint8)
    dst[src.size:] = 0xAB
    lib().sim_filter(filt, 1 if forward else 0, typesize, version, src.ctypes.data, dst.ctypes.data, int(src.size),
                     blocksize, grid)
    assert np.all(dst[src.size:] == 0xAB)
    return dst[:src.size]


def stream_decode(fmt, payload, cap):
    payload = np.ascontiguousarray(payload)
    out = np.full(cap + 64, 0xAB, dtype=np.uint8)
    r = lib().sim_stream_decode(fmt, payload.ctypes.data, int(payload.size), out.ctypes.data, cap)
    assert np.all(out[cap:] == 0xAB), "stream decoder wrote past the capacity"
    return r, out[:max(r, 0)]


def stream_encode(codec, src, cap, hlog=12):
    src = np.ascontiguousarray(src)
    out = np.full(cap + 64, 0xAB, dtype=np.uint8)
    r = lib().sim_stream_encode(codec, src.ctypes.data, int(src.size), out.ctypes.data, cap, hlog)
    assert np.all(out[cap:] == 0xAB), "stream encoder wrote past the capacity"
    return r, out[:max(r, 0)]
