"""ctypes bindings for the CHECKERS used by the test-suite (test infrastructure only):

* ``oracle()``  -> oracle/liboracle.so, our plain-C restatement of the reference path (oracle/b2o.h)
* ``ref()``     -> oracle/_ref/libblosc2_ref.so, the unmodified reference c-blosc2 compiled from
                   /root/reference by oracle/Makefile, behind oracle/ref_shim.c (None when not built)

Nothing in the product package imports this module.
"""
import ctypes as C
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

vp, i32, ci, cl = C.c_void_p, C.c_int32, C.c_int, C.c_long
pi32 = C.POINTER(C.c_int32)


def _build():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True, capture_output=True)


_oracle = None
_ref = False


def oracle():
    global _oracle
    if _oracle is None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            _build()
        L = C.CDLL(path)
        L.b2o_compress.argtypes = [ci, ci, ci, ci, i32, vp, i32, vp, i32]
        L.b2o_decompress.argtypes = [vp, i32, vp, i32]
        L.b2o_chunk_sizes.argtypes = [vp, pi32, pi32, pi32]
        for f in ("b2o_shuffle", "b2o_unshuffle", "b2o_bitshuffle"):
            getattr(L, f).argtypes = [ci, i32, vp, vp]
        L.b2o_bitunshuffle.argtypes = [ci, i32, vp, vp, ci]
        L.b2o_shuffle.restype = None
        L.b2o_unshuffle.restype = None
        L.b2o_lz4_compress.argtypes = [vp, ci, vp, ci]
        L.b2o_lz4_decompress.argtypes = [vp, ci, vp, ci]
        L.b2o_blosclz_compress.argtypes = [ci, vp, ci, vp, ci]
        L.b2o_blosclz_decompress.argtypes = [vp, ci, vp, ci]
        L.b2o_zstd_decompress.argtypes = [vp, cl, vp, cl]
        L.b2o_zstd_decompress.restype = cl
        _oracle = L
    return _oracle


def ref():
    """The reference library, or None if oracle/_ref was never built (it cannot be built on the GPU box)."""
    global _ref
    if _ref is False:
        path = os.path.join(ORACLE_DIR, "_ref", "libblosc2_ref.so")
        if not os.path.exists(path) and os.path.isdir("/root/reference"):
            _build()
        if not os.path.exists(path):
            _ref = None
            return None
        L = C.CDLL(path)
        L.ref_compress.argtypes = [ci, ci, ci, ci, ci, ci, vp, i32, vp, i32]
        L.ref_decompress.argtypes = [vp, i32, vp, i32, ci]
        L.ref_cbuffer_sizes.argtypes = [vp, pi32, pi32, pi32]
        pvp = C.POINTER(C.c_void_p)
        L.ref_compress_batch.argtypes = [ci, ci, ci, ci, ci, ci, pvp, pi32, pvp, pi32, pi32, ci]
        L.ref_decompress_batch.argtypes = [ci, pvp, pi32, pvp, pi32, pi32, ci]
        L.ref_lz4_compress.argtypes = [vp, ci, vp, ci]
        L.ref_lz4hc_compress.argtypes = [vp, ci, vp, ci, ci]
        L.ref_lz4_decompress.argtypes = [vp, ci, vp, ci]
        L.ref_blosclz_compress.argtypes = [ci, ci, ci, vp, ci, vp, ci]
        L.ref_blosclz_decompress.argtypes = [vp, ci, vp, ci]
        L.ref_zstd_compress.argtypes = [vp, cl, vp, cl, ci]
        L.ref_zstd_compress.restype = cl
        L.ref_zstd_decompress.argtypes = [vp, cl, vp, cl]
        L.ref_zstd_decompress.restype = cl
        for f in ("ref_shuffle", "ref_unshuffle", "ref_bitshuffle"):
            getattr(L, f).argtypes = [ci, ci, vp, vp]
        L.ref_bitunshuffle.argtypes = [ci, ci, vp, vp, ci]
        _ref = L
    return _ref


def ptr(a):
    """Address of a numpy array's buffer (0-size arrays get a valid dummy address)."""
    return a.ctypes.data


# ---- numpy-level helpers over the two checkers -------------------------------------------------------------
import numpy as np  # noqa: E402


def ref_sizes(chunk):
    R = ref()
    nb, cb, bs = i32(), i32(), i32()
    R.ref_cbuffer_sizes(ptr(chunk), C.byref(nb), C.byref(cb), C.byref(bs))
    return nb.value, cb.value, bs.value


def ref_compress(codec, clevel, typesize, filt, blocksize, src, nthreads=1):
    R = ref()
    src = np.ascontiguousarray(src).view(np.uint8).reshape(-1)
    dst = np.zeros(src.size + 32, dtype=np.uint8)
    r = R.ref_compress(codec, clevel, typesize, filt, blocksize, nthreads, ptr(src), int(src.size), ptr(dst),
                       int(dst.size))
    assert r > 0, r
    return dst[:r].copy()


def ref_decompress(chunk, cap=None):
    """-> (rc, bytes).  rc < 0 is the reference's error code."""
    R = ref()
    chunk = np.ascontiguousarray(chunk)
    nb = int.from_bytes(chunk[4:8].tobytes(), "little", signed=True) if cap is None else cap
    out = np.zeros(max(nb, 1), dtype=np.uint8)
    r = R.ref_decompress(ptr(chunk), int(chunk.size), ptr(out), int(nb), 1)
    return r, out[:max(r, 0)]


def oracle_compress(codec, clevel, typesize, filt, blocksize, src, cap=None):
    O = oracle()
    src = np.ascontiguousarray(src).view(np.uint8).reshape(-1)
    dst = np.zeros((src.size + 32) if cap is None else cap, dtype=np.uint8)
    r = O.b2o_compress(codec, clevel, typesize, filt, blocksize, ptr(src), int(src.size), ptr(dst), int(dst.size))
    return r, dst[:max(r, 0)].copy()


def oracle_decompress(chunk, cap=None):
    O = oracle()
    chunk = np.ascontiguousarray(chunk)
    nb = int.from_bytes(chunk[4:8].tobytes(), "little", signed=True) if cap is None else cap
    out = np.zeros(max(nb, 1), dtype=np.uint8)
    r = O.b2o_decompress(ptr(chunk), int(chunk.size), ptr(out), int(nb))
    return r, out[:max(r, 0)]
