"""Generates tests/golden/ref_cframe.npz: contiguous frames written by the UNMODIFIED reference library
(oracle/_ref/libblosc2_ref.so: to_schunk() as compressed/blosc2/schunk.h:107-121 + blosc2_schunk_to_buffer), together
with the raw data and the chunks that went in.  Run in the dev container (needs /root/reference to build oracle/_ref):
    python tests/golden/make_cframe_golden.py
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _libs  # noqa: E402


def ref_cframe(chunks):
    R = _libs.ref()
    R.ref_chunks_to_cframe.restype = C.c_int64
    R.ref_chunks_to_cframe.argtypes = [C.c_int, C.POINTER(C.c_void_p), C.c_void_p, C.c_int64]
    n = len(chunks)
    dst = np.empty(sum(c.size for c in chunks) + 4096, np.uint8)
    ln = R.ref_chunks_to_cframe(n, (C.c_void_p * max(n, 1))(*[c.ctypes.data for c in chunks]), dst.ctypes.data, dst.size)
    assert ln > 0, ln
    return dst[:ln].copy()


def main():
    rng = np.random.default_rng(20261104)
    out = {}
    cases = {"u16_lz4_5chunks": (1, 2, [6000] * 4 + [2345]), "f32_blosclz_1chunk": (0, 4, [3000]),
             "u8_lz4_12chunks": (1, 1, [4096] * 12)}
    for name, (codec, ts, sizes) in cases.items():
        dt = {1: np.uint8, 2: np.uint16, 4: np.uint32}[ts]
        raws = [(np.cumsum(rng.integers(0, 3, size=s)) & 0x3ff).astype(dt) for s in sizes]
        chunks = [_libs.ref_compress(codec, 9, ts, 1, 32768, r) for r in raws]
        out[name + "/frame"] = ref_cframe(chunks)
        out[name + "/raw"] = np.concatenate([r.view(np.uint8) for r in raws])
        out[name + "/cbytes"] = np.array([c.size for c in chunks], np.int64)
        out[name + "/nbytes"] = np.array([r.nbytes for r in raws], np.int64)
        out[name + "/chunks"] = np.concatenate(chunks)
    out["empty/frame"] = ref_cframe([])
    np.savez_compressed(os.path.join(HERE, "ref_cframe.npz"), **out)
    print({k: v.shape for k, v in out.items() if k.endswith("frame")})


if __name__ == "__main__":
    main()
