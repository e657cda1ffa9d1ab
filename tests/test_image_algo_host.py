"""Host-side argument checks of compressed_image_b200.image_algo (no GPU needed: they fire before any device call),
matching the std::invalid_argument cases of image_algo.h:37-53 and :88-98."""
import numpy as np
import pytest

from compressed_image_b200 import image_algo


def test_interleave_argument_errors():
    a = np.zeros(8, np.uint8)
    with pytest.raises(ValueError, match="No spans"):
        image_algo.interleave(np.zeros(0, np.uint8), [])
    with pytest.raises(ValueError, match="same size"):
        image_algo.interleave(np.zeros(15, np.uint8), [a, a[:7]])
    with pytest.raises(ValueError, match="not large enough"):
        image_algo.interleave(np.zeros(15, np.uint8), [a, a])
    with pytest.raises(ValueError, match="dtype"):
        image_algo.interleave(np.zeros(16, np.uint16), [a, a])


def test_deinterleave_argument_errors():
    a = np.zeros(8, np.uint8)
    with pytest.raises(ValueError, match="same size"):
        image_algo.deinterleave(np.zeros(16, np.uint8), [a.copy(), a[:7].copy()])
    with pytest.raises(ValueError, match="does not match"):
        image_algo.deinterleave(np.zeros(17, np.uint8), [a.copy(), a.copy()])
    with pytest.raises(ValueError, match="does not match"):
        image_algo.deinterleave(np.zeros(17, np.uint8), 2)
    with pytest.raises(ValueError, match="No spans"):
        image_algo.deinterleave(np.zeros(0, np.uint8), [])


def test_empty_spans_are_a_no_op():
    buf = np.zeros(0, np.float32)
    assert image_algo.interleave(buf, [np.zeros(0, np.float32)] * 3) is buf
    outs = image_algo.deinterleave(buf, [np.zeros(0, np.float32)] * 3)
    assert len(outs) == 3


def _ref_image_algo():
    """oracle/_ref/libimage_algo_ref.so = the reference's own header-only image_algo.h instantiated by oracle/ref_image_algo.cpp
    (built by oracle/Makefile when /root/reference is present), or None."""
    import ctypes as C
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    path = os.path.join(root, "oracle", "_ref", "libimage_algo_ref.so")
    if not os.path.exists(path) and os.path.isdir("/root/reference"):
        subprocess.run(["make", "-s", "-C", os.path.join(root, "oracle")], check=False, capture_output=True)
    if not os.path.exists(path):
        return None
    L = C.CDLL(path)
    L.ref_interleave.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_void_p), C.c_void_p, C.c_int64]
    L.ref_deinterleave.argtypes = [C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_void_p), C.c_int64]
    return L


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32, np.float64])
@pytest.mark.parametrize("nch", [1, 2, 3, 4, 7, 16])
def test_numpy_restatement_is_what_the_reference_computes(dtype, nch):
    """The checker of tests/test_gpu_interleave.py -- np.stack(planes, axis=-1).reshape(-1) -- against the reference's own
    compressed::image_algo::interleave / deinterleave (image_algo.h:34-65, :84-111), both directions."""
    import ctypes as C
    L = _ref_image_algo()
    if L is None:
        pytest.skip("oracle/_ref/libimage_algo_ref.so not built (no reference tree on this machine)")
    n = 1000 + nch
    rng = np.random.default_rng(nch * 17 + np.dtype(dtype).itemsize)
    planes = [rng.integers(0, 256, size=n * np.dtype(dtype).itemsize, dtype=np.uint8).view(dtype).copy() for _ in range(nch)]
    expect = np.stack(planes, axis=-1).reshape(-1)
    out = np.zeros(n * nch, dtype=dtype)
    ptrs = (C.c_void_p * nch)(*[p.ctypes.data for p in planes])
    assert L.ref_interleave(np.dtype(dtype).itemsize, nch, ptrs, out.ctypes.data, n) == 0
    assert np.array_equal(out.view(np.uint8), expect.view(np.uint8))
    back = [np.zeros(n, dtype=dtype) for _ in range(nch)]
    bptrs = (C.c_void_p * nch)(*[b.ctypes.data for b in back])
    assert L.ref_deinterleave(np.dtype(dtype).itemsize, nch, expect.ctypes.data, bptrs, n) == 0
    for b, p in zip(back, planes):
        assert np.array_equal(b.view(np.uint8), p.view(np.uint8))
