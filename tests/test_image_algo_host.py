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
