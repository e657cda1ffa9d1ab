"""interleave / deinterleave of planar channels <-> pixel-interleaved buffers on the GPU.

Mirror of compressed::image_algo (compressed_image/include/compressed/image_algo.h:34-65 interleave, :84-111 and
:131-160 deinterleave; used by image<T>::read at image.h:1880): same argument meaning, same checks, and ValueError where
the reference throws std::invalid_argument (what pybind11 maps it to).  The work is done by k_interleave_vec /
k_deinterleave_vec / the element kernels of csrc/b2g_interleave.cuh through b2g_interleave_device /
b2g_deinterleave_device; there is no CPU path.

Host arrays (numpy) go through b2g_interleave / b2g_deinterleave (staged to the device and back by the library);
`interleave_device` / `deinterleave_device` take torch CUDA tensors and move nothing.
"""
import numpy as np

from . import _b2g

_ctx = None


def _context():
    global _ctx
    if _ctx is None:
        _ctx = _b2g.Context()
    return _ctx


def _check_interleave(buffer_size, span_sizes):
    if len(span_sizes) == 0:
        raise ValueError("Interleave: No spans provided for interleaving.")
    if any(s != span_sizes[0] for s in span_sizes):
        raise ValueError("Interleave: All input spans must have the same size.")
    if buffer_size != span_sizes[0] * len(span_sizes):
        raise ValueError("Interleave: Provided buffer is not large enough to hold all the elements to interleave.")


def _check_deinterleave(interleaved_size, span_sizes):
    if len(span_sizes) == 0:
        raise ValueError("Deinterleave: No spans provided for deinterleaving.")
    if any(s != span_sizes[0] for s in span_sizes):
        raise ValueError("Deinterleave: All output spans must have the same size.")
    if interleaved_size != span_sizes[0] * len(span_sizes):
        raise ValueError("Deinterleave: Input buffer size does not match the expected size for deinterleaving.")


def interleave_device(buffer, spans, stream=None):
    """buffer[i * n + c] = spans[c][i]; torch CUDA tensors of one dtype, contiguous.  Enqueued on `stream`
    (a raw cudaStream_t; None = the context's own stream), not waited for."""
    _check_interleave(buffer.numel(), [s.numel() for s in spans])
    if any(s.dtype != buffer.dtype for s in spans):
        raise ValueError("Interleave: All spans must have the dtype of the buffer.")
    _context().interleave_device(buffer.element_size(), [s.data_ptr() for s in spans], buffer.data_ptr(),
                                 spans[0].numel(), stream)


def deinterleave_device(interleaved, channel_spans, stream=None):
    """channel_spans[c][i] = interleaved[i * n + c]; torch CUDA tensors of one dtype, contiguous."""
    _check_deinterleave(interleaved.numel(), [s.numel() for s in channel_spans])
    if any(s.dtype != interleaved.dtype for s in channel_spans):
        raise ValueError("Deinterleave: All spans must have the dtype of the buffer.")
    _context().deinterleave_device(interleaved.element_size(), interleaved.data_ptr(),
                                   [s.data_ptr() for s in channel_spans], channel_spans[0].numel(), stream)


def interleave(buffer, spans):
    """Host version: `buffer` (preallocated numpy array, exactly len(spans) * spans[0].size elements) receives the
    interleaved `spans` (numpy arrays of the same dtype and size)."""
    spans = [np.ascontiguousarray(s).reshape(-1) for s in spans]
    _check_interleave(buffer.size, [s.size for s in spans])
    if any(s.dtype != buffer.dtype for s in spans):
        raise ValueError("Interleave: All spans must have the dtype of the buffer.")
    if not buffer.flags.c_contiguous:
        raise ValueError("Interleave: the buffer must be contiguous.")
    if spans[0].size:
        _context().interleave(buffer.dtype.itemsize, [s.ctypes.data for s in spans], buffer.ctypes.data, spans[0].size)
    return buffer


def deinterleave(interleaved, channel_spans):
    """Host version: `channel_spans` (preallocated numpy arrays of equal size) receive the channels of
    `interleaved`; with an int instead of a list, that many new arrays are returned (image_algo.h:131-160)."""
    interleaved = np.ascontiguousarray(interleaved).reshape(-1)
    if isinstance(channel_spans, (int, np.integer)):
        if channel_spans <= 0 or interleaved.size % int(channel_spans):
            raise ValueError("Deinterleave: Input buffer size does not match the expected size for deinterleaving.")
        channel_spans = [np.empty(interleaved.size // int(channel_spans), dtype=interleaved.dtype)
                         for _ in range(int(channel_spans))]
    _check_deinterleave(interleaved.size, [s.size for s in channel_spans])
    if any(s.dtype != interleaved.dtype or not s.flags.c_contiguous for s in channel_spans):
        raise ValueError("Deinterleave: All spans must be contiguous and have the dtype of the buffer.")
    if channel_spans[0].size:
        _context().deinterleave(interleaved.dtype.itemsize, interleaved.ctypes.data,
                                [s.ctypes.data for s in channel_spans], channel_spans[0].size)
    return channel_spans
