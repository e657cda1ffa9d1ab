"""interleave / deinterleave of planar channels <-> pixel-interleaved buffers on the GPU.

Mirror of compressed::image_algo (compressed_image/include/compressed/image_algo.h:34-65 interleave, :84-111 and
:131-160 deinterleave; used by image<T>::read at image.h:1880): same argument meaning, same checks, and ValueError where
the reference throws std::invalid_argument (what pybind11 maps it to).  The work is done by k_interleave_vec /
k_deinterleave_vec / the element kernels of csrc/b2g_interleave.cuh through b2g_interleave_device /
b2g_deinterleave_device; there is no CPU path.

Host arrays (numpy) are copied to the device and back; `interleave_device` / `deinterleave_device` take torch CUDA
tensors and move nothing.
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


def _as_bytes_tensor(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a).reshape(-1).view(np.uint8))


def interleave(buffer, spans):
    """Host version: `buffer` (preallocated numpy array, exactly len(spans) * spans[0].size elements) receives the
    interleaved `spans` (numpy arrays of the same dtype and size)."""
    import torch
    spans = [np.asarray(s) for s in spans]
    _check_interleave(buffer.size, [s.size for s in spans])
    if any(s.dtype != buffer.dtype for s in spans):
        raise ValueError("Interleave: All spans must have the dtype of the buffer.")
    if not buffer.flags.c_contiguous:
        raise ValueError("Interleave: the buffer must be contiguous.")
    ts, n = buffer.dtype.itemsize, spans[0].size
    if n == 0:
        return buffer
    ctx = _context()
    d_planes = [_as_bytes_tensor(s).cuda() for s in spans]
    d_out = torch.empty(n * len(spans) * ts, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    ctx.interleave_device(ts, [p.data_ptr() for p in d_planes], d_out.data_ptr(), n)
    torch.cuda.synchronize()
    buffer.reshape(-1).view(np.uint8)[:] = d_out.cpu().numpy()
    return buffer


def deinterleave(interleaved, channel_spans):
    """Host version: `channel_spans` (preallocated numpy arrays of equal size) receive the channels of
    `interleaved`; with an int instead of a list, that many new arrays are returned (image_algo.h:131-160)."""
    import torch
    interleaved = np.asarray(interleaved)
    if isinstance(channel_spans, (int, np.integer)):
        if channel_spans <= 0 or interleaved.size % int(channel_spans):
            raise ValueError("Deinterleave: Input buffer size does not match the expected size for deinterleaving.")
        channel_spans = [np.empty(interleaved.size // int(channel_spans), dtype=interleaved.dtype)
                         for _ in range(int(channel_spans))]
    _check_deinterleave(interleaved.size, [s.size for s in channel_spans])
    if any(s.dtype != interleaved.dtype or not s.flags.c_contiguous for s in channel_spans):
        raise ValueError("Deinterleave: All spans must be contiguous and have the dtype of the buffer.")
    ts, n = interleaved.dtype.itemsize, channel_spans[0].size
    if n == 0:
        return channel_spans
    ctx = _context()
    d_in = _as_bytes_tensor(interleaved).cuda()
    d_planes = [torch.empty(n * ts, dtype=torch.uint8, device="cuda") for _ in channel_spans]
    torch.cuda.synchronize()
    ctx.deinterleave_device(ts, d_in.data_ptr(), [p.data_ptr() for p in d_planes], n)
    torch.cuda.synchronize()
    for s, p in zip(channel_spans, d_planes):
        s.reshape(-1).view(np.uint8)[:] = p.cpu().numpy()
    return channel_spans
