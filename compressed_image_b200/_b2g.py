"""ctypes binding of the C ABI in include/b2g.h (compressed_image_b200/libb2g.so).

This is the only way Python reaches the engine: every (de)compression below runs the sm_100a kernels of
csrc/.  There is no CPU fallback -- a missing library or a box without a CUDA device raises.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb2g.so")

BLOSCLZ, LZ4, LZ4HC, ZSTD = 0, 1, 2, 5
NOFILTER, SHUFFLE, BITSHUFFLE = 0, 1, 2
MAX_OVERHEAD = 32
NOALLOC_MAX_CHUNKS, NOALLOC_MAX_BLOCKSIZE = 1024, 65536   # b2g.h: limits of b2g_decompress_batch_device_noalloc
INT32_MAX = 2**31 - 1


class Params(C.Structure):
    _fields_ = [("codec", C.c_uint8), ("clevel", C.c_uint8), ("typesize", C.c_uint8), ("filter", C.c_uint8),
                ("blocksize", C.c_int32), ("device", C.c_int32)]


class Chunk(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("src_bytes", C.c_int32), ("dst_capacity", C.c_int32),
                ("out_bytes", C.c_int32), ("status", C.c_int32)]


class B2GError(RuntimeError):
    """A negative b2g / blosc2 status (the reference throws std::runtime_error: wrapper.h:146-149, 254-257)."""

    def __init__(self, code, what=""):
        self.code = code
        msg = lib().b2g_strerror(code).decode() if _lib is not None else "error"
        extra = lib().b2g_last_error().decode() if (_lib is not None and code <= -100) else ""
        super().__init__("b2g: %s%s (code %d)%s" % (what + ": " if what else "", msg, code, " [" + extra + "]" if extra else ""))


_lib = None


def lib():
    """The loaded libb2g.so.  Raises if it was never built: the product has no other engine."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("compressed_image_b200: %s is missing -- build it with `make` (or "
                              "__graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        pc, vp, ci = C.POINTER(Chunk), C.c_void_p, C.c_int
        L.b2g_create.argtypes = [C.POINTER(Params)]
        L.b2g_create.restype = vp
        L.b2g_destroy.argtypes = [vp]
        L.b2g_destroy.restype = None
        for f in ("b2g_compress_batch", "b2g_decompress_batch", "b2g_compress_batch_device",
                  "b2g_decompress_batch_device", "b2g_compress_batch_async", "b2g_decompress_batch_async"):
            getattr(L, f).argtypes = [vp, pc, ci, vp]
            getattr(L, f).restype = ci
        L.b2g_wait.argtypes = [vp]
        L.b2g_decompress_batch_device_noalloc.argtypes = [ci, pc, ci, vp]
        L.b2g_decompress_batch_device_noalloc.restype = ci
        L.b2g_chunk_sizes.argtypes = [vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
        L.b2g_filter_device.argtypes = [vp, ci, ci, ci, ci, vp, vp, C.c_int64, C.c_int32, vp]
        L.b2g_interleave_device.argtypes = [vp, ci, ci, C.POINTER(vp), vp, C.c_int64, vp]
        L.b2g_deinterleave_device.argtypes = [vp, ci, ci, vp, C.POINTER(vp), C.c_int64, vp]
        L.b2g_interleave_device.restype = ci
        L.b2g_interleave.argtypes = [vp, ci, ci, C.POINTER(vp), vp, C.c_int64]
        L.b2g_deinterleave.argtypes = [vp, ci, ci, vp, C.POINTER(vp), C.c_int64]
        L.b2g_interleave.restype = ci
        L.b2g_deinterleave.restype = ci
        L.b2g_deinterleave_device.restype = ci
        L.b2g_host_alloc.argtypes = [C.c_size_t]
        L.b2g_host_alloc.restype = vp
        L.b2g_host_free.argtypes = [vp]
        L.b2g_host_free.restype = None
        L.b2g_launch_count.argtypes = [vp]
        L.b2g_launch_count.restype = C.c_int64
        L.b2g_last_kernel_ms.argtypes = [vp, C.POINTER(C.c_float)]
        L.b2g_strerror.argtypes = [ci]
        L.b2g_strerror.restype = C.c_char_p
        L.b2g_last_error.restype = C.c_char_p
        L.b2g_device_count.restype = ci
        L.b2g_version.restype = C.c_char_p
        _lib = L
    return _lib


EXPORTS = ("b2g_create", "b2g_destroy", "b2g_compress_batch", "b2g_decompress_batch", "b2g_compress_batch_async",
           "b2g_decompress_batch_async", "b2g_compress_batch_device",
           "b2g_decompress_batch_device", "b2g_decompress_batch_device_noalloc", "b2g_wait", "b2g_chunk_sizes", "b2g_filter_device", "b2g_interleave_device",
           "b2g_deinterleave_device", "b2g_interleave", "b2g_deinterleave", "b2g_host_alloc", "b2g_host_free", "b2g_clone", "b2g_get_params", "b2g_launch_count",
           "b2g_last_kernel_ms", "b2g_strerror", "b2g_last_error", "b2g_device_count", "b2g_version")


def chunk_sizes(chunk):
    """(nbytes, cbytes, blocksize) of a chunk in host memory (numpy uint8 array / bytes-like, >= 16 bytes)."""
    import numpy as np
    a = np.frombuffer(chunk, dtype=np.uint8, count=16) if not isinstance(chunk, np.ndarray) else chunk
    if a.size < 16:
        raise B2GError(-5, "chunk_sizes")
    nb, cb, bs = C.c_int32(), C.c_int32(), C.c_int32()
    rc = lib().b2g_chunk_sizes(a.ctypes.data, C.byref(nb), C.byref(cb), C.byref(bs))
    if rc < 0:
        raise B2GError(rc, "chunk_sizes")
    return nb.value, cb.value, bs.value


class Context:
    """One b2g_ctx: replaces the pair of blosc2 contexts a compressed::channel owns (channel.h:575-576)."""

    def __init__(self, codec=LZ4, clevel=9, typesize=1, filter=SHUFFLE, blocksize=32768, device=-1):
        p = Params(codec, clevel, typesize, filter, blocksize, device)
        self.params = p
        self._h = lib().b2g_create(C.byref(p))
        if not self._h:
            raise B2GError(-100, "b2g_create")

    def close(self):
        if getattr(self, "_h", None):
            lib().b2g_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- batch calls ---------------------------------------------------------------------------------------
    @staticmethod
    def make_batch(srcs, dsts, src_bytes, dst_caps):
        n = len(srcs)
        arr = (Chunk * n)()
        for i in range(n):
            arr[i].src = srcs[i]
            arr[i].dst = dsts[i]
            arr[i].src_bytes = src_bytes[i]
            arr[i].dst_capacity = dst_caps[i]
        return arr

    def _call(self, fn, batch, n, stream):
        rc = fn(self._h, batch, n, stream)
        if rc < 0:
            raise B2GError(rc, fn.__name__)

    def compress_batch(self, batch, n=None, stream=None):
        self._call(lib().b2g_compress_batch, batch, len(batch) if n is None else n, stream)

    def decompress_batch(self, batch, n=None, stream=None):
        self._call(lib().b2g_decompress_batch, batch, len(batch) if n is None else n, stream)

    def compress_batch_async(self, batch, n=None):
        """Host pointers, enqueued only: `batch` and its buffers must stay alive until wait()."""
        self._call(lib().b2g_compress_batch_async, batch, len(batch) if n is None else n, None)

    def decompress_batch_async(self, batch, n=None):
        self._call(lib().b2g_decompress_batch_async, batch, len(batch) if n is None else n, None)

    def compress_batch_device(self, batch, n=None, stream=None):
        self._call(lib().b2g_compress_batch_device, batch, len(batch) if n is None else n, stream)

    def decompress_batch_device(self, batch, n=None, stream=None):
        self._call(lib().b2g_decompress_batch_device, batch, len(batch) if n is None else n, stream)

    def wait(self):
        rc = lib().b2g_wait(self._h)
        if rc < 0:
            raise B2GError(rc, "b2g_wait")

    def filter_device(self, filt, forward, typesize, src_ptr, dst_ptr, n, blocksize, version=5, stream=None):
        rc = lib().b2g_filter_device(self._h, filt, 1 if forward else 0, typesize, version, src_ptr, dst_ptr, n,
                                     blocksize, stream)
        if rc < 0:
            raise B2GError(rc, "b2g_filter_device")

    def interleave_device(self, typesize, plane_ptrs, interleaved_ptr, n_elems, stream=None):
        """interleaved[i * nch + c] = plane[c][i] on device-resident buffers (image_algo.h:34-65); stream-ordered."""
        arr = (C.c_void_p * max(len(plane_ptrs), 1))(*plane_ptrs)
        rc = lib().b2g_interleave_device(self._h, typesize, len(plane_ptrs), arr, interleaved_ptr, n_elems, stream)
        if rc < 0:
            raise B2GError(rc, "b2g_interleave_device")

    def deinterleave_device(self, typesize, interleaved_ptr, plane_ptrs, n_elems, stream=None):
        """plane[c][i] = interleaved[i * nch + c] on device-resident buffers (image_algo.h:84-111); stream-ordered."""
        arr = (C.c_void_p * max(len(plane_ptrs), 1))(*plane_ptrs)
        rc = lib().b2g_deinterleave_device(self._h, typesize, len(plane_ptrs), interleaved_ptr, arr, n_elems, stream)
        if rc < 0:
            raise B2GError(rc, "b2g_deinterleave_device")

    def interleave(self, typesize, plane_ptrs, interleaved_ptr, n_elems):
        """Host-pointer form of interleave_device (staged through the context's device buffers, synchronous)."""
        arr = (C.c_void_p * max(len(plane_ptrs), 1))(*plane_ptrs)
        rc = lib().b2g_interleave(self._h, typesize, len(plane_ptrs), arr, interleaved_ptr, n_elems)
        if rc < 0:
            raise B2GError(rc, "b2g_interleave")

    def deinterleave(self, typesize, interleaved_ptr, plane_ptrs, n_elems):
        arr = (C.c_void_p * max(len(plane_ptrs), 1))(*plane_ptrs)
        rc = lib().b2g_deinterleave(self._h, typesize, len(plane_ptrs), interleaved_ptr, arr, n_elems)
        if rc < 0:
            raise B2GError(rc, "b2g_deinterleave")

    @property
    def launches(self):
        return int(lib().b2g_launch_count(self._h))

    def last_kernel_ms(self):
        out = (C.c_float * 2)()
        lib().b2g_last_kernel_ms(self._h, out)
        return float(out[0]), float(out[1])

    # ---- numpy conveniences (host pointers) ------------------------------------------------------------------
    def compress(self, arrays):
        """list of numpy arrays (raw elements) -> list of chunks (numpy uint8), one batched launch."""
        import numpy as np
        srcs = [np.ascontiguousarray(a).reshape(-1).view(np.uint8) for a in arrays]
        # worst-case destinations come out of one arena the context keeps (fresh 4 MiB buffers per chunk would be page
        # faults on every call); the chunks are copied out at their compressed size
        offs, total = [], 0
        for s in srcs:
            offs.append(total)
            total += (int(s.size) + MAX_OVERHEAD + 63) // 64 * 64
        arena = getattr(self, "_arena", None)
        if arena is None or arena.size < total:
            arena = self._arena = np.empty(max(total, 1 << 20), dtype=np.uint8)
        base = arena.ctypes.data
        b = self.make_batch([s.ctypes.data for s in srcs], [base + o for o in offs], [int(s.size) for s in srcs],
                            [int(s.size) + MAX_OVERHEAD for s in srcs])
        self.compress_batch(b)
        out = []
        for i, o in enumerate(offs):
            if b[i].status < 0:
                raise B2GError(b[i].status, "compress chunk %d" % i)
            out.append(arena[o:o + b[i].out_bytes].copy())
        return out

    def decompress(self, chunks, outs=None, trust_header=False):
        """list of chunks (numpy uint8) -> list of numpy uint8 arrays, one batched launch."""
        import numpy as np
        chunks = [np.ascontiguousarray(c) for c in chunks]
        if outs is None:
            outs = [np.empty(max(chunk_sizes(c)[0], 0), dtype=np.uint8) for c in chunks]
        b = self.make_batch([c.ctypes.data for c in chunks], [o.ctypes.data for o in outs],
                            [INT32_MAX if trust_header else int(c.size) for c in chunks],
                            [int(o.nbytes) for o in outs])
        self.decompress_batch(b)
        for i in range(len(chunks)):
            if b[i].status < 0:
                raise B2GError(b[i].status, "decompress chunk %d" % i)
        return [o.reshape(-1).view(np.uint8)[:b[i].out_bytes] for i, o in enumerate(outs)]
