"""compressed.Channel -- host-side mirror of compressed::channel<T> (compressed_image/include/compressed/channel.h)
as exposed to Python by the reference (python/bindings/wrappers/dynamic_channel.h, bind_channel.h:17-257).

A channel is a list of independently compressed blosc2 chunks (scanline-aligned, 4 MiB by default).  Where the
reference loops over chunks calling c-blosc2 once per chunk (blosc2/schunk.h:85-104, 123-141), this class hands the
whole list to ONE batched b2g call; the chunk bytes it stores are ordinary blosc2 chunks.

Same names, argument meaning and error behaviour as the reference: size mismatches raise ValueError
(std::invalid_argument through pybind11), bad chunk indices IndexError (std::out_of_range), codec failures and
constructor size mismatches RuntimeError (std::runtime_error).
"""
import threading

import numpy as np

from . import _b2g
from .enums import Codec

DEFAULT_BLOCKSIZE = 32_768        # compressed::s_default_blocksize (constants.h:11)
DEFAULT_CHUNKSIZE = 4_194_304     # compressed::s_default_chunksize (constants.h:9)

# dtypes the reference binds (python/bindings/util/variant_t.h:92-103)
_DTYPES = tuple(np.dtype(t) for t in (np.uint8, np.int8, np.uint16, np.int16, np.uint32, np.int32, np.float16,
                                      np.float32))

_ctx_tls = threading.local()


def _context(codec, clevel, typesize, filt, blocksize, device=-1):
    """Contexts own device scratch + pinned staging, so they are pooled per parameter set rather than created per
    channel / per get_chunk as the reference does with blosc2 contexts (channel.h:511-514).  A b2g_ctx is not
    thread-safe (b2g.h) and ctypes releases the GIL during calls, so the pool is PER THREAD: two threads working on
    different channels never meet in one context -- the reference gives every channel its own blosc2 contexts, which
    makes per-channel threading safe there too."""
    pool = getattr(_ctx_tls, "pool", None)
    if pool is None:
        pool = _ctx_tls.pool = {}
    key = (int(codec), int(clevel), int(typesize), int(filt), int(blocksize), int(device))
    ctx = pool.get(key)
    if ctx is None:
        ctx = _b2g.Context(*key)
        pool[key] = ctx
    return ctx


def _check_dtype(dtype):
    dt = np.dtype(dtype)
    if dt not in _DTYPES:
        raise ValueError("Unsupported dtype %s" % dt)
    return dt


def ensure_compression_level(level):
    """compressed::util::ensure_compression_level (util.h:86-94): clamp to 0..9."""
    return max(0, min(9, int(level)))


def align_chunk_to_scanlines_bytes(width, chunk_size, itemsize):
    """compressed::util::align_chunk_to_scanlines_bytes (util.h:115-131)."""
    lines = int(chunk_size) // itemsize // width if width > 0 else 0
    if lines == 0:   # util.h:121-131 throws std::runtime_error
        raise RuntimeError("Unable to align chunk size to scanlines as the size of a scanline exceeds the chunk size. "
                           "Got a scanline size of %d x %d (sizeof(T)) while the max size of the chunks is %d"
                           % (width, itemsize, int(chunk_size)))
    return lines * width * itemsize


_iter_free = {}


def _iter_context(key):
    """A context nobody else is using, for the lifetime of one channel iterator (returned by _iter_release)."""
    free = _iter_free.setdefault(key, [])
    return free.pop() if free else _b2g.Context(*key)


def _iter_release(key, ctx):
    _iter_free.setdefault(key, []).append(ctx)


_pin_free = {}


def _pinned(nbytes):
    """Page-locked host bytes (so the iterator's copies really are asynchronous), from a free list by size: pinning
    costs about a millisecond per buffer, an iterator needs five.  Plain numpy memory without torch."""
    nbytes = int(nbytes)
    free = _pin_free.get(nbytes)
    if free:
        return free.pop()
    try:
        import torch
        return torch.empty(nbytes, dtype=torch.uint8, pin_memory=True).numpy()
    except Exception:
        return np.empty(nbytes, dtype=np.uint8)


def _pinned_release(buf):
    _pin_free.setdefault(int(buf.nbytes), []).append(buf.reshape(-1).view(np.uint8))


def special_value_chunk(value, nbytes, block_size=DEFAULT_BLOCKSIZE):
    """The chunk blosc2_chunk_zeros / blosc2_chunk_repeatval build (thirdparty/c-blosc2/blosc/blosc2.c:4304-4347,
    4444-4489) for `nbytes` of one repeated element `value` (a numpy scalar): a 32-byte extended header with
    blosc2_flags = SPECIAL_ZERO << 4 (all bytes of the value zero) or SPECIAL_VALUE << 4 followed by the value's bytes.
    Host-side metadata only; decoding it is a fill (blosc2.c:1694-1718; k_decode's special-value items)."""
    v = np.asarray(value).reshape(1)
    ts = v.dtype.itemsize
    nbytes = int(nbytes)
    if nbytes % ts:
        raise ValueError("nbytes must be a multiple of the element size")
    vb = v.view(np.uint8)
    zero = not vb.any()
    bs = int(block_size) if int(block_size) > 0 else DEFAULT_BLOCKSIZE     # blosc/stune.c:46-164 with a user blocksize
    if nbytes < ts:
        bs = 1
    else:
        bs = min(bs, nbytes)
        if bs > ts:
            bs = bs // ts * ts
        bs = max(bs, 1)
    h = np.zeros(32 + (0 if zero else ts), dtype=np.uint8)
    h[0], h[1], h[2], h[3] = 5, 1, 0x05, ts                      # format version, blosclz format, extended-header marker
    h[4:8] = np.frombuffer(np.int32(nbytes).tobytes(), np.uint8)
    h[8:12] = np.frombuffer(np.int32(bs).tobytes(), np.uint8)
    h[12:16] = np.frombuffer(np.int32(h.size).tobytes(), np.uint8)
    h[31] = (1 if zero else 3) << 4                              # BLOSC2_SPECIAL_ZERO / BLOSC2_SPECIAL_VALUE
    if not zero:
        h[32:] = vb
    return h


class _Lazy:
    """A lazy fill-value chunk (blosc2/lazyschunk.h:47-105): nbytes of one repeated element, never compressed."""
    __slots__ = ("value", "nbytes")

    def __init__(self, value, nbytes):
        self.value = value
        self.nbytes = nbytes


class Channel:
    def __init__(self, data, width, height, compression_codec=Codec.lz4, compression_level=9,
                 block_size=DEFAULT_BLOCKSIZE, chunk_size=DEFAULT_CHUNKSIZE, *, filter=_b2g.SHUFFLE, device=-1):
        data = np.asarray(data)
        self._dtype = _check_dtype(data.dtype)
        self._init_params(width, height, compression_codec, compression_level, block_size, chunk_size, filter, device)
        if data.size != self._width * self._height:   # channel.h:113-123 throws std::runtime_error
            raise RuntimeError("Invalid channel data passed. Expected its size to match up to width * height (%d) but "
                             "instead got %d" % (self._width * self._height, data.size))
        self._check_encoder()
        flat = np.ascontiguousarray(data).reshape(-1).view(np.uint8)
        cb = self._chunk_bytes
        pieces = [flat[o:o + cb] for o in range(0, flat.size, cb)]
        self._chunks = self._ctx().compress(pieces) if pieces else []

    # ---- construction helpers ------------------------------------------------------------------------------
    def _init_params(self, width, height, codec, level, block_size, chunk_size, filt, device):
        self._width, self._height = int(width), int(height)
        self._codec = Codec(codec)
        self._level = ensure_compression_level(level)
        self._block_size = int(block_size)
        self._filter = int(filt)
        self._device = int(device)
        self._nthreads = 1
        self._chunk_size_arg = int(chunk_size)
        self._chunk_bytes = align_chunk_to_scanlines_bytes(self._width, chunk_size, self._dtype.itemsize)
        self._chunks = []

    def _check_encoder(self):
        """The engine encodes blosclz, lz4 and lz4hc (as LZ4 streams); zstd chunks are decode-only.
        Fail where the channel is built, not deep inside a batch."""
        if self._codec == Codec.zstd:
            raise RuntimeError("compressed_image_b200 has no zstd encoder (zstd chunks produced elsewhere decode bit-exactly); "
                               "use Codec.lz4, Codec.lz4hc or Codec.blosclz to compress")
        ts = self._dtype.itemsize                      # split rule of stune.c:200-209
        split = (self._codec in (Codec.lz4, Codec.blosclz) and self._filter == _b2g.SHUFFLE and self._block_size // ts >= 32)
        if self._block_size // (ts if split else 1) > 192 * 1024:
            raise RuntimeError("block_size %d gives streams larger than the 192 KiB encoder window of this engine"
                               % self._block_size)

    def _ctx(self):
        return _context(self._codec.blosc2_id, self._level, self._dtype.itemsize, self._filter, self._block_size,
                        self._device)

    @classmethod
    def _lazy(cls, dtype, value, width, height, codec, level, block_size, chunk_size):
        self = cls.__new__(cls)
        self._dtype = _check_dtype(dtype)
        self._init_params(width, height, codec, level, block_size, chunk_size, _b2g.SHUFFLE, -1)
        try:
            v = np.array(value).astype(self._dtype, casting="unsafe")[()]
        except (TypeError, ValueError):
            raise RuntimeError("Could not convert fill_value to the target dtype.")
        total = self._width * self._height * self._dtype.itemsize
        cb = self._chunk_bytes
        self._chunks = [_Lazy(v, min(cb, total - o)) for o in range(0, total, cb)]
        return self

    @staticmethod
    def full(dtype, fill_value, width, height, compression_codec=Codec.lz4, compression_level=9,
             block_size=DEFAULT_BLOCKSIZE, chunk_size=DEFAULT_CHUNKSIZE):
        return Channel._lazy(dtype, fill_value, width, height, compression_codec, compression_level, block_size,
                             chunk_size)

    @staticmethod
    def zeros(dtype, width, height, compression_codec=Codec.lz4, compression_level=9, block_size=DEFAULT_BLOCKSIZE,
              chunk_size=DEFAULT_CHUNKSIZE):
        return Channel._lazy(dtype, 0, width, height, compression_codec, compression_level, block_size, chunk_size)

    @staticmethod
    def full_like(other, fill_value):
        return Channel._lazy(other._dtype, fill_value, other._width, other._height, other._codec, other._level,
                             other._block_size, other._chunk_size_arg)

    @staticmethod
    def zeros_like(other):
        return Channel.full_like(other, 0)

    # ---- attributes ------------------------------------------------------------------------------------------
    @property
    def dtype(self):
        return self._dtype

    @property
    def shape(self):
        return (self._height, self._width)

    @property
    def width(self):
        return self._width

    @property
    def height(self):
        return self._height

    def compression(self):
        return self._codec

    def compression_level(self):
        return self._level

    def num_chunks(self):
        return len(self._chunks)

    def _nbytes_of(self, i):
        c = self._chunks[i]
        return c.nbytes if isinstance(c, _Lazy) else _b2g.chunk_sizes(c)[0]

    def _index(self, i):
        i = int(i)
        if i < 0 or i >= len(self._chunks):   # schunk_mixin.h:204-214 -> std::out_of_range
            raise IndexError("Cannot access index %d in schunk. Total amount of chunks is %d" % (i, len(self._chunks)))
        return i

    def block_size(self):
        """schunk_mixin.h:191-194 max_block_size: the block size the channel was created with."""
        return self._block_size

    def chunk_size(self, chunk_index=None):
        if chunk_index is None:
            return self._chunk_bytes   # schunk_mixin.h:156-159: m_ChunkSize, also for the last chunk
        return self._nbytes_of(self._index(chunk_index))

    def chunk_elems(self, chunk_index=None):
        return self.chunk_size(chunk_index) // self._dtype.itemsize

    def compressed_bytes(self):
        """schunk::csize / lazy_schunk::csize (lazyschunk.h:312-327: a lazy chunk counts sizeof(T))."""
        return sum(self._dtype.itemsize if isinstance(c, _Lazy) else int(c.size) for c in self._chunks)

    def uncompressed_size(self):
        """In ELEMENTS (channel.h:404-420)."""
        return sum(self._nbytes_of(i) for i in range(len(self._chunks))) // self._dtype.itemsize

    def update_nthreads(self, nthreads, block_size=DEFAULT_BLOCKSIZE):
        """Kept for API compatibility (channel.h:343-348): the grid is the thread pool, so only block_size matters."""
        self._nthreads = int(nthreads)
        self._block_size = int(block_size)

    # ---- chunk access ----------------------------------------------------------------------------------------
    def _decode_into(self, indices, outs):
        """Decompress chunks[indices] into the uint8 views `outs` with one batched launch (lazy chunks are fills)."""
        self._collect_iterators()
        real_c, real_o = [], []
        for i, o in zip(indices, outs):
            c = self._chunks[i]
            if isinstance(c, _Lazy):
                o.view(self._dtype)[:] = c.value
            else:
                real_c.append(c)
                real_o.append(o)
        if real_c:
            self._ctx().decompress(real_c, real_o)

    def get_chunk(self, chunk_index, array=None):
        i = self._index(chunk_index)
        n = self._nbytes_of(i) // self._dtype.itemsize
        if array is None:
            array = np.empty(n, dtype=self._dtype)
        else:
            if not isinstance(array, np.ndarray) or array.dtype != self._dtype:
                raise ValueError("Array must have dtype matching channel element type.")
            if array.ndim != 1:
                raise ValueError("Array must be 1-dimensional.")
            if array.shape[0] != n:
                raise ValueError("Array length does not match number of chunk elements.")
            if not array.flags.c_contiguous or not array.flags.writeable:
                raise ValueError("Array must be contiguous and writeable.")
        self._decode_into([i], [array.view(np.uint8)])
        return array

    def set_chunk(self, chunk_index, array):
        array = np.asarray(array)
        if array.ndim != 1:
            raise ValueError("set_chunk requires a 1D numpy array")
        if array.dtype != self._dtype:
            raise ValueError("set_chunk requires numpy array of dtype '%s', instead got '%s'" % (self._dtype, array.dtype))
        i = self._index(chunk_index)
        if array.size * self._dtype.itemsize != self._nbytes_of(i):   # schunk_mixin.h:216-246 -> std::invalid_argument
            raise ValueError("Invalid chunk size: expected %d elements, got %d" %
                             (self._nbytes_of(i) // self._dtype.itemsize, array.size))
        self._check_encoder()
        self._collect_iterators()
        self._chunks[i] = self._ctx().compress([array])[0]
        for _, refresh in getattr(self, "_iter_hooks", ()):
            refresh(i, array)

    def set_chunks(self, chunk_indices, arrays):
        """Batched set_chunk (no reference equivalent: one launch for many chunks)."""
        idx = [self._index(i) for i in chunk_indices]
        for i, a in zip(idx, arrays):
            a = np.asarray(a)
            if a.dtype != self._dtype or a.ndim != 1 or a.size * self._dtype.itemsize != self._nbytes_of(i):
                raise ValueError("Invalid chunk passed to set_chunks")
        for i, c in zip(idx, self._ctx().compress(list(arrays))):
            self._chunks[i] = c

    def get_decompressed(self):
        out = np.empty(self._width * self._height, dtype=self._dtype)
        flat = out.view(np.uint8)
        outs, o = [], 0
        for i in range(len(self._chunks)):
            nb = self._nbytes_of(i)
            outs.append(flat[o:o + nb])
            o += nb
        self._decode_into(range(len(self._chunks)), outs)
        return out.reshape(self._height, self._width)

    # ---- chunk iteration (iterators/iterator.h:97-141: decompress -> user modifies -> recompress) --------------
    def __iter__(self):
        """Yields each chunk as a writable 1-D array; when the iterator advances (and at exhaustion, and when the
        iteration is abandoned) the chunk is recompressed into the channel, like compressed::channel_iterator, which
        compresses on every increment whether or not the caller wrote to the span (iterators/iterator.h:97-141, 270-304).
        A lazy chunk that still holds nothing but its fill value stays lazy.

        Pipelined in GROUPS: the chunks are fetched `group` at a time (about 16 MiB, at most 8 chunks) with ONE
        asynchronous decompress batch, group g + 1 while the caller works on group g; every chunk the caller leaves is
        handed to one of three compress contexts (b2g_compress_batch_async) and collected later, so the PCIe link carries
        the next group down and the finished chunks up while the caller computes.  Three group buffers rotate: a view
        stays valid until the iterator has advanced by two groups.  The contexts come from a free list for the lifetime of
        the iterator (zip() over several channels and calls on the channel from inside the loop never meet a context with
        a batch in flight); get_chunk / set_chunk / get_decompressed on the channel first collect what the iterator still
        has in flight."""
        n = len(self._chunks)
        if n == 0:
            return
        key = (self._codec.blosc2_id, self._level, self._dtype.itemsize, self._filter, self._block_size, self._device)
        itemsize = self._dtype.itemsize
        nb0 = self._nbytes_of(0)
        group = max(1, min(8, (16 << 20) // max(nb0, 1), n))
        ngroups = (n + group - 1) // group
        ncomp = 3
        dctx = _iter_context(key)
        cctxs = [_iter_context(key) for _ in range(min(ncomp, n))]
        gbufs = [_pinned(nb0 * group) for _ in range(min(ngroups, 3))]
        dsts = [_pinned(nb0 + _b2g.MAX_OVERHEAD) for _ in cctxs]
        pending = [None] * len(cctxs)     # per compress context: (index, batch, group buffer index) in flight
        state = {"fetch": None}

        def view_of(i):
            g, k = divmod(i, group)
            nb = self._nbytes_of(i)
            return gbufs[g % len(gbufs)][k * nb0:k * nb0 + nb].view(self._dtype)

        def collect(slot):
            p = pending[slot]
            if p is None:
                return
            pending[slot] = None
            i, b, _gb = p
            cctxs[slot].wait()
            if b[0].status < 0:
                raise _b2g.B2GError(b[0].status, "compress chunk %d" % i)
            self._chunks[i] = dsts[slot][:b[0].out_bytes].copy()

        def collect_all(gb=None):
            for slot in range(len(cctxs)):
                if pending[slot] is not None and (gb is None or pending[slot][2] == gb):
                    collect(slot)

        def start_fetch(g):
            gb = g % len(gbufs)
            collect_all(gb)                 # every compress that still reads this buffer has finished
            srcs, dsts_, sizes, caps, keep = [], [], [], [], []
            for i in range(g * group, min(n, (g + 1) * group)):
                c = self._chunks[i]
                v = view_of(i)
                if isinstance(c, _Lazy):
                    v[:] = c.value
                    continue
                c = np.ascontiguousarray(c)
                keep.append(c)
                srcs.append(c.ctypes.data)
                dsts_.append(v.ctypes.data)
                sizes.append(int(c.size))
                caps.append(int(v.nbytes))
            b = None
            if srcs:
                b = dctx.make_batch(srcs, dsts_, sizes, caps)
                dctx.decompress_batch_async(b)
            state["fetch"] = (g, b, keep)
            state["hi"] = min(n, (g + 1) * group)

        def finish_fetch():
            f = state["fetch"]
            state["fetch"] = None
            if f is not None and f[1] is not None:
                dctx.wait()
                for k in range(len(f[1])):
                    if f[1][k].status < 0:
                        raise _b2g.B2GError(f[1][k].status, "decompress chunk in group %d" % f[0])

        def start_compress(i, view):
            c = self._chunks[i]
            if isinstance(c, _Lazy) and bool((view == c.value).all()):
                return                      # untouched fill: stays lazy
            slot = i % len(cctxs)
            collect(slot)
            src = view.view(np.uint8)
            dst = dsts[slot]
            b = cctxs[slot].make_batch([src.ctypes.data], [dst.ctypes.data], [int(src.size)], [int(dst.size)])
            cctxs[slot].compress_batch_async(b)
            pending[slot] = (i, b, (i // group) % len(gbufs))

        def refresh(j, array):
            """set_chunk(j) from inside the loop: a chunk that is already fetched but not yet handed out gets the new data."""
            if state["cur"] < j < state["hi"]:
                finish_fetch()
                view_of(j)[:] = np.asarray(array).reshape(-1)

        state["cur"], state["hi"] = -1, 0
        hooks = (collect_all, refresh)
        self._iter_hooks = getattr(self, "_iter_hooks", [])
        self._iter_hooks.append(hooks)
        try:
            start_fetch(0)
            for g in range(ngroups):
                finish_fetch()
                if g + 1 < ngroups:
                    start_fetch(g + 1)
                for i in range(g * group, min(n, (g + 1) * group)):
                    state["cur"] = i
                    buf = view_of(i)
                    try:
                        yield buf
                    finally:    # also runs when the iteration is abandoned (zip over channels stops at the first
                        # exhausted iterator): the reference flushes the last chunk in ~channel_iterator (iterator.h:72-93)
                        start_compress(i, buf)
        finally:
            try:
                finish_fetch()
                collect_all()
            finally:
                self._iter_hooks.remove(hooks)
                _iter_release(key, dctx)
                for cx in cctxs:
                    _iter_release(key, cx)
                for b in gbufs + dsts:
                    _pinned_release(b)

    def _collect_iterators(self):
        """Chunks an active iterator is still compressing land in the channel before anything else looks at them."""
        for flush, _ in getattr(self, "_iter_hooks", ()):
            flush()

    def chunks(self, special=False):
        """The stored chunk bytes (numpy uint8 arrays).  Lazy chunks are materialised: by default the fill buffer is
        compressed, once per distinct (value, size) like lazy_schunk::to_schunk (lazyschunk.h:122-172); with
        special=True they become blosc2 SPECIAL-VALUE chunks of 32 (+ sizeof(T)) bytes (special_value_chunk below), which
        any blosc2 decoder expands -- on the GPU that is an O(1)-input fill inside k_decode, the device-resident form of a
        lazy chunk."""
        out, cache = [], {}
        for c in self._chunks:
            if isinstance(c, _Lazy):
                key = (c.value.tobytes(), c.nbytes)
                if key not in cache:
                    if special:
                        cache[key] = special_value_chunk(c.value, c.nbytes, self._block_size)
                    else:
                        cache[key] = self._ctx().compress([np.full(c.nbytes // self._dtype.itemsize, c.value, self._dtype)])[0]
                out.append(cache[key])
            else:
                out.append(c)
        return out

    # ---- contiguous frame (b2frame) export / import: to_schunk() + blosc2_schunk_to_buffer (schunk.h:107-121) -------
    def to_cframe(self):
        """The channel's chunks as one contiguous blosc2 frame (bytes) that stock c-blosc2 opens with
        blosc2_schunk_from_buffer; the offsets chunk is compressed on the GPU (b2frame.py)."""
        from . import b2frame
        enc, _ = b2frame.engine_from_context(_context(0, 5, 8, 1, DEFAULT_BLOCKSIZE, self._device))
        return b2frame.to_cframe(self.chunks(), enc)

    def set_from_cframe(self, frame):
        """Replace the channel's chunks by those of a contiguous frame; the decompressed sizes must be the channel's."""
        from . import b2frame
        _, dec = b2frame.engine_from_context(_context(0, 5, 8, 1, DEFAULT_BLOCKSIZE, self._device))
        hd, chunks = b2frame.from_cframe(frame, dec)
        if hd["special"]:
            raise ValueError("special-value frame entries are not supported")
        if len(chunks) != self.num_chunks():
            raise ValueError("frame holds %d chunks, channel has %d" % (len(chunks), self.num_chunks()))
        for i, c in enumerate(chunks):
            if _b2g.chunk_sizes(c)[0] != self._nbytes_of(i):
                raise ValueError("Invalid chunk size for chunk %d" % i)
        self._chunks = [np.array(c, dtype=np.uint8) for c in chunks]

