"""Contiguous frame ("b2frame" / cframe) export and import of a channel's chunks.

What the reference does to serialise a channel: schunk<T>::to_schunk() (compressed/blosc2/schunk.h:107-121) then
blosc2_schunk_to_buffer (thirdparty/c-blosc2/blosc/schunk.c:357), i.e. frame_from_schunk (blosc/frame.c:924-1072) with
new_header_frame (frame.c:70-352) and frame_update_trailer (frame.c:553-760); format in
thirdparty/c-blosc2/README_CFRAME_FORMAT.rst.  Layout written and read here:

    header (97 bytes without metalayers, msgpack, big-endian integers)
    chunk 0 | chunk 1 | ... | chunk N-1          the blosc2 chunks, untouched
    offsets chunk                                 a blosc2 chunk of N little-endian int64: where each chunk starts,
                                                  counted from the END of the header (frame.c:950); a set top bit
                                                  marks a special-value chunk with no bytes in the frame
    trailer (35 bytes without vl-metalayers)      msgpack, ends with the trailer length and an empty fingerprint

The header and trailer are host-side metadata (a few dozen bytes).  The one piece of data work, the offsets chunk, is
compressed / decompressed by the engine passed in as `encode` / `decode` -- on a GPU box the b2g context
(`engine_from_context`); the frame code itself contains no codec.
"""
import struct

import numpy as np

HEADER_MINLEN = 87          # frame.h:45
HEADER_LEN = 97             # + the empty metalayer section (frame.c:262-331)
TRAILER_LEN = 35            # frame.c:559-690 without vl-metalayers
FRAME_FORMAT_VERSION = 2    # BLOSC2_VERSION_FRAME_FORMAT (blosc2.h:157-159)
MAGIC = b"b2frame\x00"
SPECIAL_BIT = 1 << 63


class FrameError(ValueError):
    pass


def engine_from_context(ctx):
    """(encode, decode) over a compressed_image_b200._b2g.Context created with typesize 8."""
    def encode(array):
        return ctx.compress([array])[0]

    def decode(chunk, out):
        ctx.decompress([chunk], [out])
    return encode, decode


def _chunk_sizes(chunk):
    nbytes, blocksize, cbytes = struct.unpack_from("<iii", chunk, 4)
    return nbytes, cbytes, blocksize


def to_cframe(chunks, encode, typesize=8, codec=0, clevel=5, blocksize=0, filters=(0, 0, 0, 0, 0, 1),
              filters_meta=(0, 0, 0, 0, 0, 0), splitmode=4, nthreads=(0, 1)):
    """list of blosc2 chunks (bytes-like) -> bytes of a contiguous frame.

    The header defaults are those of the schunk compressed-image exports (wrapper.h:301-310: BLOSC2_CPARAMS_DEFAULTS =
    blosclz, level 5, typesize 8, shuffle, forward-compatible split); they describe the container only -- every chunk
    carries its own header and decodes by it."""
    chunks = [np.frombuffer(c, dtype=np.uint8) if not isinstance(c, np.ndarray) else c.reshape(-1).view(np.uint8) for c in chunks]
    n = len(chunks)
    offsets = np.zeros(n, dtype=np.int64)
    nbytes_total, cbytes_total, chunksize = 0, 0, -1
    for i, c in enumerate(chunks):
        if c.size < 16:
            raise FrameError("chunk %d is shorter than a blosc header" % i)
        nb, cb, _ = _chunk_sizes(c)
        if cb != c.size:
            raise FrameError("chunk %d: header says %d bytes, buffer has %d" % (i, cb, c.size))
        offsets[i] = cbytes_total
        cbytes_total += cb
        nbytes_total += nb
        # the header carries the size of chunk 0 and readers derive the chunk count from it (frame.c:484-497), so --
        # as for blosc2_schunk_append_chunk -- every chunk but the last must have that size and the last may be shorter
        if i == 0:
            chunksize = nb
        elif nb > chunksize or (nb < chunksize and i != n - 1):
            raise FrameError("chunk %d holds %d bytes; only the last chunk may be shorter than chunk 0 (%d)" % (i, nb, chunksize))
    off_chunk = np.zeros(0, np.uint8)
    if n > 0:
        off_chunk = np.asarray(encode(offsets)).reshape(-1).view(np.uint8)
    frame_len = HEADER_LEN + cbytes_total + off_chunk.size + TRAILER_LEN

    h = bytearray(HEADER_LEN)
    h[0] = 0x90 + 14
    h[1] = 0xa0 + 8
    h[2:10] = MAGIC
    h[10] = 0xd2
    struct.pack_into(">i", h, 11, HEADER_LEN)
    h[15] = 0xcf
    struct.pack_into(">Q", h, 16, frame_len)
    h[24] = 0xa0 + 4
    h[25] = FRAME_FORMAT_VERSION + 0x10                   # 64-bit offsets
    h[26] = 0                                             # contiguous
    h[27] = (codec & 0xf) | ((clevel & 0xf) << 4)
    h[28] = splitmode - 1
    h[29] = 0xd3
    struct.pack_into(">q", h, 30, nbytes_total)
    h[38] = 0xd3
    struct.pack_into(">q", h, 39, cbytes_total)
    h[47] = 0xd2
    struct.pack_into(">i", h, 48, typesize)
    h[52] = 0xd2
    struct.pack_into(">i", h, 53, blocksize)
    h[57] = 0xd2
    struct.pack_into(">i", h, 58, chunksize)
    h[62] = 0xd1
    struct.pack_into(">h", h, 63, nthreads[0])
    h[65] = 0xd1
    struct.pack_into(">h", h, 66, nthreads[1])
    h[68] = 0xc2                                          # no vl-metalayers
    h[69] = 0xd8
    h[70] = 6                                             # BLOSC2_MAX_FILTERS
    h[71:77] = bytes(filters)
    h[77] = codec                                         # udcodec slot holds the compcode (frame.c:243-244)
    h[78] = 0
    h[79:85] = bytes(filters_meta)
    h[87] = 0x93
    h[88] = 0xcd
    struct.pack_into(">H", h, 89, 7)                      # size of the (empty) metalayer index (frame.c:298-304)
    h[91] = 0xde
    struct.pack_into(">H", h, 92, 0)
    h[94] = 0xdc
    struct.pack_into(">H", h, 95, 0)

    t = bytearray(TRAILER_LEN)
    t[0] = 0x90 + 4
    t[1] = 1                                              # FRAME_TRAILER_VERSION
    t[2] = 0x93
    t[3] = 0xcd
    struct.pack_into(">H", t, 4, 6)
    t[6] = 0xde
    t[9] = 0xdc
    t[12] = 0xce
    struct.pack_into(">I", t, 13, TRAILER_LEN)
    t[17] = 0xd8                                          # fixext 16: fingerprint type 0 + 16 zero bytes
    return b"".join([bytes(h)] + [c.tobytes() for c in chunks] + [off_chunk.tobytes(), bytes(t)])


def read_header(frame):
    """-> dict of the header fields (get_header_info, frame.c:355-505); raises FrameError on a malformed header."""
    f = memoryview(frame)
    if len(f) < HEADER_MINLEN:
        raise FrameError("frame shorter than the minimum header")
    if f[0] != 0x9e or f[1] != 0xa8 or bytes(f[2:10]) != MAGIC:
        raise FrameError("not a b2frame")
    header_len = struct.unpack_from(">i", f, 11)[0]
    frame_len = struct.unpack_from(">q", f, 16)[0]
    if header_len < HEADER_MINLEN or header_len > frame_len or frame_len > len(f):
        raise FrameError("inconsistent header / frame length")
    if f[25] & 0x0f > FRAME_FORMAT_VERSION:
        raise FrameError("frame format version %d is newer than this reader" % (f[25] & 0x0f))
    if f[26] & 0x0f != 0:
        raise FrameError("only contiguous frames are supported")
    nbytes = struct.unpack_from(">q", f, 30)[0]
    cbytes = struct.unpack_from(">q", f, 39)[0]
    typesize, blocksize, chunksize = (struct.unpack_from(">i", f, o)[0] for o in (48, 53, 58))
    if typesize <= 0 or cbytes < 0:
        raise FrameError("invalid typesize / cbytes")
    nfilters = f[70]
    if nfilters > 6:
        raise FrameError("too many filters")
    nchunks = 0
    if nbytes > 0 and chunksize > 0:
        nchunks = -(-nbytes // chunksize)
    return {"header_len": header_len, "frame_len": frame_len, "nbytes": nbytes, "cbytes": cbytes, "typesize": typesize,
            "blocksize": blocksize, "chunksize": chunksize, "nchunks": nchunks, "codec": f[27] & 0xf, "clevel": f[27] >> 4,
            "splitmode": f[28] + 1, "filters": bytes(f[71:71 + nfilters]), "filters_meta": bytes(f[79:79 + nfilters]),
            "has_vlmetalayers": f[68] == 0xc3}


def from_cframe(frame, decode):
    """bytes of a contiguous frame -> (header dict, list of chunks as numpy uint8 views into the frame).

    Chunks of variable size (header chunksize 0) are counted from the offsets chunk itself.  Special-value entries
    (top bit of the offset set) have no bytes in the frame and are returned as None with their code in
    header["special"][i]."""
    buf = np.frombuffer(frame, dtype=np.uint8)
    hd = read_header(buf)
    hl, fl = hd["header_len"], hd["frame_len"]
    tail = buf[:fl]
    if fl < hl + TRAILER_LEN - 10:
        raise FrameError("no room for a trailer")
    trailer_len = struct.unpack_from(">I", tail, fl - 22)[0] if tail[fl - 23] == 0xce else -1   # FRAME_TRAILER_LEN_OFFSET (frame.h:51)
    if trailer_len < 25 or hl + hd["cbytes"] + trailer_len > fl:
        raise FrameError("invalid trailer length")
    hd["trailer_len"] = trailer_len
    off_start = hl + hd["cbytes"]
    off_end = fl - trailer_len
    chunks, special = [], {}
    if off_end > off_start:
        off_chunk = buf[off_start:off_end]
        if off_chunk.size < 16:
            raise FrameError("offsets chunk truncated")
        onb, ocb, _ = _chunk_sizes(off_chunk)
        if ocb != off_chunk.size or onb % 8:
            raise FrameError("offsets chunk does not fill the space between the chunks and the trailer")
        offsets = np.empty(onb // 8, dtype=np.int64)
        decode(off_chunk, offsets.view(np.uint8))
        if hd["nchunks"] and hd["nchunks"] != offsets.size:
            raise FrameError("header announces %d chunks, offsets chunk holds %d" % (hd["nchunks"], offsets.size))
        hd["nchunks"] = int(offsets.size)
        for i, o in enumerate(offsets.view(np.uint64)):
            o = int(o)
            if o & SPECIAL_BIT:
                special[i] = (o >> 56) & 0x7
                chunks.append(None)
                continue
            pos = hl + o
            if o > hd["cbytes"] - 16 or pos + 16 > off_start:
                raise FrameError("chunk %d starts outside the chunk section" % i)
            _, cb, _ = _chunk_sizes(buf[pos:pos + 16])
            if cb < 16 or pos + cb > off_start:
                raise FrameError("chunk %d runs past the chunk section" % i)
            chunks.append(buf[pos:pos + cb])
    elif hd["nchunks"]:
        raise FrameError("header announces chunks but the frame has no offsets chunk")
    hd["special"] = special
    return hd, chunks
