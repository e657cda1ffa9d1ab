"""compressed.Image -- host-side mirror of compressed::image<T> (compressed_image/include/compressed/image.h) as the
reference exposes it to Python (python/bindings/bind_image.h:18-479, wrappers/dynamic_image.h).

An image is a list of Channels + names + a metadata dict.  Construction and get_decompressed() batch the chunks of ALL
channels into one b2g call where the reference compresses channel after channel, chunk after chunk
(image.h:119-159, blosc2/schunk.h:85-104).  File reading (`read`, `dtype_from_file`) is OpenImageIO territory and is
out of scope (OUT OF SCOPE rows 5/15).
"""
import json

import numpy as np

from . import _b2g
from .channel import Channel, DEFAULT_BLOCKSIZE, DEFAULT_CHUNKSIZE, _check_dtype
from .enums import Codec


class Image:
    def __init__(self, dtype, channels, width, height, channel_names=(), compression_codec=Codec.lz4,
                 compression_level=9, block_size=DEFAULT_BLOCKSIZE, chunk_size=DEFAULT_CHUNKSIZE):
        self._dtype = _check_dtype(dtype)
        self._width, self._height = int(width), int(height)
        self._chunk_size = int(chunk_size)
        self._metadata = {}
        channel_names = list(channel_names)
        if channel_names and len(channel_names) != len(channels):   # image.h:100-107
            raise ValueError("Invalid number of channel names passed: got %d names for %d channels" %
                             (len(channel_names), len(channels)))
        self._names = channel_names
        arrs = []
        for a in channels:
            a = np.asarray(a)
            if a.dtype != self._dtype:
                raise ValueError("Channel dtype %s does not match image dtype %s" % (a.dtype, self._dtype))
            arrs.append(a)
        self._channels = Image._build_channels(arrs, self._width, self._height, compression_codec, compression_level,
                                               block_size, chunk_size)

    @staticmethod
    def _build_channels(arrs, width, height, codec, level, block_size, chunk_size):
        """All chunks of all channels in ONE compress batch."""
        chans, pieces, owner = [], [], []
        for k, a in enumerate(arrs):
            ch = Channel.__new__(Channel)
            ch._dtype = _check_dtype(a.dtype)
            ch._init_params(width, height, codec, level, block_size, chunk_size, _b2g.SHUFFLE, -1)
            if a.size != ch._width * ch._height:
                raise RuntimeError("Invalid channel data passed. Expected its size to match up to width * height (%d) "
                                 "but instead got %d" % (ch._width * ch._height, a.size))
            flat = np.ascontiguousarray(a).reshape(-1).view(np.uint8)
            for o in range(0, flat.size, ch._chunk_bytes):
                pieces.append(flat[o:o + ch._chunk_bytes])
                owner.append(k)
            chans.append(ch)
        if pieces:
            out = chans[0]._ctx().compress(pieces)
            for k, c in zip(owner, out):
                chans[k]._chunks.append(c)
        return chans

    # ---- reading files: OpenImageIO path, out of scope ---------------------------------------------------------
    @staticmethod
    def read(*args, **kwargs):
        raise NotImplementedError("Image.read needs OpenImageIO (image.h:360-896); file I/O is outside the chunk "
                                  "compress/decompress path this package implements")

    @staticmethod
    def dtype_from_file(filepath):
        raise NotImplementedError("Image.dtype_from_file needs OpenImageIO")

    # ---- channels ------------------------------------------------------------------------------------------------
    def add_channel(self, data, width, height, name=None, compression_codec=Codec.lz4, compression_level=9,
                    block_size=DEFAULT_BLOCKSIZE, chunk_size=DEFAULT_CHUNKSIZE):
        data = np.asarray(data)
        if data.dtype != self._dtype:
            raise ValueError("Channel dtype %s does not match image dtype %s" % (data.dtype, self._dtype))
        if int(width) != self._width or int(height) != self._height:   # image.h:960-968
            raise ValueError("Invalid channel dimensions: expected %dx%d, got %dx%d" %
                             (self._width, self._height, int(width), int(height)))
        if data.size != self._width * self._height:
            raise ValueError("Invalid channel data: expected %d elements, got %d" %
                             (self._width * self._height, data.size))
        if self._names and name is None:
            raise ValueError("The image has channel names: a name is required for the new channel")
        self._channels.append(Channel(data, width, height, compression_codec, compression_level, block_size,
                                      chunk_size))
        if name is not None:
            self._names.append(str(name))

    def get_channel_index(self, channelname):
        try:
            return self._names.index(channelname)
        except ValueError:
            raise ValueError("Unknown channel name '%s' encountered" % channelname)   # image.h:1216-1232

    def _resolve(self, key):
        if isinstance(key, str):
            return self.get_channel_index(key)
        i = int(key)
        if i < 0 or i >= len(self._channels):
            raise IndexError("Cannot access channel index %d, total number of channels is %d" % (i, len(self._channels)))
        return i

    def remove_channel(self, name_or_index):
        i = self._resolve(name_or_index)
        del self._channels[i]
        if i < len(self._names):
            del self._names[i]

    def channel(self, key):
        return self._channels[self._resolve(key)]

    __getitem__ = channel

    def __len__(self):
        return len(self._channels)

    def __iter__(self):
        return iter(self._channels)

    def channels(self):
        return list(self._channels)

    # ---- attributes ----------------------------------------------------------------------------------------------
    @property
    def dtype(self):
        return self._dtype

    @property
    def shape(self):
        return (len(self._channels), self._height, self._width)

    @property
    def width(self):
        return self._width

    @property
    def height(self):
        return self._height

    @property
    def num_channels(self):
        return len(self._channels)

    def get_channel_names(self):
        return list(self._names)

    def set_channel_names(self, names):
        names = list(names)
        if len(names) != len(self._channels):
            raise ValueError("Invalid number of channel names passed")
        self._names = names

    def get_metadata(self):
        return json.loads(json.dumps(self._metadata))

    def set_metadata(self, metadata):
        self._metadata = json.loads(json.dumps(metadata))

    def chunk_size(self):
        return self._chunk_size

    def update_nthreads(self, nthreads):
        for c in self._channels:
            c.update_nthreads(nthreads, c.block_size())

    def compression_ratio(self):
        """image.h:1160-1170 -- note the reference seeds both sums with 1 and mixes elements with bytes."""
        unc, comp = 1, 1
        for c in self._channels:
            comp += c.compressed_bytes()
            unc += c.uncompressed_size()
        return unc / comp

    def print_statistics(self):
        comp = sum(c.compressed_bytes() for c in self._channels)
        unc = sum(c.uncompressed_size() for c in self._channels)
        print("Statistics for image buffer:")
        print(" Width:             %d" % self._width)
        print(" Height:            %d" % self._height)
        print(" Channels:          %d" % len(self._channels))
        print(" Channelnames:      [%s]" % ", ".join(self._names))
        print(" --------------     ")
        print(" Compressed Size:   %d" % comp)
        print(" Uncompressed Size: %d" % unc)
        print(" Compression ratio: %gx" % (unc / comp if comp else float("inf")))
        print(" Num Chunks:        %d" % sum(c.num_chunks() for c in self._channels))
        print(" Metadata:          \n %s" % json.dumps(self._metadata, indent=4))

    # ---- decompression ---------------------------------------------------------------------------------------------
    def get_decompressed(self):
        """list of (height, width) arrays; every chunk of every channel goes through one decompress batch."""
        outs = [np.empty(self._width * self._height, dtype=self._dtype) for _ in self._channels]
        chunks, views = [], []
        for ch, out in zip(self._channels, outs):
            flat, o, lazy_i, lazy_o = out.view(np.uint8), 0, [], []
            for i in range(ch.num_chunks()):
                nb = ch._nbytes_of(i)
                c = ch._chunks[i]
                if isinstance(c, np.ndarray):
                    chunks.append(c)
                    views.append(flat[o:o + nb])
                else:
                    lazy_i.append(i)
                    lazy_o.append(flat[o:o + nb])
                o += nb
            if lazy_i:
                ch._decode_into(lazy_i, lazy_o)
        if chunks:
            self._channels[0]._ctx().decompress(chunks, views)
        return [o.reshape(self._height, self._width) for o in outs]
