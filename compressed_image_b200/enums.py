"""compressed.Codec -- mirror of compressed::enums::codec (compressed_image/include/compressed/enums.h) with the
values the pybind11 module exposes (python/bindings/bind_enums.h:15-24; stubs: zstd = 3)."""
import enum


class Codec(enum.IntEnum):
    blosclz = 0
    lz4 = 1
    lz4hc = 2
    zstd = 3

    @property
    def blosc2_id(self):
        """compressed::blosc2::codec_to_blosc2 (blosc2/wrapper.h:74-96): BLOSC_BLOSCLZ 0, LZ4 1, LZ4HC 2, ZSTD 5."""
        return (0, 1, 2, 5)[int(self)]

    @staticmethod
    def from_blosc2(compcode):
        """compressed::blosc2::blosc2_to_codec (blosc2/wrapper.h:98-119)."""
        try:
            return {0: Codec.blosclz, 1: Codec.lz4, 2: Codec.lz4hc, 5: Codec.zstd}[int(compcode)]
        except KeyError:
            raise ValueError("Unknown blosc2 compression codec %r" % (compcode,))
