"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol include/b2g.h declares,
its host-only entry points agree with the checker, and the host-side mirror of the reference API applies the
reference's size / index / dtype rules.  No kernels are launched here (no GPU in the dev container)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import _golden
import _libs

from compressed_image_b200 import _b2g
from compressed_image_b200 import channel as chmod
from compressed_image_b200.enums import Codec

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_match_header():
    hdr = open(os.path.join(ROOT, "include", "b2g.h")).read()
    declared = re.findall(r"B2G_API\s+[\w\s\*]+?\b(b2g_\w+)\s*\(", hdr)
    assert len(declared) >= 15
    L = _b2g.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert sorted(declared) == sorted(_b2g.EXPORTS)


def test_struct_layout_matches_header():
    assert C.sizeof(_b2g.Params) == 12
    assert C.sizeof(_b2g.Chunk) == 32
    assert _b2g.Chunk.src_bytes.offset == 16 and _b2g.Chunk.status.offset == 28


@pytest.mark.parametrize("key", _golden.REF_KEYS[::3])
def test_chunk_sizes_matches_checker(key):
    chunk = _golden.ref_chunk(key)
    nb, cb, bs = _b2g.chunk_sizes(chunk)
    O = _libs.oracle()
    a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
    O.b2o_chunk_sizes(_libs.ptr(chunk), C.byref(a), C.byref(b), C.byref(c))
    assert (nb, cb, bs) == (a.value, b.value, c.value)
    assert cb == chunk.size
    if _libs.ref() is not None:
        assert (nb, cb, bs) == _libs.ref_sizes(chunk)


def test_chunk_sizes_rejects_bad_headers():
    chunk = _golden.ref_chunk(_golden.REF_KEYS[0]).copy()
    chunk[0] = 9
    with pytest.raises(_b2g.B2GError) as e:
        _b2g.chunk_sizes(chunk)
    assert e.value.code == -10
    chunk[0] = 5
    chunk[3] = 0
    with pytest.raises(_b2g.B2GError):
        _b2g.chunk_sizes(chunk)


def test_no_device_fails_loudly():
    """Without a CUDA device a context cannot be created: there is no CPU fallback to fall into."""
    if _b2g.lib().b2g_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(_b2g.B2GError) as e:
        _b2g.Context()
    assert e.value.code == -100


def test_strerror_and_version():
    L = _b2g.lib()
    assert b"sm_100a" in L.b2g_version()
    assert L.b2g_strerror(-11) == b"invalid chunk header"
    assert L.b2g_strerror(0) == b"ok"


def test_codec_enum_mapping():
    """wrapper.h:74-119 codec_to_blosc2 / blosc2_to_codec; bind_enums.h:15-24 python values."""
    assert [int(c) for c in (Codec.blosclz, Codec.lz4, Codec.lz4hc, Codec.zstd)] == [0, 1, 2, 3]
    assert [c.blosc2_id for c in Codec] == [0, 1, 2, 5]
    for c in Codec:
        assert Codec.from_blosc2(c.blosc2_id) == c
    with pytest.raises(ValueError):
        Codec.from_blosc2(4)


def test_chunk_alignment_rule():
    """util.h:115-131 / 180-184."""
    assert chmod.align_chunk_to_scanlines_bytes(4096, 4 * 1024 * 1024, 1) == 4 * 1024 * 1024
    assert chmod.align_chunk_to_scanlines_bytes(1920, 4 * 1024 * 1024, 2) == (4 * 1024 * 1024 // 3840) * 3840
    assert chmod.align_chunk_to_scanlines_bytes(123, 123 * 4, 4) == 123 * 4
    with pytest.raises(RuntimeError):
        chmod.align_chunk_to_scanlines_bytes(4096, 4096, 2)
    assert chmod.ensure_compression_level(12) == 9 and chmod.ensure_compression_level(0) == 0


def test_lazy_channel_needs_no_device():
    """full / zeros channels are lazy fill-value chunks (lazyschunk.h): pure host metadata until decompressed."""
    ch = chmod.Channel.full(np.float32, 25, 1920, 1080, compression_codec=Codec.blosclz, compression_level=2)
    assert ch.shape == (1080, 1920) and ch.uncompressed_size() == 1920 * 1080
    assert ch.num_chunks() == -(-1920 * 1080 * 4 // ch.chunk_size())
    assert ch.compressed_bytes() == 4 * ch.num_chunks()          # lazyschunk.h:312-327
    assert ch.chunk_elems(ch.num_chunks() - 1) * 4 == 1920 * 1080 * 4 - (ch.num_chunks() - 1) * ch.chunk_size()
    assert np.array_equal(ch.get_decompressed(), np.full((1080, 1920), 25, np.float32))
    with pytest.raises(IndexError):
        ch.get_chunk(ch.num_chunks())
    with pytest.raises(ValueError):
        chmod.Channel.zeros(np.bool_, 1, 1)
