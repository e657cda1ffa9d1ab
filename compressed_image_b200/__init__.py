"""compressed_image_b200 -- B200-native chunk compress / decompress engine behind the compressed-image API.

`Channel`, `Image`, `Codec` mirror the reference's Python module (python/py_module/compressed_image); the engine
underneath is libb2g.so (C ABI: include/b2g.h), hand-written sm_100a CUDA.  No CPU fallback exists.
"""
from .enums import Codec  # noqa: F401
from .channel import Channel  # noqa: F401
from .image import Image  # noqa: F401
from . import _b2g  # noqa: F401

__all__ = ["Codec", "Channel", "Image"]
