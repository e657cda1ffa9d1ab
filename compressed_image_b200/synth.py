"""Deterministic synthetic image channels for the benchmark configs (SURVEY.md section 8d / BASELINE.json configs).

Pure numpy; used by bench.py and the tests.  Every generator returns a C-contiguous 2-D array (height, width)
of the channel's dtype; `half` data is carried as float16.
"""
import numpy as np


def _rng(seed, image=0, channel=0):
    return np.random.Generator(np.random.PCG64(0xC0FFEE + 1000 * image + channel + (seed << 20)))


def u8_rgba_channel(width, height, channel, seed=0, image=0):
    """cfg 1: R,G,B = ((x>>3) + (y>>3)*k_c + 2 noise bits) & 255 ; A = 255."""
    if channel == 3:
        return np.full((height, width), 255, dtype=np.uint8)
    r = _rng(seed, image, channel)
    x = np.arange(width, dtype=np.uint32)[None, :] >> 3
    y = np.arange(height, dtype=np.uint32)[:, None] >> 3
    k = (1, 2, 3)[channel]
    noise = r.integers(0, 4, size=(height, width), dtype=np.uint32)
    return ((x + y * k + noise) & 255).astype(np.uint8)


def half_channel(width, height, channel, seed=0, image=0):
    """cfg 2: smooth radial gradient in [0,1] + 2^-9 uniform noise as IEEE half; channel 3 = elliptical alpha."""
    x = (np.arange(width, dtype=np.float32)[None, :] / max(width - 1, 1)) * 2 - 1
    y = (np.arange(height, dtype=np.float32)[:, None] / max(height - 1, 1)) * 2 - 1
    if channel == 3:
        return ((x / 0.8) ** 2 + (y / 0.6) ** 2 <= 1.0).astype(np.float16)
    r = _rng(seed, image, channel)
    cx, cy = (0.1, -0.2, 0.3)[channel % 3], (-0.15, 0.25, 0.05)[channel % 3]
    g = np.clip(1.0 - np.sqrt((x - cx) ** 2 + (y - cy) ** 2) / 1.6, 0, 1)
    g = g + r.random((height, width), dtype=np.float32) * (2.0 ** -9)
    return np.clip(g, 0, 1).astype(np.float16)


def f32_channel(width, height, kind, seed=0, image=0, channel=0):
    """cfg 3: kind in {'beauty','sparse','id','depth'}."""
    r = _rng(seed, image, channel)
    x = np.arange(width, dtype=np.float32)[None, :] / max(width - 1, 1)
    y = np.arange(height, dtype=np.float32)[:, None] / max(height - 1, 1)
    if kind == "beauty":  # gradient with noise in the low 10 mantissa bits
        g = (0.25 + 0.5 * x + 0.25 * y).astype(np.float32)
        bits = g.view(np.uint32) & np.uint32(0xFFFFFC00)
        bits = bits | r.integers(0, 1024, size=(height, width), dtype=np.uint32)
        return bits.view(np.float32)
    if kind == "sparse":  # 92 % exact zeros
        v = r.random((height, width), dtype=np.float32)
        m = r.random((height, width), dtype=np.float32) < 0.92
        v[m] = 0.0
        return v
    if kind == "id":  # piece-wise constant per 64x64 tile, float-encoded hash
        tx = (np.arange(width, dtype=np.uint32)[None, :] >> 6)
        ty = (np.arange(height, dtype=np.uint32)[:, None] >> 6)
        h = (tx * np.uint32(2654435761) ^ (ty * np.uint32(40503) + np.uint32(channel * 977))) & np.uint32(0xFFFFF)
        return h.astype(np.float32)
    if kind == "depth":  # planar ramp + 1e-4 noise
        return (1.0 + 3.0 * x + 2.0 * y + 1e-4 * r.random((height, width), dtype=np.float32)).astype(np.float32)
    raise ValueError(kind)


F32_KINDS = ("beauty", "sparse", "id", "depth")


def u16_sensor_channel(width, height, channel, seed=0, image=0):
    """cfg 4: 12-bit sensor-like clip(gradient + triangular noise sigma~4, 0, 4095)."""
    r = _rng(seed, image, channel)
    x = np.arange(width, dtype=np.float32)[None, :] / max(width - 1, 1)
    y = np.arange(height, dtype=np.float32)[:, None] / max(height - 1, 1)
    g = 200 + 3000 * (0.6 * x + 0.4 * y) * (1.0 + 0.1 * channel)
    n = (r.random((height, width), dtype=np.float32) - r.random((height, width), dtype=np.float32)) * 9.8
    return np.clip(g + n, 0, 4095).astype(np.uint16)


def aligned_chunk_bytes(chunk_size, width, itemsize):
    """compressed::util::align_chunk_to_scanlines_bytes (compressed_image/include/compressed/util.h:115-131)."""
    scan = width * itemsize
    n = (chunk_size // scan) * scan
    return n if n > 0 else scan


def split_chunks(arr, chunk_size=4 * 1024 * 1024):
    """Split a 2-D channel into scanline-aligned chunk views (bytes), as compressed::channel does."""
    h, w = arr.shape
    flat = arr.reshape(-1).view(np.uint8)
    cb = aligned_chunk_bytes(chunk_size, w, arr.dtype.itemsize)
    return [flat[o:o + cb] for o in range(0, flat.size, cb)]
