"""Seeded test inputs + the parameter grids shared by the oracle tests, the simulator tests and the GPU tests."""
import numpy as np

from compressed_image_b200 import synth

BLOSCLZ, LZ4, LZ4HC, ZSTD = 0, 1, 2, 5
NOFILTER, SHUFFLE, BITSHUFFLE = 0, 1, 2


def make_data(kind, nbytes, typesize, seed=1):
    """nbytes of test data of a given character, viewed as bytes."""
    r = np.random.Generator(np.random.PCG64(seed * 7919 + nbytes + typesize * 131))
    if kind == "zeros":
        return np.zeros(nbytes, dtype=np.uint8)
    if kind == "const":
        pat = (np.arange(typesize, dtype=np.uint8) * 37 + 5).astype(np.uint8)
        return np.resize(pat, nbytes).copy()
    if kind == "random":
        return r.integers(0, 256, size=nbytes, dtype=np.uint8)
    if kind == "iota":  # the reference's own test input (test_schunk.cpp:39-74, compat/filegen.c)
        ne = nbytes // typesize
        if typesize in (1, 2, 4, 8):
            v = np.arange(ne, dtype=np.uint64).astype({1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[typesize])
            b = v.view(np.uint8)
        else:
            b = np.zeros((ne, typesize), dtype=np.uint8)
            idx = np.arange(ne, dtype=np.uint64)
            for j in range(min(typesize, 8)):
                b[:, j] = (idx >> np.uint64(8 * j)).astype(np.uint8)
            b = b.reshape(-1)
        return np.concatenate([b, np.arange(nbytes - b.size, dtype=np.uint8)])
    if kind == "ramp":  # gradient + 2 noise bits per element (cfg 1 / 2 flavour)
        ne = nbytes // typesize + 1
        base = (np.arange(ne, dtype=np.uint64) >> np.uint64(3)) + r.integers(0, 4, size=ne, dtype=np.uint64)
        b = np.zeros((ne, typesize), dtype=np.uint8)
        for j in range(min(typesize, 8)):
            b[:, j] = (base >> np.uint64(8 * j)).astype(np.uint8)
        return b.reshape(-1)[:nbytes].copy()
    if kind == "text":  # repetitive, many short matches at mixed distances
        words = [bytes(r.integers(97, 123, size=int(k), dtype=np.uint8)) for k in r.integers(2, 9, size=64)]
        out = bytearray()
        while len(out) < nbytes:
            out += words[int(r.integers(0, 64))] + b" "
        return np.frombuffer(bytes(out[:nbytes]), dtype=np.uint8).copy()
    if kind == "sparse":  # mostly zeros with isolated values
        b = np.zeros(nbytes, dtype=np.uint8)
        k = max(1, nbytes // 50)
        b[r.integers(0, nbytes, size=k)] = r.integers(1, 256, size=k, dtype=np.uint8)
        return b
    if kind == "half":
        w = 512
        h = max(1, (nbytes // 2 + w - 1) // w)
        return synth.half_channel(w, h, 0, seed=seed).reshape(-1).view(np.uint8)[:nbytes].copy()
    if kind == "f32":
        w = 256
        h = max(1, (nbytes // 4 + w - 1) // w)
        return synth.f32_channel(w, h, "beauty", seed=seed).reshape(-1).view(np.uint8)[:nbytes].copy()
    if kind == "sensor":
        w = 512
        h = max(1, (nbytes // 2 + w - 1) // w)
        return synth.u16_sensor_channel(w, h, 0, seed=seed).reshape(-1).view(np.uint8)[:nbytes].copy()
    if kind == "longrun":  # long runs of a few values with random run lengths (overlapping matches)
        out = np.empty(nbytes, dtype=np.uint8)
        i = 0
        while i < nbytes:
            n = int(r.integers(1, 700))
            out[i:i + n] = r.integers(0, 4)
            i += n
        return out
    raise ValueError(kind)


KINDS = ("zeros", "const", "random", "iota", "ramp", "text", "sparse", "half", "f32", "sensor", "longrun")


def case_key(c):
    return "{kind}_n{nbytes}_ts{typesize}_c{codec}_l{clevel}_f{filter}_b{blocksize}".format(**c)


def make_case(c):
    return make_data(c["kind"], c["nbytes"], c["typesize"], seed=c.get("seed", 1))


def _grid():
    g = []

    def add(kind, nbytes, typesize, codec, clevel=9, filt=SHUFFLE, blocksize=32768):
        g.append(dict(kind=kind, nbytes=nbytes, typesize=typesize, codec=codec, clevel=clevel, filter=filt,
                      blocksize=blocksize))

    for codec in (LZ4, BLOSCLZ, ZSTD, LZ4HC):
        for ts, kind in ((1, "ramp"), (2, "half"), (4, "f32"), (2, "sensor"), (4, "iota"), (8, "iota"), (1, "text")):
            add(kind, 100000, ts, codec)
        add("zeros", 70000, 4, codec)
        add("const", 70000, 4, codec)
        add("random", 70000, 2, codec)
        add("sparse", 131072, 4, codec)
        add("longrun", 65536, 1, codec)
        add("f32", 131072 + 36, 4, codec, filt=BITSHUFFLE)
        add("half", 65536 + 6, 2, codec, filt=BITSHUFFLE)
        add("text", 50000, 1, codec, filt=NOFILTER)
        add("iota", 20, 4, codec)           # < 32 bytes: memcpyed chunk (b2c:2400-2407)
        add("iota", 7 * 13, 7, codec)       # odd typesize
        add("ramp", 3 * 1000 + 2, 3, codec, blocksize=999)
        add("iota", 16 * 5000, 16, codec)
        add("iota", 24 * 3000, 24, codec)   # typesize > 16: never split
        add("text", 300000, 1, codec, blocksize=131072)
    for lvl in (1, 5):
        add("sensor", 100000, 2, ZSTD, clevel=lvl)      # zstd level <= 5 -> split streams
        add("f32", 100000, 4, BLOSCLZ, clevel=lvl)
    add("ramp", 100000, 2, LZ4, clevel=0)               # clevel 0 -> memcpyed
    return g


REF_CHUNK_GRID = _grid()
