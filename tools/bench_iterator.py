"""The reference's "modify" usage (benchmark/main.cpp:139-157: iterate the chunks of a channel, change every value,
let the iterator recompress) on one 8192 x 4096 half channel (64 MiB, 16 chunks of 4 MiB): the double-buffered iterator
against the same loop written with blocking get_chunk / set_chunk.  Host numpy work included; median of 5."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import compressed_image_b200 as c  # noqa: E402
from compressed_image_b200 import synth  # noqa: E402

data = synth.half_channel(8192, 4096, 0)
ch = c.Channel(data, 8192, 4096)


def modify(a):
    v = a.view(np.uint16)
    v ^= 1


def run_iter():
    for chunk in ch:
        modify(chunk)


def run_blocking():
    for i in range(ch.num_chunks()):
        a = ch.get_chunk(i)
        modify(a)
        ch.set_chunk(i, a)


res = {}
for name, fn in (("iterator_double_buffered", run_iter), ("blocking_get_set", run_blocking)):
    fn()
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    ms = sorted(ts)[2] * 1e3
    res[name] = {"ms": round(ms, 2), "GBps_raw": round(data.nbytes / ms / 1e6, 2)}
t0 = time.perf_counter()
for _ in range(5):
    for i in range(16):
        modify(data.reshape(-1)[i * (1 << 21):(i + 1) * (1 << 21)])
res["uncompressed_loop_ms"] = round((time.perf_counter() - t0) / 5 * 1e3, 2)
print(json.dumps(res))
