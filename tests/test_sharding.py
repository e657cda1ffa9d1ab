"""Multi-GPU sharding logic on CPU: partition properties, and the N > 1 path with world_size 2 over gloo.

There is no GPU here, so the per-rank ENGINE of these tests is the oracle (test infrastructure: tests/_libs.py); what
is under test is compressed_image_b200/sharding.py -- ranges, ownership, the no-collective operations and the one
all-gather behind get_decompressed().  On the GPU box the same class runs over the b2g context (test_gpu_api.py)."""
import os
import socket
import sys

import numpy as np
import pytest

from compressed_image_b200.sharding import ShardedChannel, shard_of, shard_ranges


def test_ranges_cover_and_balance():
    rng = np.random.default_rng(0)
    for n in (0, 1, 2, 3, 7, 64, 1000):
        for world in (1, 2, 3, 4, 8):
            w = rng.integers(1, 1 << 22, size=n)
            r = shard_ranges(w, world)
            assert len(r) == world
            assert r[0][0] == 0 and r[-1][1] == n
            for a, b in zip(r, r[1:]):
                assert a[1] == b[0] and a[0] <= a[1]
            if n >= 8 * world:
                loads = [int(w[lo:hi].sum()) for lo, hi in r]
                assert max(loads) <= w.sum() / world + w.max()
    assert shard_ranges([0, 0, 0, 0], 2) == [(0, 2), (2, 4)]
    assert shard_ranges([5], 4)[-1][1] == 1


def test_equal_chunks_split_evenly():
    assert shard_ranges([4 << 20] * 64, 8) == [(8 * r, 8 * r + 8) for r in range(8)]
    r = shard_ranges([4 << 20] * 64, 8)
    assert [shard_of(i, r) for i in (0, 7, 8, 63)] == [0, 0, 1, 7]
    with pytest.raises(IndexError):
        shard_of(64, r)


def _engine():
    import _libs
    LZ4, SHUFFLE = 1, 1

    def enc(arrays):
        out = []
        for a in arrays:
            r, c = _libs.oracle_compress(LZ4, 9, 2, SHUFFLE, 32768, a)
            assert r > 0
            out.append(c)
        return out

    def dec(chunks, outs):
        for c, o in zip(chunks, outs):
            r, b = _libs.oracle_decompress(c)
            assert r == o.size
            o[:] = b
    return enc, dec


def _channel_data(nchunks=7, chunk=40000):
    rng = np.random.default_rng(42)
    sizes = [chunk] * (nchunks - 1) + [chunk // 2 + 6]
    raw = [(np.cumsum(rng.integers(0, 3, size=s // 2)) & 0x3ff).astype(np.uint16).view(np.uint8) for s in sizes]
    return sizes, raw


def _worker(rank, world, port, q):
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        enc, dec = _engine()
        sizes, raw = _channel_data()
        sc = ShardedChannel(sizes, rank, world, decode=dec, encode=enc)
        sc.set_local(raw[sc.lo:sc.hi])                      # compress: no collective
        mine = sc.lo
        got = sc.get_chunk(mine)                            # random access on the owner: no collective
        ok = np.array_equal(got, raw[mine])
        other = (sc.hi) % len(sizes) if world > 1 else None
        raised = False
        if other is not None and sc.owner(other) != rank:
            try:
                sc.get_chunk(other)
            except IndexError:
                raised = True
        else:
            raised = True
        mod = (raw[mine].astype(np.uint16) + 1).astype(np.uint8)
        sc.set_chunk(mine, mod)
        raw[mine] = mod
        whole = sc.get_decompressed()                       # the one exchange step
        # every rank modified its first chunk: rebuild the expectation from all ranks' edits
        exp = list(_channel_data()[1])
        for lo, hi in sc.ranges:
            if hi > lo:
                exp[lo] = (exp[lo].astype(np.uint16) + 1).astype(np.uint8)
        q.put((rank, bool(ok), raised, bool(np.array_equal(whole, np.concatenate(exp)))))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2])
def test_sharded_channel_gloo(world):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert sorted(r[0] for r in res) == list(range(world))
    for r in res:
        assert r[1] and r[2] and r[3], r


def test_single_rank_no_process_group():
    enc, dec = _engine()
    sizes, raw = _channel_data(3)
    sc = ShardedChannel(sizes, 0, 1, decode=dec, encode=enc)
    sc.set_local(raw)
    assert np.array_equal(sc.get_decompressed(), np.concatenate(raw))
    with pytest.raises(ValueError):
        sc.set_chunk(0, raw[0][:10])
