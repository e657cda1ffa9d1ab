"""Multi-GPU sharding of the chunk path: one process per GPU, chunks dealt over ranks, no collective on
the data path.

Every blosc2 chunk is self-describing and independent, so compress / decompress / get_chunk / set_chunk need no exchange:
a batch is flattened to chunks, every rank takes one contiguous range balanced by bytes and runs its own b2g context on
its own GPU.  The single exchange step is `get_decompressed()` of a channel whose chunks live on several GPUs: each rank
decodes its range into its slice of a flat buffer -- on a GPU box a DEVICE buffer, decoded in place by a device-resident batch --
and ONE all-gather (NCCL over NVLink on the GPU box; any torch.distributed backend works, the CPU tests use gloo)
assembles the channel.  The reference has nothing comparable
(single process, compressed_image/include/compressed/channel.h:539-566 decodes the chunks in a loop).
"""
import numpy as np


def shard_ranges(weights, world):
    """Contiguous ranges [lo, hi) per rank over len(weights) chunks, balanced by the weights (bytes per chunk:
    nbytes for compress, nbytes + cbytes for decompress).  Every chunk belongs to exactly one rank; ranks may be empty
    when there are fewer chunks than ranks.  Greedy on the prefix sum: rank r ends at the first chunk whose inclusive
    prefix reaches (r + 1) / world of the total."""
    n = len(weights)
    world = max(int(world), 1)
    pre = np.concatenate(([0], np.cumsum(np.asarray(weights, dtype=np.float64))))
    total = pre[-1]
    bounds = [0]
    for r in range(1, world):
        if total <= 0:
            cut = (n * r) // world
        else:
            cut = int(np.searchsorted(pre, total * r / world, side="left"))
        cut = min(max(cut, bounds[-1]), n)
        bounds.append(cut)
    bounds.append(n)
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def shard_of(index, ranges):
    """Rank that owns chunk `index`."""
    for r, (lo, hi) in enumerate(ranges):
        if lo <= index < hi:
            return r
    raise IndexError("chunk index %d is not in any shard" % index)


class ShardedChannel:
    """The chunks of one channel, spread over the ranks of a torch.distributed process group.

    `chunk_nbytes[i]` is the decompressed size of chunk i (known on every rank: it follows from width, height and the
    chunk size, channel.h:101-104); the chunk BYTES exist only on the owning rank.  `decode(chunks, outs)` is the batched
    decompress of the local engine (Channel._ctx().decompress on a GPU box); `encode(arrays)` the batched compress.
    """

    def __init__(self, chunk_nbytes, rank, world, decode, encode=None, group=None, decode_device=None):
        self.chunk_nbytes = [int(b) for b in chunk_nbytes]
        self.rank, self.world = int(rank), int(world)
        self.ranges = shard_ranges(self.chunk_nbytes, self.world)
        self.lo, self.hi = self.ranges[self.rank]
        self._decode, self._encode = decode, encode
        self._group = group
        self._decode_device = decode_device
        self.local_chunks = [None] * (self.hi - self.lo)

    # ---- no-collective operations ------------------------------------------------------------------------------
    def owner(self, index):
        return shard_of(index, self.ranges)

    def set_local(self, arrays):
        """Compress this rank's range (one batched launch); `arrays` are the raw chunks lo..hi-1."""
        if len(arrays) != self.hi - self.lo:
            raise ValueError("expected %d chunks for rank %d, got %d" % (self.hi - self.lo, self.rank, len(arrays)))
        for k, a in enumerate(arrays):
            if np.asarray(a).nbytes != self.chunk_nbytes[self.lo + k]:
                raise ValueError("Invalid chunk size for chunk %d" % (self.lo + k))
        self.local_chunks = list(self._encode(list(arrays))) if arrays else []

    def set_chunk(self, index, array):
        if self.owner(index) != self.rank:
            raise IndexError("chunk %d lives on rank %d" % (index, self.owner(index)))
        if np.asarray(array).nbytes != self.chunk_nbytes[index]:
            raise ValueError("Invalid chunk size for chunk %d" % index)
        self.local_chunks[index - self.lo] = self._encode([array])[0]

    def get_chunk(self, index):
        if self.owner(index) != self.rank:
            raise IndexError("chunk %d lives on rank %d" % (index, self.owner(index)))
        out = np.empty(self.chunk_nbytes[index], dtype=np.uint8)
        self._decode([self.local_chunks[index - self.lo]], [out])
        return out

    def decode_local(self, out=None):
        """This rank's range decoded into one flat uint8 buffer (one batched launch)."""
        nb = self.chunk_nbytes[self.lo:self.hi]
        if out is None:
            out = np.empty(sum(nb), dtype=np.uint8)
        views, o = [], 0
        for b in nb:
            views.append(out[o:o + b])
            o += b
        if views:
            self._decode(self.local_chunks, views)
        return out

    # ---- the one exchange step -----------------------------------------------------------------------------------
    def get_decompressed(self, device=None, decode_device=None, as_tensor=False):
        """Every rank returns the whole channel as a flat uint8 array: local decode, then ONE all-gather of the per-rank
        segments (padded to the longest one so a single all_gather_into_tensor moves everything).

        On a GPU box pass `decode_device(chunks, device_ptrs)` (shard_channel does): the rank's chunks are decoded
        straight INTO its slice of a device buffer and the all-gather (NCCL over NVLink) runs on device pointers; nothing
        of the decoded data touches host memory before the gather.  Without it (CPU tests over gloo) the local decode
        lands in host memory first.  `as_tensor` returns the gathered torch tensor (on `device`) instead of numpy."""
        import torch
        import torch.distributed as dist

        seg = [sum(self.chunk_nbytes[lo:hi]) for lo, hi in self.ranges]
        pad = max(max(seg), 1)
        dev = torch.device(device) if device is not None else torch.device("cpu")
        gathered = torch.empty(pad * self.world, dtype=torch.uint8, device=dev)
        mine = gathered[self.rank * pad:(self.rank + 1) * pad]       # decode in place: rank r's slice of the result
        if decode_device is None:
            decode_device = self._decode_device
        if decode_device is not None and dev.type == "cuda" and self.hi > self.lo:
            ptrs, o = [], 0
            for b in self.chunk_nbytes[self.lo:self.hi]:
                ptrs.append(mine.data_ptr() + o)
                o += b
            decode_device(self.local_chunks, ptrs, self.chunk_nbytes[self.lo:self.hi])
        else:
            local = self.decode_local()
            if local.size:
                mine[:local.size] = torch.from_numpy(local).to(dev)
        if self.world > 1:
            dist.all_gather_into_tensor(gathered, mine.clone() if dist.get_backend(self._group) == "gloo" else mine,
                                        group=self._group)
        if as_tensor:
            if all(s == pad for s in seg):
                return gathered
            return torch.cat([gathered[r * pad:r * pad + seg[r]] for r in range(self.world)])
        g = gathered.cpu().numpy()
        return np.concatenate([g[r * pad:r * pad + seg[r]] for r in range(self.world)])


def device_decoder(ctx):
    """decode_device(chunks, device_ptrs, nbytes) on a _b2g.Context: the compressed chunks (host numpy) are uploaded into
    one device buffer ONCE -- a sharded channel keeps its range of chunk bytes resident in HBM until a chunk changes --
    and decoded by ONE device-resident batch (b2g_decompress_batch_device) into the given device addresses."""
    from . import _b2g
    cache = {}

    def decode(chunks, ptrs, nbytes):
        import torch
        key = tuple(id(c) for c in chunks)
        if cache.get("key") != key:
            sizes = [int(c.size) for c in chunks]
            offs, total = [], 0
            for sz in sizes:
                offs.append(total)
                total += (sz + 255) // 256 * 256
            host = torch.empty(max(total, 1), dtype=torch.uint8).pin_memory()
            hv = host.numpy()
            for c, o, sz in zip(chunks, offs, sizes):
                hv[o:o + sz] = c
            dev = host.cuda(non_blocking=True)
            torch.cuda.current_stream().synchronize()  # the batch runs on the context's own stream: the upload has landed
            cache.update(key=key, keep=list(chunks), dev=dev, offs=offs, sizes=sizes)
        dev, offs, sizes = cache["dev"], cache["offs"], cache["sizes"]
        torch.cuda.current_stream().synchronize()      # whatever the caller enqueued on the destination is done
        b = ctx.make_batch([dev.data_ptr() + o for o in offs], list(ptrs), sizes, [int(x) for x in nbytes])
        ctx.decompress_batch_device(b)
        ctx.wait()                                     # ... and the decoded bytes are there before the caller's all-gather
        for i in range(len(chunks)):
            if b[i].status < 0:
                raise _b2g.B2GError(b[i].status, "decompress chunk %d" % i)
    return decode


def shard_channel(channel, rank, world, group=None):
    """ShardedChannel over the chunks of a compressed_image_b200.Channel that exists on every rank (e.g. built from the
    same file): each rank keeps only its own range of chunk bytes and decodes / encodes it on its own GPU."""
    ctx = channel._ctx()
    nbytes = [channel._nbytes_of(i) for i in range(channel.num_chunks())]
    sc = ShardedChannel(nbytes, rank, world, decode=lambda cs, outs: ctx.decompress(cs, outs),
                        encode=lambda arrs: ctx.compress(arrs), group=group, decode_device=device_decoder(ctx))
    sc.local_chunks = channel.chunks(special=True)[sc.lo:sc.hi]   # lazy chunks travel as 32-byte special-value chunks
    return sc
