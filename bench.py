#!/usr/bin/env python
"""bench.py -- chunk decompress / compress throughput of the b2g engine on B200 (contract: DESIGN.md section 6).

HEADLINE (value, e2e, roofline, cpu_baseline): BASELINE.json configs[1], the configuration the metric is quoted on:
compressed::image<half>, synthetic 8192x4096, 4 channels, lz4 + byte-shuffle, default 4 MiB chunks / 32 KiB blocks ->
64 chunks, 256 MiB.  One STEP is the random-access pass of that config: get_chunk over all 64 chunks in a seeded
shuffled order (one batched decompress launch) followed by set_chunk over all of them (one batched compress launch).

  value          uncompressed GB/s over the step (bytes decompressed + bytes compressed, / step time), chunks and raw
                 image resident in HBM, CUDA events on the launching stream.  Inputs exceed the 126 MB L2.
  e2e            the same step through the C ABI with HOST (pinned) buffers: H2D of chunks / raw data and D2H of the
                 results inside the timed region; e2e.link_gbs is this box's measured pinned H2D || D2H ceiling in the same
                 run (all ranks copying at once at N > 1), e2e.frac_of_link the step's PCIe traffic against it.
  e2e_pageable   the same step with plain numpy buffers -- what compressed::channel<T> / Channel really hand over.
  configs        one entry per BASELINE.json config (cfg 1 ... cfg 5): kernel-scope GB/s per direction, roofline fraction
                 per kernel, end-to-end GB/s with host buffers, the reference on this box's host cores, both ratios.
  strong         N > 1: cfg 4's batch of 256 images split over the ranks by chunk (shard_ranges), max-over-ranks time.
  gather         N > 1: one 256 MiB channel decoded on N GPUs into device memory + ONE NCCL all-gather on device pointers.

`--impl reference` times the reference's own CPU implementation (oracle/_ref/libblosc2_ref.so = unmodified c-blosc2, one
nthreads=1 context per host thread, chunks dealt over all cores) on the same workload, the same `config`, and the same
per-config samples.
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BLOSCLZ, LZ4, ZSTD = 0, 1, 5
NOFILTER, SHUFFLE, BITSHUFFLE = 0, 1, 2
WIDTH, HEIGHT, NCHAN = 8192, 4096, 4
CHUNK, BLOCK, CLEVEL, TYPESIZE = 4 * 1024 * 1024, 32768, 9, 2
WORKLOAD = ("compressed::image<half> synthetic 8192x4096 4-channel, lz4 + byte-shuffle, chunk 4 MiB / block 32 KiB: "
            "get_chunk + set_chunk over all 64 chunks in seeded random order")
METRIC = "chunk decompress+compress GB/s (uncompressed)"


def headline_config(gpus):
    """The `config` object of the JSON line: the same for both arms (the reference arm runs it on the host cores)."""
    return {"workload": WORKLOAD, "chunks_per_gpu": 64, "bytes_per_step_per_gpu": 2 * WIDTH * HEIGHT * NCHAN * TYPESIZE,
            "l2": "inputs (256 MiB raw + chunks per direction) exceed the 126 MB L2; no explicit flush",
            "parallelism": "chunks sharded by image across %d GPU(s), no collective on the data path" % gpus}


def make_image(image_index):
    from compressed_image_b200 import synth
    chans = [synth.half_channel(WIDTH, HEIGHT, c, image=image_index) for c in range(NCHAN)]
    pieces = []
    for a in chans:
        pieces += synth.split_chunks(a, CHUNK)
    return pieces   # 64 uint8 views of 4 MiB


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (NVML; nvidia-smi's clocks line carries the same fields)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def sample(self):
        if not self.nv:
            return
        nv = self.nv
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                     0x80: "hw_power_brake_slowdown"}
            for bit, name in names.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def run(self):
        while not self._stop_evt.is_set():
            self.sample()
            time.sleep(0.004)

    def stop(self):
        self._stop_evt.set()
        self.join()
        self.sample()
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ---------------------------------------------------------------------------------------------------------------------
# The workloads of BASELINE.json configs[0..4] (SURVEY.md section 8d generators).  Every case is a list of raw chunk
# views (`pieces`, numpy uint8) + codec parameters; `replicas` says how many copies of the list make up the configured
# batch on the device (each copy is its own HBM buffer, so the kernels stream the full configured size).
# ---------------------------------------------------------------------------------------------------------------------
def case_cfg1():
    from compressed_image_b200 import synth
    p = []
    for k in range(4):
        p += synth.split_chunks(synth.u8_rgba_channel(4096, 4096, k), CHUNK)
    return {"name": "cfg1 image<uint8> 4096x4096 RGBA lz4+shuffle, compress then get_decompressed", "pieces": p,
            "codec": LZ4, "clevel": 9, "ts": 1, "filter": SHUFFLE, "replicas": 1}


def case_cfg2():
    return {"name": "cfg2 image<half> 8192x4096x4 lz4+shuffle, get_chunk/set_chunk over all chunks", "pieces": make_image(0),
            "codec": LZ4, "clevel": 9, "ts": 2, "filter": SHUFFLE, "replicas": 1}


def case_cfg3(filt=BITSHUFFLE, nchan=16):
    from compressed_image_b200 import synth
    p = []
    for k in range(nchan):
        p += synth.split_chunks(synth.f32_channel(8192, 8192, synth.F32_KINDS[k // (nchan // 4)], channel=k), CHUNK)
    return {"name": "cfg3 image<float> 8192x8192 x%d (beauty/sparse/id/depth) blosclz+%s" %
                    (nchan, "bitshuffle" if filt == BITSHUFFLE else "shuffle"),
            "pieces": p, "codec": BLOSCLZ, "clevel": 9, "ts": 4, "filter": filt, "replicas": 1}


CFG4_IMAGES, CFG4_DISTINCT = 256, 32


def cfg4_pieces(images):
    from compressed_image_b200 import synth
    p = []
    for im in images:
        for ch in range(3):
            p += synth.split_chunks(synth.u16_sensor_channel(2048, 2048, ch, image=im), CHUNK)
    return p   # 6 chunks of 4 MiB per image


def case_cfg4_lz4(distinct=CFG4_DISTINCT, images=CFG4_IMAGES):
    return {"name": "cfg4 batch of %d uint16 RGB 2048x2048 images (%d distinct synthetic images x %d copies, each copy its "
                    "own HBM buffer) lz4+shuffle re-encode" % (images, distinct, images // distinct),
            "pieces": cfg4_pieces(range(distinct)), "codec": LZ4, "clevel": 9, "ts": 2, "filter": SHUFFLE,
            "replicas": images // distinct}


def ref_lib():
    path = os.path.join(ROOT, "oracle", "_ref", "libblosc2_ref.so")
    if not os.path.exists(path):
        return None
    L = C.CDLL(path)
    pvp, pi32, ci = C.POINTER(C.c_void_p), C.POINTER(C.c_int32), C.c_int
    L.ref_compress_batch.argtypes = [ci, ci, ci, ci, ci, ci, pvp, pi32, pvp, pi32, pi32, ci]
    L.ref_decompress_batch.argtypes = [ci, pvp, pi32, pvp, pi32, pi32, ci]
    return L


def ref_compress(L, pieces, codec, clevel, ts, filt, nthreads):
    """Chunks of `pieces` produced by the reference on the host cores -> (list of numpy uint8 chunks, seconds)."""
    n, vp = len(pieces), C.c_void_p
    srcs = [np.ascontiguousarray(p) for p in pieces]
    dsts = [np.empty(s.size + 32, np.uint8) for s in srcs]
    out_c = (C.c_int32 * n)()
    a = ((vp * n)(*[s.ctypes.data for s in srcs]), (C.c_int32 * n)(*[int(s.size) for s in srcs]),
         (vp * n)(*[d.ctypes.data for d in dsts]), (C.c_int32 * n)(*[int(d.size) for d in dsts]))
    t0 = time.perf_counter()
    L.ref_compress_batch(codec, clevel, ts, filt, BLOCK, n, a[0], a[1], a[2], a[3], out_c, nthreads)
    dt = time.perf_counter() - t0
    assert all(v > 0 for v in out_c), "reference compress failed"
    return [d[:out_c[i]] for i, d in enumerate(dsts)], dt


def ref_decompress(L, chunks, nbytes, nthreads, repeat=2):
    """Best-of-`repeat` seconds for the reference to decompress `chunks` on the host cores; also the outputs."""
    n, vp = len(chunks), C.c_void_p
    outs = [np.empty(nb, np.uint8) for nb in nbytes]
    res = (C.c_int32 * n)()
    a = ((vp * n)(*[c.ctypes.data for c in chunks]), (C.c_int32 * n)(*[int(c.size) for c in chunks]),
         (vp * n)(*[o.ctypes.data for o in outs]), (C.c_int32 * n)(*[int(nb) for nb in nbytes]))
    best = None
    for _ in range(repeat):
        t0 = time.perf_counter()
        L.ref_decompress_batch(n, a[0], a[1], a[2], a[3], res, nthreads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    assert all(res[i] == nbytes[i] for i in range(n)), "reference decompress failed"
    return best, outs


def sample_of(pieces, max_bytes):
    """An evenly spread sample of the chunk list of at most ~max_bytes (the CPU legs are bounded, the GPU ones are not)."""
    total = sum(int(p.size) for p in pieces)
    if total <= max_bytes:
        return list(range(len(pieces)))
    stride = -(-total // max_bytes)
    return list(range(0, len(pieces), stride))


def cpu_case(L, case, cores, max_bytes=256 << 20, chunks=None):
    """The reference on this box's host cores on a bounded sample of the case -> dict (GB/s per direction, ratio)."""
    idx = sample_of(case["pieces"], max_bytes)
    pieces = [case["pieces"][i] for i in idx]
    nb = [int(p.size) for p in pieces]
    total = sum(nb)
    out = {"cores": cores, "kind": "reference",
           "sample": "%d of the case's %d distinct chunks (%.0f MiB), best of 2; unmodified c-blosc2 (oracle/_ref), one "
                     "nthreads=1 context per host thread" % (len(idx), len(case["pieces"]), total / 2 ** 20)}
    if chunks is None:
        best = None
        for _ in range(2):
            ch, dt = ref_compress(L, pieces, case["codec"], case["clevel"], case["ts"], case["filter"], cores)
            best = dt if best is None else min(best, dt)
        out["compress_gbs"] = total / best / 1e9
    else:
        ch = [chunks[i] for i in idx]
    dt, outs = ref_decompress(L, ch, nb, cores)
    assert np.array_equal(outs[0], pieces[0].reshape(-1).view(np.uint8))
    out["decompress_gbs"] = total / dt / 1e9
    out["ratio"] = total / sum(int(c.size) for c in ch)
    return out


# ---------------------------------------------------------------------------------------------------------------------
def gpu_case(case, peak, reps=5, chunks=None, e2e_bytes=512 << 20, check=True):
    """One config on the GPU.  Kernel scope: the whole configured batch (pieces x replicas) device-resident, median of
    `reps` launches per direction, CUDA events around the kernels inside the library.  End to end: a bounded sample of the
    batch through the host-pointer C ABI with plain (pageable) numpy buffers, as Channel / Image hand them over.
    chunks: reference-produced chunks (zstd) -> decode only."""
    import torch
    from compressed_image_b200 import _b2g

    pieces, R = case["pieces"], case["replicas"]
    n1 = len(pieces)
    n = n1 * R
    sizes1 = [int(p.size) for p in pieces]
    total1 = sum(sizes1)
    total = total1 * R
    raw = torch.empty(total1, dtype=torch.uint8)
    offs1, o = [], 0
    for p in pieces:
        raw[o:o + p.size] = torch.from_numpy(np.ascontiguousarray(p).reshape(-1).view(np.uint8))
        offs1.append(o)
        o += int(p.size)
    d_raw1 = raw.cuda()
    d_raw = d_raw1.repeat(R) if R > 1 else d_raw1
    offs = [r * total1 + o for r in range(R) for o in offs1]
    sizes = sizes1 * R
    slot = max(sizes1) + 256
    d_chunks = torch.zeros(n * slot, dtype=torch.uint8, device="cuda")
    d_out = torch.zeros(total, dtype=torch.uint8, device="cuda")
    ctx = _b2g.Context(case["codec"], case["clevel"], case["ts"], case["filter"], BLOCK)
    torch.cuda.synchronize()        # uploads and fills above ran on torch's stream, the batches below run on the context's
    res = {"name": case["name"], "chunks": n, "nbytes": total}
    comp_ms = None
    if chunks is None:
        comp = ctx.make_batch([d_raw.data_ptr() + offs[i] for i in range(n)], [d_chunks.data_ptr() + i * slot for i in range(n)],
                              sizes, [s + 32 for s in sizes])
        t = []
        for _ in range(reps + 1):
            ctx.compress_batch_device(comp)
            ctx.wait()
            t.append(sum(ctx.last_kernel_ms()))
        comp_ms = float(np.median(t[1:]))
        assert all(comp[i].status == 0 and comp[i].out_bytes > 0 for i in range(n)), "compress failed: " + case["name"]
        cb = [comp[i].out_bytes for i in range(n)]
    else:
        cb1 = [int(c.size) for c in chunks]
        stage = torch.zeros(n1 * slot, dtype=torch.uint8)
        for i, c in enumerate(chunks):
            stage[i * slot:i * slot + c.size] = torch.from_numpy(c)
        d1 = stage.cuda()
        for r in range(R):
            d_chunks[r * n1 * slot:(r + 1) * n1 * slot] = d1
        cb = cb1 * R
    dec = ctx.make_batch([d_chunks.data_ptr() + i * slot for i in range(n)], [d_out.data_ptr() + offs[i] for i in range(n)], cb,
                         sizes)
    torch.cuda.synchronize()
    t = []
    for _ in range(reps + 1):
        ctx.decompress_batch_device(dec)
        ctx.wait()
        t.append(sum(ctx.last_kernel_ms()))
    dec_ms = float(np.median(t[1:]))
    torch.cuda.synchronize()
    assert all(dec[i].status == 0 for i in range(n)), "decompress failed: " + case["name"]
    if check:
        assert torch.equal(d_out, d_raw), "round trip mismatch: " + case["name"]
    cbytes = sum(cb)
    res["ratio"] = total / cbytes
    res["decompress_gbs_kernel"] = total / dec_ms / 1e6
    res["decompress_ms"] = dec_ms
    res["roofline"] = {"k_decode": {"achieved": (total + cbytes) / dec_ms / 1e6, "frac": (total + cbytes) / dec_ms / 1e6 / peak}}
    if comp_ms:
        res["compress_gbs_kernel"] = total / comp_ms / 1e6
        res["compress_ms"] = comp_ms
        res["roofline"]["k_compress"] = {"achieved": (total + cbytes) / comp_ms / 1e6, "frac": (total + cbytes) / comp_ms / 1e6 / peak}
    res["launches"] = ctx.launches

    # ---- end to end with the buffers the API hands over (pageable numpy), bounded sample ------------------------------
    if e2e_bytes <= 0:
        ctx.close()
        del d_raw, d_chunks, d_out
        torch.cuda.empty_cache()
        return res
    idx = sample_of(pieces, e2e_bytes)
    h_chunks = [d_chunks[i * slot:i * slot + cb[i]].cpu().numpy() for i in idx]     # the first copy's chunks
    srcs = [np.ascontiguousarray(pieces[i]).reshape(-1).view(np.uint8) for i in idx]
    sb = sum(int(s.size) for s in srcs)
    outs = [np.empty(s.size, np.uint8) for s in srcs]
    best_d = best_c = None
    for _ in range(3):
        t0 = time.perf_counter()
        ctx.decompress(h_chunks, outs)
        dt = time.perf_counter() - t0
        best_d = dt if best_d is None else min(best_d, dt)
    assert np.array_equal(outs[-1], srcs[-1])
    e2e = {"decompress_gbs": sb / best_d / 1e9, "host_buffers": "pageable numpy", "sample_bytes": sb,
           "h2d_bytes": sum(int(c.size) for c in h_chunks), "d2h_bytes": sb,
           "calls": "b2g_decompress_batch / b2g_compress_batch (what Channel / Image call), best of 3"}
    if chunks is None:
        dsts = [np.zeros(s.size + 32, np.uint8) for s in srcs]      # allocated (and touched) once, like the CPU leg's
        cb_ = ctx.make_batch([s.ctypes.data for s in srcs], [d.ctypes.data for d in dsts], [int(s.size) for s in srcs],
                             [int(d.size) for d in dsts])
        for _ in range(3):
            t0 = time.perf_counter()
            ctx.compress_batch(cb_)
            dt = time.perf_counter() - t0
            best_c = dt if best_c is None else min(best_c, dt)
        assert all(cb_[i].status == 0 and cb_[i].out_bytes > 0 for i in range(len(srcs)))
        e2e["compress_gbs"] = sb / best_c / 1e9
        e2e["value"] = 2 * sb / (best_d + best_c) / 1e9
    else:
        e2e["value"] = e2e["decompress_gbs"]
    e2e["unit"] = "GB/s"
    res["e2e"] = e2e
    ctx.close()
    del d_raw, d_chunks, d_out
    torch.cuda.empty_cache()
    return res


def api_cfg1():
    """cfg 1 through the public API: Image ctor (compress all channels) + get_decompressed, numpy in / numpy out."""
    from compressed_image_b200 import Image, synth
    chans = [synth.u8_rgba_channel(4096, 4096, k) for k in range(4)]
    nbytes = sum(c.nbytes for c in chans)
    best = [None, None]
    for _ in range(3):
        t0 = time.perf_counter()
        img = Image(np.uint8, chans, 4096, 4096, ["R", "G", "B", "A"])
        t1 = time.perf_counter()
        out = img.get_decompressed()
        t2 = time.perf_counter()
        best[0] = t1 - t0 if best[0] is None else min(best[0], t1 - t0)
        best[1] = t2 - t1 if best[1] is None else min(best[1], t2 - t1)
    assert all(np.array_equal(a, b) for a, b in zip(chans, out))
    return {"calls": "Image(np.uint8, channels, 4096, 4096) + Image.get_decompressed(), numpy buffers, best of 3",
            "compress_gbs": nbytes / best[0] / 1e9, "decompress_gbs": nbytes / best[1] / 1e9,
            "value": 2 * nbytes / (best[0] + best[1]) / 1e9, "unit": "GB/s", "ratio": img.compression_ratio()}


def sweep_cfg5(peak, L, cores, reps=3, target=128 << 20, gpu=True):
    """cfg 5: chunk size x codec x element type.  Generators of cfg 1/2/3 reshaped to width 2048; one batch of `target`
    bytes per cell (SURVEY.md asks for >= 1 GiB per GPU; bounded here so that the default run stays within minutes)."""
    from compressed_image_b200 import synth
    out = []
    data = {}
    w = 2048
    for ts, gen in ((1, lambda h: synth.u8_rgba_channel(w, h, 1)), (2, lambda h: synth.half_channel(w, h, 1).view(np.uint16)),
                    (4, lambda h: synth.f32_channel(w, h, "depth", channel=12))):
        data[ts] = gen(target // (w * ts))
    for ts in (1, 2, 4):
        for kib in (64, 256, 1024, 4096, 16384):
            pieces = synth.split_chunks(data[ts], kib * 1024)
            for codec, label in ((LZ4, "lz4"), (BLOSCLZ, "blosclz"), (ZSTD, "zstd-decode")):
                case = {"name": "cfg5 %s %s %d KiB chunks" % ({1: "uint8", 2: "half", 4: "float"}[ts], label, kib),
                        "pieces": pieces, "codec": codec, "clevel": 5 if codec == ZSTD else 9, "ts": ts, "filter": SHUFFLE,
                        "replicas": 1}
                chunks = None
                cell = {"name": case["name"], "typesize": ts, "chunk_kib": kib, "codec": label}
                if codec == ZSTD:
                    if L is None:
                        continue
                    chunks, _ = ref_compress(L, pieces, ZSTD, 5, ts, SHUFFLE, cores)
                if gpu:
                    r = gpu_case(case, peak, reps=reps, chunks=chunks, e2e_bytes=0)
                    cell.update({k: r[k] for k in ("chunks", "ratio", "decompress_gbs_kernel", "compress_gbs_kernel") if k in r})
                    cell["decode_roofline_frac"] = r["roofline"]["k_decode"]["frac"]
                if L is not None:
                    c = cpu_case(L, case, cores, max_bytes=64 << 20, chunks=chunks)
                    cell["cpu_decompress_gbs"] = c["decompress_gbs"]
                    if "compress_gbs" in c:
                        cell["cpu_compress_gbs"] = c["compress_gbs"]
                    cell["cpu_ratio"] = c["ratio"]
                out.append(cell)
    return out


def round_floats(x, nd=4):
    if isinstance(x, float):
        return float("%.*g" % (nd + 2, x))
    if isinstance(x, dict):
        return {k: round_floats(v, nd) for k, v in x.items()}
    if isinstance(x, list):
        return [round_floats(v, nd) for v in x]
    return x


# ---------------------------------------------------------------------------------------------------------------------
def cpu_reference_pass(L, pieces, order, nthreads, repeat=1):
    """One step of the headline workload on the host cores with the reference library -> (seconds per step, chunks)."""
    n = len(order)
    srcs = [pieces[i] for i in order]
    dsts = [np.empty(s.size + 32, np.uint8) for s in srcs]
    vp = C.c_void_p
    a_src = (vp * n)(*[s.ctypes.data for s in srcs])
    a_dst = (vp * n)(*[d.ctypes.data for d in dsts])
    a_nb = (C.c_int32 * n)(*[int(s.size) for s in srcs])
    a_cap = (C.c_int32 * n)(*[int(d.size) for d in dsts])
    out_c = (C.c_int32 * n)()
    L.ref_compress_batch(LZ4, CLEVEL, TYPESIZE, SHUFFLE, BLOCK, n, a_src, a_nb, a_dst, a_cap, out_c, nthreads)
    assert all(v > 0 for v in out_c)
    chunks = [d[:out_c[i]] for i, d in enumerate(dsts)]
    outs = [np.empty(s.size, np.uint8) for s in srcs]
    a_out = (vp * n)(*[o.ctypes.data for o in outs])
    a_cb = (C.c_int32 * n)(*[int(c.size) for c in chunks])
    out_n = (C.c_int32 * n)()
    best = None
    for _ in range(repeat):
        t0 = time.perf_counter()
        L.ref_decompress_batch(n, a_dst, a_cb, a_out, a_nb, out_n, nthreads)
        L.ref_compress_batch(LZ4, CLEVEL, TYPESIZE, SHUFFLE, BLOCK, n, a_src, a_nb, a_dst, a_cap, out_c, nthreads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    assert all(out_n[i] == srcs[i].size for i in range(n))
    assert np.array_equal(outs[0], srcs[0])
    return best, [d[:out_c[i]] for i, d in enumerate(dsts)]


def reference_configs(L, cores, quick):
    """The reference on the host cores for every config (bounded samples): the CPU side of `configs`."""
    out = []
    for make in (case_cfg1, case_cfg2):
        case = make()
        out.append({"name": case["name"], "cpu_baseline": cpu_case(L, case, cores)})
    if quick:
        return out
    for filt in (BITSHUFFLE, SHUFFLE):
        case = case_cfg3(filt, nchan=4)
        out.append({"name": case["name"], "cpu_baseline": cpu_case(L, case, cores)})
    case = case_cfg4_lz4(distinct=8, images=8)
    out.append({"name": case["name"], "cpu_baseline": cpu_case(L, case, cores)})
    for clevel in (5, 9):
        zc = dict(case, codec=ZSTD, clevel=clevel, pieces=case["pieces"][:48 if clevel == 5 else 12],
                  name="cfg4 zstd clevel %d chunks, decode" % clevel)
        chunks, _ = ref_compress(L, zc["pieces"], ZSTD, clevel, 2, SHUFFLE, cores)
        out.append({"name": zc["name"], "cpu_baseline": cpu_case(L, zc, cores, chunks=chunks)})
    return out


def run_reference(args, rank, world):
    if rank != 0:
        return
    L = ref_lib()
    cores = os.cpu_count() or 1
    if L is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libblosc2_ref.so was not built"}))
        return
    pieces = make_image(0)
    order = list(np.random.default_rng(1234).permutation(len(pieces)))
    times = []
    for i in range(args.warmup + args.steps):
        dt, chunks = cpu_reference_pass(L, pieces, order, cores)
        if i >= args.warmup:
            times.append(dt)
    step = sum(times) / len(times)
    nbytes = sum(int(p.size) for p in pieces)
    gbs = 2 * nbytes / step / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": headline_config(args.gpus),
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "reference",
                             "sample": "the full step (64 chunks decompressed + 64 compressed) per timed step; unmodified "
                                       "c-blosc2, one nthreads=1 context per host thread, chunks dealt over all threads"},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "ratio": nbytes / sum(int(c.size) for c in chunks)}
    if not args.headline_only:
        line["configs"] = reference_configs(L, cores, args.quick)
        if not args.quick:
            line["cfg5_sweep"] = sweep_cfg5(0.0, L, cores, gpu=False)
    print(json.dumps(round_floats(line)))


# ---------------------------------------------------------------------------------------------------------------------
def link_ceiling(torch, dist, nbytes=256 << 20, reps=3):
    """Pinned H2D || D2H bandwidth of this rank (GB/s, both directions summed) with every rank copying at once."""
    h1 = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    h2 = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d1 = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    d2 = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def once():
        with torch.cuda.stream(s1):
            d1.copy_(h1, non_blocking=True)
        with torch.cuda.stream(s2):
            h2.copy_(d2, non_blocking=True)
    once()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    if dist:
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return 2 * nbytes / dt / 1e9


def host_memcpy_gbs(threads, nbytes=64 << 20, reps=3):
    """What `threads` host threads copy per second between plain numpy buffers (GB/s of payload): the ceiling of any path
    that has to move pageable caller memory through pinned staging."""
    srcs = [np.ones(nbytes, np.uint8) for _ in range(threads)]
    dsts = [np.ones(nbytes, np.uint8) for _ in range(threads)]
    best = None
    for _ in range(reps):
        ts = [threading.Thread(target=np.copyto, args=(dsts[i], srcs[i])) for i in range(threads)]
        t0 = time.perf_counter()
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return threads * nbytes / best / 1e9


def strong_cfg4(torch, dist, rank, world, local_rank, L, cores):
    """cfg 4 strong scaling: ONE fixed batch of 256 images (1536 chunks) split over the ranks by chunk with shard_ranges;
    every rank decodes its zstd chunks (clevel 5, produced by the reference on the host cores) and re-encodes with lz4.
    Time = max over ranks of the device time of the rank's two launches."""
    from compressed_image_b200 import _b2g
    from compressed_image_b200.sharding import shard_ranges
    nchunks = CFG4_IMAGES * 6
    lo, hi = shard_ranges([CHUNK] * nchunks, world)[rank]
    pieces = cfg4_pieces(range(CFG4_DISTINCT))                  # chunk k of the batch = distinct chunk k % (32 * 6)
    nd = len(pieces)
    zchunks, _ = ref_compress(L, pieces, ZSTD, 5, 2, SHUFFLE, cores)
    slot = CHUNK + 256
    stage = torch.zeros(nd * slot, dtype=torch.uint8)
    for i, c in enumerate(zchunks):
        stage[i * slot:i * slot + c.size] = torch.from_numpy(c)
    d_z1 = stage.cuda()
    n = hi - lo
    d_z = torch.empty(n * slot, dtype=torch.uint8, device="cuda")
    for k in range(n):                                          # every chunk of the shard its own HBM copy
        j = (lo + k) % nd
        d_z[k * slot:(k + 1) * slot] = d_z1[j * slot:(j + 1) * slot]
    d_raw = torch.empty(n * CHUNK, dtype=torch.uint8, device="cuda")
    d_lz = torch.empty(n * slot, dtype=torch.uint8, device="cuda")
    zctx = _b2g.Context(ZSTD, 5, 2, SHUFFLE, BLOCK, local_rank)
    lctx = _b2g.Context(LZ4, 9, 2, SHUFFLE, BLOCK, local_rank)
    dec = zctx.make_batch([d_z.data_ptr() + k * slot for k in range(n)], [d_raw.data_ptr() + k * CHUNK for k in range(n)],
                          [int(zchunks[(lo + k) % nd].size) for k in range(n)], [CHUNK] * n)
    enc = lctx.make_batch([d_raw.data_ptr() + k * CHUNK for k in range(n)], [d_lz.data_ptr() + k * slot for k in range(n)],
                          [CHUNK] * n, [CHUNK + 32] * n)
    stream = torch.cuda.current_stream()
    sh = C.c_void_p(stream.cuda_stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def step():
        zctx.decompress_batch_device(dec, stream=sh)
        zctx.wait()
        lctx.compress_batch_device(enc, stream=sh)
        lctx.wait()
    step()
    assert all(dec[k].status == 0 for k in range(n)) and all(enc[k].status == 0 for k in range(n))
    ref0 = torch.from_numpy(np.ascontiguousarray(pieces[lo % nd]).reshape(-1).view(np.uint8)).cuda()
    assert torch.equal(d_raw[:CHUNK], ref0), "strong cfg4: decoded chunk differs from the source"
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    reps = 3
    e0.record(stream)
    for _ in range(reps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    if dist:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    total = nchunks * CHUNK
    return {"workload": "cfg4: one batch of 256 uint16 RGB 2048x2048 images (1536 chunks of 4 MiB; 32 distinct synthetic "
                        "images x 8 copies), sharded by chunk over %d rank(s) with shard_ranges: zstd clevel 5 chunks "
                        "(reference-produced) decoded + lz4 re-encoded, device-resident" % world,
            "scaling": "strong", "chunks_total": nchunks, "chunks_this_rank": n, "ms_max_over_ranks": ms,
            "value": 2 * total / ms / 1e6, "unit": "GB/s (uncompressed, decode + re-encode, whole job)"}


def gather_leg(torch, dist, rank, world, local_rank):
    """get_decompressed() of ONE 256 MiB channel whose chunks live on N GPUs: every rank decodes its range into its
    slice of a device buffer (device-resident batch), one NCCL all-gather on device pointers assembles the channel."""
    from compressed_image_b200 import Channel, synth
    from compressed_image_b200.sharding import shard_channel
    a = synth.f32_channel(8192, 8192, "depth", channel=12)
    ch = Channel(a, 8192, 8192)                                  # lz4 + shuffle, 64 chunks of 4 MiB
    sc = shard_channel(ch, rank, world)
    dev = "cuda:%d" % local_rank
    out = sc.get_decompressed(device=dev, as_tensor=True)
    torch.cuda.synchronize()
    assert torch.equal(out.cpu(), torch.from_numpy(a.reshape(-1).view(np.uint8))), "gathered channel differs"
    if dist:
        dist.barrier()
    reps, best = 3, None
    for _ in range(reps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = sc.get_decompressed(device=dev, as_tensor=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    # the collective alone, same buffers
    pad = out.numel() // world
    g = torch.empty(pad * world, dtype=torch.uint8, device=dev)
    mine = g[rank * pad:(rank + 1) * pad]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if dist:
        dist.all_gather_into_tensor(g, mine)
        torch.cuda.synchronize()
        dist.barrier()
        e0.record()
        for _ in range(5):
            dist.all_gather_into_tensor(g, mine)
        e1.record()
        torch.cuda.synchronize()
        ag_ms = e0.elapsed_time(e1) / 5
        t = torch.tensor([best, ag_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best, ag_ms = float(t[0].item()), float(t[1].item())
    else:
        ag_ms = 0.0
    nbytes = a.nbytes
    recv = nbytes * (world - 1) / world                          # bytes every rank receives over NVLink
    return {"workload": "Channel.get_decompressed() of one 256 MiB float channel (64 chunks, lz4+shuffle) sharded over %d "
                        "rank(s): device-resident decode into the rank's slice + one all_gather_into_tensor (NCCL) on "
                        "device pointers" % world,
            "ms_max_over_ranks": best * 1e3, "value": nbytes / best / 1e9, "unit": "GB/s (channel bytes / wall time)",
            "all_gather_ms": ag_ms, "all_gather_recv_gbs_per_gpu": (recv / ag_ms / 1e6) if ag_ms else None,
            "nvlink_frac_of_900": (recv / ag_ms / 1e6 / 900.0) if ag_ms else None}


# ---------------------------------------------------------------------------------------------------------------------
def run_b2g(args, rank, world, local_rank):
    import torch
    from compressed_image_b200 import _b2g

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the b2g engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    pieces = make_image(rank)
    n = len(pieces)
    nbytes_total = sum(int(p.size) for p in pieces)
    order = [int(i) for i in np.random.default_rng(1234 + rank).permutation(n)]
    ctx = _b2g.Context(LZ4, CLEVEL, TYPESIZE, SHUFFLE, BLOCK, local_rank)
    # a stream of the bench's own (torch's default stream has handle 0, which the C ABI reads as "the context's own
    # stream"): every device-resident batch of the timed region is launched on it, and so are the timing events
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sh = C.c_void_p(stream.cuda_stream)
    assert stream.cuda_stream != 0

    # ---- device-resident buffers -----------------------------------------------------------------------------------
    slot = CHUNK + 256
    d_raw = torch.empty(nbytes_total, dtype=torch.uint8, device="cuda")
    h_raw = torch.empty(nbytes_total, dtype=torch.uint8).pin_memory()
    o = 0
    raw_off = []
    for p in pieces:
        h_raw[o:o + p.size] = torch.from_numpy(p)
        raw_off.append(o)
        o += int(p.size)
    d_raw.copy_(h_raw, non_blocking=True)
    d_chunks = torch.zeros(n * slot, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(nbytes_total, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()

    comp = ctx.make_batch([d_raw.data_ptr() + raw_off[i] for i in order], [d_chunks.data_ptr() + i * slot for i in order],
                          [int(pieces[i].size) for i in order], [CHUNK + 32] * n)
    ctx.compress_batch_device(comp, stream=sh)
    ctx.wait()
    cbytes = {order[k]: comp[k].out_bytes for k in range(n)}
    assert all(comp[k].status == 0 and comp[k].out_bytes > 0 for k in range(n)), "compress failed"
    cbytes_total = sum(cbytes.values())
    dec = ctx.make_batch([d_chunks.data_ptr() + i * slot for i in order], [d_out.data_ptr() + raw_off[i] for i in order],
                         [cbytes[i] for i in order], [int(pieces[i].size) for i in order])

    # correctness gate before any timing counts: GPU decode of GPU chunks == input, bit-exact
    ctx.decompress_batch_device(dec, stream=sh)
    ctx.wait()
    assert all(dec[k].status == 0 for k in range(n)), "decompress failed"
    assert torch.equal(d_out, d_raw), "round trip mismatch"

    k_ms = {"decode": 0.0, "compress": 0.0}

    # the two batches of a step go to two contexts on the SAME stream (a context carries one batch at a time): the compress
    # launches queue behind the decode kernels without a host round trip in between
    ctx_c = _b2g.Context(LZ4, CLEVEL, TYPESIZE, SHUFFLE, BLOCK, local_rank)

    def step(collect=False):
        ctx.decompress_batch_device(dec, stream=sh)
        ctx_c.compress_batch_device(comp, stream=sh)
        ctx.wait()
        ctx_c.wait()
        if collect:
            k_ms["decode"] += ctx.last_kernel_ms()[1]
            k_ms["compress"] += ctx_c.last_kernel_ms()[1]

    for _ in range(args.warmup):
        step()
    launches0 = ctx.launches + ctx_c.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0.record(stream)
    for _ in range(args.steps):
        step(collect=True)
    e1.record(stream)
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    clocks = sampler.stop()
    launches = ctx.launches + ctx_c.launches - launches0
    ms = e0.elapsed_time(e1)
    if dist:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_step = ms / args.steps
    value = world * 2 * nbytes_total / (ms_step * 1e-3) / 1e9

    # ---- e2e: the same step through the C ABI with host buffers ---------------------------------------------------------
    h_chunks = torch.empty(n * slot, dtype=torch.uint8).pin_memory()
    h_chunks.copy_(d_chunks)
    h_out = torch.empty(nbytes_total, dtype=torch.uint8).pin_memory()
    h_cout = torch.empty(n * slot, dtype=torch.uint8).pin_memory()
    hdec = ctx.make_batch([h_chunks.data_ptr() + i * slot for i in order], [h_out.data_ptr() + raw_off[i] for i in order],
                          [cbytes[i] for i in order], [int(pieces[i].size) for i in order])
    hcomp = ctx.make_batch([h_raw.data_ptr() + raw_off[i] for i in order], [h_cout.data_ptr() + i * slot for i in order],
                           [int(pieces[i].size) for i in order], [CHUNK + 32] * n)

    # The get_chunk pass and the set_chunk pass are independent batches (decode reads the stored chunks, encode reads the
    # raw image): they go to two contexts through the asynchronous C-ABI calls, so the D2H-heavy decode and the H2D-heavy
    # encode share the PCIe link in both directions.  B2G_E2E_SERIAL=1 runs them one after the other instead.
    ctx2 = _b2g.Context(LZ4, CLEVEL, TYPESIZE, SHUFFLE, BLOCK, local_rank)
    serial = os.environ.get("B2G_E2E_SERIAL") == "1"
    only = os.environ.get("B2G_E2E_ONLY", "")

    def make_e2e_step(bd, bc):
        def e2e_step():
            if serial:
                ctx.decompress_batch(bd, stream=sh)
                ctx.compress_batch(bc, stream=sh)
                return
            if only:                       # diagnosis: one direction alone (the JSON line is then not a bench line)
                if only == "dec":
                    ctx.decompress_batch_async(bd)
                    ctx.wait()
                else:
                    ctx2.compress_batch_async(bc)
                    ctx2.wait()
                return
            ctx.decompress_batch_async(bd)
            ctx2.compress_batch_async(bc)
            ctx2.wait()
            ctx.wait()
        return e2e_step

    def time_e2e(fn, steps, warm):
        for _ in range(warm):
            fn()
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        t_ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / steps
        if dist:
            tt = torch.tensor([t_ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t_ms = float(tt.item())
        return t_ms

    e2e_steps = max(3, min(args.steps, 10))
    if args.kernels_only:
        e2e_steps = 1
    e2e_ms = time_e2e(make_e2e_step(hdec, hcomp), e2e_steps, 1 if args.kernels_only else 3)
    if not only:
        assert all(hdec[k].status == 0 for k in range(n)) and all(hcomp[k].status == 0 for k in range(n))
        assert torch.equal(h_out, h_raw), "e2e round trip mismatch"
    cout_total = sum(hcomp[k].out_bytes for k in range(n))
    link = None if args.kernels_only else link_ceiling(torch, dist)
    traffic = (cbytes_total + nbytes_total) + (nbytes_total + cout_total)
    e2e = {"value": world * 2 * nbytes_total / (e2e_ms * 1e-3) / 1e9, "unit": "GB/s",
           "h2d_bytes_per_step": cbytes_total + nbytes_total, "d2h_bytes_per_step": nbytes_total + cout_total,
           "ms_per_step": e2e_ms, "steps": e2e_steps, "host_buffers": "pinned",
           "link_gbs": link, "link": "pinned 256 MiB H2D || D2H on this rank, both directions summed, all %d rank(s) "
                                     "copying at once (max over ranks), measured in this run" % world,
           "frac_of_link": (traffic / (e2e_ms * 1e-3) / 1e9 / link) if link else None,
           "calls": "b2g_decompress_batch + b2g_compress_batch, one after the other" if serial else
                    "b2g_decompress_batch_async (ctx A) + b2g_compress_batch_async (ctx B) + b2g_wait x2"}

    # ---- the same step with the memory the API really hands over: plain (pageable) numpy buffers -----------------------
    e2e_pageable = None
    if not args.kernels_only:
        p_chunks = h_chunks.numpy().copy()
        p_raw = h_raw.numpy().copy()
        p_out = np.empty(nbytes_total, np.uint8)
        p_cout = np.empty(n * slot, np.uint8)
        pdec = ctx.make_batch([p_chunks.ctypes.data + i * slot for i in order], [p_out.ctypes.data + raw_off[i] for i in order],
                              [cbytes[i] for i in order], [int(pieces[i].size) for i in order])
        pcomp = ctx.make_batch([p_raw.ctypes.data + raw_off[i] for i in order], [p_cout.ctypes.data + i * slot for i in order],
                               [int(pieces[i].size) for i in order], [CHUNK + 32] * n)
        p_ms = time_e2e(make_e2e_step(pdec, pcomp), max(3, e2e_steps // 2), 2)
        if not only:
            assert all(pdec[k].status == 0 for k in range(n)) and all(pcomp[k].status == 0 for k in range(n))
            assert np.array_equal(p_out, p_raw), "pageable e2e round trip mismatch"
        hm = host_memcpy_gbs(12)
        e2e_pageable = {"value": world * 2 * nbytes_total / (p_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": p_ms,
                        "host_buffers": "pageable (numpy), as compressed::channel<T> / Channel pass them; staged through "
                                        "per-lane pinned buffers by the library's worker threads",
                        "frac_of_pinned": e2e_ms / p_ms, "host_memcpy_gbs": hm,
                        "host_memcpy": "12 host threads copying numpy -> numpy (the two contexts of the step run 6 worker "
                                       "threads each); every byte of the step crosses host memory once more than with "
                                       "pinned buffers",
                        "frac_of_host_memcpy": (traffic / (p_ms * 1e-3) / 1e9) / hm}

    # ---- multi-GPU legs (all ranks take part) -----------------------------------------------------------------------------
    L = ref_lib()
    cores = os.cpu_count() or 1
    strong = gather = None
    if world > 1 and not args.kernels_only and not args.headline_only:
        del d_out, h_out, h_cout
        torch.cuda.empty_cache()
        if L is not None:
            strong = strong_cfg4(torch, dist, rank, world, local_rank, L, max(1, cores // world))
        gather = gather_leg(torch, dist, rank, world, local_rank)

    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel --------------------------------------------------------------------------------
    peak, peak_src = measured_peaks()
    alg_bytes = nbytes_total + cbytes_total    # per launch, either direction: nbytes + cbytes
    dec_ms, comp_ms = k_ms["decode"] / args.steps, k_ms["compress"] / args.steps
    kern = {"k_decode": dec_ms, "k_compress": comp_ms}
    dom = max(kern, key=kern.get)

    # dram__bytes_read.sum + dram__bytes_write.sum per launch of the same workload, from the committed ncu --set full
    # capture (profiles/traffic.json names the report it was read from); never measured inside this run
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            traffic_j = json.load(fh)
    except Exception:
        traffic_j = {}

    def roof(name):
        t = kern[name]
        a = alg_bytes / (t * 1e-3) / 1e9 if t > 0 else 0.0
        tr = traffic_j.get(name)
        launched = {"k_decode": "k_plan_decode + k_decode<false>",
                    "k_compress": "k_compress_team<lz4, 256> (two-warp teams on the 16 KiB planes) + k_place + k_finalize"}
        return {"kernel": name, "launched_as": launched.get(name), "bound": "hbm", "achieved": a, "peak": peak, "unit": "GB/s",
                "frac": a / peak,
                "traffic": (tr["dram_bytes_read"] + tr["dram_bytes_write"]) if tr else None,
                "traffic_source": traffic_j.get("source") if tr else None,
                "ms_per_launch": t, "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src}

    roofline = roof(dom)
    roofline["kernels"] = {k: roof(k) for k in kern}

    # ---- CPU baseline on this box's host cores (bounded sample; oracle/_ref = the unmodified reference) ---------------
    cpu = None
    if L is not None and not args.kernels_only:
        sample = list(range(n)) if cores >= 16 else list(range(0, n, 4))
        dt, ref_chunks = cpu_reference_pass(L, pieces, [order[i] for i in sample], cores, repeat=2)
        sb = sum(int(pieces[order[i]].size) for i in sample)
        cpu = {"value": 2 * sb / dt / 1e9, "unit": "GB/s", "cores": cores, "kind": "reference",
               "sample": "%d of the %d chunks of the step, decompress + compress, best of 2; unmodified c-blosc2 "
                         "(oracle/_ref), one nthreads=1 context per host thread" % (len(sample), n),
               "ratio": sb / sum(int(c.size) for c in ref_chunks)}

    cfg = headline_config(world)
    line = {"metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg,
            "decompress_gbs_kernel": nbytes_total / (dec_ms * 1e-3) / 1e9 if dec_ms else None,
            "compress_gbs_kernel": nbytes_total / (comp_ms * 1e-3) / 1e9 if comp_ms else None,
            "ratio": nbytes_total / cbytes_total, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "e2e_pageable": e2e_pageable, "gpu_launches": launches, "clocks": clocks}
    if strong:
        line["strong"] = strong
    if gather:
        line["gather"] = gather

    # ---- every BASELINE.json config (N = 1 only: the per-N runs of the scaling series carry `strong` and `gather`) ------
    if world == 1 and not args.kernels_only and not args.headline_only:
        del d_raw, d_chunks, h_raw, h_chunks
        torch.cuda.empty_cache()
        configs = []
        t_cfg = time.perf_counter()

        def add(case, chunks=None, cpu_bytes=256 << 20, **kw):
            r = gpu_case(case, peak, chunks=chunks, **kw)
            if L is not None:
                r["cpu_baseline"] = cpu_case(L, case, cores, max_bytes=cpu_bytes, chunks=chunks)
                c = r["cpu_baseline"]
                r["gpu_over_cpu"] = {"decompress_kernel": r["decompress_gbs_kernel"] / c["decompress_gbs"],
                                     "e2e_decompress": r["e2e"]["decompress_gbs"] / c["decompress_gbs"]}
                if "compress_gbs" in c and "compress_gbs_kernel" in r:
                    r["gpu_over_cpu"]["compress_kernel"] = r["compress_gbs_kernel"] / c["compress_gbs"]
                    r["gpu_over_cpu"]["e2e_compress"] = r["e2e"]["compress_gbs"] / c["compress_gbs"]
            configs.append(r)
            return r

        r1 = add(case_cfg1())
        r1["api"] = api_cfg1()
        add(case_cfg2())
        # the same two images at compression level 5: the reference's LZ4 path ignores clevel (accel forced to 1,
        # blosc2.c:413), the GPU encoder trades ~1 % (half) / ~6 % (u8) of ratio for speed on BOTH sides of the codec
        for base in (case_cfg1(), case_cfg2()):
            add(dict(base, clevel=5, name=base["name"] + " -- at compression level 5 (not the reference default 9)"))
        if not args.quick:
            add(case_cfg3(BITSHUFFLE), cpu_bytes=256 << 20, e2e_bytes=256 << 20)
            add(case_cfg3(SHUFFLE, nchan=4), cpu_bytes=256 << 20, e2e_bytes=256 << 20)
            c4 = case_cfg4_lz4()
            add(c4, e2e_bytes=256 << 20)
            if L is not None:
                for clevel, distinct in ((5, CFG4_DISTINCT), (9, 4)):
                    zp = c4["pieces"][:distinct * 6]
                    zc = {"name": "cfg4 batch of %d uint16 RGB images (%d distinct x %d copies), zstd clevel %d chunks produced "
                                  "by the reference, GPU decode" % (CFG4_IMAGES, distinct, CFG4_IMAGES // distinct, clevel),
                          "pieces": zp, "codec": ZSTD, "clevel": clevel, "ts": 2, "filter": SHUFFLE,
                          "replicas": CFG4_IMAGES // distinct}
                    chunks, _ = ref_compress(L, zp, ZSTD, clevel, 2, SHUFFLE, cores)
                    add(zc, chunks=chunks, cpu_bytes=(256 << 20) if clevel == 5 else (48 << 20), e2e_bytes=96 << 20)
            line["cfg5_sweep"] = sweep_cfg5(peak, L, cores)
        line["configs"] = configs
        line["configs_seconds"] = time.perf_counter() - t_cfg
    print(json.dumps(round_floats(line)))
    if dist:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b2g", choices=["b2g", "reference"])
    ap.add_argument("--kernels-only", action="store_true", help="profiling runs: skip the e2e, cpu_baseline and configs legs")
    ap.add_argument("--headline-only", action="store_true", help="skip the per-config, strong and gather legs")
    ap.add_argument("--quick", action="store_true", help="configs: cfg 1 and cfg 2 only")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b2g(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
