#!/usr/bin/env python
"""Kernel-scope throughput of the other BASELINE.json configs (bench.py measures cfg 2, the headline).  Developer /
profiling tool: device-resident chunks, CUDA events inside the library (b2g_last_kernel_ms), median of `reps` launches.
Prints one JSON line per case: uncompressed GB/s per direction, ratio, fraction of the measured HBM roofline
((nbytes + cbytes) / kernel time / MEASURED_PEAKS.json hbm_gbs).   usage: python tools/bench_configs.py [reps]"""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

from compressed_image_b200 import _b2g, synth  # noqa: E402

CHUNK, BLOCK = 4 << 20, 32768
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6650.0


def run_case(name, pieces, codec, clevel, ts, filt, reps, chunks=None):
    """chunks given (numpy, e.g. reference zstd chunks): decode only.  B2G_CASE=<substring> runs the matching cases only."""
    if os.environ.get("B2G_CASE") and os.environ["B2G_CASE"] not in name:
        return
    n = len(pieces)
    total = sum(int(p.size) for p in pieces)
    raw = torch.empty(total, dtype=torch.uint8)
    offs, o = [], 0
    for p in pieces:
        raw[o:o + p.size] = torch.from_numpy(np.ascontiguousarray(p))
        offs.append(o)
        o += int(p.size)
    d_raw = raw.cuda()
    slot = max(int(p.size) for p in pieces) + 256
    d_chunks = torch.zeros(n * slot, dtype=torch.uint8, device="cuda")
    d_out = torch.zeros(total, dtype=torch.uint8, device="cuda")
    ctx = _b2g.Context(codec, clevel, ts, filt, BLOCK)
    comp_ms = None
    if chunks is None:
        comp = ctx.make_batch([d_raw.data_ptr() + offs[i] for i in range(n)], [d_chunks.data_ptr() + i * slot for i in range(n)],
                              [int(p.size) for p in pieces], [int(p.size) + 32 for p in pieces])
        t = []
        for _ in range(reps + 1):
            ctx.compress_batch_device(comp)
            ctx.wait()
            t.append(sum(ctx.last_kernel_ms()))
        comp_ms = float(np.median(t[1:]))
        cb = [comp[i].out_bytes for i in range(n)]
        assert all(comp[i].status == 0 for i in range(n))
    else:
        for i, c in enumerate(chunks):
            d_chunks[i * slot:i * slot + c.size] = torch.from_numpy(c).cuda()
        cb = [int(c.size) for c in chunks]
    dec = ctx.make_batch([d_chunks.data_ptr() + i * slot for i in range(n)], [d_out.data_ptr() + offs[i] for i in range(n)], cb,
                         [int(p.size) for p in pieces])
    t = []
    for _ in range(reps + 1):
        ctx.decompress_batch_device(dec)
        ctx.wait()
        t.append(sum(ctx.last_kernel_ms()))
    dec_ms = float(np.median(t[1:]))
    torch.cuda.synchronize()
    assert all(dec[i].status == 0 for i in range(n)) and torch.equal(d_out, d_raw), name
    cbytes = sum(cb)
    line = {"case": name, "chunks": n, "nbytes": total, "ratio": round(total / cbytes, 3),
            "decompress_gbs": round(total / dec_ms / 1e6, 1), "decompress_ms": round(dec_ms, 3),
            "decompress_roofline_frac": round((total + cbytes) / dec_ms / 1e6 / PEAK, 4)}
    if comp_ms:
        line.update({"compress_gbs": round(total / comp_ms / 1e6, 1), "compress_ms": round(comp_ms, 3),
                     "compress_roofline_frac": round((total + cbytes) / comp_ms / 1e6 / PEAK, 4)})
    line.update(cpu_reference(pieces, codec, clevel, ts, filt, chunks is None,
                              [d_chunks[i * slot:i * slot + cb[i]].cpu().numpy() for i in range(n)]))
    print(json.dumps(line), flush=True)
    ctx.close()


def cpu_reference(pieces, codec, clevel, ts, filt, do_compress, chunks):
    """The unmodified c-blosc2 (oracle/_ref) on all host cores, one nthreads=1 context per thread, chunks dealt over the
    threads: decompress of the very chunks the GPU decoded, compress with the same parameters."""
    import time
    try:
        import _libs
        R = _libs.ref()
    except Exception:
        R = None
    if R is None:
        return {}
    n, cores, vp = len(pieces), os.cpu_count() or 1, C.c_void_p
    srcs = [np.ascontiguousarray(p) for p in pieces]
    outs = [np.empty(p.size, np.uint8) for p in srcs]
    a_ch = (vp * n)(*[c.ctypes.data for c in chunks])
    a_cb = (C.c_int32 * n)(*[int(c.size) for c in chunks])
    a_out = (vp * n)(*[o.ctypes.data for o in outs])
    a_nb = (C.c_int32 * n)(*[int(p.size) for p in srcs])
    res = (C.c_int32 * n)()
    best = None
    for _ in range(2):
        t0 = time.perf_counter()
        R.ref_decompress_batch(n, a_ch, a_cb, a_out, a_nb, res, cores)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    assert all(res[i] == srcs[i].size for i in range(n)) and np.array_equal(outs[n // 2], srcs[n // 2].view(np.uint8).reshape(-1))
    total = sum(int(p.size) for p in srcs)
    out = {"cpu_cores": cores, "cpu_decompress_gbs": round(total / best / 1e9, 2)}
    if do_compress:
        dsts = [np.empty(p.size + 32, np.uint8) for p in srcs]
        a_src = (vp * n)(*[p.ctypes.data for p in srcs])
        a_dst = (vp * n)(*[d.ctypes.data for d in dsts])
        a_cap = (C.c_int32 * n)(*[int(d.size) for d in dsts])
        best = None
        for _ in range(2):
            t0 = time.perf_counter()
            R.ref_compress_batch(codec, clevel, ts, filt, BLOCK, n, a_src, a_nb, a_dst, a_cap, res, cores)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        out["cpu_compress_gbs"] = round(total / best / 1e9, 2)
        out["cpu_ratio"] = round(total / sum(int(r) for r in res), 3)
    return out


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    LZ4, BLZ, ZSTD, SH, BSH = _b2g.LZ4, _b2g.BLOSCLZ, _b2g.ZSTD, _b2g.SHUFFLE, _b2g.BITSHUFFLE
    p = []
    for k in range(4):
        p += synth.split_chunks(synth.u8_rgba_channel(4096, 4096, k), CHUNK)
    run_case("cfg1 u8 RGBA 4096x4096 lz4+shuffle", p, LZ4, 9, 1, SH, reps)
    p = []
    for k in range(4):
        p += synth.split_chunks(synth.half_channel(8192, 4096, k).view(np.uint16), CHUNK)
    run_case("cfg2 half 8192x4096x4 lz4+shuffle", p, LZ4, 9, 2, SH, reps)
    p = []
    for k, kind in enumerate(synth.F32_KINDS):
        p += synth.split_chunks(synth.f32_channel(8192, 8192, kind, channel=k), CHUNK)
    run_case("cfg3 f32 8192x8192 x4 kinds blosclz+bitshuffle", p, BLZ, 9, 4, BSH, reps)
    run_case("cfg3 f32 8192x8192 x4 kinds blosclz+shuffle", p, BLZ, 9, 4, SH, reps)
    p = []
    for im in range(12):
        for ch in range(3):
            p += synth.split_chunks(synth.u16_sensor_channel(2048, 2048, ch, image=im), CHUNK)
    run_case("cfg4 u16 RGB 2048x2048 x12 lz4+shuffle (re-encode)", p, LZ4, 9, 2, SH, reps)
    try:
        import _libs
        R = _libs.ref()
    except Exception:
        R = None
    if R is not None:
        for clevel in (5, 9):
            q = p if clevel == 5 else p[:12]
            n = len(q)
            vp = C.c_void_p
            dsts = [np.empty(x.size + 32, np.uint8) for x in q]
            out_c = (C.c_int32 * n)()
            R.ref_compress_batch(ZSTD, clevel, 2, SH, BLOCK, n, (vp * n)(*[x.ctypes.data for x in q]),
                                 (C.c_int32 * n)(*[int(x.size) for x in q]), (vp * n)(*[d.ctypes.data for d in dsts]),
                                 (C.c_int32 * n)(*[int(d.size) for d in dsts]), out_c, os.cpu_count() or 1)
            run_case("cfg4 zstd clevel %d chunks from the reference, GPU decode" % clevel, q, ZSTD, 9, 2, SH, reps,
                     chunks=[d[:out_c[i]] for i, d in enumerate(dsts)])
    for kib in (64, 1024, 16384):
        h = (64 << 20) // (2048 * 2)
        q = synth.split_chunks(synth.half_channel(2048, h, 1).view(np.uint16), kib * 1024)
        run_case("cfg5 half 64 MiB in %d KiB chunks lz4+shuffle" % kib, q, LZ4, 9, 2, SH, reps)


if __name__ == "__main__":
    main()
