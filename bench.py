#!/usr/bin/env python
"""bench.py -- chunk decompress / compress throughput of the b2g engine on B200 (contract: see DESIGN.md section 6).

Workload (BASELINE.json configs[1], the configuration the metric is quoted on): compressed::image<half>, synthetic
8192x4096, 4 channels, lz4 + byte-shuffle, default 4 MiB chunks / 32 KiB blocks -> 64 chunks, 256 MiB.  One STEP is the
random-access pass of that config: get_chunk over all 64 chunks in a seeded shuffled order (one batched decompress
launch) followed by set_chunk over all of them (one batched compress launch).

  value   uncompressed GB/s over the step (bytes decompressed + bytes compressed, / step time), chunks and raw image
          resident in HBM, timed with CUDA events on the launching stream.  Inputs (256 MiB raw + chunks) exceed the
          126 MB L2, so every step streams from HBM.
  e2e     the same step through the C ABI with HOST (pinned) buffers: H2D of chunks / raw data and D2H of the results
          inside the timed region.
  roofline / cpu_baseline / clocks / gpu_launches: as the contract asks.

`--impl reference` times the reference's own CPU implementation (oracle/_ref/libblosc2_ref.so = unmodified c-blosc2,
one nthreads=1 context per host thread, chunks dealt over all cores) on the same workload.

N > 1 (torchrun): every rank runs the workload on its own image (weak scaling, chunks are independent: no collective
on the data path); value = all ranks' bytes / max-over-ranks time.
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

LZ4, SHUFFLE = 1, 1
WIDTH, HEIGHT, NCHAN = 8192, 4096, 4
CHUNK, BLOCK, CLEVEL, TYPESIZE = 4 * 1024 * 1024, 32768, 9, 2
WORKLOAD = ("compressed::image<half> synthetic 8192x4096 4-channel, lz4 + byte-shuffle, chunk 4 MiB / block 32 KiB: "
            "get_chunk + set_chunk over all 64 chunks in seeded random order")


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
def ref_lib():
    path = os.path.join(ROOT, "oracle", "_ref", "libblosc2_ref.so")
    if not os.path.exists(path):
        return None
    L = C.CDLL(path)
    pvp, pi32, ci = C.POINTER(C.c_void_p), C.POINTER(C.c_int32), C.c_int
    L.ref_compress_batch.argtypes = [ci, ci, ci, ci, ci, ci, pvp, pi32, pvp, pi32, pi32, ci]
    L.ref_decompress_batch.argtypes = [ci, pvp, pi32, pvp, pi32, pi32, ci]
    return L


def cpu_reference_pass(L, pieces, order, nthreads, repeat=1):
    """One step of the workload on the host cores with the reference library.  -> (seconds per step, chunks)."""
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
    line = {"impl": "reference", "metric": "chunk decompress+compress GB/s (uncompressed)", "value": gbs, "unit": "GB/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": step * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "chunks": len(pieces), "bytes_per_step": 2 * nbytes},
            "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "reference",
                             "sample": "the full step (64 chunks decompressed + 64 compressed) per timed step; unmodified "
                                       "c-blosc2, one nthreads=1 context per host thread, chunks dealt over all threads"},
            "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "ratio": nbytes / sum(int(c.size) for c in chunks)}
    print(json.dumps(line))


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
    stream = torch.cuda.current_stream()
    sh = C.c_void_p(stream.cuda_stream)

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

    def step(collect=False):
        ctx.decompress_batch_device(dec, stream=sh)
        ctx.wait()
        if collect:
            k_ms["decode"] += ctx.last_kernel_ms()[1]
        ctx.compress_batch_device(comp, stream=sh)
        ctx.wait()
        if collect:
            k_ms["compress"] += ctx.last_kernel_ms()[1]

    for _ in range(args.warmup):
        step()
    launches0 = ctx.launches
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
    launches = ctx.launches - launches0
    ms = e0.elapsed_time(e1)
    if dist:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_step = ms / args.steps
    value = world * 2 * nbytes_total / (ms_step * 1e-3) / 1e9

    # ---- e2e: the same step through the C ABI with host (pinned) buffers ----------------------------------------------
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

    def e2e_step():
        if serial:
            ctx.decompress_batch(hdec, stream=sh)
            ctx.compress_batch(hcomp, stream=sh)
            return
        ctx.decompress_batch_async(hdec)
        ctx2.compress_batch_async(hcomp)
        ctx2.wait()
        ctx.wait()

    e2e_steps = max(3, min(args.steps, 10))
    if args.kernels_only:
        e2e_steps = 1
    for _ in range(1 if args.kernels_only else 3):
        e2e_step()
    assert all(hdec[k].status == 0 for k in range(n)) and all(hcomp[k].status == 0 for k in range(n))
    assert torch.equal(h_out, h_raw), "e2e round trip mismatch"
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record(stream)
    for _ in range(e2e_steps):
        e2e_step()
    e1.record(stream)
    torch.cuda.synchronize()
    e2e_ms = max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / e2e_steps
    if dist:
        t = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    cout_total = sum(hcomp[k].out_bytes for k in range(n))
    e2e = {"value": world * 2 * nbytes_total / (e2e_ms * 1e-3) / 1e9, "unit": "GB/s",
           "h2d_bytes_per_step": cbytes_total + nbytes_total, "d2h_bytes_per_step": nbytes_total + cout_total,
           "ms_per_step": e2e_ms, "steps": e2e_steps, "host_buffers": "pinned",
           "calls": "b2g_decompress_batch + b2g_compress_batch, one after the other" if serial else
                    "b2g_decompress_batch_async (ctx A) + b2g_compress_batch_async (ctx B) + b2g_wait x2"}

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
            traffic = json.load(fh)
    except Exception:
        traffic = {}

    def roof(name):
        t = kern[name]
        a = alg_bytes / (t * 1e-3) / 1e9 if t > 0 else 0.0
        tr = traffic.get(name)
        return {"kernel": name, "bound": "hbm", "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak,
                "traffic": (tr["dram_bytes_read"] + tr["dram_bytes_write"]) if tr else None,
                "traffic_source": traffic.get("source") if tr else None,
                "ms_per_launch": t, "algorithmic_bytes_per_launch": alg_bytes, "peak_source": peak_src}

    roofline = roof(dom)
    roofline["kernels"] = {k: roof(k) for k in kern}

    # ---- CPU baseline on this box's host cores (bounded sample; oracle/_ref = the unmodified reference) ---------------
    cpu = None
    L = ref_lib()
    cores = os.cpu_count() or 1
    if L is not None and not args.kernels_only:
        sample = list(range(n)) if cores >= 16 else list(range(0, n, 4))
        dt, ref_chunks = cpu_reference_pass(L, pieces, [order[i] for i in sample], cores, repeat=2)
        sb = sum(int(pieces[order[i]].size) for i in sample)
        cpu = {"value": 2 * sb / dt / 1e9, "unit": "GB/s", "cores": cores, "kind": "reference",
               "sample": "%d of the %d chunks of the step, decompress + compress, best of 2; unmodified c-blosc2 "
                         "(oracle/_ref), one nthreads=1 context per host thread" % (len(sample), n),
               "ratio": sb / sum(int(c.size) for c in ref_chunks)}

    line = {"metric": "chunk decompress+compress GB/s (uncompressed)", "value": value, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "chunks_per_gpu": n, "bytes_per_step_per_gpu": 2 * nbytes_total,
                       "l2": "inputs (256 MiB raw + chunks per direction) exceed the 126 MB L2; no explicit flush",
                       "parallelism": "chunks sharded by image across %d GPU(s), no collective" % world},
            "decompress_gbs_kernel": nbytes_total / (dec_ms * 1e-3) / 1e9 if dec_ms else None,
            "compress_gbs_kernel": nbytes_total / (comp_ms * 1e-3) / 1e9 if comp_ms else None,
            "ratio": nbytes_total / cbytes_total, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": launches, "clocks": clocks}
    print(json.dumps(line))
    if dist:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b2g", choices=["b2g", "reference"])
    ap.add_argument("--kernels-only", action="store_true", help="profiling runs: skip the e2e and cpu_baseline legs")
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
