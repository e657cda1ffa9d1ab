"""Device-resident bandwidth of the interleave / deinterleave kernels on BASELINE-sized images,
CUDA events around 20 launches each, buffers larger than L2.  One JSON line per case; algorithmic bytes = 2 x image."""
import json
import sys

import torch

sys.path.insert(0, ".")
from compressed_image_b200 import _b2g  # noqa: E402

CASES = [("cfg1 u8 4096x4096x4", 1, 4, 4096 * 4096), ("cfg2 half 8192x4096x4", 2, 4, 8192 * 4096),
         ("f32 8192x8192x4", 4, 4, 8192 * 8192), ("cfg4 u16 2048x2048x3 x32", 2, 3, 2048 * 2048 * 32),
         ("f32 rgb 8192x4096x3", 4, 3, 8192 * 4096), ("u8 2ch 8192x8192", 1, 2, 8192 * 8192),
         ("cfg3 f32 16ch 4096x4096 (tile kernel)", 4, 16, 4096 * 4096), ("half 8ch 8192x4096 (tile kernel)", 2, 8, 8192 * 4096),
         ("u8 8ch 8192x8192 (tile kernel)", 1, 8, 8192 * 8192), ("u16 5ch 4096x4096 (element kernel)", 2, 5, 4096 * 4096)]
peak = json.load(open("MEASURED_PEAKS.json")) if len(sys.argv) > 1 else None
with _b2g.Context() as ctx:
    st = torch.cuda.Stream()
    for name, ts, nch, n in CASES:
        inter = torch.randint(0, 256, (n * nch * ts,), dtype=torch.uint8, device="cuda")
        planes = [torch.empty(n * ts, dtype=torch.uint8, device="cuda") for _ in range(nch)]
        back = torch.empty_like(inter)
        torch.cuda.synchronize()
        res = {"case": name, "typesize": ts, "channels": nch, "MiB": inter.numel() / 2**20}
        for label, fn in (("deinterleave", lambda: ctx.deinterleave_device(ts, inter.data_ptr(), [p.data_ptr() for p in planes], n, st.cuda_stream)),
                          ("interleave", lambda: ctx.interleave_device(ts, [p.data_ptr() for p in planes], back.data_ptr(), n, st.cuda_stream))):
            with torch.cuda.stream(st):
                for _ in range(3):
                    fn()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(20):
                    fn()
                e1.record(st)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 20
            res[label + "_ms"] = round(ms, 4)
            res[label + "_GBps"] = round(2 * inter.numel() / ms / 1e6, 1)
        assert torch.equal(back, inter)
        print(json.dumps(res), flush=True)
        del inter, planes, back
