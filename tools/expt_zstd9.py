"""Developer tooling: one kernel-scope decode of reference-produced zstd clevel 9 chunks (for ncu)."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
peak, _ = bench.measured_peaks()
L = bench.ref_lib(); cores = os.cpu_count()
zp = bench.cfg4_pieces(range(4))
chunks, _ = bench.ref_compress(L, zp, bench.ZSTD, 9, 2, bench.SHUFFLE, cores)
case = {"name": "zstd clevel 9", "pieces": zp, "codec": bench.ZSTD, "clevel": 9, "ts": 2, "filter": bench.SHUFFLE, "replicas": 8}
r = bench.gpu_case(case, peak, chunks=chunks, e2e_bytes=0, reps=2)
print(r["decompress_gbs_kernel"], r["decompress_ms"])
