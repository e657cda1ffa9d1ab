"""Developer tooling: kernel-scope decode of reference-produced zstd chunks (cfg 4 sample), clevel 5 and 9."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
peak, _ = bench.measured_peaks()
L = bench.ref_lib(); cores = os.cpu_count()
p = bench.cfg4_pieces(range(8))
for clevel, cnt in ((5, 48), (9, 24)):
    zp = p[:cnt]
    chunks, _ = bench.ref_compress(L, zp, bench.ZSTD, clevel, 2, bench.SHUFFLE, cores)
    case = {"name": "zstd clevel %d" % clevel, "pieces": zp, "codec": bench.ZSTD, "clevel": clevel, "ts": 2, "filter": bench.SHUFFLE, "replicas": 256 * 6 // cnt}
    r = bench.gpu_case(case, peak, chunks=chunks, e2e_bytes=0, reps=3)
    print(case["name"], "chunks", r["chunks"], "decode GB/s", round(r["decompress_gbs_kernel"], 1), "ms", round(r["decompress_ms"], 3), "frac", round(r["roofline"]["k_decode"]["frac"], 4))
