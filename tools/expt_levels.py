"""Developer tooling: kernel-scope cfg 2 and cfg 4 (lz4) at compression levels 9 / 7 / 5 (ratio against speed)."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
peak, _ = bench.measured_peaks()
for base in (bench.case_cfg2(), bench.case_cfg4_lz4(distinct=8, images=32), bench.case_cfg1()):
    for lvl in (9, 7, 5, 2):
        case = dict(base, clevel=lvl)
        r = bench.gpu_case(case, peak, e2e_bytes=0, reps=3)
        print(case["name"][:44], "clevel", lvl, "ratio", round(r["ratio"], 4), "dec GB/s", round(r["decompress_gbs_kernel"], 1), "(%.3f)" % r["roofline"]["k_decode"]["frac"], "comp GB/s", round(r["compress_gbs_kernel"], 1), "(%.3f)" % r["roofline"]["k_compress"]["frac"])
