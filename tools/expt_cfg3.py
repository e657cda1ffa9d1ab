"""Developer tooling: kernel-scope cfg 3 (4 channels, one per kind) blosclz + bitshuffle / shuffle, and cfg 1."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
peak, _ = bench.measured_peaks()
for case in (bench.case_cfg3(bench.BITSHUFFLE, nchan=4), bench.case_cfg3(bench.SHUFFLE, nchan=4), bench.case_cfg1()):
    r = bench.gpu_case(case, peak, e2e_bytes=0, reps=3)
    print(case["name"][:60], "ratio", round(r["ratio"], 3), "dec GB/s", round(r["decompress_gbs_kernel"], 1), "comp GB/s", round(r["compress_gbs_kernel"], 1))
