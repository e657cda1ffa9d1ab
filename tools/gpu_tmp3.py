import sys, os, numpy as np
sys.path.insert(0, "/root/repo/tools"); sys.path.insert(0, "/root/repo")
import bench_configs as B
from compressed_image_b200 import _b2g, synth
for kind in synth.F32_KINDS:
    p = synth.split_chunks(synth.f32_channel(8192, 4096, kind), B.CHUNK)
    B.run_case("blosclz+bitshuffle " + kind, p, _b2g.BLOSCLZ, 9, 4, _b2g.BITSHUFFLE, 3)
    B.run_case("lz4+bitshuffle " + kind, p, _b2g.LZ4, 9, 4, _b2g.BITSHUFFLE, 3)
    B.run_case("lz4+noshuffle " + kind, p, _b2g.LZ4, 9, 4, _b2g.NOFILTER, 3)
