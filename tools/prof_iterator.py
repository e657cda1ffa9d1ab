"""Developer tooling: cProfile of the iterator "modify" loop (tools/bench_iterator.py)."""
import cProfile, pstats, sys
import numpy as np
sys.path.insert(0, ".")
import compressed_image_b200 as c
from compressed_image_b200 import synth
data = synth.half_channel(8192, 4096, 0)
ch = c.Channel(data, 8192, 4096)
def run_iter():
    for chunk in ch:
        v = chunk.view(np.uint16); v ^= 1
run_iter(); run_iter()
pr = cProfile.Profile(); pr.enable()
for _ in range(5): run_iter()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
