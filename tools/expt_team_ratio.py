"""Ratio of the team encoder (k_compress_team on the CPU simulator) against one warp per stream, on the bench data."""
import sys, time
import numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import _sim, _libs, _data
from compressed_image_b200 import synth

kind = sys.argv[1] if len(sys.argv) > 1 else "half"
nbytes = int(sys.argv[2]) if len(sys.argv) > 2 else 512 * 1024
if kind == "half":
    data = synth.half_channel(8192, 4096, 0).view(np.uint8).reshape(-1); ts, codec, filt = 2, _data.LZ4, _data.SHUFFLE
elif kind == "u8":
    data = synth.u8_rgba_channel(4096, 4096, 0).view(np.uint8).reshape(-1); ts, codec, filt = 1, _data.LZ4, _data.SHUFFLE
elif kind == "f32bit":
    data = synth.f32_channel(8192, 8192, "depth").view(np.uint8).reshape(-1); ts, codec, filt = 4, _data.BLOSCLZ, _data.BITSHUFFLE
elif kind == "f32":
    data = synth.f32_channel(8192, 8192, "depth").view(np.uint8).reshape(-1); ts, codec, filt = 4, _data.BLOSCLZ, _data.SHUFFLE
off = (data.size // 3) // 32768 * 32768
src = data[off:off + nbytes]
L = _sim.lib()
import ctypes as C
L.sim_set_team.argtypes = [C.c_int] * 3
rows = []
for team, prime, hlog in [(1, 0, 0)] + [(-1, pr, hl) for hl in (9, 10, 11) for pr in (0, 1024, 2048, 4096, 65536)]:
    L.sim_set_team(team, prime, hlog)
    t0 = time.time()
    (st, chunk), = _sim.compress_batch([src], codec, 9, ts, filt, 32768)
    assert st == 0
    rc, back = _libs.ref_decompress(chunk)
    assert rc == src.size and np.array_equal(back, src), (team, prime)
    print(f"team={team} prime={prime} hlog={hlog}: {chunk.size} bytes ratio {src.size / chunk.size:.4f}  ({time.time() - t0:.1f}s)", flush=True)
L.sim_set_team(-1, -1, 0)
