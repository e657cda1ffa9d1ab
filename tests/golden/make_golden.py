"""Regenerates tests/golden/ from the reference tree (run in the dev container only; /root/reference does not
exist on the GPU box).

* compat/*.cdata : the c-blosc2 backward-compatibility chunks (thirdparty/c-blosc2/compat/, generator
  compat/filegen.c:21-57: int32 i for i < 1_000_000 compressed by historic Blosc releases with blosclz / lz4 /
  lz4hc / zstd, plus the lz4 + bitshuffle files).  These are DATA vectors, copied byte for byte; zlib files are
  skipped (the reference is built with DEACTIVATE_ZLIB, CMakeLists.txt:48).
* ref_chunks.npz : chunks produced HERE by the unmodified reference library (oracle/_ref/libblosc2_ref.so) for
  a grid of codec x filter x typesize x size, with the seeds that regenerate their inputs (tests/_data.py).
* manifest.json  : sha256 of every decoded payload, decoded with the reference library itself.
"""
import glob
import hashlib
import json
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import _data  # noqa: E402
import _libs  # noqa: E402

REF_COMPAT = "/root/reference/thirdparty/c-blosc2/compat"


def main():
    R = _libs.ref()
    assert R is not None, "build oracle/_ref first (make -C oracle)"
    man = {"compat": {}, "ref_chunks": {}}
    os.makedirs(os.path.join(HERE, "compat"), exist_ok=True)
    for f in sorted(glob.glob(os.path.join(REF_COMPAT, "*.cdata"))):
        name = os.path.basename(f)
        if "zlib" in name:
            continue
        shutil.copyfile(f, os.path.join(HERE, "compat", name))
        chunk = np.fromfile(f, dtype=np.uint8)
        nb, cb, bs = _libs.ref_sizes(chunk)
        out = np.zeros(nb, dtype=np.uint8)
        r = R.ref_decompress(_libs.ptr(chunk), int(chunk.size), _libs.ptr(out), int(nb), 1)
        assert r == nb, (name, r)
        man["compat"][name] = {"nbytes": int(nb), "cbytes": int(cb), "blocksize": int(bs),
                               "version": int(chunk[0]), "flags": int(chunk[2]), "typesize": int(chunk[3]),
                               "sha256": hashlib.sha256(out.tobytes()).hexdigest()}
    chunks = {}
    for case in _data.REF_CHUNK_GRID:
        key = _data.case_key(case)
        src = _data.make_case(case)
        chunk = _libs.ref_compress(case["codec"], case["clevel"], case["typesize"], case["filter"],
                                   case["blocksize"], src)
        chunks[key] = chunk
        man["ref_chunks"][key] = {"case": case, "cbytes": int(chunk.size),
                                  "sha256": hashlib.sha256(src.tobytes()).hexdigest()}
    np.savez_compressed(os.path.join(HERE, "ref_chunks.npz"), **chunks)
    with open(os.path.join(HERE, "manifest.json"), "w") as fh:
        json.dump(man, fh, indent=1, sort_keys=True)
    print("compat files:", len(man["compat"]), " ref chunks:", len(chunks))


if __name__ == "__main__":
    main()
