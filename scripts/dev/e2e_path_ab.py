#!/usr/bin/env python
"""Dev tool: host -> host time of b200jpeg_decode_batch on the C2 batch with the entropy path forced either way,
alternating in one process (PCIe-bound: differences are first-chunk latency and the tail)."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
from jpeg_decoder_b200 import abi  # noqa: E402
import bench  # noqa: E402


def main():
    lib = jd.load()
    dec = jd.Decoder(0)

    class A:
        workload, images, quality = sys.argv[1] if len(sys.argv) > 1 else "C2", 0, 0
    wl = bench.workload(A)
    arena, files = bench.pack(bench.make_files(wl, 0, wl["n"]), alloc=lib.b200jpeg_host_alloc)
    n = len(files)
    rgb_bytes = 3 * wl["w"] * wl["h"]
    ptr = lib.b200jpeg_host_alloc(n * rgb_bytes)
    host = np.ctypeslib.as_array((C.c_uint8 * (n * rgb_bytes)).from_address(ptr))
    rgbs = [host[i * rgb_bytes:(i + 1) * rgb_bytes] for i in range(n)]
    descs, _ = jd.parse_headers(files, wl["flags"])
    res = {"auto": [], "segment": [], "sync": []}
    modes = {"auto": abi.ENTROPY_AUTO, "segment": abi.ENTROPY_SEGMENT, "sync": abi.ENTROPY_SYNC}
    for rep in range(8):
        for name, m in modes.items():
            dec.set_entropy_path(m)
            t0 = time.perf_counter()
            st = dec.decode_batch(descs, files, rgbs)
            dt = (time.perf_counter() - t0) * 1e3
            assert not any(st)
            if rep >= 2:
                res[name].append(dt)
    fl = [dec.copy_floor(descs, files, rgbs) for _ in range(4)]
    print(json.dumps({"workload": A.workload, **{k: round(float(np.median(v)), 3) for k, v in res.items()}, "copy_floor_ms": round(float(np.median(fl)), 3),
                      "path_auto": dec.stats()["entropy_path"]}))
    dec.close()


if __name__ == "__main__":
    main()
