#!/usr/bin/env python
"""Measurement: host -> host throughput of the PRODUCT's multi-device entry point (b200jpeg_multi_decode_batch: ONE process,
one host thread + context per GPU) on N = 1, 2, 4, 8 GPUs, weak scaling (256 C2 images per GPU), pinned buffers, header
parse inside the timed region -- the one-process counterpart of `torchrun ... bench.py --gpus N` (N processes).
    python scripts/dev/multi_e2e.py [max_gpus] [images_per_gpu]     -> one JSON line per N"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
import bench  # noqa: E402


def main():
    import torch
    ndev = torch.cuda.device_count()
    max_gpus = min(int(sys.argv[1]) if len(sys.argv) > 1 else ndev, ndev)
    per = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    lib = jd.load()

    class A:
        workload, images, quality = "C2", per * max_gpus, 0
    wl = bench.workload(A)
    arena, files = bench.pack(bench.make_files(wl, 0, per * max_gpus), alloc=lib.b200jpeg_host_alloc)
    rgb_bytes = 3 * wl["w"] * wl["h"]
    rgb_ptr = lib.b200jpeg_host_alloc(per * max_gpus * rgb_bytes)
    rgb_host = np.ctypeslib.as_array((C.c_uint8 * (per * max_gpus * rgb_bytes)).from_address(rgb_ptr))
    n = 1
    while n <= max_gpus:
        m = jd.MultiDecoder(list(range(n)))
        cnt = per * n
        fs = files[:cnt]
        rgbs = [rgb_host[i * rgb_bytes:(i + 1) * rgb_bytes] for i in range(cnt)]
        descs = (jd.ImageDesc * cnt)()
        ts = []
        for k in range(6):
            t0 = time.perf_counter()
            d, rcs = jd.parse_headers(fs, 0, descs=descs)
            st = m.decode_batch(d, fs, rgbs)
            if k >= 2:
                ts.append(time.perf_counter() - t0)
        assert not any(st)
        t = float(np.median(ts))
        print(json.dumps({"gpus": n, "images": cnt, "ms": t * 1e3, "e2e_mp_s": cnt * wl["w"] * wl["h"] / t / 1e6,
                          "d2h_gbs_total": cnt * rgb_bytes / t / 1e9, "form": "one process, one host thread per GPU"}), flush=True)
        m.close()
        n *= 2
    lib.b200jpeg_host_free(rgb_ptr)


if __name__ == "__main__":
    main()
