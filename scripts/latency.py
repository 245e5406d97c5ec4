#!/usr/bin/env python
"""Single-image latency of BASELINE.json config C1 (images/img.jpg, copied to tests/golden/) and of the other reference
fixtures: GPU time of the kernels (CUDA events), host -> host time of b200jpeg_decode_batch with n = 1 (header parse +
H2D + kernels + D2H), next to the reference binary's own time on one host core (ctor + readJPEG(), jpeg.cpp:797-907).
    python scripts/latency.py [files...]      -> one JSON line per file"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
from oracle import oracle  # noqa: E402


def main():
    names = sys.argv[1:] or [os.path.join(ROOT, "tests", "golden", n) for n in ("img.jpg", "img2.jpg", "img4.jpg", "img5.jpg")]
    lib = jd.load()
    dec = jd.Decoder(0)
    dec.set_timing(True)
    for path in names:
        data = open(path, "rb").read()
        ptr = lib.b200jpeg_host_alloc(len(data))
        buf = np.ctypeslib.as_array((C.c_uint8 * len(data)).from_address(ptr))
        buf[:] = np.frombuffer(data, np.uint8)
        descs, rcs = jd.parse_headers([buf])
        assert rcs == [0]
        w, h = descs[0].image_width, descs[0].image_height
        rptr = lib.b200jpeg_host_alloc(3 * w * h)
        rgb = np.ctypeslib.as_array((C.c_uint8 * (3 * w * h)).from_address(rptr))
        dec.prepare(descs, [buf])
        gpu = []
        for k in range(30):
            dec.run(); st = dec.fetch(None, None)
            if k >= 10:
                gpu.append(dec.last_timing())
        assert st == [0]
        e2e = []
        for k in range(30):
            t0 = time.perf_counter()
            d, _ = jd.parse_headers([buf], descs=descs)
            st = dec.decode_batch(d, [buf], [rgb])
            if k >= 10:
                e2e.append((time.perf_counter() - t0) * 1e3)
        assert st == [0]
        line = {"file": os.path.basename(path), "w": w, "h": h, "bytes": len(data),
                "gpu_ms": {k: float(np.median([g[k] for g in gpu])) for k in gpu[0]},
                "host_to_host_ms": float(np.median(e2e)), "launches": dec.stats()["kernel_launches"],
                "entropy_path": dec.stats()["entropy_path"]}
        if oracle.have_ref():
            r = oracle.ref_bench([path], reps=5)
            line["reference_1core_ms"] = r["seconds"] / r["images"] * 1e3
        print(json.dumps(line), flush=True)
        lib.b200jpeg_host_free(ptr); lib.b200jpeg_host_free(rptr)
    dec.close()


if __name__ == "__main__":
    main()
