#!/usr/bin/env python
"""Timing of the progressive (SOF2) path on the reference's own fixtures, next to the Annex-G port on one host core.
    python scripts/dev/prog_time.py   -> one JSON line per case"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
from oracle import oracle  # noqa: E402


def main():
    dec = jd.Decoder(0)
    g = os.path.join(ROOT, "tests", "golden")
    big, small = open(os.path.join(g, "prograssive-sample.jpg"), "rb").read(), open(os.path.join(g, "prograssive-sample-2.jpg"), "rb").read()
    for name, datas in (("prograssive-sample-2.jpg x1", [small]), ("prograssive-sample.jpg x1", [big]), ("prograssive-sample.jpg x32", [big] * 32),
                        ("prograssive-sample-2.jpg x256", [small] * 256)):
        ts = []
        for k in range(4):
            t0 = time.perf_counter()
            st, rgbs, _ = dec.decode_progressive(datas)
            ts.append(time.perf_counter() - t0)
        assert not any(st)
        mp = sum(r.shape[0] * r.shape[1] for r in rgbs) / 1e6
        t = min(ts[1:])
        line = {"case": name, "megapixels": mp, "host_to_host_ms": t * 1e3, "mp_s": mp / t, "gpu_stage_ms": None}
        t0 = time.perf_counter()
        rc, _, _, _ = oracle.port_decode_progressive(datas[0], want_coef=False)
        line["port_1core_ms_per_image"] = (time.perf_counter() - t0) * 1e3
        print(json.dumps(line), flush=True)
    dec.close()


if __name__ == "__main__":
    main()
