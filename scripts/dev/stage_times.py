#!/usr/bin/env python
"""Dev tool: per-stage CUDA-event times (median of N device-resident steps) of one workload with the library
B200JPEG_LIB names, after the golden-digest parity check through both entropy decoders.
    B200JPEG_LIB=build/variants/x.so python scripts/dev/stage_times.py C2 [images] [entropy]  -> one JSON line"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
from jpeg_decoder_b200 import abi  # noqa: E402
import bench  # noqa: E402


def main():
    key = sys.argv[1] if len(sys.argv) > 1 else "C2"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    ent = sys.argv[3] if len(sys.argv) > 3 else "auto"

    class A:
        workload, images, quality = key, n, 0
    wl = bench.workload(A)
    dec = jd.Decoder(0)
    ok = None if os.environ.get("B2_DIAG_BUILD") else bench.rank_parity_check(dec, abi)   # a diagnostic build decodes wrongly on purpose
    dec.set_entropy_path({"auto": abi.ENTROPY_AUTO, "segment": abi.ENTROPY_SEGMENT, "sync": abi.ENTROPY_SYNC}[ent])
    dec.set_timing(True)
    files = bench.make_files(wl, 0, wl["n"])
    descs, rcs = jd.parse_headers(files, wl["flags"])
    dec.prepare(descs, files)
    ts = []
    for k in range(14):
        dec.run(); st = dec.fetch(None, None)
        if k >= 4:
            ts.append(dec.last_timing())
    assert os.environ.get("B2_DIAG_BUILD") or all(s == 0 for s in st)
    med = {k: round(float(np.median([t[k] for t in ts])), 4) for k in ts[0]}
    print(json.dumps({"lib": os.path.basename(jd.LIB_SO), "workload": key, "images": wl["n"], "entropy": ent, "parity": ok, **med}), flush=True)
    dec.close()


if __name__ == "__main__":
    main()
