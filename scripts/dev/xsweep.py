#!/usr/bin/env python
"""Dev tool (needs the B200JPEG_TUNING build: scripts/build_variant.sh tuning -DB200JPEG_TUNING, B200JPEG_LIB=build/variants/tuning.so):
sub size / unit size / run-in of the self-synchronising path, per-stage CUDA-event times on fixed inputs.
    B200JPEG_LIB=build/variants/tuning.so python scripts/dev/xsweep.py [C2|C3|C5|C1 ...]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
from jpeg_decoder_b200 import abi  # noqa: E402
import bench  # noqa: E402

CASES = {"C2": ("C2", 256), "C3": ("C3", 1), "C5": ("C5", 2048), "C4": ("C4-422", 4), "C1": ("C1", 1)}


def main():
    which = sys.argv[1:] or ["C2", "C3", "C5"]
    dec = jd.Decoder(0)
    dec.set_timing(True)
    dec.set_entropy_path(abi.ENTROPY_SYNC)
    for w in which:
        key, n = CASES[w]

        class A:
            workload, images, quality = key, n, 0
        if key == "C1":                                # BASELINE config 1: the reference's own fixture, one small picture
            files = [np.frombuffer(open(os.path.join(ROOT, "tests", "golden", "img.jpg"), "rb").read(), np.uint8)]
            descs, rcs = jd.parse_headers(files, 0)
            subs, units, ovs = (9, 10, 11, 12), (9, 10, 11, 12), (512, 1024, 2048, 4096)
        else:
            wl = bench.workload(A)
            files = bench.make_files(wl, 0, n)
            descs, rcs = jd.parse_headers(files, wl["flags"])
            subs, units, ovs = (11, 12, 13, 14), (11, 12, 13, 14, 15), (2048, 4096)
        for sub in subs:
            for unit_bits in units:
                if unit_bits < sub:
                    continue
                for ov in ovs:
                    os.environ["B200JPEG_XSUB"] = str(sub)
                    os.environ["B200JPEG_XUNIT"] = str(1 << (unit_bits - sub))
                    os.environ["B200JPEG_XOVERLAP"] = str(ov)
                    dec.prepare(descs, files)
                    ts = []
                    for k in range(6):
                        dec.run(); st = dec.fetch(None, None)
                        if k >= 2:
                            ts.append(dec.last_timing())
                    assert all(s == 0 for s in st), st[:4]
                    med = {k: float(np.median([t[k] for t in ts])) for k in ts[0]}
                    print(json.dumps({"case": w, "sub": sub, "unit_bits": unit_bits, "overlap": ov, "sync_rounds": dec.stats()["sync_rounds"],
                                      **{k: round(v, 4) for k, v in med.items()}}), flush=True)
    dec.close()


if __name__ == "__main__":
    main()
