#!/usr/bin/env python
"""Dev tool (CPU only): random corruption of the HEADER bytes of synthetic JPEGs; the host parser of the library
(b200jpeg_parse_header) against the oracle's restatement of the reference's parser, in both modes.
Checks: whatever the oracle decodes, the parser accepts; whatever the parser refuses, the oracle refuses.
    python scripts/dev/header_fuzz.py [cases] [seed]  -> one JSON line"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
from jpeg_decoder_b200 import harness  # noqa: E402
from oracle import oracle  # noqa: E402

cases = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
harness.build_harness(); oracle.build()
stat = {"cases": 0, "both_accept": 0, "both_refuse": 0, "parser_accepts_oracle_fails_later": 0, "mismatch": 0}
bad, later = [], {}
for c in range(cases):
    w, h = int(rng.integers(1, 200)), int(rng.integers(1, 150))
    hs, vs = [(1, 1), (2, 1), (1, 2), (2, 2)][int(rng.integers(0, 4))]
    gray = bool(rng.integers(0, 5) == 0)
    if gray:
        hs = vs = 1
    d = bytearray(harness.synth_jpeg(int(rng.integers(0, 1 << 30)), w, h, int(rng.integers(30, 99)), hs, vs, gray,
                                     int([0, 0, -1, 1, 3][int(rng.integers(0, 5))])))
    head = d.find(b"\xff\xda") + 14
    kind = int(rng.integers(0, 4))
    if kind == 0:
        for p in rng.integers(2, head, size=int(rng.integers(1, 4))):
            d[p] = int(rng.integers(0, 256))
    elif kind == 1:
        p = int(rng.integers(2, head)); d[p] = (d[p] + int(rng.integers(1, 4))) & 255
    elif kind == 2:
        p = int(rng.integers(2, head)); del d[p:p + int(rng.integers(1, 6))]
    else:
        p = int(rng.integers(2, head)); d[p:p] = bytes(rng.integers(0, 256, size=int(rng.integers(1, 6)), dtype=np.uint8))
    d = bytes(d)
    for flags in (0, 1):
        stat["cases"] += 1
        rc_o = oracle.port_decode(d, flags=flags, want_coef=False, want_rgb=False)[0]
        _, rcs = jd.parse_headers([np.frombuffer(d, np.uint8).copy()], flags)
        rc_p = int(rcs[0])
        if rc_o == 0 and rc_p == 0:
            stat["both_accept"] += 1
        elif rc_o != 0 and rc_p != 0:
            stat["both_refuse"] += 1
        elif rc_o != 0 and rc_p == 0:
            stat["parser_accepts_oracle_fails_later"] += 1     # allowed: the failure may lie in the scan
            later[rc_o] = later.get(rc_o, 0) + 1
        else:
            stat["mismatch"] += 1
            bad.append((c, flags, rc_o, rc_p))
            os.makedirs(os.path.join(ROOT, "gpurun_out", "hfuzz"), exist_ok=True)
            open(os.path.join(ROOT, "gpurun_out", "hfuzz", "case_%d_%d.jpg" % (c, flags)), "wb").write(d)
print(json.dumps(stat), bad[:10], "oracle codes where the parser accepted:", later)
