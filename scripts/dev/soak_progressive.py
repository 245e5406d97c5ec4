#!/usr/bin/env python
"""Dev tool: the progressive (SOF2) path under random corruption -- the fixtures of tests/golden/prog with bytes of their
scans flipped, zeroed, filled with FF 00 or FF, truncated, or with chunks copied around, in batches next to intact copies.
Checks: no hang (run under `timeout`), intact pictures equal to the Annex-G port, a corrupted one either flagged or equal
to what the port makes of it.
    timeout 300 python scripts/dev/soak_progressive.py [rounds] [seed]  -> one JSON line"""
import glob
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
from jpeg_decoder_b200 import B200JpegError  # noqa: E402
from oracle import oracle  # noqa: E402

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 50
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
oracle.build()
fixtures = [open(f, "rb").read() for f in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "prog", "*.jpg"))) if "_twin" not in f]
ref = [oracle.port_decode_progressive(d) for d in fixtures]
assert all(r[0] == 0 for r in ref)


def corrupt(d):
    m = bytearray(d)
    first = d.find(b"\xff\xda")
    at = int(rng.integers(first + 14, len(d) - 4))
    ln = int(rng.integers(1, min(len(d) - 2 - at, 300)))
    kind = int(rng.integers(0, 6))
    if kind == 0:
        for p in rng.integers(first + 14, len(d) - 2, size=int(rng.integers(1, 20))):
            m[p] = int(rng.integers(0, 255))
    elif kind == 1:
        m = m[:at] + b"\xff\xd9"
    elif kind == 2:
        m[at:at + ln] = (b"\xff\x00" * ln)[:ln]
    elif kind == 3:
        m[at:at + ln] = bytes(ln)
    elif kind == 4:
        m[at:at + ln] = b"\xff" * ln
    else:
        src = int(rng.integers(first + 14, len(d) - ln - 2))
        m[at:at + ln] = m[src:src + ln]
    return bytes(m)


dec = jd.Decoder(0)
stat = {"batches": 0, "intact": 0, "corrupt": 0, "refused_by_parser": 0, "flagged": 0, "unflagged_same_as_port": 0,
        "unflagged_port_rejects": 0, "unflagged_differs_from_port": 0}
for r in range(rounds):
    datas, which = [], []
    for _ in range(int(rng.integers(2, 10))):
        k = int(rng.integers(0, len(fixtures)))
        if rng.integers(0, 3) == 0:
            datas.append(fixtures[k]); which.append((k, True))
        else:
            datas.append(corrupt(fixtures[k])); which.append((k, False))
    keep = []
    for d, w in zip(datas, which):                       # what the progressive parser refuses, the port must refuse too
        try:
            dec.decode_progressive([d])
            keep.append((d, w))
        except B200JpegError:
            stat["refused_by_parser"] += 1
            assert oracle.port_decode_progressive(d)[0] != 0
    if not keep:
        continue
    st, rgbs, coefs = dec.decode_progressive([d for d, _ in keep], want_coef=True)
    stat["batches"] += 1
    for i, (d, (k, intact)) in enumerate(keep):
        if intact:
            stat["intact"] += 1
            assert st[i] == 0 and np.array_equal(coefs[i].astype(np.int32), ref[k][2]) and np.array_equal(rgbs[i], ref[k][3])
            continue
        stat["corrupt"] += 1
        if st[i] != 0:
            stat["flagged"] += 1
            continue
        rc, _, oc, orgb = oracle.port_decode_progressive(d)
        if rc == 0 and not np.array_equal(coefs[i].astype(np.int32), oc):
            stat["unflagged_differs_from_port"] += 1
            os.makedirs(os.path.join(ROOT, "gpurun_out", "soak"), exist_ok=True)
            open(os.path.join(ROOT, "gpurun_out", "soak", "progdiff_s%s_r%d_i%d.jpg" % (sys.argv[2] if len(sys.argv) > 2 else "1", r, i)), "wb").write(d)
        elif rc == 0:
            stat["unflagged_same_as_port"] += 1
        else:
            stat["unflagged_port_rejects"] += 1
            os.makedirs(os.path.join(ROOT, "gpurun_out", "soak"), exist_ok=True)
            open(os.path.join(ROOT, "gpurun_out", "soak", "prog_r%d_i%d.jpg" % (r, i)), "wb").write(d)
            print(json.dumps({"port_rc": rc, "round": r, "image": i}), flush=True)
print(json.dumps(stat))
assert stat["unflagged_differs_from_port"] == 0 and stat["unflagged_port_rejects"] == 0
