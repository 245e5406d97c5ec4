#!/usr/bin/env python
"""Dev tool: where the decoder warps of k_entropy_seg spend their cycles (a -DB2_SEG_DIAG=7 build only).
    B200JPEG_LIB=build/variants/diag7.so python scripts/dev/seg_diag.py [C2]  -> one JSON line: cycles per round by part"""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
from jpeg_decoder_b200 import abi  # noqa: E402
import bench  # noqa: E402

key = sys.argv[1] if len(sys.argv) > 1 else "C2"


class A:
    workload, images, quality = key, 0, 0


wl = bench.workload(A)
dec = jd.Decoder(0)
dec.set_timing(True)
files = bench.make_files(wl, 0, wl["n"])
descs, rcs = jd.parse_headers(files, wl["flags"])
dec.prepare(descs, files)
lib = ctypes.CDLL(jd.LIB_SO)
buf = (ctypes.c_ulonglong * 24)()
for k in range(4):
    dec.run(); dec.fetch(None, None)
lib.b200jpeg_debug_seg_diag(buf)
dec.run(); dec.fetch(None, None)
t = dec.last_timing()
lib.b200jpeg_debug_seg_diag(buf)
v = list(buf)
names = ["topup", "wait_room", "cp_wait", "steps", "ninth_step", "publish"]
rounds, warps = v[6], v[7]
print(json.dumps({"lib": os.path.basename(jd.LIB_SO), "write_ms": round(t["write"], 4), "decoder_warps": warps, "rounds_per_warp": round(rounds / warps, 1),
                  "cycles_per_round": {n: round(v[i] / rounds, 1) for i, n in enumerate(names)}, "total_per_round": round(sum(v[:6]) / rounds, 1),
                  "rounds_without_room": round(v[8] / rounds, 3),
                  "lanes_busy": round(v[9] / (32.0 * rounds), 3),   # share of (lane, round) slots in which the lane still had work
                  "writer_per_own_round": {n: round(v[10 + i] / max(v[13], 1), 1) for i, n in enumerate(["wait_full", "wait_asm_room", "read+compute"])},
                  "flusher_per_round": {n: round(v[16 + i] / max(v[19], 1), 1) for i, n in [(0, "wait_full"), (1, "wait_worked"), (4, "masks+dc"), (5, "list"), (6, "copy"), (2, "release"), (7, "blocks")]}}))
dec.close()
