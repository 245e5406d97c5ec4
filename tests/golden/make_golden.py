"""Regenerates tests/golden/golden.json by running the UNTOUCHED reference (oracle/_ref/ref_decode,
built from /root/reference by oracle/Makefile) on the reference's own fixture images (copied here
as data, since /root/reference does not exist on the GPU box) and on seeded synthetic streams from
the bundled encoder.  Stored per case: SHA-256 of the MCU-array dump after decodeHuffman()
(int32 [pos][3][64]) and of the gathered RGB (uint8 [H][W][3]), plus the geometry.

Run here (needs /root/reference):  python tests/golden/make_golden.py [--named]
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle  # noqa: E402
from jpeg_decoder_b200.harness import synth_jpeg  # noqa: E402

FIXTURES = ["img.jpg", "img2.jpg", "img3.jpg", "img4.jpg", "img5.jpg", "img6.jpg"]
# name -> synth_jpeg kwargs; all inside the subset the reference decodes correctly
SYNTH = {
    "s420_rstrow_200x120": dict(seed=11, w=200, h=120, quality=90, hs=2, vs=2, restart=-1),
    "s420_norst_333x211": dict(seed=12, w=333, h=211, quality=85, hs=2, vs=2, restart=0),
    "s422_rst7_200x120": dict(seed=13, w=200, h=120, quality=90, hs=2, vs=1, restart=7),
    "s440_norst_200x120": dict(seed=14, w=200, h=120, quality=75, hs=1, vs=2, restart=0),
    "s444_norst_123x77": dict(seed=15, w=123, h=77, quality=95, hs=1, vs=1, restart=0),
    "s444_rst3_96x40": dict(seed=16, w=96, h=40, quality=50, hs=1, vs=1, restart=3),
    "s420_1080p_rstrow": dict(seed=17, w=1920, h=1080, quality=90, hs=2, vs=2, restart=-1),
    "s420_512_q85": dict(seed=18, w=512, h=512, quality=85, hs=2, vs=2, restart=0),
    "s444_1024_q95": dict(seed=19, w=1024, h=1024, quality=95, hs=1, vs=1, restart=0),
    "s420_1x1": dict(seed=20, w=1, h=1, quality=90, hs=2, vs=2, restart=0),
}
# BASELINE.json configs at their NAMED sizes; minutes of reference time each, so they are
# only (re)made with --named and otherwise carried over from the existing golden.json.
#   C3: one 8192x8192 4:4:4 q95 without restart markers (the reference needs ~805 MB for MCU[] and ~1 heap
#       chunk per scan byte)
#   C4: 4096x4096 x {4:2:2, 4:4:4} x {q50, q90} (grey is rejected by the reference, jpeg.cpp:83-87: port only)
NAMED = {
    "c3_444_8192_q95": dict(seed=3001, w=8192, h=8192, quality=95, hs=1, vs=1, restart=0),
    "c4_422_4096_q50": dict(seed=4150, w=4096, h=4096, quality=50, hs=2, vs=1, restart=0),
    "c4_422_4096_q90": dict(seed=4190, w=4096, h=4096, quality=90, hs=2, vs=1, restart=0),
    "c4_444_4096_q50": dict(seed=4250, w=4096, h=4096, quality=50, hs=1, vs=1, restart=0),
    "c4_444_4096_q90": dict(seed=4290, w=4096, h=4096, quality=90, hs=1, vs=1, restart=0),
    "c2_420_1080p_rstrow_255": dict(seed=1255, w=1920, h=1080, quality=90, hs=2, vs=2, restart=-1),
    "c5_420_512_q85_16383": dict(seed=5000 + 16383, w=512, h=512, quality=85, hs=2, vs=2, restart=0),
}


def digest(a):
    return hashlib.sha256(a.tobytes()).hexdigest()


def entry(kw):
    data = synth_jpeg(**kw)
    info, coef, rgb = oracle.ref_dump(data)
    return dict(args=kw, jpeg_sha256=hashlib.sha256(data).hexdigest(), jpeg_bytes=len(data),
                coef_sha256=digest(coef), rgb_sha256=digest(rgb))


def main():
    oracle.build()
    out = {"fixtures": {}, "synthetic": {}, "named_size": {}}
    path = os.path.join(HERE, "golden.json")
    if os.path.exists(path):
        with open(path) as f:
            out["named_size"] = json.load(f).get("named_size", {})
    if "--named" in sys.argv:
        for name, kw in NAMED.items():
            out["named_size"][name] = entry(kw)
            print("named", name, out["named_size"][name]["jpeg_bytes"], flush=True)
    for f in FIXTURES:
        info, coef, rgb = oracle.ref_dump(os.path.join(HERE, f))
        out["fixtures"][f] = {k: info[k] for k in ("w", "h", "bwr", "bhr", "hs", "vs", "ri", "scan_unstuffed")}
        out["fixtures"][f].update(coef_sha256=digest(coef), rgb_sha256=digest(rgb))
    for name, kw in SYNTH.items():
        out["synthetic"][name] = entry(kw)
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote golden.json:", len(out["fixtures"]), "fixtures,", len(out["synthetic"]), "synthetic")


if __name__ == "__main__":
    main()
