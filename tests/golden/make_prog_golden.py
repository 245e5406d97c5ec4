"""Progressive (SOF2) fixtures.  Run here (needs Pillow and /root/reference): python tests/golden/make_prog_golden.py

The reference rejects progressive files (jpeg.cpp:69-73), so it cannot witness them directly.  But an encoder computes
the same quantised coefficients whether it then writes them as one baseline scan or as progressive scans: for every
fixture this script lets Pillow (libjpeg) write BOTH from the same pixels, runs the REFERENCE BINARY on the baseline twin,
and stores the digests of its coefficient dump and RGB -- which the progressive file must reproduce exactly.
Stored: tests/golden/prog/<name>.jpg (progressive), <name>_twin.jpg (baseline), golden_prog.json."""
import hashlib
import io
import json
import os
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle  # noqa: E402
from jpeg_decoder_b200.harness import synth_rgb  # noqa: E402

# name -> (seed, w, h, quality, Pillow subsampling 0=4:4:4 1=4:2:2 2=4:2:0, restart rows, grey)
CASES = {
    "p420_200x120_q85": (1, 200, 120, 85, 2, 0, False),
    "p444_333x211_q75": (2, 333, 211, 75, 0, 0, False),
    "p422_640x480_q92": (3, 640, 480, 92, 1, 0, False),
    "p420_97x55_q50": (4, 97, 55, 50, 2, 0, False),
    "p444_64x64_q30": (6, 64, 64, 30, 0, 0, False),
    "p420_320x240_q90_rst": (7, 320, 240, 90, 2, 1, False),
    "pgray_160x100_q80": (9, 160, 100, 80, 0, 0, True),
}


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    oracle.build()
    out = {}
    for name, (seed, w, h, q, ss, rst, grey) in CASES.items():
        img = Image.fromarray(synth_rgb(seed, w, h))
        if grey:
            img = img.convert("L")
        kw = dict(quality=q) if grey else dict(quality=q, subsampling=ss)
        if rst:
            kw["restart_marker_rows"] = rst
        b = io.BytesIO(); img.save(b, "JPEG", progressive=True, **kw); prog = b.getvalue()
        b = io.BytesIO(); img.save(b, "JPEG", progressive=False, optimize=False, **kw); twin = b.getvalue()
        assert b"\xff\xc2" in prog and b"\xff\xc2" not in twin
        if grey:                      # the reference rejects 1-component files too (jpeg.cpp:83-87): the port under SPEC rules
            rc, info, coef, rgb = oracle.port_decode(twin, flags=oracle.F_SPEC)
            assert rc == 0
            witness = "port (grey is rejected by the reference)"
        else:
            info, coef, rgb = oracle.ref_dump(twin)
            witness = "reference binary on the baseline twin"
        rc, pi, pc, pr = oracle.port_decode_progressive(prog)
        assert rc == 0 and np.array_equal(pc, coef) and np.array_equal(pr, rgb), name
        open(os.path.join(HERE, "prog", name + ".jpg"), "wb").write(prog)
        open(os.path.join(HERE, "prog", name + "_twin.jpg"), "wb").write(twin)
        out[name] = dict(w=w, h=h, scans=pi["scan_unstuffed"], witness=witness, coef_sha256=digest(coef), rgb_sha256=digest(rgb),
                         prog_sha256=hashlib.sha256(prog).hexdigest(), twin_sha256=hashlib.sha256(twin).hexdigest())
        print(name, len(prog), len(twin), pi["scan_unstuffed"], "scans")
    json.dump(out, open(os.path.join(HERE, "golden_prog.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
