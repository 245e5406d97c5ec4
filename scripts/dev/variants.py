"""Dev tool (not part of the product): per-stage device times of several builds of the library
(build/variants/*.so, same sources, different -D flags) on the bench workload, in ONE process --
the batch is generated once.  Optionally an e2e sweep over B200JPEG_CHUNK.

    python scripts/dev/variants.py [--images 256] [--steps 10] [--e2e 8,16,32,64] [names...]
"""
import argparse, ctypes as C, glob, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402  (device sync only)
import jpeg_decoder_b200 as jd  # noqa: E402
from jpeg_decoder_b200 import abi  # noqa: E402
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--images", type=int, default=256)
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--e2e", default="")
ap.add_argument("--entropy", default="auto")
ap.add_argument("names", nargs="*")
a = ap.parse_args()

base = jd.load()
n = a.images
arena, files = bench.make_batch(n, 1000, pinned_alloc=base.b200jpeg_host_alloc)
descs = (abi.ImageDesc * n)()
for i, b in enumerate(files):
    assert base.b200jpeg_parse_header(b.ctypes.data, b.size, 0, C.byref(descs[i])) == 0
rb = 3 * bench.W * bench.H
ptr = base.b200jpeg_host_alloc(n * rb)
host = np.ctypeslib.as_array((C.c_uint8 * (n * rb)).from_address(ptr))
rgbs = [host[i * rb:(i + 1) * rb] for i in range(n)]

paths = sorted(glob.glob(os.path.join(ROOT, "build", "variants", "*.so")))
if a.names:
    paths = [p for p in paths if os.path.basename(p)[:-3] in a.names]
mode = {"auto": abi.ENTROPY_AUTO, "segment": abi.ENTROPY_SEGMENT, "sync": abi.ENTROPY_SYNC}[a.entropy]
for p in paths:
    name = os.path.basename(p)[:-3]
    try:
        jd._lib = abi.bind(C.CDLL(p))
        dec = jd.Decoder(0)
        dec.set_entropy_path(mode)
        dec.set_timing(True)
        dec.prepare(descs, files)
        for _ in range(3):
            dec.run(); st = dec.fetch(None, None)
        bad = sum(1 for s in st if s)
        acc = {}
        t0 = time.perf_counter()
        for _ in range(a.steps):
            dec.run(); dec.fetch(None, None)
            for k, v in dec.last_timing().items():
                acc[k] = acc.get(k, 0.0) + v
        wall = (time.perf_counter() - t0) / a.steps * 1e3
        out = {"variant": name, "wall_ms": round(wall, 3), "bad_status": bad}
        out.update({k: round(v / a.steps, 3) for k, v in acc.items()})
        for spec in [x for x in a.e2e.split(",") if x]:
            ch, _, sl = spec.partition(":")
            ch = int(ch)
            os.environ["B200JPEG_CHUNK"] = str(ch)
            if sl:
                os.environ["B200JPEG_SLOTS"] = sl
            dec.decode_batch(descs, files, rgbs)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5):
                dec.decode_batch(descs, files, rgbs)
            out["e2e_chunk%s_ms" % spec] = round((time.perf_counter() - t0) / 5 * 1e3, 2)
        os.environ.pop("B200JPEG_CHUNK", None)
        os.environ.pop("B200JPEG_SLOTS", None)
        print(json.dumps(out), flush=True)
        dec.close()
    except Exception as e:  # keep going: a diagnostic build may fail on purpose
        print(json.dumps({"variant": name, "error": repr(e)}), flush=True)
