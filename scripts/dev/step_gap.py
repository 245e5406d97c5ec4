"""Dev tool: wall-clock ms per device-resident step with and without the per-stage CUDA events (what the events cost)."""
import sys, time, os
sys.path.insert(0, os.getcwd())
import jpeg_decoder_b200 as jd
from jpeg_decoder_b200 import abi
import bench
for key in ("C2", "C3"):
    class A:
        workload, images, quality = key, 0, 0
    wl = bench.workload(A)
    dec = jd.Decoder(0)
    files = bench.make_files(wl, 0, wl["n"])
    descs, rcs = jd.parse_headers(files, wl["flags"])
    dec.prepare(descs, files)
    for timing in (True, False, True, False):
        dec.set_timing(timing)
        for _ in range(3):
            dec.run(); dec.fetch(None, None)
        t0 = time.perf_counter()
        for _ in range(20):
            dec.run(); dec.fetch(None, None)
        dt = (time.perf_counter() - t0) / 20 * 1e3
        print(key, "timing", timing, "ms/step %.4f" % dt, ("events %.4f" % dec.last_timing()["total"]) if timing else "")
    dec.close()
