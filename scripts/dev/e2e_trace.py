"""Dev probe: CUPTI timeline (via torch.profiler) of one b200jpeg_decode_batch call."""
import ctypes as C, sys, os, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import jpeg_decoder_b200 as jd
from jpeg_decoder_b200 import abi
import bench
lib = jd.load(); dec = jd.Decoder(0)
n = 256
arena, files = bench.make_batch(n, 1000, pinned_alloc=lib.b200jpeg_host_alloc)
descs = (abi.ImageDesc * n)()
for i, b in enumerate(files): assert lib.b200jpeg_parse_header(b.ctypes.data, b.size, 0, C.byref(descs[i])) == 0
rb = 3 * 1920 * 1080
ptr = lib.b200jpeg_host_alloc(n * rb)
host = np.ctypeslib.as_array((C.c_uint8 * (n * rb)).from_address(ptr))
rgbs = [host[i * rb:(i + 1) * rb] for i in range(n)]
for _ in range(2): dec.decode_batch(descs, files, rgbs)
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    dec.decode_batch(descs, files, rgbs)
    torch.cuda.synchronize()
prof.export_chrome_trace("gpurun_out/e2e_trace.json")
ev = json.load(open("gpurun_out/e2e_trace.json"))["traceEvents"]
gpu = [e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "ts" in e]
gpu.sort(key=lambda e: e["ts"])
t0 = gpu[0]["ts"]
# merge consecutive same-category events per stream into spans
spans = []
for e in gpu:
    kind = "D2H" if "DtoH" in e["name"] else "H2D" if "HtoD" in e["name"] else "set" if e["cat"] == "gpu_memset" else "K"
    st = e["args"].get("stream")
    s, f = e["ts"] - t0, e["ts"] - t0 + e["dur"]
    if spans and spans[-1][0] == kind and spans[-1][1] == st and s - spans[-1][3] < 300: spans[-1][3] = f; spans[-1][4] += e["dur"]
    else: spans.append([kind, st, s, f, e["dur"]])
for sp in spans:
    if sp[3] - sp[2] > 200: print("%-4s stream %s  %8.2f -> %8.2f ms (busy %.2f)" % (sp[0], sp[1], sp[2] / 1e3, sp[3] / 1e3, sp[4] / 1e3))
