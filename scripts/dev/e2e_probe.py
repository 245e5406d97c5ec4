"""Dev probe: where does the time of b200jpeg_decode_batch go?  (not part of the product)"""
import ctypes as C, sys, time, os
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
def T(f, k=3):
    f(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(k): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / k * 1e3
print("decode_batch ms", T(lambda: dec.decode_batch(descs, files, rgbs)))
def staged():
    dec.prepare(descs, files); dec.run(); dec.fetch(rgbs, None)
print("staged ms", T(staged))
def prep_only():
    dec.prepare(descs, files); dec.sync()
print("prepare(H2D) ms", T(prep_only))
dec.prepare(descs, files)
def run_only():
    dec.run(); dec.fetch(None, None)
print("run ms", T(run_only))
def fetch_only():
    dec.fetch(rgbs, None)
print("fetch(D2H) ms", T(fetch_only))
# raw torch copies for comparison
d = torch.empty(n * rb, dtype=torch.uint8, device="cuda"); h = torch.from_numpy(host)
print("torch D2H one copy ms", T(lambda: h.copy_(d, non_blocking=True)))
for chunk in (8, 32, 64):
    sub = [(descs.__class__._type_ * chunk)(*[descs[i] for i in range(chunk)])]
    print("decode_batch n=%d ms" % chunk, T(lambda: dec.decode_batch(sub[0], files[:chunk], rgbs[:chunk])))
