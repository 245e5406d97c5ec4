#!/usr/bin/env python
"""bench.py -- decoded megapixels/s of the scan-to-pixels path (BASELINE.json metric).

Workload (config.workload): BASELINE.json configs[1] -- a batch of 256 synthetic 1920x1080 YCbCr
4:2:0 q90 baseline JPEGs with a restart interval of one MCU row, made by the bundled encoder.
A "step" is one pass of the whole path (unstuff -> entropy decode -> fused pixel stage) over that
batch.  `value`: inputs already resident in HBM.  `e2e`: the same through the C ABI with HOST
buffers (pinned), H2D of the files and D2H of the RGB inside the timed region.
Multi-GPU (torchrun, one process per GPU): every rank decodes its own batch (weak scaling), no
collective on the data path; time = max over ranks.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--images 256]
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, QUALITY = 1920, 1080, 90
METRIC = "decoded megapixels/s (baseline 4:2:0 JPEG)"
UNIT = "MP/s"


def workload_name(n):
    return "C2: %d x 1920x1080 4:2:0 q90 baseline JPEG, DRI = 1 MCU row (bundled encoder, synthetic)" % n


def make_batch(n, seed0, pinned_alloc=None):
    """n synthetic JPEGs packed back to back (256-byte aligned) in one buffer -> (arena, views)."""
    from jpeg_decoder_b200.harness import synth_jpeg, build_harness
    build_harness()
    cap = W * H  # generous per-image bound (~0.29 B/px expected)
    tmp = [np.empty(cap, np.uint8) for _ in range(n)]
    with ThreadPoolExecutor(max_workers=max(1, min(32, os.cpu_count() or 1))) as ex:
        lens = list(ex.map(lambda i: synth_jpeg(seed0 + i, W, H, QUALITY, 2, 2, False, -1, out=tmp[i]), range(n)))
    total = sum((l + 255) // 256 * 256 for l in lens) + 4096
    if pinned_alloc is not None:
        ptr = pinned_alloc(total)
        if not ptr:
            raise MemoryError("pinned allocation of %d bytes failed" % total)
        arena = np.ctypeslib.as_array((C.c_uint8 * total).from_address(ptr))
    else:
        arena = np.empty(total, np.uint8)
    views, off = [], 0
    for i in range(n):
        arena[off:off + lens[i]] = tmp[i][:lens[i]]
        views.append(arena[off:off + lens[i]])
        off += (lens[i] + 255) // 256 * 256
    return arena, views


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel, images):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/traffic.json);
    None when the capture was taken on another batch size or does not hold the kernel."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        k = t["kernels"][kernel]
        return k["read"] + k["write"] if t["images"] == images else None
    except (OSError, KeyError, ValueError):
        return None


# ------------------------------------------------------------------ CPU side (oracle as baseline)
def cpu_reference_run(files, per_worker, workers):
    """Time the reference's own CPU path on `workers` processes, `per_worker` images each.

    kind "reference": oracle/_ref/ref_decode = the UNTOUCHED jpeg.cpp compiled headless; the timed
    region per image is ctor + readJPEG() (parse + unstuff + Huffman + dequant + IDCT + colour).
    Falls back to the C restatement (kind "port", 1 thread) when the binary is absent.
    Returns (MP/s, cores, kind, sample description)."""
    from oracle import oracle
    oracle.build()
    with tempfile.TemporaryDirectory() as td:
        paths = []
        for i, v in enumerate(files):
            p = os.path.join(td, "f%04d.jpg" % i)
            with open(p, "wb") as f:
                f.write(v.tobytes())
            paths.append(p)
        if oracle.have_ref():
            procs = []
            for w in range(workers):
                mine = [paths[(w * per_worker + k) % len(paths)] for k in range(per_worker)]
                procs.append(subprocess.Popen([oracle.REF_BIN, "bench", "1"] + mine, stdout=subprocess.PIPE, text=True))
            outs = [json.loads(p.communicate()[0].strip().splitlines()[-1]) for p in procs]
            if any(p.returncode != 0 for p in procs):
                raise RuntimeError("reference bench worker failed")
            pixels = sum(o["pixels"] for o in outs)
            secs = max(o["seconds"] for o in outs)
            return pixels / secs / 1e6, workers, "reference", "%d worker processes x %d images of the workload" % (workers, per_worker)
        t0 = time.time()
        pixels = 0
        for k in range(per_worker):
            rc, info, _, _ = oracle.port_decode(files[k % len(files)].tobytes(), want_coef=False)
            assert rc == 0
            pixels += info["width"] * info["height"]
        return pixels / (time.time() - t0) / 1e6, 1, "port", "%d images of the workload, 1 thread" % per_worker


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    ncpu = os.cpu_count() or 1
    workers = max(1, ncpu)
    per_worker = 2
    _, files = make_batch(min(args.images, max(8, min(64, workers))), 1000)
    vals = []
    for s in range(args.warmup + args.steps):
        v, cores, kind, sample = cpu_reference_run(files, per_worker, workers)
        if s >= args.warmup:
            vals.append(v)
    value = float(np.mean(vals))
    pixels_step = workers * per_worker * W * H
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": pixels_step / value / 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32+int32", "data": "synthetic",
            "config": {"workload": workload_name(args.images), "step": "bounded sample: " + sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample + " per step"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ------------------------------------------------------------------ GPU arm
def run_gpu_arm(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import jpeg_decoder_b200 as jd
    from jpeg_decoder_b200 import abi, sharding

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = jd.load()
    dec = jd.Decoder(local_rank)
    dec.set_entropy_path({"auto": abi.ENTROPY_AUTO, "segment": abi.ENTROPY_SEGMENT, "sync": abi.ENTROPY_SYNC}[args.entropy])
    n = args.images
    seed0, _ = sharding.weak_shard(n, rank)
    arena, files = make_batch(n, seed0, pinned_alloc=lib.b200jpeg_host_alloc)   # pinned host files
    descs = (abi.ImageDesc * n)()
    for i, b in enumerate(files):
        rc = lib.b200jpeg_parse_header(b.ctypes.data, b.size, 0, C.byref(descs[i]))
        assert rc == 0, rc
    pixels = n * W * H
    rgb_bytes = 3 * W * H
    rgb_ptr = lib.b200jpeg_host_alloc(n * rgb_bytes)                                          # pinned host RGB
    rgb_host = np.ctypeslib.as_array((C.c_uint8 * (n * rgb_bytes)).from_address(rgb_ptr))
    rgbs = [rgb_host[i * rgb_bytes:(i + 1) * rgb_bytes] for i in range(n)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        dec.sync()

    def max_over_ranks(x):
        return sharding.max_over_ranks(x, device="cuda")

    # ---- device-resident: value ----
    dec.set_timing(True)
    dec.prepare(descs, files)
    for _ in range(args.warmup):
        dec.run(); st = dec.fetch(None, None)
    assert all(s == 0 for s in st), "decode status not clean: %r" % st[:8]
    stage = {k: 0.0 for k in ("unstuff", "sync", "write", "pixel", "total")}
    barrier()
    with ClockSampler(local_rank) as clocks:
        t0 = time.perf_counter()
        for _ in range(args.steps):
            dec.run()
            dec.fetch(None, None)          # waits on the stream, reads the convergence flag + per-image status
            for k, v in dec.last_timing().items():
                stage[k] += v
        barrier()
        t_res = time.perf_counter() - t0
    t_res = max_over_ranks(t_res)
    stats = dec.stats()
    launches = stats["kernel_launches"] * args.steps
    ev_total_ms = stage["total"] / args.steps

    # ---- end to end through the C ABI with host buffers: e2e ----
    for _ in range(min(2, args.warmup)):
        dec.decode_batch(descs, files, rgbs)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        # the reference-facing call: H2D of every scan + descriptors, all kernels, D2H of every RGB image +
        # status, chunked and double-buffered inside the library
        st = dec.decode_batch(descs, files, rgbs)
    barrier()
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    stats_e = dec.stats()
    assert all(s == 0 for s in st)

    if rank == 0:
        peak, peak_src = measured_peaks()
        blocks = stats["coded_blocks"]
        a_pix = 128 * blocks + 3 * pixels                      # int16 coefficients in + RGB8 out
        a_ent = stats["scan_bytes"] + 128 * blocks              # unstuffed scan in + coefficients out
        a_e2e = stats["scan_bytes"] + 3 * pixels
        ms = {k: v / args.steps for k, v in stage.items()}
        seg = stats["entropy_path"] == 1
        ent_name = "k_entropy_seg" if seg else "k_entropy_write"
        kernels = {
            "pixel": {"name": "k_pixel<2,2,3>", "ms": ms["pixel"], "alg_bytes": a_pix, "gbs": a_pix / ms["pixel"] / 1e6,
                      "frac": a_pix / ms["pixel"] / 1e6 / peak},
            "entropy_decode": {"name": ent_name, "ms": ms["write"], "alg_bytes": a_ent, "gbs": a_ent / ms["write"] / 1e6,
                               "frac": a_ent / ms["write"] / 1e6 / peak},
            "unstuff": {"name": "k_unstuff_count+scan+scatter", "ms": ms["unstuff"], "alg_bytes": 3 * stats["scan_bytes"],
                        "gbs": 3 * stats["scan_bytes"] / ms["unstuff"] / 1e6},
        }
        if not seg:
            kernels["entropy_sync"] = {"name": "k_entropy_sync<1>+<0>x2+scan", "ms": ms["sync"], "alg_bytes": 3 * stats["scan_bytes"],
                                       "gbs": 3 * stats["scan_bytes"] / ms["sync"] / 1e6}
        dom = max(("pixel", "entropy_decode"), key=lambda k: kernels[k]["ms"])
        roofline = {"kernel": kernels[dom]["name"], "bound": "hbm",
                    "achieved": kernels[dom]["gbs"], "peak": peak, "unit": "GB/s", "frac": kernels[dom]["frac"],
                    "traffic": ncu_traffic(kernels[dom]["name"], n), "peak_source": peak_src,
                    "note": "entropy decoding is latency/issue bound (one dependent chain per stream), see DESIGN.md; "
                            "the HBM-bound kernel of the path is the pixel kernel",
                    "pixel_kernel": {"achieved": kernels["pixel"]["gbs"], "frac": kernels["pixel"]["frac"]},
                    "whole_path": {"alg_bytes": a_e2e, "achieved": a_e2e / ev_total_ms / 1e6, "frac": a_e2e / ev_total_ms / 1e6 / peak}}
        ncpu = os.cpu_count() or 1
        cpu = None
        if world == 1:                              # the CPU baseline is a 1-GPU-run item (rank 0 at N=1 only)
            v, cores, kind, sample = cpu_reference_run(files[:min(n, max(8, min(64, ncpu)))], 2, ncpu)
            cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample}
        value = world * pixels * args.steps / t_res / 1e6
        e2e = world * pixels * args.steps / t_e2e / 1e6
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": t_res / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "int32 entropy + f32 IDCT/colour", "data": "synthetic",
                "config": {"workload": workload_name(n), "images_per_gpu": n, "parallelism": "images sharded, no collective",
                           "l2": "working set per step (scan %.0f MB, coefficients %.0f MB, RGB %.0f MB) exceeds the 126 MB L2"
                                 % (stats["scan_bytes"] / 1e6, 128 * blocks / 1e6, 3 * pixels / 1e6),
                           "timed": "K steps bracketed by barrier + stream sync; each step = all kernels + status readback"},
                "event_ms_per_step": ev_total_ms, "kernels": kernels, "roofline": roofline, "cpu_baseline": cpu,
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": stats_e["h2d_bytes"] * world,
                        "d2h_bytes_per_step": stats_e["d2h_bytes"] * world, "ms_per_step": t_e2e / args.steps * 1e3},
                "gpu_launches": launches, "sync_rounds": stats["sync_rounds"],
                "entropy_path": "segment-parallel" if seg else "self-synchronising", "clocks": clocks.summary()}
        emit(line)
    lib.b200jpeg_host_free(rgb_ptr)
    dec.close()
    if world > 1:
        dist.destroy_process_group()


_JSON_FD = None


def emit(line):
    """The ONE JSON line, on the process's real stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    # stdout carries the JSON line and nothing else: whatever libraries print there (NCCL announces its version on
    # stdout under torchrun) goes to stderr instead -- file descriptor 1 is pointed at stderr, the line is written to a
    # duplicate of the original
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--images", type=int, default=256, help="images per GPU per step")
    ap.add_argument("--entropy", default="auto", choices=["auto", "segment", "sync"],
                    help="entropy decoder (auto = the library's own choice; the others are for measurements)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
    else:
        run_gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
