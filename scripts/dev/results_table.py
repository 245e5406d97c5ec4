#!/usr/bin/env python
"""Dev tool: DESIGN.md's results table from the committed bench lines (profiles/r02_bench_*.json, r02_latency.jsonl).
    python scripts/dev/results_table.py  -> markdown rows on stdout"""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ROWS = [("C2", "C2 256 × 1080p 4:2:0 q90, DRI = row"), ("C2_sync", "C2 forced `--entropy sync` (r01: 6.45 ms)"),
        ("C3", "C3 1 × 8192² 4:4:4 q95, no RST"), ("C4-gray", "C4 8 × 4096² grey q90"), ("C4-gray_q50", "… q50"),
        ("C4-422", "C4 8 × 4096² 4:2:2 q90"), ("C4-422_q50", "… q50"), ("C4-444", "C4 8 × 4096² 4:4:4 q90"),
        ("C4-444_q50", "… q50"), ("C5", "C5 16,384 × 512² 4:2:0 q85, no RST")]


def line(path):
    return json.loads(open(path).read().strip().split("\n")[-1])


print("| Workload | entropy path | U / sync / write / P (ms) | step | device-resident | host → host (floor) | reference, all cores / 1 core |")
print("|---|---|---|---|---|---|---|")
for key, name in ROWS:
    d = line(os.path.join(ROOT, "profiles", "r02_bench_%s.json" % key))
    k = d["kernels"]
    seg = d["entropy_path"] in (1, "segment", "segment-parallel")
    st = "%.3f / %s / %.3f / %.3f" % (k["unstuff"]["ms"], "–" if seg else "%.3f" % k["entropy_sync"]["ms"],
                                      k["entropy_decode"]["ms"], k["pixel"]["ms"])
    e, c = d["e2e"], d["cpu_baseline"]
    if c.get("kind") == "reference":
        cpu = "%.0f MP/s on %d cores / %.1f on one" % (c["value"], c["cores"], c["one_core"]["value"])
    else:
        cpu = "port, 1 core: %.0f MP/s (the reference rejects grey)" % c["value"]
    print("| %s | %s | %s | **%.2f ms** | **%.1f GP/s** | %.1f ms (%.1f) = %.1f GP/s | %s |" % (
        name, "segments" if seg else "self-sync", st, d["event_ms_per_step"], d["value"] / 1e3, e["ms_per_step"], e["copy_floor_ms"],
        e["value"] / 1e3, cpu))
for ln in open(os.path.join(ROOT, "profiles", "r02_latency.jsonl")):
    d = json.loads(ln)
    if d["file"] == "img.jpg":
        g = d["gpu_ms"]
        ref = d.get("reference_1core_ms") or d.get("ref_ms")
        print("| C1 `img.jpg` %d×%d, one image | self-sync | %.3f / %.3f / %.3f / %.3f | **%.2f ms** | | host → host %.2f ms | %s |" % (
            d["w"], d["h"], g["unstuff"], g["sync"], g["write"], g["pixel"], g["total"], d["host_to_host_ms"],
            ("%.1f ms on one core" % ref) if ref else ""))
