"""profiles/traffic.json from `ncu --set full` raw exports: per workload and kernel the DRAM bytes per launch and the issue-side
counters bench.py copies into `roofline.traffic` / `roofline.issue`.
usage: python profiles/make_traffic.py C2:256:raw_C2.csv C2-sync:256:raw_C2sync.csv C3:1:raw_C3.csv > profiles/traffic.json"""
import json
import sys

from ncu_extract import rows_of

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}
out = {"source": "ncu --set full --clock-control none, first launch of each kernel; see profiles/README.md", "workloads": {}}
for arg in sys.argv[1:]:
    key, images, path = arg.split(":")
    ks = {}
    for d in rows_of(path):
        name = d["kernel"].replace(" ", "")
        if name in ks:
            continue
        ks[name] = {"read": int(d["rd"] * UNIT[d["rd_unit"]]), "write": int(d["wr"] * UNIT[d["wr_unit"]]),
                    "ncu_ms": round(d["time"] * UNIT[d["time_unit"]], 4), "ipc": round(d["ipc"], 3),
                    "issue_active_pct": round(d.get("issue%", 0.0), 1), "alu_pipe_pct": round(d["alu%"], 1),
                    "fma_pipe_pct": round(d["fma%"], 1), "xu_pipe_pct": round(d["xu%"], 1), "lsu_pipe_pct": round(d["lsu%"], 1),
                    "lanes_per_inst": round(d["lanes"], 2), "warp_inst": int(d["inst"]), "dram_pct": round(d["dram%"], 1),
                    "smem_wavefronts": int(d["wavef"]), "smem_conflict_wavefronts": int(d["conf"]), "registers": int(d["regs"])}
    out["workloads"][key] = {"images": int(images), "kernels": ks}
json.dump(out, sys.stdout, indent=1, sort_keys=True)
