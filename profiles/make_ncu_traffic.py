"""profiles/make_ncu_traffic.py REPORT.ncu-rep WORKLOAD COMMIT [SOURCE-NOTE] -> updates profiles/ncu_traffic.json

The static `traffic` (dram__bytes_read.sum + dram__bytes_write.sum per launch) and `ncu` (issue-slot / pipe
percentages) fields bench.py prints next to its live per-kernel rows come from here; the file records the
commit and the capture they were taken at (`_commit`, `_source`), and bench.py repeats both in `ncu_static`.
"""
import csv
import json
import os
import subprocess
import sys

rep, wl, commit = sys.argv[1], sys.argv[2], sys.argv[3]
note = sys.argv[4] if len(sys.argv) > 4 else ""
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
NAMES = {"preprocess_fwd": "preprocess_fwd", "tile_scan": "tile_scan", "sort_pack_small": "sort_pack",
         "blend_fwd": "blend_fwd", "blend_bwd": "blend_bwd", "preprocess_bwd": "preprocess_bwd"}
PCT = {"issue_active_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
       "pipe_fma_pct": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
       "pipe_alu_pct": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
       "pipe_xu_pct": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
       "pipe_lsu_pct": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
       "dram_throughput_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
       "warps_active_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
       "registers": "launch__registers_per_thread", "warp_instructions": "smsp__inst_executed.sum",
       "duration_us_under_ncu": "gpu__time_duration.sum"}
units = rows[1]


def num(r, name):
    i = hdr.index(name)
    v = float(r[i].replace(",", ""))
    u = units[i]
    if name.startswith("dram__bytes"):
        v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}.get(u, 1.0)
    if name == "gpu__time_duration.sum":
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u, 1.0)
    return v


traffic, pct = {}, {}
for r in rows[2:]:
    kn = r[hdr.index("Kernel Name")]
    key = next((v for k, v in NAMES.items() if k in kn), None)
    if key is None or key in traffic:
        continue
    traffic[key] = num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum")
    pct[key] = {k: num(r, m) for k, m in PCT.items() if m in hdr}
# the record gather of sort_pack is its own kernel (pack_flat_kernel): its DRAM bytes belong to the sort_pack row
for r in rows[2:]:
    if "pack_flat" in r[hdr.index("Kernel Name")] and "sort_pack" in traffic:
        traffic["sort_pack"] += num(r, "dram__bytes_read.sum") + num(r, "dram__bytes_write.sum")
        pct["sort_pack"]["pack_flat_duration_us_under_ncu"] = num(r, "gpu__time_duration.sum")
        break
path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
d = json.load(open(path)) if os.path.exists(path) else {}
d[wl] = traffic
d[wl + "_ncu"] = pct
d["_commit"] = commit
d["_source"] = (f"{os.path.basename(rep)} (ncu --set full --clock-control none --import-source on, profiles/run_step.py, "
                f"{wl}, bucketed forward) captured at commit {commit}; dram__bytes_read.sum + dram__bytes_write.sum per launch. {note}")
json.dump(d, open(path, "w"), indent=1)
print("updated", path, "kernels:", sorted(traffic))
