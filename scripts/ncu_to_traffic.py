#!/usr/bin/env python3
"""profiles/traffic.json <- one `ncu --set full` capture: per workload and kernel the figures bench.py quotes beside its roofline
(per-launch DRAM bytes, DRAM GB/s, issue-slot utilisation, lanes per instruction, spill instructions, L2 hit rate).
usage: ncu_to_traffic.py <workload> <file.ncu-rep> <source label> [launch index per kernel, default 0 = first captured]"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
workload, rep, label = sys.argv[1], sys.argv[2], sys.argv[3]
pick = int(sys.argv[4]) if len(sys.argv) > 4 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}


def num(r, name):
    v = r[col[name]].replace(",", "")
    return float(v) if v not in ("", "n/a") else 0.0


def unit_scale(name):
    u = rows[1][col[name]]
    return {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3,
            "second": 1}.get(u, 1)


seen = {}
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    key = "k_shade" if "k_shade" in name else ("k_connect" if "k_trace<1" in name or "k_trace<true" in name else ("k_extend" if "k_trace" in name else None))
    if not key:
        continue
    seen.setdefault(key, []).append(r)
path = os.path.join(ROOT, "profiles", "traffic.json")
data = json.load(open(path)) if os.path.exists(path) else {}
entry = data.setdefault(workload, {})
if not isinstance(entry, dict):
    entry = data[workload] = {}
for key, rs in seen.items():
    r = rs[min(pick, len(rs) - 1)]
    t = num(r, "gpu__time_duration.sum") * unit_scale("gpu__time_duration.sum")
    rd = num(r, "dram__bytes_read.sum") * unit_scale("dram__bytes_read.sum")
    wr = num(r, "dram__bytes_write.sum") * unit_scale("dram__bytes_write.sum")
    entry[key] = {"dram_bytes_per_launch": rd + wr, "launch_ms_under_ncu": t * 1e3, "dram_gbs": (rd + wr) / t / 1e9,
                  "issue_active_pct": num(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                  "lanes_per_inst": num(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
                  "spill_insts": int(num(r, "sass__inst_executed_register_spilling")),
                  "l2_hit_pct": num(r, "lts__t_sector_hit_rate.pct"), "l1_hit_pct": num(r, "l1tex__t_sector_hit_rate.pct"),
                  "registers": int(num(r, "launch__registers_per_thread")), "launch_of_capture": min(pick, len(rs) - 1), "source": label}
data["_comment"] = ("per workload and kernel: dram__bytes_read.sum + dram__bytes_write.sum per launch and the issue / SIMT figures of one `ncu --set full "
                    "--clock-control none` capture (scripts/ncu_to_traffic.py; summaries of the same captures: profiles/r02_ncu_*_metrics.txt). Times under ncu are "
                    "cold-cache and serialised: bench.py uses the bytes and ratios, not the times.")
json.dump(data, open(path, "w"), indent=1)
print(json.dumps(entry, indent=1))
