#!/usr/bin/env python3
"""Sums DRAM traffic and duration of the k_trace_queue launches in an `ncu --set full` report of `bench.py --profile`, per step,
and records it in profiles/traffic.json (read by bench.py for roofline.traffic).
Usage: python tools/ncu_traffic.py report.ncu-rep <workload>_<W>x<H> <launches captured = whole steps> <steps captured>"""
import csv
import io
import json
import os
import subprocess
import sys

rep, key, steps = sys.argv[1], sys.argv[2], int(sys.argv[3])
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
units = rows[1]


def scale(name):
    u = units[col[name]].lower()
    return {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0, "s": 1.0}.get(u, 1.0)


# time-weighted means over the kernel's launches: what bounds it (L2 / L1 hit rates, issue rate, lanes per instruction, memory stalls)
WEIGHTED = {"lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2_throughput_pct", "lts__t_sector_hit_rate.pct": "l2_hit_pct", "l1tex__t_sector_hit_rate.pct": "l1_hit_pct",
            "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct", "smsp__thread_inst_executed_per_inst_executed.ratio": "threads_per_inst",
            "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio": "long_scoreboard_per_issue",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct", "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct"}
weighted = {v: 0.0 for v in WEIGHTED.values()}
tot = {"read": 0.0, "write": 0.0, "seconds": 0.0, "launches": 0}
for r in rows[2:]:
    if "k_trace_queue" not in r[col["Kernel Name"]]:
        continue
    seconds = float(r[col["gpu__time_duration.sum"]].replace(",", "")) * scale("gpu__time_duration.sum")
    for name, short in WEIGHTED.items():
        if name in col and r[col[name]] not in ("", "n/a"):
            weighted[short] += float(r[col[name]].replace(",", "")) * seconds
    tot["read"] += float(r[col["dram__bytes_read.sum"]].replace(",", "")) * scale("dram__bytes_read.sum")
    tot["write"] += float(r[col["dram__bytes_write.sum"]].replace(",", "")) * scale("dram__bytes_write.sum")
    tot["seconds"] += float(r[col["gpu__time_duration.sum"]].replace(",", "")) * scale("gpu__time_duration.sum")
    tot["launches"] += 1
entry = {"trace_queue_dram_bytes_per_step": (tot["read"] + tot["write"]) / steps, "dram_read_bytes_per_step": tot["read"] / steps, "dram_write_bytes_per_step": tot["write"] / steps,
         "trace_queue_launches_per_step": tot["launches"] / steps, "ncu_serialised_ms_per_step": tot["seconds"] / steps * 1e3, "source": os.path.basename(rep),
         "ncu_time_weighted": {k: round(v / max(tot["seconds"], 1e-12), 3) for k, v in weighted.items()}}
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "traffic.json")
data = json.load(open(path)) if os.path.exists(path) else {}
data[key] = entry
json.dump(data, open(path, "w"), indent=1, sort_keys=True)
print(key, json.dumps(entry))
