#!/usr/bin/env python3
"""Tabulates an `ncu --csv --metrics ...` launch list (one row per launch). Usage: python tools/ncu_launches.py file.csv [--sum]"""
import csv
import sys
from collections import OrderedDict

M = [("gpu__time_duration.sum", "us", 1e-3), ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst", 1), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%", 1),
     ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 1), ("launch__registers_per_thread", "regs", 1),
     ("smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "no_inst", 1), ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "long_sb", 1),
     ("smsp__inst_executed.sum", "Minst", 1e-6)]
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ik, im, iv, iid = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
d = OrderedDict()
for r in rows[1:]:
    d.setdefault((int(r[iid]), r[ik].split("(")[0][:28]), {})[r[im]] = float(r[iv].replace(",", "")) if r[iv] not in ("", "n/a") else float("nan")
print("id kernel " + " ".join(n for _, n, _ in M))
tot = {}
for (i, k), m in d.items():
    vals = [m.get(name, float("nan")) * scale for name, _, scale in M]
    if "--sum" not in sys.argv:
        print(i, k, " ".join(f"{v:.2f}" for v in vals))
    t = tot.setdefault(k, [0, 0.0, 0.0])
    t[0] += 1; t[1] += vals[0]; t[2] += vals[-1]
print("-- per kernel: launches, total us, total Minst")
for k, t in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:30s} {t[0]:4d} {t[1]:10.1f} {t[2]:10.1f}")
print("total us", sum(t[1] for t in tot.values()))
