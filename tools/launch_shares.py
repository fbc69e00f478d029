#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: launch_shares.py ncu_launches.csv list_out.csv shares_out.csv
  list_out:   one line per launch (id, kernel, block, grid, time_ns) -- the compact form committed under profiles/
  shares_out: launches / total / share / mean per kernel (PyTorch's own fill / copy kernels are left out of the shares)"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 8]
hdr = rows[0]
col = {k: hdr.index(k) for k in ("ID", "Kernel Name", "Block Size", "Grid Size", "Metric Name", "Metric Unit", "Metric Value")}
scale = {"ns": 1.0, "us": 1e3, "ms": 1e6, "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6}
launches = []
for r in rows[1:]:
    if r[col["Metric Name"]] != "gpu__time_duration.sum":
        continue
    ns = float(r[col["Metric Value"]].replace(",", "")) * scale[r[col["Metric Unit"]]]
    launches.append((r[col["ID"]], r[col["Kernel Name"]][:110], r[col["Block Size"]], r[col["Grid Size"]], int(round(ns))))
with open(sys.argv[2], "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["id", "kernel", "block", "grid", "time_ns"])
    w.writerows(launches)
tot, cnt = collections.Counter(), collections.Counter()
for _, name, _, _, ns in launches:
    if name.startswith("void at::"):
        continue
    name = re.sub(r"\(.*$", "", name).replace("void ", "").replace("<unnamed>::", "").strip()
    tot[name] += ns * 1e-3
    cnt[name] += 1
total = sum(tot.values())
with open(sys.argv[3], "w", newline="") as f:
    w = csv.writer(f)
    w.writerow(["kernel", "launches", "total_us", "share", "mean_us"])
    for name, us in tot.most_common():
        w.writerow([name, cnt[name], f"{us:.1f}", f"{us / total:.4f}", f"{us / cnt[name]:.2f}"])
