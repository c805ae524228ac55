#!/usr/bin/env python3
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum --csv) of scripts/bench_decode.py:
the LAST decode step in the capture (its launches come last), grouped by kernel and grid.
Usage: ncu_decode_summary.py LAUNCHES.csv LAUNCHES_PER_STEP"""
import collections
import csv
import re
import sys

path, per_step = sys.argv[1], int(sys.argv[2])
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
byid = {}
for x in csv.DictReader(lines):
    d = byid.setdefault(int(x["ID"]), {"k": x["Kernel Name"], "g": x["Grid Size"], "b": x["Block Size"], "t": 0.0, "r": 0.0})
    v = float(x["Metric Value"].replace(",", ""))
    if x["Metric Name"] == "gpu__time_duration.sum":
        d["t"] = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(x["Metric Unit"], 1e-3)
    else:
        d["r"] = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(x["Metric Unit"], 1)
rows = [byid[i] for i in sorted(byid)][-per_step:]
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for d in rows:
    k = (re.sub(r"\(.*", "", d["k"]).replace("void ", "").replace("<unnamed>::", "")[:44], d["g"], d["b"])
    agg[k][0] += 1; agg[k][1] += d["t"]; agg[k][2] += d["r"]
tot = sum(v[1] for v in agg.values())
print(f"# {path}: last {per_step} launches = one decode step; ncu times are cold-cache and serialised (shares, not absolutes)")
print(f"{'kernel':44s} {'grid':14s} {'block':12s} {'n':>4s} {'total us':>9s} {'avg us':>8s} {'share':>6s} {'DRAM read/launch MB':>20s}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[0]:44s} {k[1]:14s} {k[2]:12s} {v[0]:4d} {v[1]:9.1f} {v[1] / v[0]:8.2f} {100 * v[1] / tot:5.1f}% {v[2] / v[0] / 1e6:20.1f}")
print(f"{'sum':44s} {'':14s} {'':12s} {len(rows):4d} {tot:9.1f}")
