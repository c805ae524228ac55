#!/usr/bin/env python3
"""Summarise an `ncu --csv` launch list (gpu__time_duration.sum, optionally dram__bytes_{read,write}.sum).

    ncu_summary.py LAUNCHES.csv [--skip N] [--take M] [--traffic-key medium_b1 --traffic-json profiles/traffic.json]

Prints, per kernel family: launches, total / mean duration, share of the summed GPU time and (when present) mean
DRAM bytes per launch.  With --traffic-key it also records the Q4 GEMM kernel's mean DRAM bytes per launch in
profiles/traffic.json, which bench.py reports as roofline.traffic.
"""
import argparse
import csv
import json
import re
import sys
from collections import OrderedDict, defaultdict
from pathlib import Path


def family(name: str) -> str:
    m = re.search(r"(?:<unnamed>::)?(\w+)(<[^>]*>)?\(", name)
    if not m:
        return name[:60]
    return m.group(1) + (m.group(2) or "")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--skip", type=int, default=0, help="ignore the first N launches (warm-up)")
    ap.add_argument("--take", type=int, default=0, help="use only M launches after the skip (0 = all)")
    ap.add_argument("--last", type=int, default=0, help="use only the last M launches (one steady-state step)")
    ap.add_argument("--traffic-key", default="")
    ap.add_argument("--traffic-json", default="profiles/traffic.json")
    ap.add_argument("--gemm-regex", default=r"q4_(gemm2|chain)_kernel")
    a = ap.parse_args()

    rows = defaultdict(dict)
    names = {}
    with open(a.csv, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        try:
            i = int(r["ID"])
            rows[i][r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
            names[i] = r["Kernel Name"]
        except (KeyError, ValueError):
            continue
    ids = sorted(rows)[a.skip:]
    if a.take:
        ids = ids[: a.take]
    if a.last:
        ids = ids[-a.last:]
    fam = OrderedDict()
    for i in ids:
        d = rows[i]
        e = fam.setdefault(family(names[i]), {"n": 0, "ns": 0.0, "rd": 0.0, "wr": 0.0, "has_dram": False})
        e["n"] += 1
        e["ns"] += d.get("gpu__time_duration.sum", 0.0)
        if "dram__bytes_read.sum" in d:
            e["has_dram"] = True
            e["rd"] += d["dram__bytes_read.sum"]
            e["wr"] += d.get("dram__bytes_write.sum", 0.0)
    total = sum(e["ns"] for e in fam.values()) or 1.0
    print(f"# {a.csv}: launches {len(ids)} (skip {a.skip}), summed GPU time {total / 1e3:.1f} us")
    print(f"{'kernel':58s} {'n':>5s} {'total us':>10s} {'mean us':>9s} {'share':>7s} {'dram rd MB/launch':>18s} {'dram wr MB/launch':>18s}")
    for k, e in sorted(fam.items(), key=lambda kv: -kv[1]["ns"]):
        rd = f"{e['rd'] / e['n'] / 1e6:18.3f}" if e["has_dram"] else f"{'-':>18s}"
        wr = f"{e['wr'] / e['n'] / 1e6:18.3f}" if e["has_dram"] else f"{'-':>18s}"
        print(f"{k[:58]:58s} {e['n']:5d} {e['ns'] / 1e3:10.1f} {e['ns'] / e['n'] / 1e3:9.2f} {e['ns'] / total * 100:6.1f}% {rd} {wr}")
    if a.traffic_key:
        g = [e for k, e in fam.items() if re.search(a.gemm_regex, k) and e["has_dram"]]
        if not g:
            print("no DRAM metrics for the GEMM kernel in this CSV", file=sys.stderr)
            return 1
        n = sum(e["n"] for e in g)
        by = sum(e["rd"] + e["wr"] for e in g) / n
        p = Path(a.traffic_json)
        d = json.loads(p.read_text()) if p.exists() else {}
        d[a.traffic_key] = {"gemm_bytes_per_launch": by, "gemm_launches": n,
                            "gemm_share_of_gpu_time": sum(e["ns"] for e in g) / total,
                            "source": f"ncu dram__bytes_read.sum+dram__bytes_write.sum, mean over {n} launches ({Path(a.csv).name})"}
        p.write_text(json.dumps(d, indent=1) + "\n")
        print(f"wrote {p}: {a.traffic_key} -> {by / 1e6:.3f} MB per GEMM launch")
    return 0


if __name__ == "__main__":
    sys.exit(main())
