#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch log of one training step:
per-kernel totals, or (--list) every launch in order.  Usage: timeline_summary.py file.csv [--list] [--grep PAT]"""
import collections
import csv
import re
import sys


def load(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    out = []
    for row in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void kdl::", "").replace("kdl::", "")
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        out.append((name, row["Grid Size"], v))
    return out


def main():
    rows = load(sys.argv[1])
    pat = sys.argv[sys.argv.index("--grep") + 1] if "--grep" in sys.argv else None
    if "--list" in sys.argv:
        for i, (n, g, v) in enumerate(rows):
            if pat is None or re.search(pat, n):
                print(f"{i:4d} {v:8.1f}us {g:14s} {n[:70]}")
        return
    agg = collections.defaultdict(lambda: [0, 0.0])
    for n, g, v in rows:
        agg[n][0] += 1
        agg[n][1] += v
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{t / 1e3:8.3f} ms {n:4d}  {k[:100]}")
    print(f"{sum(v for _, _, v in rows) / 1e3:8.3f} ms total, {len(rows)} launches")


if __name__ == "__main__":
    main()
