#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: launches, total time and share per kernel.
usage: python scripts/summarise_launches.py launches.csv "<command line that was profiled>" > summary.txt"""
import csv
import re
import sys
from collections import defaultdict


def main():
    path, cmd = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    rows = [l for l in open(path) if l.startswith('"')]
    agg = defaultdict(lambda: [0, 0.0])
    for r in csv.DictReader(rows):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*$", "", r["Kernel Name"])[:110]
        a = agg[name]
        a[0] += 1
        a[1] += float(r["Metric Value"]) / 1e6
    tot = sum(a[1] for a in agg.values())
    print(cmd)
    print("(cold-cache, serialised: compare shares)  total %.3f ms over %d launches" % (tot, sum(a[0] for a in agg.values())))
    for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name:<112}{n:>6}{ms:>10.3f} ms{100 * ms / tot:>6.1f}%")


if __name__ == "__main__":
    main()
