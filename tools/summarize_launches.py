#!/usr/bin/env python
"""Summarise an ncu launch list (`--metrics gpu__time_duration.sum --csv --log-file launches.csv`) into a
per-kernel table: launches, average duration, share of the captured region.

    python tools/summarize_launches.py gpurun_out/launches.csv > profiles/r01_launches_cfg2a.md
"""
import collections
import csv
import re
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.OrderedDict()
    n = 0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        n += 1
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(row["Metric Unit"], 1.0)
        k = (name, row["Grid Size"], row["Block Size"])
        c, t = agg.get(k, (0, 0.0))
        agg[k] = (c + 1, t + v)
    tot = sum(t for _, t in agg.values())
    print("| kernel | grid | block | launches | avg us | share |")
    print("|---|---|---|---|---|---|")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| `%s` | %s | %s | %d | %.1f | %.3f |" % (k[0][:70], k[1], k[2], c, t / c, t / tot))
    print()
    print("%d launches, %.1f us in total (ncu: cold-cache, serialised -- compare shares, not absolutes)" % (n, tot))


if __name__ == "__main__":
    main(sys.argv[1])
