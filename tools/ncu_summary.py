#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total time and share."""
import collections
import csv
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    h = rows[hdr]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[hdr + 1:]:
        if len(r) < len(h):
            continue
        d = dict(zip(h, r))
        if d["Metric Name"] != "gpu__time_duration.sum":
            continue
        v = float(d["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(d["Metric Unit"], 1.0)
        k = d["Kernel Name"].split("(")[0][-60:]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    print("%-62s %6s %12s %7s %10s" % ("kernel", "n", "total_us", "share", "avg_us"))
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-62s %6d %12.1f %7.3f %10.1f" % (k, v[0], v[1], v[1] / tot, v[1] / v[0]))


if __name__ == "__main__":
    main(sys.argv[1])
