#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: share of time per kernel."""
import collections
import csv
import io
import sys


def main(path):
    lines = open(path).read().splitlines()
    start = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
    rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
    agg = collections.OrderedDict()
    for r in rows:
        v = agg.setdefault(r["Kernel Name"], [0, 0.0, r["Grid Size"], r["Block Size"]])
        v[0] += 1
        v[1] += float(r["Metric Value"].replace(",", ""))
    tot = sum(v[1] for v in agg.values())
    print(f"{len(rows)} launches, {tot / 1e6:.3f} ms total (cold-cache, serialised: compare shares, not absolutes)")
    print(f"{'share':>7} {'n':>5} {'avg_us':>11}  grid / block  kernel")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1] / tot * 100:6.2f}% {v[0]:5d} {v[1] / v[0] / 1e3:11.1f}  {v[2]} {v[3]}  {k[:100]}")


if __name__ == "__main__":
    main(sys.argv[1])
