#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (share of step)."""
import collections
import csv
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    tot = collections.defaultdict(float)
    cnt = collections.Counter()
    for row in csv.DictReader(lines):
        name = row["Kernel Name"].split("(")[0]
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v *= {"ns": 1.0, "nsecond": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6}.get(u, 1.0)
        tot[name] += v
        cnt[name] += 1
    T = sum(tot.values())
    print(f"| kernel | launches | total ms | share | avg us |\n|---|---|---|---|---|")
    for k, v in sorted(tot.items(), key=lambda x: -x[1]):
        print(f"| `{k}` | {cnt[k]} | {v / 1e6:.3f} | {v / T:.3f} | {v / cnt[k] / 1e3:.1f} |")


if __name__ == "__main__":
    main(sys.argv[1])
