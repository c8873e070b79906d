#!/usr/bin/env python
"""summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel count / mean / total."""
import collections
import csv
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        k = row["Kernel Name"][:72]
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        agg.setdefault(k, []).append(v)
    tot = sum(sum(v) for v in agg.values())
    for k, v in agg.items():
        print(f"{k:72s} n={len(v):4d} mean={sum(v) / len(v):10.1f} us  total={sum(v) / 1e3:9.3f} ms  {100 * sum(v) / tot:5.1f}%")
    print(f"total {tot / 1e3:.3f} ms")


if __name__ == "__main__":
    main(sys.argv[1])
