#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
Usage: python scripts/summarize_launches.py gpurun_out/<tag>_launches.csv [steps] > profiles/<tag>_launches.md"""
import collections
import csv
import sys


def main():
    path = sys.argv[1]
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    with open(path) as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v * 1e6 if u == "s" else v
        name = row["Kernel Name"].split("(")[0][-90:]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"# ncu launch list: {path}\n")
    print(f"total {tot/1e3:.2f} ms over {sum(a[0] for a in agg.values())} launches"
          + (f" ({steps} pipeline steps -> {tot/1e3/steps:.2f} ms/step serialised, cold-cache)" if steps else "") + "\n")
    print("| kernel | launches | total us | us/launch | share |\n|---|---:|---:|---:|---:|")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"| `{k}` | {a[0]} | {a[1]:.1f} | {a[1]/a[0]:.1f} | {100*a[1]/tot:.1f}% |")


if __name__ == "__main__":
    main()
