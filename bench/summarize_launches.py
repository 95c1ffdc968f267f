#!/usr/bin/env python
"""Turn an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel table (markdown)."""
import collections
import csv
import sys


def main(path, steps=1):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0, 0.0, 1e30, 0.0])
    total = 0.0
    for r in csv.DictReader(lines):
        v = float(r["Metric Value"].replace(",", ""))
        u = r["Metric Unit"]
        v = v / 1000 if u == "ns" else v * 1000 if u == "ms" else v
        name = r["Kernel Name"].split("(")[0].replace("void ", "")
        a = agg[name]
        a[0] += 1; a[1] += v; a[2] = min(a[2], v); a[3] = max(a[3], v)
        total += v
    print(f"| kernel | launches/step | total us/step | avg us | min | max | share |")
    print("|---|---:|---:|---:|---:|---:|---:|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {a[0] / steps:.0f} | {a[1] / steps:.1f} | {a[1] / a[0]:.1f} | {a[2]:.1f} | {a[3]:.1f} | {a[1] / total:.1%} |")
    print(f"| **total** | {sum(a[0] for a in agg.values()) / steps:.0f} | {total / steps:.1f} | | | | |")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 1)
