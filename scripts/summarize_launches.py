#!/usr/bin/env python
"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections
import csv
import re
import sys


def main(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.defaultdict(list)
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        v = {"ns": v / 1e3, "us": v, "ms": v * 1e3, "s": v * 1e6}.get(r[ui], v)
        name = re.sub(r"\(.*", "", r[ki]).replace("<unnamed>::", "")
        agg[name].append(v)
    tot = sum(sum(v) for v in agg.values())
    print(f"{'kernel':42s} {'launches':>8s} {'avg us':>10s} {'total ms':>10s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k:42s} {len(v):8d} {sum(v) / len(v):10.1f} {sum(v) / 1e3:10.2f} {100 * sum(v) / tot:6.1f}%")
    print(f"{'total':42s} {sum(len(v) for v in agg.values()):8d} {'':10s} {tot / 1e3:10.2f}")


if __name__ == "__main__":
    main(sys.argv[1])
