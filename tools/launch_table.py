#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list as a markdown table (per kernel: launches,
total, mean, max, share).  Times under ncu are per-launch, cold-cache and serialised: shares are meaningful, absolute
values are not bench numbers.
    python tools/launch_table.py gpurun_out/launches.csv "title" > profiles/x.md"""
import collections
import csv
import sys


def main(path, title):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        v = v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        a = agg.setdefault(row["Kernel Name"], [0, 0.0, 0.0])
        a[0] += 1; a[1] += v; a[2] = max(a[2], v)
    tot = sum(a[1] for a in agg.values())
    print("# %s\n" % title)
    print("Source: `%s` (`ncu --metrics gpu__time_duration.sum --clock-control none`; per-launch, cold-cache, serialised).\n" % path)
    print("| kernel | launches | total µs | mean µs | max µs | share |")
    print("|---|---|---|---|---|---|")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        if a[1] / tot < 0.001:
            continue
        print("| `%s` | %d | %.1f | %.1f | %.1f | %.1f %% |" % (k[:90], a[0], a[1], a[1] / a[0], a[2], 100 * a[1] / tot))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "kernel launch list")
