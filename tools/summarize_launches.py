"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel table.
usage: python tools/summarize_launches.py gpurun_out/launches.csv > profiles/<name>.md"""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
agg, tot = collections.OrderedDict(), 0.0
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(row["Metric Unit"], 1.0)
    n, t = agg.get(row["Kernel Name"], (0, 0.0))
    agg[row["Kernel Name"]] = (n + 1, t + v)
    tot += v
print("| kernel | launches | total us | share | avg us |\n|---|---|---|---|---|")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("| `%s` | %d | %.1f | %.3f | %.1f |" % (k.split("(")[0][-70:], n, t, t / tot, t / n))
print("\ntotal %.1f us over %d launches (cold-cache, serialised under ncu: compare SHARES, not absolutes)" % (tot, sum(n for n, _ in agg.values())))
