"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel totals and shares.
Usage: python scripts/summarize_launches.py gpurun_out/launches.csv > profiles/rNN_launches.txt"""
import collections
import csv
import re
import sys

lines = open(sys.argv[1]).readlines()
start = [i for i, l in enumerate(lines) if l.startswith('"ID"')][0]
agg = collections.defaultdict(lambda: [0, 0.0])
by_grid = collections.defaultdict(lambda: [0, 0.0])
tot = 0.0
for x in csv.DictReader(lines[start:]):
    if x["Metric Name"] != "gpu__time_duration.sum":
        continue
    t = float(x["Metric Value"].replace(",", ""))
    t = t / 1e3 if x["Metric Unit"] == "ns" else (t * 1e3 if x["Metric Unit"] == "ms" else t)
    name = re.sub(r"\(.*", "", x["Kernel Name"]).replace("void ", "")[:100]
    agg[name][0] += 1
    agg[name][1] += t
    if "fod::" in name:
        k = (name, x["Grid Size"], x["Block Size"])
        by_grid[k][0] += 1
        by_grid[k][1] += t
    tot += t
ours = sum(v[1] for k, v in agg.items() if "fod::" in k)
print("# ncu launch list summary (cold-cache, serialised launches; shares, not absolutes, are comparable with bench.py)")
print("total %.1f us over %d launches; fod:: kernels %.1f us (%.1f%%)" % (tot, sum(v[0] for v in agg.values()), ours, 100 * ours / tot))
print("%-100s %6s %10s %6s %8s" % ("kernel", "n", "us", "share", "avg us"))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:60]:
    print("%-100s %6d %10.1f %5.1f%% %8.1f" % (k, v[0], v[1], 100 * v[1] / tot, v[1] / v[0]))
print("\n# fod:: kernels by grid")
for k, v in sorted(by_grid.items(), key=lambda kv: -kv[1][1])[:60]:
    print("%-60s %-14s %-12s %5d %9.1f us avg %6.1f" % (k[0][:60], k[1], k[2], v[0], v[1], v[1] / v[0]))
