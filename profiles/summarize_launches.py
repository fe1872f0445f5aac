#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel totals.

    python profiles/summarize_launches.py gpurun_out/launches.csv > profiles/launches_rNN.md

ncu serialises launches and runs them cold-cache, so only the SHARES are comparable with the
CUDA-event numbers bench.py prints (B200_PROFILING.md).
"""
import csv
import re
import sys
from collections import defaultdict


def main(path):
    rows = []
    with open(path, newline="") as f:
        lines = [l for l in f if l.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            rows.append((r["Kernel Name"], r["Grid Size"], r["Block Size"], float(r["Metric Value"].replace(",", ""))))
    tot = sum(r[3] for r in rows)
    agg = defaultdict(lambda: [0, 0.0])
    for name, grid, block, ns in rows:
        short = re.sub(r"\(.*", "", name)
        agg[short][0] += 1
        agg[short][1] += ns
    print("| kernel | launches | total ms | share | mean us |")
    print("|---|---:|---:|---:|---:|")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| `%s` | %d | %.3f | %.1f%% | %.1f |" % (k, n, ns / 1e6, 100 * ns / tot, ns / n / 1e3))
    print("\n%d launches, %.3f ms of device time in total (serialised, cold-cache: compare shares, not absolutes)" % (len(rows), tot / 1e6))


if __name__ == "__main__":
    main(sys.argv[1])
