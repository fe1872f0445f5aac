#!/usr/bin/env python
"""Per-layer table from bench.py --launch-csv (CUDA-event time of every launch of one step).

    python profiles/layer_table.py gpurun_out/launch_events.csv [top_n]
"""
import collections
import csv
import sys


def main(path, top=20):
    agg = collections.OrderedDict()
    for r in csv.DictReader(open(path)):
        k = r["label"] or ("kind%s" % r["kind"])
        a = agg.setdefault(k, [0, 0.0, 0.0])
        a[0] += 1
        a[1] += float(r["ms"])
        a[2] += float(r["flops"])
    tot = sum(a[1] for a in agg.values())
    print("total %.3f ms" % tot)
    print("| launch | n | ms | share | TFLOP/s |")
    print("|---|---:|---:|---:|---:|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print("| %s | %d | %.3f | %.1f%% | %.0f |" % (k, a[0], a[1], 100 * a[1] / tot, a[2] / a[1] / 1e9 if a[1] else 0))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 20)
