#!/usr/bin/env python
"""ncu CSV (dram__bytes_read.sum, dram__bytes_write.sum, gpu__time_duration.sum per conv launch of one warm denoiser
evaluation) -> profiles/conv_traffic.json, which bench.py reports as roofline.traffic (bytes per launch, averaged like
roofline.achieved).      python profiles/traffic_from_ncu.py gpurun_out/conv_traffic_TAG.csv TAG"""
import csv
import json
import os
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "ms": 1e-3, "nsecond": 1e-9, "usecond": 1e-6, "msecond": 1e-3}


def main(path, tag):
    lines = [l for l in open(path, newline="") if l.startswith('"')]
    per = {}
    for r in csv.DictReader(lines):
        v = float(r["Metric Value"].replace(",", "")) * UNIT.get(r["Metric Unit"], 1.0)
        per.setdefault(r["ID"], {})[r["Metric Name"]] = v
    n = len(per)
    rd = sum(p.get("dram__bytes_read.sum", 0.0) for p in per.values())
    wr = sum(p.get("dram__bytes_write.sum", 0.0) for p in per.values())
    t = sum(p.get("gpu__time_duration.sum", 0.0) for p in per.values())
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from bench import kernel_source_hash
    out = {"tag": tag, "kernel_source_sha16": kernel_source_hash(), "launches": n, "dram_read_bytes_per_launch": rd / n, "dram_write_bytes_per_launch": wr / n,
           "traffic_bytes_per_launch": (rd + wr) / n, "mean_duration_us_under_ncu": 1e6 * t / n,
           "what": "all conv_umma launches of one warm denoiser evaluation (B=64, 4 bands, 64x64, bf16), ncu --clock-control none"}
    json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "conv_traffic.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "")
