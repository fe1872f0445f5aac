#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page raw --csv` (one row per captured launch, `--set full`) into a table.

    python profiles/summarize_ncu_raw.py gpurun_out/conv_umma_rNN_raw.csv > profiles/ncu_full_rNN.md

Columns: duration, DRAM read/write per launch (the `traffic` of bench.py's roofline object), tensor-pipe
activity, achieved bf16 tensor op rate, SM / L2 / DRAM throughput percentages, registers.
"""
import csv
import re
import sys

COLS = [
    ("us", "gpu__time_duration.sum"),
    ("dram rd MB", "dram__bytes_read.sum"),
    ("dram wr MB", "dram__bytes_write.sum"),
    ("tensor pipe % of active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("SM thr %", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("L2 thr %", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("DRAM thr %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("regs", "launch__registers_per_thread"),
]
SCALE = {"": 1.0, "K": 1e3, "M": 1e6, "G": 1e9, "T": 1e12, "k": 1e3, "n": 1e-9, "u": 1e-6, "m": 1e-3}


def to_base(value, unit):
    """value in `unit` -> value in the unprefixed base unit (bytes, seconds, ops/s)."""
    f = float(value.replace(",", ""))
    u = unit.strip()
    for base in ("byte", "second", "ops/s", "/s", "s"):
        if u.endswith(base) or u == base:
            pre = u[: len(u) - len(base)]
            if pre in SCALE:
                return f * SCALE[pre]
    return f


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [(n, m) for n, m in COLS if m in idx]
    print("| # | kernel | grid | " + " | ".join(n for n, _ in cols) + " |")
    print("|---|---|---|" + "---:|" * len(cols))
    for k, r in enumerate(data):
        name = r[idx["Kernel Name"]].replace("unnamed>::", "")
        m = re.search(r"(\w+)<([^>]*)>", name)
        short = "%s<%s>" % (m.group(1), m.group(2)) if m else name[:40]
        vals = []
        for n, met in cols:
            try:
                f = to_base(r[idx[met]], units[idx[met]])
            except ValueError:
                vals.append(r[idx[met]])
                continue
            if n == "us": f *= 1e6
            elif n.endswith("MB"): f /= 1e6
            elif n.endswith("TFLOP/s"): f /= 1e12
            vals.append("%.1f" % f)
        print("| %d | `%s` | %s | " % (k, short, r[idx["Grid Size"]]) + " | ".join(vals) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
