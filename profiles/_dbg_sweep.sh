#!/bin/bash
# Role sweep of the tcgen05 conv kernel on the bench workload (TMD_UMMA_DEBUG bits: 1 no MMA, 2 no activation loads, 4 no weight
# loads, 8 no epilogue, 16 no tanh, 32 no residual loads, 64 no raw stores, 128 no activated stores): per-layer CUDA-event times.
mkdir -p gpurun_out
for dbg in ${DBG_LIST:-0 1 192 193 7 16 224}; do
  TMD_UMMA_DEBUG=$dbg timeout 200 python bench.py --steps 2 --warmup 3 --no-parity --no-eager-baseline --no-cpu-baseline --launch-csv gpurun_out/ev_dbg$dbg.csv > gpurun_out/bench_dbg$dbg.json 2>/dev/null
  python - <<PY
import csv, collections, json
rows = list(csv.DictReader(open("gpurun_out/ev_dbg$dbg.csv")))
agg = collections.OrderedDict()
for r in rows:
    k = r["label"]
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += float(r["ms"])
tot = sum(v[1] for v in agg.values())
d = json.load(open("gpurun_out/bench_dbg$dbg.json"))
print("dbg=$dbg  patches/s %.1f  sum-of-launch ms %.2f" % (d["value"], tot))
for k in ("conv2.conv21", "final.conv20.conv20", "final.conv21.conv20", "final.conv21.conv21", "final.conv23.conv21", "up3.up1.Conv_1", "down1.conv20.conv20", "down1.conv20.conv21", "down1.down.Conv_0", "up3.up1.Conv_0", "down3.conv20.conv21"):
    for kk, v in agg.items():
        if kk.startswith(k + " "):
            print("   %-70s %7.1f us" % (kk, 1e3 * v[1] / v[0]))
PY
done
