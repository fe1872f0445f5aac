#!/bin/bash
# A/B runs of the bench workload on ONE box: each argument is a tag=ENV1=v1,ENV2=v2 (or just a tag for the default build);
# prints patches/s and the per-launch time of the layers that matter.   bash profiles/_ab.sh base generic=TMD_UMMA_GENERIC=1
mkdir -p gpurun_out
for spec in "$@"; do
  tag=${spec%%=*}; envs=""
  if [[ "$spec" == *=* ]]; then envs=$(echo "${spec#*=}" | tr ',' ' '); fi
  env $envs timeout 300 python bench.py --steps ${AB_STEPS:-3} --warmup 3 --no-parity --no-eager-baseline --no-cpu-baseline --launch-csv gpurun_out/ev_$tag.csv > gpurun_out/bench_$tag.json 2>gpurun_out/bench_$tag.err || tail -3 gpurun_out/bench_$tag.err
  python - <<PY
import csv, collections, json
rows = list(csv.DictReader(open("gpurun_out/ev_$tag.csv")))
agg = collections.OrderedDict()
for r in rows:
    a = agg.setdefault(r["label"], [0, 0.0, 0.0]); a[0] += 1; a[1] += float(r["ms"]); a[2] += float(r["flops"])
tot = sum(v[1] for v in agg.values())
d = json.load(open("gpurun_out/bench_$tag.json"))
print("== $tag ($envs): %.1f patches/s, e2e %.1f, conv3 %.1f TFLOP/s, launches %.2f ms, gap %.2f ms, clocks %s" % (
    d["value"], d["e2e"]["value"], d["roofline"]["achieved"], tot, d["launch_gap_ms"], d["clocks"]["sm_mhz"]))
if "$AB_LAYERS" != "none":
    for kk, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:${AB_TOP:-28}]:
        print("   %-72s x%-3d %7.1f us  %6.0f TFLOP/s  %5.2f ms" % (kk[:72], v[0], 1e3 * v[1] / v[0], v[2] / (v[1] * 1e-3) / 1e12 if v[1] else 0, v[1]))
PY
done
