#!/bin/bash
# A/B of an environment switch on the default bench workload: profiles/_ab2.sh TAG VAR=a VAR=b [reps]
tag=$1; a=$2; b=$3; reps=${4:-2}
mkdir -p gpurun_out
run() { t=$1; shift; env "$@" timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-parity > gpurun_out/ab_${tag}_$t.json 2> gpurun_out/ab_${tag}_$t.err || tail -3 gpurun_out/ab_${tag}_$t.err; python -c "
import json;d=json.loads(open('gpurun_out/ab_${tag}_$t.json').read().strip().splitlines()[-1]);print('$t', '$*', round(d['value'],1), round(d['e2e']['value'],1), d['kernel_kinds']['conv3x3x3']['ms'], d['clocks']['sm_mhz'])"; }
for i in $(seq 1 $reps); do run a$i $a; run b$i $b; done 2>&1 | tee gpurun_out/ab_${tag}.txt
