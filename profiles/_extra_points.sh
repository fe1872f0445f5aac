# other workloads on one GPU with the final library (values quoted in DESIGN.md section 6)
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 400 python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/extra_$tag.json 2> gpurun_out/extra_$tag.err || tail -3 gpurun_out/extra_$tag.err; python -c "
import json;d=json.loads(open('gpurun_out/extra_$tag.json').read().strip().splitlines()[-1]);print('$tag', round(d['value'],2), round(d['e2e']['value'],2), round(d['ms_per_step'],2), d['config'].get('workload'))"; }
run wv3_b64 --bands 8 --sensor WV3
run qb_b256 --batch 256
run cfg1_bf16 --batch 1 --bands 8 --sensor WV3 --solver-steps 10
run cfg1_fp32 --batch 1 --bands 8 --sensor WV3 --solver-steps 10 --dtype fp32
run qb_b16 --batch 16
