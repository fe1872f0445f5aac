mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --launch-csv gpurun_out/e14_ev_$tag.csv > gpurun_out/e14_b_$tag.json 2> gpurun_out/e14_b_$tag.err || tail -3 gpurun_out/e14_b_$tag.err; python -c "
import json;d=json.load(open('gpurun_out/e14_b_$tag.json'));print('$tag', round(d['value'],1), round(d['e2e']['value'],1), d['kernel_kinds']['conv3x3x3']['ms'], d['clocks']['sm_mhz'])"; }
run bt3 X=0
run bt9 TMD_UMMA_BTAPS=9
run bt3b X=0
run bt9b TMD_UMMA_BTAPS=9
run bt3c X=0
run bt9c TMD_UMMA_BTAPS=9
