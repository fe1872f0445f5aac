mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/e8_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/e8_tests.log
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --launch-csv gpurun_out/e8_ev_$tag.csv > gpurun_out/e8_b_$tag.json 2> gpurun_out/e8_b_$tag.err || tail -3 gpurun_out/e8_b_$tag.err; python -c "
import json;d=json.load(open('gpurun_out/e8_b_$tag.json'));print('$tag', round(d['value'],1), round(d['e2e']['value'],1), d['kernel_kinds']['conv3x3x3']['ms'], d['kernel_kinds']['elementwise']['ms'], d['clocks']['sm_mhz'])"; }
run acc2 TMD_UMMA_ACC4=0
run acc4 X=0
run acc2b TMD_UMMA_ACC4=0
run acc4b X=0
