mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --launch-csv gpurun_out/e13_ev_$tag.csv > gpurun_out/e13_b_$tag.json 2> gpurun_out/e13_b_$tag.err || tail -3 gpurun_out/e13_b_$tag.err; python -c "
import json;d=json.load(open('gpurun_out/e13_b_$tag.json'));print('$tag', round(d['value'],1), round(d['e2e']['value'],1), d['kernel_kinds']['conv3x3x3']['ms'], d['clocks']['sm_mhz'])"; }
V=build/variants
run v5 TMDIFF_B200_LIB=$V/libv5.so
run new X=0
run nofold TMD_UMMA_NO_FOLD=1
run v5b TMDIFF_B200_LIB=$V/libv5.so
run newb X=0
