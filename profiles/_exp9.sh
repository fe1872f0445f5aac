mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/e9_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/e9_tests.log
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --launch-csv gpurun_out/e9_ev_$tag.csv > gpurun_out/e9_b_$tag.json 2> gpurun_out/e9_b_$tag.err || tail -3 gpurun_out/e9_b_$tag.err; python -c "
import json;d=json.load(open('gpurun_out/e9_b_$tag.json'));print('$tag', round(d['value'],1), round(d['e2e']['value'],1), d['kernel_kinds']['conv3x3x3']['ms'], d['kernel_kinds']['elementwise']['ms'], d['clocks']['sm_mhz'])"; }
V=build/variants
run nohint TMDIFF_B200_LIB=$V/libnohint.so
run hint20k X=0
run hint1k TMDIFF_B200_LIB=$V/libhint1k.so
run nohintb TMDIFF_B200_LIB=$V/libnohint.so
run hint20kb X=0
