mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --launch-csv gpurun_out/e10_ev_$tag.csv > gpurun_out/e10_b_$tag.json 2> gpurun_out/e10_b_$tag.err || tail -3 gpurun_out/e10_b_$tag.err; python -c "
import json;d=json.load(open('gpurun_out/e10_b_$tag.json'));print('$tag', round(d['value'],1), round(d['e2e']['value'],1), d['kernel_kinds']['conv3x3x3']['ms'], d['clocks']['sm_mhz'])"; }
V=build/variants
run base X=0
run leader TMDIFF_B200_LIB=$V/libleader.so
run sleep32 TMDIFF_B200_LIB=$V/libsleep32.so
run both TMDIFF_B200_LIB=$V/libboth.so
run baseb X=0
run leaderb TMDIFF_B200_LIB=$V/libleader.so
TMDIFF_B200_LIB=$V/libleader.so timeout 600 python -m pytest tests/test_gpu_denoiser.py tests/test_gpu_ops.py -m gpu -x -q 2>&1 | tail -2
