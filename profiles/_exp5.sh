mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/e5_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/e5_tests.log
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --launch-csv gpurun_out/e5_ev_$tag.csv > gpurun_out/e5_b_$tag.json 2> gpurun_out/e5_b_$tag.err || tail -3 gpurun_out/e5_b_$tag.err; python -c "
import json;d=json.load(open('gpurun_out/e5_b_$tag.json'));print('$tag', round(d['value'],1), d['kernel_kinds']['conv3x3x3']['ms'], d['kernel_kinds']['conv1x1x1']['ms'])"; }
V=build/variants
run old TMDIFF_B200_LIB=$V/libold.so
run new X=0
run nosplit TMD_UMMA_COSPLIT=0
run sleep TMDIFF_B200_LIB=$V/libsleep.so
run nopref TMDIFF_B200_LIB=$V/libnopref.so
run bt9 TMD_UMMA_BTAPS=9
run old2 TMDIFF_B200_LIB=$V/libold.so
run new2 X=0
