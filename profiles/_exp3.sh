# epilogue-vector staging + multi-tap weight stages: parity first, then role sweeps (scratch script)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/e3_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/e3_tests.log
run() { tag=$1; shift; env "$@" timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --launch-csv gpurun_out/e3_ev_$tag.csv > gpurun_out/e3_b_$tag.json 2> gpurun_out/e3_b_$tag.err || tail -3 gpurun_out/e3_b_$tag.err; python -c "
import json;d=json.load(open('gpurun_out/e3_b_$tag.json'));print('$tag', round(d['value'],1), d['kernel_kinds']['conv3x3x3']['ms'], d['kernel_kinds']['conv1x1x1']['ms'])"; }
run base X=0
run bt1 TMD_UMMA_BTAPS=1
run bt9 TMD_UMMA_BTAPS=9
run mma_only TMD_UMMA_DEBUG=14
run epi_only TMD_UMMA_DEBUG=7
run epi_nostore TMD_UMMA_DEBUG=199
run epi_bare TMD_UMMA_DEBUG=247
run mma_only_bt1 TMD_UMMA_DEBUG=14 TMD_UMMA_BTAPS=1
TMD_UMMA_TRACE=1 TMD_NO_GRAPH=1 timeout 300 python profiles/_one_eval.py > gpurun_out/e3_trace.txt 2>&1; grep -c "umma trace" gpurun_out/e3_trace.txt
