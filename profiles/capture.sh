#!/bin/bash
# Round evidence capture on one B200 (run under gpurun): tests, bench, ncu launch list, ncu --set full.
# Usage: bash profiles/capture.sh <tag>
tag=${1:-r01}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > gpurun_out/smi_$tag.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$tag.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_$tag.log
tail -3 gpurun_out/pytest_$tag.log
timeout 900 python bench.py --steps 5 --warmup 3 --launch-csv gpurun_out/ev_$tag.csv > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench exit $?"
tail -c 1500 gpurun_out/bench_$tag.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err; echo "ref exit $?"
# launch list of the same command (short run: 1 timed step), cold-cache + serialised: shares only
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_$tag.csv \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-parity > gpurun_out/ncu_bench_$tag.log 2>&1; echo "ncu list exit $?"
# full capture of the conv kernel on one denoiser evaluation (second evaluation: warm)
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:conv_umma --launch-skip ${SKIP:-64} -c 38 \
  -o gpurun_out/conv_umma_$tag -f python profiles/_one_eval.py 64 > gpurun_out/ncu_full_$tag.log 2>&1; echo "ncu full exit $?"
# DRAM traffic of every conv launch of one warm evaluation (the `traffic` of bench.py's roofline object)
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:conv_umma --launch-skip ${SKIP:-64} -c 38 --csv \
  --log-file gpurun_out/conv_traffic_$tag.csv python profiles/_one_eval.py 64 > gpurun_out/ncu_traffic_$tag.log 2>&1; echo "ncu traffic exit $?"
ncu -i gpurun_out/conv_umma_$tag.ncu-rep --page raw --csv > gpurun_out/conv_umma_${tag}_raw.csv 2>/dev/null
# the report itself is scratch (the raw CSV above is what gets summarised); keep the merge-back under its size limit
rm -f gpurun_out/conv_umma_$tag.ncu-rep
ls -la gpurun_out | tail -20
