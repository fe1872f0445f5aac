#!/bin/bash
# Attention kernel evidence on one B200 (run under gpurun): parity tests, sweep, ncu --set full of attn_umma_kernel.
# Usage: bash profiles/capture_attn.sh <tag>
tag=${1:-r01}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_operators.py -m gpu -q > gpurun_out/attn_test_$tag.log 2>&1; echo "operator tests exit $?"
tail -5 gpurun_out/attn_test_$tag.log
timeout 300 python profiles/attn_bench.py gpurun_out/attn_sweep_$tag.md > gpurun_out/attn_bench_$tag.log 2>&1; echo "attn bench exit $?"
tail -14 gpurun_out/attn_bench_$tag.log
timeout 300 python profiles/_attn_one.py; echo "one exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_umma -c 5 -o gpurun_out/attn_umma_$tag -f \
  python profiles/_attn_one.py > gpurun_out/ncu_attn_$tag.log 2>&1; echo "ncu full exit $?"
ncu -i gpurun_out/attn_umma_$tag.ncu-rep --page raw --csv > gpurun_out/attn_umma_${tag}_raw.csv 2>/dev/null
rm -f gpurun_out/attn_umma_$tag.ncu-rep
ls -la gpurun_out | tail
