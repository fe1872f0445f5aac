#!/bin/bash
# compute-sanitizer pass over every kernel of the library on a reduced matrix (small shapes, one case per kernel template
# instantiation): memcheck on the whole selection, racecheck / synccheck on the tcgen05 kernels.  Summaries -> gpurun_out/sanitizer_*.txt
mkdir -p gpurun_out
SEL_ALL='test_forward_bf16_golden or test_forward_fp32_golden or test_haar_golden or test_dynamic_threshold_boundary_counts or test_dynamic_threshold_degenerate or test_sampler_variants or test_scene_front_and_back_end_kernels or test_mixed_sensor_two_band_buckets or test_ancestral_steps_match_reference or test_small_channels_and_odd_band_count or (test_modulated_conv3d and not benchmark and not linearity and bf16)'
SEL_OPS='tests/test_gpu_operators.py'
SEL_TC='test_forward_bf16_golden or (test_modulated_conv3d and not benchmark and not linearity and bf16)'
run() {  # tool, timeout, tag, pytest args...
  tool=$1; to=$2; tag=$3; shift 3
  timeout $to compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 99 python -m pytest -q -x -p no:cacheprovider "$@" > gpurun_out/sanitizer_${tag}.log 2>&1
  rc=$?
  echo "== $tool / $tag: exit $rc" | tee -a gpurun_out/sanitizer_summary.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error" gpurun_out/sanitizer_${tag}.log | tail -4 | tee -a gpurun_out/sanitizer_summary.txt
  grep -E "=========     at |Invalid|hazard|Race" gpurun_out/sanitizer_${tag}.log | sort | uniq -c | sort -rn | head -12 | tee -a gpurun_out/sanitizer_summary.txt
}
: > gpurun_out/sanitizer_summary.txt
run memcheck ${T_MEM:-700} memcheck_path tests -m gpu -k "$SEL_ALL"
run memcheck ${T_MEM2:-500} memcheck_operators $SEL_OPS -m gpu
run racecheck ${T_RACE:-500} racecheck_tcgen05 tests -m gpu -k "$SEL_TC"
run synccheck ${T_SYNC:-400} synccheck_tcgen05 tests -m gpu -k "$SEL_TC"
run racecheck ${T_RACE2:-400} racecheck_attention tests/test_gpu_operators.py -m gpu -k "attention or SpatialSelf or Cross"
