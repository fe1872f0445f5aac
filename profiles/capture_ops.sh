#!/bin/bash
# ncu --set full of the memory-bound kernels (Haar, GroupNorm cluster kernel, per-evaluation glue) for the HBM roofline.
tag=${1:-r01}
mkdir -p gpurun_out
cat > /tmp/_ops_probe.py <<'PY'
import os, sys, torch
sys.path.insert(0, os.getcwd())
from tmdiff_b200 import ops, operators
dwt, idwt = ops.DWT_2D("haar"), ops.IDWT_2D("haar")
x = torch.randn(256, 128, 64, 64, device="cuda")
for _ in range(2):
    b = dwt(x); y = idwt(*b); l = dwt(x, ll_only=True)
g = torch.randn(256, 32, 4, 64, 64, device="cuda")
ga, be = torch.randn(256, 32, device="cuda"), torch.randn(256, 32, device="cuda")
for _ in range(2):
    operators.group_norm(g, 8, ga, be, swish=True)
torch.cuda.synchronize()
PY
timeout 900 ncu --set full --clock-control none -k regex:"dwt2d|idwt2d|gn_cluster" --launch-skip 4 -c 4 -o gpurun_out/ops_$tag -f python /tmp/_ops_probe.py > gpurun_out/ncu_ops_$tag.log 2>&1; echo "ncu ops exit $?"
ncu -i gpurun_out/ops_$tag.ncu-rep --page raw --csv > gpurun_out/ops_${tag}_raw.csv 2>/dev/null
timeout 900 ncu --set full --clock-control none -k regex:"slice_activate|stem_expand|project_to_band|x0_threshold|solver_update" --launch-skip 8 -c 7 -o gpurun_out/glue_$tag -f python profiles/_one_eval.py 64 > gpurun_out/ncu_glue_$tag.log 2>&1; echo "ncu glue exit $?"
ncu -i gpurun_out/glue_$tag.ncu-rep --page raw --csv > gpurun_out/glue_${tag}_raw.csv 2>/dev/null
rm -f gpurun_out/ops_$tag.ncu-rep gpurun_out/glue_$tag.ncu-rep
ls -la gpurun_out | tail -8
