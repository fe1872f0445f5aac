import os, sys, torch
sys.path.insert(0, os.getcwd())
from tmdiff_b200 import operators
g = torch.randn(256, 32, 4, 64, 64, device="cuda")
ga, be = torch.randn(256, 32, device="cuda"), torch.randn(256, 32, device="cuda")
for _ in range(3): operators.group_norm(g, 8, ga, be, swish=True)
torch.cuda.synchronize()
ts=[]
for _ in range(10):
    a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    a.record(); operators.group_norm(g, 8, ga, be, swish=True); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
ms=sorted(ts)[5]; print(os.environ.get("TMD_GN_SLICE"), ms, 2*g.numel()*4/ms/1e6, "GB/s")
