import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tmdiff_b200 import ops
for (B, Ci, Co, N, H, W) in [(64, 32, 32, 4, 64, 64), (64, 64, 64, 4, 64, 64), (64, 128, 128, 4, 32, 32), (64, 256, 256, 4, 16, 16)]:
    x = torch.randn(B, Ci, N, H, W, device="cuda")
    w = torch.randn(Co, Ci, 3, 3, 3, device="cuda") / (27 * Ci) ** 0.5
    for _ in range(2):
        y = ops.modulated_conv3d(x, w, None, padding=1, dtype="bf16")
    torch.cuda.synchronize()
