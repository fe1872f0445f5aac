"""Seeded synthetic inputs shared by the oracle, the tests and bench.py.
TEST INFRASTRUCTURE ONLY."""
import torch
import torch.nn.functional as F


def make_inputs(B, N, H=64, W=64, seed=3407):
    """PAN ~ U[0,1) [B,1,H,W]; MS = bicubic x4 of U[0,1) [B,N,H/4,W/4] (a smooth `lms`
    stand-in, data/LRHR_dataset.py:134-151 scales real data to [0,1]); x_T ~ N(0,1)."""
    g = torch.Generator().manual_seed(seed)
    pan = torch.rand(B, 1, H, W, generator=g)
    lr = torch.rand(B, N, H // 4, W // 4, generator=g)
    ms = F.interpolate(lr, scale_factor=4, mode="bicubic", align_corners=False)
    x_T = torch.randn(B, N, H, W, generator=g)
    return pan, ms, x_T
