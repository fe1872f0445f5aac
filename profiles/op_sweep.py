#!/usr/bin/env python
"""Per-operator sweep (BASELINE config 5): the memory-bound operators at the UNet's level shapes against the measured
HBM copy bandwidth, the attention core against its FLOPs.  CUDA events on the launching stream, inputs larger than L2
or an L2 flush between iterations.

    python profiles/op_sweep.py > gpurun_out/op_sweep.md
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tmdiff_b200 import ops, operators  # noqa: E402

PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6650.0
FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def _median_ms(run, iters):
    ts = []
    for _ in range(iters):
        FLUSH.zero_()                       # evict L2 (126 MB) between iterations
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def timed(fn, iters=5):
    """(call, device): one call through the Python wrapper (its allocations, ctypes marshalling and launch latency included --
    about 30 us, which is all a small operator costs) and the same call replayed from a CUDA graph (device time only)."""
    fn(); fn()
    call = _median_ms(fn, iters)
    dev = None
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            fn()
        g.replay()
        dev = _median_ms(g.replay, iters)
    except Exception:                       # an operator that cannot be captured keeps the call time only
        torch.cuda.synchronize()
    return call, dev


def row(name, shape, t, nbytes=None, flops=None):
    call, dev = t if isinstance(t, tuple) else (t, None)
    ms = dev if dev else call
    gbs = nbytes / ms / 1e6 if nbytes else None
    print("| %s | %s | %.3f | %s | %s | %s | %s |" % (name, shape, call, "%.3f" % dev if dev else "", "%.0f" % gbs if gbs else "",
                                                      "%.0f%%" % (100 * gbs / PEAK) if gbs else "", "%.1f" % (flops / ms / 1e9) if flops else ""))


def main():
    print("# per-operator sweep (median of 5, L2 flushed between iterations); HBM peak = %.0f GB/s (measured copy)\n" % PEAK)
    print("`call ms` = one call through the Python wrapper; `device ms` = the same call replayed from a CUDA graph; GB/s, the fraction and")
    print("TFLOP/s are of the device time (of the call time where the operator could not be captured).\n")
    print("| operator | shape | call ms | device ms | GB/s (algorithmic) | of HBM peak | TFLOP/s |")
    print("|---|---|---:|---:|---:|---:|---:|")
    dwt, idwt = ops.DWT_2D("haar"), ops.IDWT_2D("haar")
    for B in (1, 16, 64, 256, 1024):
        for (CN, HW) in ((128, 64), (256, 32), (512, 16)):         # 4-band level shapes (C*N, H)
            x = torch.randn(B, CN, HW, HW, device="cuda")
            row("DWT_2D", tuple(x.shape), timed(lambda: dwt(x)), 2 * x.numel() * 4)
            row("DWT_2D LL-only", tuple(x.shape), timed(lambda: dwt(x, ll_only=True)), 1.25 * x.numel() * 4)
            bands = dwt(x)
            row("IDWT_2D", tuple(x.shape), timed(lambda: idwt(*bands)), 2 * x.numel() * 4)
    for B in (1, 16, 64, 256, 1024):
        for (C, HW) in ((32, 64), (64, 32), (128, 16), (256, 8)):
            x = torch.randn(B, C, 4, HW, HW, device="cuda")
            ga, be = torch.randn(B, C, device="cuda"), torch.randn(B, C, device="cuda")
            row("GroupNorm(8)+FiLM+Swish", tuple(x.shape), timed(lambda: operators.group_norm(x, 8, ga, be, swish=True)), 2 * x.numel() * 4)
    for B in (16, 64, 256):
        x0 = torch.randn(B, 4, 64, 64, device="cuda") * 1.2
        row("dynamic_thresholding", tuple(x0.shape), timed(lambda: ops.dynamic_thresholding(x0)), 2 * x0.numel() * 4)
    from tmdiff_b200 import metrics
    for (rows, C) in ((64 * 4096, 64), (64 * 1024, 320), (16 * 4096, 512), (64 * 256, 1280)):
        x, w, b = torch.randn(rows, C, device="cuda"), torch.randn(C, device="cuda"), torch.randn(C, device="cuda")
        row("LayerNorm", (rows, C), timed(lambda: operators.layer_norm(x, w, b)), 2 * x.numel() * 4)
        p = torch.randn(rows, 2 * C, device="cuda")
        row("GEGLU gate", (rows, 2 * C), timed(lambda: operators.geglu(p)), 1.5 * p.numel() * 4)
    for B in (16, 64, 256):
        a, b = torch.rand(B, 4, 256, 256, device="cuda"), torch.rand(B, 4, 256, 256, device="cuda")
        row("band moments (4x4 pairs)", tuple(a.shape), timed(lambda: metrics.band_moments(a, b)), 2 * a.numel() * 4)
        row("band moments + 3x3 high-pass", tuple(a.shape), timed(lambda: metrics.band_moments(a, b, highpass=True)), 2 * a.numel() * 4)
        row("quality_metrics (SAM/ERGAS/Q)", tuple(a.shape), timed(lambda: ops.quality_metrics(a, b)), 2 * a.numel() * 4)
        blur = operators.GaussianKernel(kernel_size=8, bands=4, scale_factor=4).cuda()
        row("Gaussian PSF blur /4", tuple(a.shape), timed(lambda: blur(a)), (1 + 1 / 16) * a.numel() * 4)
        e = torch.randn_like(a)
        row("ancestral step (Philox in kernel)", tuple(a.shape), timed(lambda: ops.ancestral_update(a, e, None, 1.1, 0.4, 0.3, 0.7, 0.5, True, 1, 2)),
            3 * a.numel() * 4)
    # Dynamic_conv3d (core/dynamic_conv.py:244-268): expert attention (pool + 2 FCs + softmax), expert mix, ONE grouped conv launch
    for (B, Ci, Co, N, HW, dt) in ((64, 32, 32, 4, 64, "bf16"), (64, 64, 64, 4, 32, "bf16"), (64, 128, 128, 4, 16, "bf16"), (16, 32, 32, 4, 64, "bf16"),
                                   (64, 32, 32, 4, 32, "fp32")):
        m = operators.Dynamic_conv3d(Ci, Co, 3, padding=1, K=4, dtype=dt).cuda()
        x = torch.randn(B, Ci, N, HW, HW, device="cuda")
        row("Dynamic_conv3d k=3 K=4 %s (head + mix + one conv launch)" % dt, tuple(x.shape), timed(lambda: m(x), 3),
            nbytes=2 * x.numel() * 4, flops=2.0 * 27 * Ci * Co * B * N * HW * HW)
    # scene front / back end (data/LRHR_dataset.py:55-139): DN scaling + bilinear x4 + tiling, and the way back
    from tmdiff_b200 import tiling
    pan, lrms = torch.rand(1, 2048, 2048, device="cuda") * 2047, torch.rand(8, 512, 512, device="cuda") * 2047
    row("scene_to_tiles 2048^2, 8 bands -> 1024 tiles", (8, 2048, 2048), timed(lambda: tiling.scene_to_tiles(pan, lrms, 64, 0, 1024, 2047.0)),
        nbytes=(pan.numel() + lrms.numel() + 9 * 2048 * 2048) * 4)
    tiles = torch.rand(1024, 8, 64, 64, device="cuda")
    scene = torch.empty(8, 2048, 2048, device="cuda")
    row("tiles_to_scene 1024 tiles -> 2048^2, 8 bands", (1024, 8, 64, 64), timed(lambda: tiling.tiles_to_scene(tiles, scene, 0, 2047.0, True)),
        nbytes=2 * tiles.numel() * 4)
    for (BH, Lq, Lk, D) in ((64, 64, 64, 1024), (16, 256, 256, 512), (8, 1024, 1024, 256), (2, 4096, 4096, 128), (128, 1024, 77, 64),
                            (512, 256, 77, 64)):
        q, k, v = (torch.randn(BH, L, D, device="cuda") for L in (Lq, Lk, Lk))
        row("attention core (fp32)", (BH, Lq, Lk, D), timed(lambda: operators.attention(q, k, v, D ** -0.5), 3), flops=4.0 * BH * Lq * Lk * D)


if __name__ == "__main__":
    main()
