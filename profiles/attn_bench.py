"""Attention-core sweep on one B200 (BASELINE config 5): the tcgen05 kernel (tmd_attention_bf16) against the
fp32 CUDA-core kernel (tmd_attention), CUDA events on the launching stream, operands already in the kernel's layout.
Algorithmic FLOPs = 4 * BH * Lq * Lk * D (QK^T and PV); algorithmic bytes = bf16 q, k, v read once + fp32 out written once.  A shape
is judged against the roofline that bounds it: tensor when FLOPs / bytes exceeds the machine balance (measured bf16 peak / measured
HBM copy bandwidth), HBM otherwise.  Usage: python profiles/attn_bench.py [out.md]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tmdiff_b200 import _lib  # noqa: E402

PEAKS = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))


def _peak():
    for k in ("bf16_tflops_sustained", "bf16_tflops"):
        if k in PEAKS:
            return float(PEAKS[k]), k
    return 1394.3, "fallback"


def time_ms(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    lib = _lib.load()
    peak, src = _peak()
    rows = []
    shapes = [  # (what, BH, Lq, Lk, D)
        ("self-attn L0 tokens 64x64, C*N=256", 16, 4096, 4096, 256),
        ("self-attn L0, B=64", 64, 4096, 4096, 256),
        ("self-attn L1 tokens 32x32, C*N=512", 64, 1024, 1024, 512),
        ("self-attn L2 tokens 16x16, C*N=1024", 256, 256, 256, 1024),
        ("self-attn L3 tokens 8x8, C*N=2048", 1024, 64, 64, 2048),
        ("head 128", 64, 4096, 4096, 128),
        ("cross-attn 8x64 heads, 77 text keys, 4096 tokens, B=64", 512, 4096, 77, 64),
        ("cross-attn, 1024 tokens, B=256", 2048, 1024, 77, 64),
    ]
    for what, BH, Lq, Lk, D in shapes:
        g = torch.Generator(device="cuda").manual_seed(0)
        q = torch.randn(BH, Lq, D, device="cuda", generator=g)
        k = torch.randn(BH, Lk, D, device="cuda", generator=g)
        v = torch.randn(BH, Lk, D, device="cuda", generator=g)
        pitch = (Lk + 7) // 8 * 8
        qb, kb = q.bfloat16(), k.bfloat16()
        vt = torch.nn.functional.pad(v.transpose(1, 2), (0, pitch - Lk)).bfloat16().contiguous()
        out = torch.empty(BH, Lq, D, device="cuda")
        scale = D ** -0.5
        st = _lib.stream_ptr()

        def bf():
            _lib.check(lib.tmd_attention_bf16(_lib.ptr(qb), _lib.ptr(kb), _lib.ptr(vt), _lib.ptr(out), BH, Lq, Lk, D, pitch, scale, st))

        def f32():
            _lib.check(lib.tmd_attention(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), BH, Lq, Lk, D, scale, st))

        flops = 4.0 * BH * Lq * Lk * D
        t_bf = time_ms(bf, 10)
        # the fp32 kernel is ~100x slower: time it on a slice of the batch
        sub = max(1, BH // 16)
        q, k, v = q[:sub].contiguous(), k[:sub].contiguous(), v[:sub].contiguous()
        BH_full, BH = BH, sub
        t_f32 = time_ms(f32, 2) * BH_full / sub
        BH = BH_full
        nbytes = 2.0 * BH * D * (Lq + 2 * Lk) + 4.0 * BH * Lq * D
        hbm = float(PEAKS.get("hbm_gbs", 6537.6))
        tf, gbs = flops / t_bf / 1e9, nbytes / t_bf / 1e6
        bound = "tensor" if flops / nbytes > peak * 1e3 / hbm else "hbm"
        rows.append((what, BH, Lq, Lk, D, t_bf, tf, gbs, bound, tf / peak if bound == "tensor" else gbs / hbm, flops / t_f32 / 1e9))
        print(rows[-1], flush=True)
    lines = ["# Attention core sweep (tcgen05 `attn_umma_kernel` vs fp32 CUDA-core kernel)", "",
             "Peaks (MEASURED_PEAKS.json): %.1f TFLOP/s (%s), %.1f GB/s HBM copy. FLOPs = 4*BH*Lq*Lk*D; bytes = bf16 q,k,v + fp32 out."
             % (peak, src, float(PEAKS.get("hbm_gbs", 6537.6))), "",
             "| shape | BH | Lq | Lk | D | ms | TFLOP/s | GB/s | bound | frac of that roofline | fp32 kernel TFLOP/s |", "|---|---|---|---|---|---|---|---|---|---|---|"]
    for r in rows:
        lines.append("| %s | %d | %d | %d | %d | %.3f | %.1f | %.0f | %s | %.2f | %.1f |" % r)
    text = "\n".join(lines) + "\n"
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(text)
    print(text)


if __name__ == "__main__":
    main()
