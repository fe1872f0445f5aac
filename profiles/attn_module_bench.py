"""Module-level throughput of the attention operators: the whole `forward` of SpatialSelfAttention
(core/Attention.py:140-159), AttnBlockpp (Hyper_unet_general.py:487-512) and CrossAttention (core/Attention.py:185-214) in bf16
mode -- GroupNorm, the q / k / v / output projections, layout changes and the attention core -- next to the core alone on
operands already in its layout, and next to the same module written with torch library kernels (F.group_norm, F.conv2d /
F.linear, F.scaled_dot_product_attention in bf16).  CUDA events on the launching stream, 3 warm-up calls.
FLOPs = 4 projections (2*tokens*C*C each) + 4*B*heads*Lq*Lk*Dh.    Usage: python profiles/attn_module_bench.py [out.md]"""
import json
import os
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tmdiff_b200 import operators  # noqa: E402

PEAKS = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
PEAK = float(PEAKS.get("bf16_tflops_sustained", PEAKS.get("bf16_tflops", 1394.3)))


def time_ms(fn, iters=10):
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


def torch_self_attention(m, x):
    """The reference's SpatialSelfAttention.forward with library kernels in bf16 (autocast-like)."""
    B, C, H, W = x.shape
    h = F.group_norm(x, 32, m.norm.weight, m.norm.bias, eps=1e-6).bfloat16()
    q, k, v = (F.conv2d(h, c.weight.bfloat16(), c.bias.bfloat16()) for c in (m.q, m.k, m.v))
    tok = lambda t: t.reshape(B, 1, C, H * W).transpose(2, 3)
    o = F.scaled_dot_product_attention(tok(q), tok(k), tok(v), scale=C ** -0.5)
    o = o.transpose(2, 3).reshape(B, C, H, W)
    return x + F.conv2d(o, m.proj_out.weight.bfloat16(), m.proj_out.bias.bfloat16()).float()


def torch_cross_attention(m, x, ctx):
    B, L, _ = x.shape
    h = m.heads
    xb, cb = x.bfloat16(), ctx.bfloat16()
    q, k, v = F.linear(xb, m.to_q.weight.bfloat16()), F.linear(cb, m.to_k.weight.bfloat16()), F.linear(cb, m.to_v.weight.bfloat16())
    sp = lambda t: t.reshape(B, t.shape[1], h, -1).transpose(1, 2)
    o = F.scaled_dot_product_attention(sp(q), sp(k), sp(v), scale=m.scale).transpose(1, 2).reshape(B, L, -1)
    return F.linear(o, m.to_out[0].weight.bfloat16(), m.to_out[0].bias.bfloat16()).float()


def main():
    out = ["# Attention modules end to end (bf16 mode), one B200", "",
           "Peak for the fraction: %.1f TFLOP/s (MEASURED_PEAKS.json, sustained bf16).  `module` = the whole forward through"
           " `tmdiff_b200.operators`; `core` = `tmd_attention_bf16` alone on operands already in its layout; `torch` = the same forward"
           " written with F.group_norm / F.conv2d / F.scaled_dot_product_attention in bf16." % PEAK, "",
           "| module | shape | module ms | core ms | torch ms | module TFLOP/s | frac of peak | module / torch |", "|---|---|---:|---:|---:|---:|---:|---:|"]
    torch.manual_seed(0)
    for (B, C, HW) in ((16, 256, 64), (64, 256, 64), (64, 512, 32), (64, 1024, 16)):
        m = operators.SpatialSelfAttention(C, dtype="bf16").cuda()
        x = torch.randn(B, C, HW, HW, device="cuda")
        L = HW * HW
        flops = 4 * 2.0 * B * L * C * C + 4.0 * B * L * L * C
        t_mod = time_ms(lambda: m(x))
        q = torch.randn(B, L, C, device="cuda")
        v = torch.randn(B, C, L, device="cuda")
        qb, vb = q.bfloat16(), v.bfloat16()
        t_core = time_ms(lambda: operators.attention_bf16(qb, qb, vb, C ** -0.5, v_channel_major=True))
        with torch.no_grad():
            t_ref = time_ms(lambda: torch_self_attention(m, x))
            err = float((m(x) - torch_self_attention(m, x)).abs().max() / torch_self_attention(m, x).abs().max())
        out.append("| SpatialSelfAttention | B=%d C=%d %dx%d (err vs torch %.1e) | %.3f | %.3f | %.3f | %.1f | %.2f | %.2f |" % (
            B, C, HW, HW, err, t_mod, t_core, t_ref, flops / t_mod / 1e9, flops / t_mod / 1e9 / PEAK, t_ref / t_mod))
    for (B, Cf, N, HW) in ((64, 64, 4, 64), (64, 128, 4, 32)):
        m = operators.AttnBlockpp(Cf * N, dtype="bf16").cuda()
        x = torch.randn(B, Cf, N, HW, HW, device="cuda")
        L, C = HW * HW, Cf * N
        flops = 4 * 2.0 * B * L * C * C + 4.0 * B * L * L * C
        t_mod = time_ms(lambda: m(x))
        out.append("| AttnBlockpp | B=%d C*N=%d %dx%d | %.3f | | | %.1f | %.2f | |" % (B, C, HW, HW, t_mod, flops / t_mod / 1e9, flops / t_mod / 1e9 / PEAK))
    for (B, L, qd) in ((64, 4096, 320), (256, 1024, 640)):
        m = operators.CrossAttention(qd, context_dim=768, heads=8, dim_head=64, dtype="bf16").cuda()
        x, ctx = torch.randn(B, L, qd, device="cuda"), torch.randn(B, 77, 768, device="cuda")
        inner = 512
        flops = 2.0 * B * L * qd * inner * 2 + 2.0 * B * 77 * 768 * inner * 2 + 4.0 * B * 8 * L * 77 * 64
        t_mod = time_ms(lambda: m(x, ctx))
        qh = torch.randn(B * 8, L, 64, device="cuda").bfloat16()
        kh = torch.randn(B * 8, 77, 64, device="cuda").bfloat16()
        t_core = time_ms(lambda: operators.attention_bf16(qh, kh, kh, 0.125))
        with torch.no_grad():
            t_ref = time_ms(lambda: torch_cross_attention(m, x, ctx))
        out.append("| CrossAttention 8x64, 77 keys | B=%d L=%d dim=%d | %.3f | %.3f | %.3f | %.1f | %.2f | %.2f |" % (
            B, L, qd, t_mod, t_core, t_ref, flops / t_mod / 1e9, flops / t_mod / 1e9 / PEAK, t_ref / t_mod))
    out += ["", "Projection alone (`tmd_linear`, bf16 in / bf16 out, weight pack included):", "",
            "| tokens | Ci | Co | ms | TFLOP/s | GB/s (x + y) |", "|---:|---:|---:|---:|---:|---:|"]
    for (M, Ci, Co) in ((262144, 256, 256), (262144, 320, 512), (65536, 512, 512), (16384, 1024, 1024), (262144, 512, 320)):
        x = torch.randn(M, Ci, device="cuda").bfloat16()
        w = torch.randn(Co, Ci, device="cuda") / Ci ** 0.5
        t = time_ms(lambda: operators._linear_tokens(x, w, None, "bf16", out_bf16=True))
        out.append("| %d | %d | %d | %.3f | %.1f | %.0f |" % (M, Ci, Co, t, 2.0 * M * Ci * Co / t / 1e9, (M * Ci + M * Co) * 2 / t / 1e6))
    text = "\n".join(out) + "\n"
    print(text)
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(text)


if __name__ == "__main__":
    main()
