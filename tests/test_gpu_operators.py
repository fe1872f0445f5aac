"""Stand-alone operators against outputs of the reference's own classes (tests/golden/
make_golden_operators.py): AdaptiveGroupNorm, AttnBlockpp, SpatialSelfAttention, CrossAttention, Dynamic_conv3d."""
import os

import numpy as np
import pytest
import torch

from util import rel_l2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FP32_TOL, BF16_TOL = 1e-5, 2e-2      # north star: relative L2 in fp32 / bf16 mode


@pytest.fixture(scope="module")
def vec():
    return np.load(os.path.join(ROOT, "tests", "golden", "operator_vectors.npz"))


def _sd(vec, tag):
    pre = tag + "/sd/"
    return {k[len(pre):]: torch.from_numpy(vec[k]) for k in vec.files if k.startswith(pre)}


def _inp(vec, tag, name, cuda=True):
    t = torch.from_numpy(vec["%s/in/%s" % (tag, name)])
    return t.cuda() if cuda else t


def _build(tag, dtype="fp32"):
    from tmdiff_b200 import operators as O
    kw = {} if tag == "adagn" else {"dtype": dtype}
    return {"adagn": lambda: O.AdaptiveGroupNorm(8, 32, 128),
            "attnpp": lambda: O.AttnBlockpp(64, skip_rescale=True, init_scale=1.0, **kw),
            "ssa": lambda: O.SpatialSelfAttention(64, **kw),
            "xattn": lambda: O.CrossAttention(64, context_dim=128, heads=8, dim_head=64, **kw),
            "dyn3d": lambda: O.Dynamic_conv3d(16, 32, 3, padding=1, K=4, temperature=34, **kw)}[tag]()


@pytest.mark.parametrize("tag", ["adagn", "attnpp", "ssa", "xattn", "dyn3d"])
def test_state_dict_keys_match_reference(vec, tag):
    m = _build(tag)
    ref = _sd(vec, tag)
    mine = m.state_dict()
    assert set(mine) == set(ref)
    assert all(tuple(mine[k].shape) == tuple(ref[k].shape) for k in ref)


@pytest.mark.gpu
@pytest.mark.parametrize("tag,dtype,tol", [("adagn", "fp32", FP32_TOL), ("attnpp", "fp32", FP32_TOL), ("attnpp", "bf16", BF16_TOL),
                                           ("ssa", "fp32", FP32_TOL), ("ssa", "bf16", BF16_TOL), ("xattn", "fp32", FP32_TOL),
                                           ("xattn", "bf16", BF16_TOL), ("dyn3d", "fp32", FP32_TOL), ("dyn3d", "bf16", BF16_TOL)])
def test_operator_against_reference(vec, tag, dtype, tol):
    m = _build(tag, dtype)
    m.load_state_dict(_sd(vec, tag), strict=True)
    m = m.cuda()
    args = {"adagn": ("x", "style"), "attnpp": ("x",), "ssa": ("x",), "xattn": ("x", "context"), "dyn3d": ("x",)}[tag]
    ins = [_inp(vec, tag, a) for a in args]
    got = m(ins[0], context=ins[1]) if tag == "xattn" else m(*ins)
    assert rel_l2(got, torch.from_numpy(vec[tag + "/out"])) < tol
    if tag == "adagn":
        assert rel_l2(m(*ins, swish=True), torch.from_numpy(vec["adagn/out_swish"])) < tol


@pytest.mark.gpu
def test_group_norm_properties_full_size():
    """Level-0 map of the bench workload (B=16 here): every group of the output has zero mean and unit variance, the
    per-sample affine is applied per channel, and odd sizes take the scalar path."""
    from tmdiff_b200.operators import group_norm
    g = torch.Generator().manual_seed(0)
    x = (torch.randn(16, 32, 4, 64, 64, generator=g) * 3 + 10).cuda()
    y = group_norm(x, 8)
    grp = y.reshape(16, 8, -1)
    assert float(grp.mean(-1).abs().max()) < 1e-4 and float((grp.var(-1, unbiased=False) - 1).abs().max()) < 1e-4
    want = torch.nn.functional.group_norm(x, 8, eps=1e-6)
    assert rel_l2(y, want) < FP32_TOL
    ga, be = torch.randn(16, 32, generator=g).cuda(), torch.randn(16, 32, generator=g).cuda()
    assert rel_l2(group_norm(x, 8, ga, be), want * ga[:, :, None, None, None] + be[:, :, None, None, None]) < FP32_TOL
    xo = torch.randn(3, 6, 5, 7, generator=g).cuda()
    assert rel_l2(group_norm(xo, 3, swish=True), torch.nn.functional.silu(torch.nn.functional.group_norm(xo, 3, eps=1e-6))) < FP32_TOL
    with pytest.raises(RuntimeError):
        group_norm(xo, 4)          # 6 channels do not split into 4 groups (torch raises too)


@pytest.mark.gpu
@pytest.mark.parametrize("BH,Lq,Lk,D", [(4, 1024, 1024, 256), (16, 256, 77, 64), (2, 37, 50, 48), (1, 64, 64, 2048), (3, 9, 1, 8)])
def test_attention_core(BH, Lq, Lk, D):
    """Level shapes of the sweep (self-attention over HW tokens with C*N-wide heads, 77-token cross attention), ragged
    sizes, a single key; plus the property that constant values come back unchanged (softmax rows sum to one)."""
    from tmdiff_b200.operators import attention
    g = torch.Generator().manual_seed(1)
    q, k, v = (torch.randn(BH, L, D, generator=g).cuda() for L in (Lq, Lk, Lk))
    scale = D ** -0.5
    want = torch.softmax(torch.einsum("bid,bjd->bij", q.double(), k.double()) * scale, -1) @ v.double()
    assert rel_l2(attention(q, k, v, scale), want) < FP32_TOL
    ones = torch.ones_like(v)
    assert float((attention(q, k, ones, scale) - 1).abs().max()) < 1e-5
    with pytest.raises(RuntimeError):
        attention(q[:, :, :4].contiguous().repeat(1, 1, 1024), k[:, :, :4].contiguous().repeat(1, 1, 1024), v[:, :, :4].contiguous().repeat(1, 1, 1024), 1.0)


@pytest.mark.gpu
@pytest.mark.parametrize("BH,Lq,Lk,D", [(4, 1024, 1024, 256), (16, 256, 77, 64), (2, 300, 333, 128), (1, 64, 64, 2048), (2, 4096, 4096, 512),
                                        (3, 9, 1, 64), (2, 130, 200, 192), (2, 200, 100, 256), (2, 130, 128, 192), (3, 70, 40, 640), (4, 256, 256, 1024), (2, 140, 190, 320)])
def test_attention_core_bf16(BH, Lq, Lk, D):
    """The tcgen05 attention kernel (bf16 operands, fp32 logits / statistics / accumulation) against float64 softmax attention
    of the SAME bf16-rounded operands: level shapes of the sweep, ragged query / key counts (partial tiles, a partial key
    block, a single key), heads wider than one 256-column accumulator.  Logits up to +-30 so that the running maximum is
    raised (and the accumulator rescaled) in mid-sequence; constant values must come back unchanged."""
    from tmdiff_b200.operators import attention
    g = torch.Generator().manual_seed(2)
    q, k, v = (torch.randn(BH, L, D, generator=g).cuda() for L in (Lq, Lk, Lk))
    k[:, Lk // 2:] *= 3.0                  # later key blocks carry the larger logits
    scale = 2.0 * D ** -0.5
    r = lambda t: t.to(torch.bfloat16).double()
    want = torch.softmax(torch.einsum("bid,bjd->bij", r(q), r(k)) * scale, -1) @ r(v)
    got = attention(q, k, v, scale, "bf16")
    assert got.dtype == torch.float32 and got.shape == q.shape
    assert rel_l2(got, want) < 5e-3                       # only P's bf16 rounding and the fp32 accumulation order remain
    exact = torch.softmax(torch.einsum("bid,bjd->bij", q.double(), k.double()) * scale, -1) @ v.double()
    assert rel_l2(got, exact) < BF16_TOL                  # north-star bf16 gate against the unrounded fp32 inputs
    got_cm = attention(q, k, v.transpose(1, 2).contiguous(), scale, "bf16", v_channel_major=True)
    assert torch.equal(got_cm, got)
    ones = torch.ones_like(v)
    assert float((attention(q, k, ones, scale, "bf16") - 1).abs().max()) < 1e-4     # the row sum is taken over the rounded P
    # widths the tensor-core kernel does not take run the fp32 kernel; the raw entry refuses them
    from tmdiff_b200 import _lib
    with pytest.raises(RuntimeError):
        _lib.check(_lib.load().tmd_attention_bf16(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(got), BH, Lq, Lk, 48, Lk, 1.0, _lib.stream_ptr()))


@pytest.mark.gpu
@pytest.mark.parametrize("BH,Lq,Lk,D", [(16, 2048, 2048, 256), (16, 2048, 2048, 128), (8, 1024, 1024, 512), (64, 256, 256, 1024), (64, 1024, 77, 64)])
def test_attention_bf16_bit_stable(BH, Lq, Lk, D):
    """The same launch repeated gives the same bits (logits scaled so that the accumulator is rescaled often): a hazard between
    the TMA producers, the MMA thread and the softmax warpgroups of attn_umma_kernel shows up as run-to-run differences long
    before it exceeds a parity tolerance (profiles/_attn_stress.py is the long version of this test)."""
    from tmdiff_b200.operators import attention_bf16
    g = torch.Generator().manual_seed(3)
    q, k, v = (torch.randn(BH, L, D, generator=g).cuda() for L in (Lq, Lk, Lk))
    k *= 2.0
    first = attention_bf16(q, k, v, D ** -0.5)
    for _ in range(25):
        assert torch.equal(attention_bf16(q, k, v, D ** -0.5), first)


# ---- second batch: BasicTransformerBlock / SpatialTransformer, Dynamic_conv1d/2d, learned-PSF kernels, torch metrics
# (tests/golden/make_golden_operators2.py) ----
@pytest.fixture(scope="module")
def vec2():
    return np.load(os.path.join(ROOT, "tests", "golden", "operator_vectors2.npz"))


def _build2(tag, dtype="fp32"):
    from tmdiff_b200 import operators as O
    return {"btb": lambda: O.BasicTransformerBlock(64, 4, 16, context_dim=96, checkpoint=False, dtype=dtype),
            "st": lambda: O.SpatialTransformer(64, 4, 16, depth=2, context_dim=[96, 96], use_checkpoint=False, dtype=dtype),
            "dyn2d": lambda: O.Dynamic_conv2d(16, 32, 3, padding=1, K=4, temperature=34, dtype=dtype),
            "dyn1d": lambda: O.Dynamic_conv1d(16, 32, 3, padding=1, K=4, temperature=34, dtype=dtype),
            "psf_entire": lambda: O.EntireKernel(kernel_size=4, bands=4, scale_factor=4),
            "psf_gauss": lambda: O.GaussianKernel(kernel_size=5, bands=4, scale_factor=4)}[tag]()


@pytest.mark.parametrize("tag", ["btb", "st", "dyn2d", "dyn1d", "psf_entire", "psf_gauss"])
def test_state_dict_keys_match_reference_2(vec2, tag):
    m = _build2(tag)
    ref = _sd(vec2, tag)
    mine = m.state_dict()
    assert set(mine) == set(ref)
    assert all(tuple(mine[k].shape) == tuple(ref[k].shape) for k in ref)


@pytest.mark.gpu
@pytest.mark.parametrize("tag,dtype,tol", [("btb", "fp32", FP32_TOL), ("btb", "bf16", BF16_TOL), ("st", "fp32", FP32_TOL), ("st", "bf16", BF16_TOL),
                                           ("dyn2d", "fp32", FP32_TOL), ("dyn2d", "bf16", BF16_TOL), ("dyn1d", "fp32", FP32_TOL),
                                           ("psf_entire", "fp32", FP32_TOL), ("psf_gauss", "fp32", FP32_TOL)])
def test_operator_against_reference_2(vec2, tag, dtype, tol):
    m = _build2(tag, dtype)
    m.load_state_dict(_sd(vec2, tag), strict=True)
    m = m.cuda()
    x = _inp(vec2, tag, "x")
    if tag == "btb":
        got = m(x, context=_inp(vec2, tag, "context"))
    elif tag == "st":
        got = m(x, context=[_inp(vec2, tag, "context"), _inp(vec2, tag, "context2")])
    else:
        got = m(x)
    ref = torch.from_numpy(vec2[tag + "/out"])
    assert got.shape == ref.shape
    assert rel_l2(got, ref) < tol
    if tag == "psf_entire":
        assert rel_l2(m.psf, torch.from_numpy(vec2["psf_entire/psf"])) < FP32_TOL
    if tag == "psf_gauss":
        assert rel_l2(m.KernelAdaption[0, 0], torch.from_numpy(vec2["psf_gauss/psf"])) < FP32_TOL


@pytest.mark.gpu
def test_layer_norm_and_geglu_against_torch():
    """plain fp32 torch references of the two row kernels, including widths off the vector path and rows that do not fill a block"""
    from tmdiff_b200.operators import geglu, layer_norm
    g = torch.Generator().manual_seed(5)
    for rows, C in [(7, 64), (300, 320), (33, 1024), (5, 1280), (9, 30)]:
        x = (torch.randn(rows, C, generator=g) * 3 + 1).cuda()
        w, b = torch.randn(C, generator=g).cuda(), torch.randn(C, generator=g).cuda()
        ref = torch.nn.functional.layer_norm(x, (C,), w, b, 1e-5)
        assert rel_l2(layer_norm(x, w, b, 1e-5), ref) < FP32_TOL
    assert rel_l2(layer_norm(x), torch.nn.functional.layer_norm(x, (C,))) < FP32_TOL
    for rows, D in [(3, 64), (77, 256), (1000, 1280)]:
        p = torch.randn(rows, 2 * D, generator=g).cuda()
        a, gate = p.chunk(2, dim=-1)
        assert rel_l2(geglu(p), a * torch.nn.functional.gelu(gate)) < FP32_TOL


@pytest.mark.gpu
def test_torch_metrics_match_reference(vec2):
    """QIndex_torch / D_lambda_torch / D_s_torch / SCC_torch / cc on the device against the reference's values (north star: 0.5 %)."""
    from tmdiff_b200 import metrics as M
    t = {k: torch.from_numpy(vec2["metrics/in/" + k]).cuda() for k in ("ps", "lms", "pan", "lpan", "gt")}
    close = lambda got, key: np.testing.assert_allclose(got.cpu().numpy(), vec2["metrics/" + key], rtol=5e-3, atol=1e-5)
    close(M.QIndex_torch(t["ps"][:, 0], t["gt"][:, 0]), "QIndex")
    close(M.D_lambda_torch(t["lms"], t["ps"]), "D_lambda")
    close(M.D_s_torch(t["lms"], t["pan"], t["lpan"], t["ps"]), "D_s")
    close(M.cc(t["ps"], t["gt"]), "cc")
    np.testing.assert_allclose(np.array([float(M.SCC_torch(t["gt"][:, c:c + 1], t["ps"][:, c:c + 1])) for c in range(4)]), vec2["metrics/SCC"], rtol=5e-3)
    np.testing.assert_allclose(float(M.SCC_torch(t["gt"], t["ps"])), float(vec2["metrics/SCC"].mean()), rtol=5e-3)
    # full-size property: a fused image against itself is perfect (Q = 1, SCC = 1, cc = 1, D_lambda of identical inputs = 0)
    big = torch.rand(8, 8, 256, 256, generator=torch.Generator().manual_seed(1)).cuda()
    assert abs(float(M.SCC_torch(big, big)) - 1) < 1e-5 and float((M.cc(big, big) - 1).abs().max()) < 1e-5
    assert float(M.D_lambda_torch(big, big)) < 1e-6
    q = M.q_index_matrix(big, big)
    assert float((q.diagonal(dim1=1, dim2=2) - 1).abs().max()) < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(5, 32, 64, 4, 16, 16, 3), (3, 64, 32, 2, 8, 8, 3), (4, 64, 128, 1, 32, 16, 1), (2, 32, 32, 8, 16, 24, 3)])
def test_dynamic_conv3d_one_launch_tensor_path(lib_built, shape):
    """Dynamic_conv3d (core/dynamic_conv.py:244-268: per-sample expert mix, groups=B conv) on the tcgen05 kernel with per-sample
    weight rows -- REUSE and SIMPLE schedules, 1x1x1 and 3x3x3, with the mixed bias -- against the fp32 CUDA-core mode of the same
    layer and a torch restatement of the reference formulation."""
    import torch.nn.functional as F
    from tmdiff_b200 import operators as O
    B, Ci, Co, N, H, W, k = shape
    g = torch.Generator().manual_seed(sum(shape))
    layers = {}
    for dtype in ("fp32", "bf16"):
        m = O.Dynamic_conv3d(Ci, Co, k, padding=k // 2, K=4, temperature=34, dtype=dtype)
        if dtype == "fp32":
            m.weight.data = torch.randn(m.weight.shape, generator=g) / (Ci * k ** 3) ** 0.5
            m.bias.data = torch.randn(m.bias.shape, generator=g) * 0.3
            m.attention.fc1.weight.data = torch.randn(m.attention.fc1.weight.shape, generator=g) * 2.0
            m.attention.fc2.weight.data = torch.randn(m.attention.fc2.weight.shape, generator=g) * 30.0      # distinct mixes per sample
        else:
            m.load_state_dict(layers["fp32"].state_dict())
        layers[dtype] = m.cuda()
    x = torch.randn(B, Ci, N, H, W, generator=g)
    y32, y16 = layers["fp32"](x.cuda()).cpu(), layers["bf16"](x.cuda()).cpu()
    m = layers["fp32"].cpu()
    pooled = x.mean(dim=(2, 3, 4))
    hid = torch.relu(pooled @ m.attention.fc1.weight.reshape(-1, Ci).t())
    alpha = torch.softmax((hid @ m.attention.fc2.weight.reshape(4, -1).t()) / 34.0, dim=1)
    assert float((alpha[0] - alpha[1]).abs().max()) > 1e-3                       # the samples really get different weights
    wmix = (alpha @ m.weight.reshape(4, -1)).reshape(B * Co, Ci, k, k, k)
    bmix = (alpha @ m.bias).reshape(-1)
    want = F.conv3d(x.reshape(1, B * Ci, N, H, W), wmix, bmix, padding=k // 2, groups=B).reshape(B, Co, N, H, W)
    assert rel_l2(y32, want) < 1e-5
    assert rel_l2(y16, want) < 2e-2
    per = ((y16 - want).flatten(1).norm(dim=1) / want.flatten(1).norm(dim=1)).max().item()
    assert per < 2e-2, per                                                       # every sample (a wrong weight row would show here)


@pytest.mark.gpu
@pytest.mark.parametrize("M,Ci,Co", [(4096, 256, 768), (77 * 3, 768, 512), (1000, 320, 320), (8, 64, 16), (262144, 64, 64)])
@pytest.mark.parametrize("dtype,tol", [("fp32", 1e-5), ("bf16", 1e-2)])
def test_linear_tokens(lib_built, M, Ci, Co, dtype, tol):
    """tmd_linear (nn.Linear / 1x1 convolution over token-major activations; the projections of core/Attention.py:131-134,
    176-183, 213) against F.linear: wide outputs run as channel groups of 256, an odd token count takes the CUDA-core kernel,
    bf16 operands go in and out without a cast pass."""
    from tmdiff_b200 import operators as O
    g = torch.Generator(device="cuda").manual_seed(M + Ci)
    x = torch.randn(3, M // 3, Ci, generator=g, device="cuda") if M % 3 == 0 else torch.randn(M, Ci, generator=g, device="cuda")
    w = torch.randn(Co, Ci, generator=g, device="cuda") / Ci ** 0.5
    b = torch.randn(Co, generator=g, device="cuda")
    want = torch.nn.functional.linear(x.double(), w.double(), b.double()).float()
    got = O._linear_tokens(x, w, b, dtype)
    assert got.shape == want.shape and got.dtype == torch.float32
    assert rel_l2(got, want) < tol
    assert rel_l2(O._linear_tokens(x, w, None, dtype), want - b) < tol
    if dtype == "bf16":
        got16 = O._linear_tokens(x.bfloat16(), w, b, dtype, out_bf16=True)
        assert got16.dtype == torch.bfloat16 and rel_l2(got16.float(), want) < tol


@pytest.mark.gpu
@pytest.mark.parametrize("B,h,Lq,Lk,d", [(3, 8, 300, 77, 64), (2, 4, 256, 333, 128), (2, 1, 200, 200, 256), (1, 2, 130, 1024, 512)])
@pytest.mark.parametrize("out_bf16", [False, True])
def test_attention_bf16_heads_in_place_layout(lib_built, B, h, Lq, Lk, d, out_bf16):
    """tmd_attention_bf16_strided: heads side by side in the rows of q / k / out (the projections' own layout,
    core/Attention.py:189-191, 211) against softmax(q k^T * scale) v per head in fp64 on the same bf16-rounded operands, and
    against the split-heads entry point."""
    from tmdiff_b200 import operators as O
    g = torch.Generator(device="cuda").manual_seed(B * 100 + h)
    q, k, v = (torch.randn(B, L, h * d, generator=g, device="cuda").bfloat16() for L in (Lq, Lk, Lk))
    scale = d ** -0.5
    got = O.attention_bf16_heads(q, k, v, h, scale, out_bf16=out_bf16)
    assert got.shape == (B, Lq, h * d) and got.dtype == (torch.bfloat16 if out_bf16 else torch.float32)
    sp = lambda t: t.double().reshape(B, t.shape[1], h, d).permute(0, 2, 1, 3)
    want = (torch.softmax(sp(q) @ sp(k).transpose(-1, -2) * scale, -1) @ sp(v)).permute(0, 2, 1, 3).reshape(B, Lq, h * d).float()
    assert rel_l2(got.float(), want) < (8e-3 if out_bf16 else 5e-3)
    split = lambda t: t.reshape(B, t.shape[1], h, d).permute(0, 2, 1, 3).reshape(B * h, t.shape[1], d).contiguous()
    old = O.attention_bf16(split(q), split(k), split(v), scale).view(B, h, Lq, d).permute(0, 2, 1, 3).reshape(B, Lq, h * d)
    if not out_bf16:
        assert torch.equal(got, old)          # same kernel, same operands: bit-identical
    # v read in place (MN-major operand) or copied into the channel-major layout first: the same products in the same order
    assert torch.equal(got, O.attention_bf16_heads(q, k, v, h, scale, out_bf16=out_bf16, v_in_place=False))
