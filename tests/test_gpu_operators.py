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
