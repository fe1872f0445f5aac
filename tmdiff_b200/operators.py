"""The operators the north star names that the reference's sampling path never executes, with
the reference's class names, constructor arguments and state-dict keys, backed by the CUDA library:

    AdaptiveGroupNorm       GeneralModel/Hyper_unet_general.py:311-331  (+ optional fused Swish)
    AttnBlockpp, NIN        GeneralModel/Hyper_unet_general.py:471-515
    SpatialSelfAttention    core/Attention.py:112-162
    CrossAttention          core/Attention.py:165-214
    Dynamic_conv1d/2d/3d    core/dynamic_conv.py:45-104, 140-190, 221-268
    GEGLU, FeedForward, BasicTransformerBlock, SpatialTransformer     core/Attention.py:69-96, 266-362
    EntireKernel, GaussianKernel                                      core/kernelAdaption.py:117-180

Normalisation, the softmax-attention core and the expert mixing are hand-written kernels (csrc/operators.cu); the
1x1 projections run through the library's convolution operator (tcgen05 in bf16 mode, fp32 CUDA cores otherwise).
Forward only, CUDA tensors only.  torch is used for parameter storage and layout views.
"""
import math

import torch
import torch.nn as nn

from . import _lib
from .ops import _cuda_f32, modulated_conv3d


def group_norm(x, num_groups, weight=None, bias=None, eps=1e-6, swish=False):
    """F.group_norm(x, G, weight, bias, eps) [+ Swish]; weight/bias [C] shared or [B, C] per sample."""
    x = _cuda_f32(x, "x")
    B, Cc = x.shape[0], x.shape[1]
    S = x[0, 0].numel()
    per_sample = weight is not None and weight.dim() == 2
    w = None if weight is None else _cuda_f32(weight, "weight")
    b = None if bias is None else _cuda_f32(bias, "bias")
    y = torch.empty_like(x)
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().tmd_groupnorm(_lib.ptr(x), _lib.ptr(y), _lib.ptr(w), _lib.ptr(b), int(per_sample), B, Cc, S,
                                             int(num_groups), float(eps), int(bool(swish)), _lib.stream_ptr()))
    return y


def attention(q, k, v, scale, dtype="fp32", v_channel_major=False):
    """softmax(scale * q k^T) v for q [BH, Lq, D], k/v [BH, Lk, D] in one fused pass (v [BH, D, Lk] when v_channel_major).

    dtype "fp32": the CUDA-core kernel (1e-5 parity mode).  dtype "bf16": the tcgen05 kernel (bf16 operands, fp32 logits and
    accumulation) whenever the head width is a multiple of 64; other widths run the fp32 kernel."""
    if dtype == "bf16" and q.shape[-1] % 64 == 0:
        return attention_bf16(q, k, v, scale, v_channel_major)
    if v_channel_major:
        v = v.transpose(1, 2)
    q, k, v = _cuda_f32(q, "q"), _cuda_f32(k, "k"), _cuda_f32(v, "v")
    BH, Lq, D = q.shape
    if k.shape[0] != BH or k.shape[2] != D or v.shape != k.shape:
        raise ValueError("attention: q [BH,Lq,D], k and v [BH,Lk,D] expected")
    out = torch.empty_like(q)
    with torch.cuda.device(q.device):
        _lib.check(_lib.load().tmd_attention(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), BH, Lq, k.shape[1], D,
                                             float(scale), _lib.stream_ptr()))
    return out


def attention_bf16(q, k, v, scale, v_channel_major=False, out_bf16=False):
    """The tensor-core attention core (tmd_attention_bf16).  q [BH, Lq, D], k [BH, Lk, D]; v [BH, Lk, D], or [BH, D, Lk] when
    v_channel_major (the layout the reference's 1x1 convs produce, which is the one the kernel reads).  Returns fp32, or bf16 when
    `out_bf16` (for a consumer that takes bf16: the output projection)."""
    if not (q.is_cuda and k.is_cuda and v.is_cuda):
        raise RuntimeError("attention_bf16: CUDA tensors required (no CPU fallback)")
    BH, Lq, D = q.shape
    Lk = k.shape[1]
    vt = v if v_channel_major else v.transpose(1, 2)
    if k.shape[0] != BH or k.shape[2] != D or tuple(vt.shape) != (BH, D, Lk):
        raise ValueError("attention: q [BH,Lq,D], k and v [BH,Lk,D] expected")
    pitch = (Lk + 7) // 8 * 8
    if pitch != Lk:
        vt = torch.nn.functional.pad(vt, (0, pitch - Lk))
    # one pass per operand: a strided (transposed / sliced) fp32 view is cast and laid out by the same copy
    qb, kb, vb = (t if t.dtype == torch.bfloat16 and t.is_contiguous() else torch.empty(t.shape, dtype=torch.bfloat16, device=t.device).copy_(t)
                  for t in (q, k, vt))
    out = torch.empty(BH, Lq, D, device=q.device, dtype=torch.bfloat16 if out_bf16 else torch.float32)
    with torch.cuda.device(q.device):
        if out_bf16:
            _lib.check(_lib.load().tmd_attention_bf16_strided(_lib.ptr(qb), _lib.ptr(kb), _lib.ptr(vb), _lib.ptr(out), BH, 1, Lq, Lk, D, D, D,
                                                              pitch, D, 1, float(scale), _lib.stream_ptr()))
        else:
            _lib.check(_lib.load().tmd_attention_bf16(_lib.ptr(qb), _lib.ptr(kb), _lib.ptr(vb), _lib.ptr(out), BH, Lq, Lk, D, pitch,
                                                      float(scale), _lib.stream_ptr()))
    return out


def attention_bf16_heads(q, k, v, heads, scale, out_bf16=True, v_in_place=True):
    """Multi-head attention on the projections' own layout: q [B, Lq, heads*d], k and v [B, Lk, heads*d] bf16, head h in columns
    [h*d, (h+1)*d) -- no `(b h) n d` split and no merge of the result (core/Attention.py:189-191, 211).  v_in_place: v is read
    as it is too (tmd_attention_bf16_tokens: the MN-major B operand of P V); otherwise it is copied into the channel-major
    per-head layout first (tmd_attention_bf16_strided).  Returns [B, Lq, heads*d], bf16 or fp32."""
    if not (q.is_cuda and k.is_cuda and v.is_cuda):
        raise RuntimeError("attention_bf16_heads: CUDA tensors required (no CPU fallback)")
    B, Lq, inner = q.shape
    Lk, d = k.shape[1], inner // heads
    if inner % heads or d % 64 or k.shape != (B, Lk, inner) or v.shape != k.shape:
        raise ValueError("attention_bf16_heads: q [B,Lq,h*d], k and v [B,Lk,h*d] with d a multiple of 64 expected")
    qb, kb = (t if t.dtype == torch.bfloat16 and t.is_contiguous() else torch.empty(t.shape, dtype=torch.bfloat16, device=t.device).copy_(t)
              for t in (q, k))
    if v_in_place:
        vb = v if v.dtype == torch.bfloat16 and v.is_contiguous() else torch.empty(v.shape, dtype=torch.bfloat16, device=v.device).copy_(v)
        out = torch.empty(B, Lq, inner, device=q.device, dtype=torch.bfloat16 if out_bf16 else torch.float32)
        with torch.cuda.device(q.device):
            _lib.check(_lib.load().tmd_attention_bf16_tokens(_lib.ptr(qb), _lib.ptr(kb), _lib.ptr(vb), _lib.ptr(out), B, heads, Lq, Lk, d, inner,
                                                             inner, inner, inner, int(bool(out_bf16)), float(scale), _lib.stream_ptr()))
        return out
    pitch = (Lk + 7) // 8 * 8
    vt = torch.zeros(B, heads, d, pitch, dtype=torch.bfloat16, device=q.device) if pitch != Lk else \
        torch.empty(B, heads, d, pitch, dtype=torch.bfloat16, device=q.device)
    vt[..., :Lk].copy_(v.reshape(B, Lk, heads, d).permute(0, 2, 3, 1))
    out = torch.empty(B, Lq, inner, device=q.device, dtype=torch.bfloat16 if out_bf16 else torch.float32)
    with torch.cuda.device(q.device):
        _lib.check(_lib.load().tmd_attention_bf16_strided(_lib.ptr(qb), _lib.ptr(kb), _lib.ptr(vt), _lib.ptr(out), B, heads, Lq, Lk, d, inner,
                                                          inner, pitch, inner, int(bool(out_bf16)), float(scale), _lib.stream_ptr()))
    return out


def _conv1x1(x, weight, bias, dtype):
    """1x1 conv of x [B, C, *spatial] with weight [Co, Ci] through the library's conv operator."""
    B, Ci = x.shape[:2]
    sp = x.shape[2:]
    n = int(torch.tensor(sp).prod())
    w2 = next(d for d in (8, 4, 2, 1) if n % d == 0)
    y = modulated_conv3d(x.reshape(B, Ci, 1, n // w2, w2), weight.reshape(weight.shape[0], Ci, 1, 1, 1), bias=bias, dtype=dtype)
    return y.reshape(B, weight.shape[0], *sp)


def _linear_tokens(x, weight, bias, dtype, out_bf16=False):
    """nn.Linear over the last dim of TOKEN-MAJOR x [..., Cin] (tmd_linear: the rows are the voxels of the channels-last conv
    kernels, nothing is re-laid-out).  x fp32 or bf16; returns [..., Co] fp32, or bf16 when `out_bf16` in bf16 mode."""
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor (tmdiff_b200 has no CPU path)")
    code = {"fp32": _lib.F32, "bf16": _lib.BF16}[dtype]
    x_bf16 = code == _lib.BF16 and x.dtype == torch.bfloat16
    x = x.contiguous() if x_bf16 else x.contiguous().float()
    Cin, Co = x.shape[-1], weight.shape[0]
    M = x.numel() // Cin
    w = _cuda_f32(weight, "weight").reshape(Co, Cin)
    b = None if bias is None else _cuda_f32(bias, "bias")
    y_bf16 = bool(out_bf16) and code == _lib.BF16
    y = torch.empty(*x.shape[:-1], Co, device=x.device, dtype=torch.bfloat16 if y_bf16 else torch.float32)
    if M == 0:
        return y
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().tmd_linear(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(y), M, Cin, Co, int(x_bf16), int(y_bf16), code,
                                          _lib.stream_ptr()))
    return y


def _relayout(t, shape, dtype):
    """`t` (any strides) as a contiguous tensor of `dtype`: layout change and cast in ONE copy."""
    return torch.empty(shape, dtype=dtype, device=t.device).copy_(t)


def _token_self_attention(h, qkv, out_w, out_b, scale, dtype):
    """Single-head self-attention over the HW tokens of a normalised feature map h [B, C, HW] (core/Attention.py:140-159,
    Hyper_unet_general.py:494-508): q / k / v / output projections as token-major GEMMs (tmd_linear), the attention core, and the
    result back as [B, C, HW] (a view).  bf16 mode: the operands stay bf16 between the kernels; what is left in torch is the
    layout change of h into tokens, of v into the core's channel-major operand and of the result back (three copies)."""
    B, C, L = h.shape
    bf = dtype == "bf16" and C % 64 == 0
    tdt = torch.bfloat16 if bf else torch.float32
    ht = _relayout(h.transpose(1, 2), (B, L, C), tdt)
    q, k, v = (_linear_tokens(ht, w, b, dtype, out_bf16=bf) for (w, b) in qkv)
    if bf:          # q, k, v as the projections wrote them; the result bf16 for the output projection
        o = attention_bf16_heads(q, k, v, 1, scale, out_bf16=True)
    else:
        o = attention(q, k, v, scale, "fp32")
    return _linear_tokens(o, out_w, out_b, dtype).transpose(1, 2)


class AdaptiveGroupNorm(nn.Module):
    def __init__(self, num_groups, in_channel, style_dim):
        super().__init__()
        self.num_groups, self.in_channel = num_groups, in_channel
        self.style = nn.Linear(style_dim, in_channel * 2)
        nn.init.zeros_(self.style.bias)
        self.style.bias.data[:in_channel] = 1
        self.style.bias.data[in_channel:] = 0

    @torch.no_grad()
    def forward(self, input, style, swish=False):
        """input [B, C, N, H, W], style [B, style_dim]; swish=True fuses the activation that follows the norm."""
        gb = _linear_tokens(style[None].float(), self.style.weight, self.style.bias, "fp32")[0]     # [B, 2C]
        gamma, beta = gb[:, :self.in_channel].contiguous(), gb[:, self.in_channel:].contiguous()
        return group_norm(input, self.num_groups, gamma, beta, eps=1e-6, swish=swish)


class NIN(nn.Module):
    def __init__(self, in_dim, num_units, init_scale=0.1):
        super().__init__()
        bound = math.sqrt(3.0 * max(init_scale, 1e-10) / ((in_dim + num_units) / 2.0))
        self.W = nn.Parameter(torch.empty(in_dim, num_units).uniform_(-bound, bound), requires_grad=False)
        self.b = nn.Parameter(torch.zeros(num_units), requires_grad=False)


class AttnBlockpp(nn.Module):
    def __init__(self, channels, skip_rescale=True, init_scale=0.0, dtype="fp32"):
        super().__init__()
        self.channels, self.skip_rescale, self.compute_dtype = channels, skip_rescale, dtype
        self.GroupNorm_0 = nn.GroupNorm(num_groups=min(channels // 4, 32), num_channels=channels, eps=1e-6)
        self.NIN_0, self.NIN_1, self.NIN_2 = NIN(channels, channels), NIN(channels, channels), NIN(channels, channels)
        self.NIN_3 = NIN(channels, channels, init_scale=init_scale)

    @torch.no_grad()
    def forward(self, x):
        """x [B, C, N, H, W] with C*N == channels; the logits are scaled by C^-0.5 (the feature C, :500)."""
        B, Cf, N, H, W = x.shape
        x2 = x.reshape(B, Cf * N, H, W)
        h = group_norm(x2, self.GroupNorm_0.num_groups, self.GroupNorm_0.weight, self.GroupNorm_0.bias, eps=1e-6)
        Cc = Cf * N
        nin = lambda n: (n.W.t(), n.b)
        o = _token_self_attention(h.reshape(B, Cc, H * W), [nin(self.NIN_0), nin(self.NIN_1), nin(self.NIN_2)], self.NIN_3.W.t(), self.NIN_3.b,
                                  int(Cf) ** (-0.5), self.compute_dtype).reshape(B, Cc, H, W)
        out = (x2 + o).reshape(B, Cf, N, H, W)
        return out / math.sqrt(2.0) if self.skip_rescale else out


class SpatialSelfAttention(nn.Module):
    def __init__(self, in_channels, dtype="fp32"):
        super().__init__()
        self.in_channels, self.compute_dtype = in_channels, dtype
        self.norm = nn.GroupNorm(num_groups=32, num_channels=in_channels, eps=1e-6, affine=True)
        self.q, self.k, self.v, self.proj_out = (nn.Conv2d(in_channels, in_channels, 1) for _ in range(4))

    @torch.no_grad()
    def forward(self, x):
        B, Cc, H, W = x.shape
        h = group_norm(x, 32, self.norm.weight, self.norm.bias, eps=1e-6)
        conv = lambda m: (m.weight.reshape(Cc, Cc), m.bias)
        o = _token_self_attention(h.reshape(B, Cc, H * W), [conv(self.q), conv(self.k), conv(self.v)], self.proj_out.weight.reshape(Cc, Cc),
                                  self.proj_out.bias, int(Cc) ** (-0.5), self.compute_dtype)
        return x + o.reshape(B, Cc, H, W)


class CrossAttention(nn.Module):
    def __init__(self, query_dim, context_dim=None, heads=8, dim_head=64, dropout=0.0, dtype="fp32"):
        super().__init__()
        inner = dim_head * heads
        context_dim = query_dim if context_dim is None else context_dim
        self.scale, self.heads, self.compute_dtype = dim_head ** -0.5, heads, dtype
        self.to_q = nn.Linear(query_dim, inner, bias=False)
        self.to_k = nn.Linear(context_dim, inner, bias=False)
        self.to_v = nn.Linear(context_dim, inner, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner, query_dim), nn.Dropout(dropout))

    @torch.no_grad()
    def forward(self, x, context=None, mask=None):
        if mask is not None:
            raise NotImplementedError("attention masks are not used anywhere in the reference")
        context = x if context is None else context
        B, L, _ = x.shape
        h = self.heads
        d = self.to_q.weight.shape[0] // h
        bf = self.compute_dtype == "bf16" and d % 64 == 0
        tdt = torch.bfloat16 if bf else torch.float32
        q = _linear_tokens(x, self.to_q.weight, None, self.compute_dtype, out_bf16=bf)            # [B, L, h*d], bf16 in bf16 mode
        k = _linear_tokens(context, self.to_k.weight, None, self.compute_dtype, out_bf16=bf)
        v = _linear_tokens(context, self.to_v.weight, None, self.compute_dtype, out_bf16=bf)
        if bf:          # the heads stay side by side in the rows: no split / merge copies
            o = attention_bf16_heads(q, k, v, h, self.scale, out_bf16=True)
        else:
            split = lambda t: _relayout(t.reshape(B, t.shape[1], h, d).permute(0, 2, 1, 3), (B, h, t.shape[1], d), tdt).view(B * h, t.shape[1], d)
            o = attention(split(q), split(k), split(v), self.scale, self.compute_dtype)          # logits always in fp32 (:194-197)
            o = _relayout(o.view(B, h, L, d).permute(0, 2, 1, 3), (B, L, h, d), tdt).view(B, L, h * d)
        return _linear_tokens(o, self.to_out[0].weight, self.to_out[0].bias, self.compute_dtype)


class _Attention3d(nn.Module):
    def __init__(self, in_planes, ratios, K, temperature):
        super().__init__()
        assert temperature % 3 == 1
        hidden = int(in_planes * ratios) + 1 if in_planes != 3 else K
        self.fc1 = nn.Conv3d(in_planes, hidden, 1, bias=False)
        self.fc2 = nn.Conv3d(hidden, K, 1, bias=False)
        self.temperature = temperature


def _dynconv_workspace(module, device, B, Ci, Co, N, H, W, k, K, code):
    """The layer's scratch block (tmd_dynamic_conv_workspace_bytes), kept on the module and grown on demand: no allocation
    inside the call."""
    need = int(_lib.load().tmd_dynamic_conv_workspace_bytes(B, Ci, Co, N, H, W, k, K, code))
    if need < 0:
        raise ValueError("Dynamic_conv: unsupported problem")
    ws = getattr(module, "_workspace", None)
    if ws is None or ws.numel() < need or ws.device != device:
        ws = torch.empty(need, dtype=torch.uint8, device=device)
        object.__setattr__(module, "_workspace", ws)      # scratch, not a buffer of the state dict
    return ws


class Dynamic_conv3d(nn.Module):
    def __init__(self, in_planes, out_planes, kernel_size, ratio=0.25, stride=1, padding=0, dilation=1, groups=1, bias=True, K=4,
                 temperature=34, dtype="fp32"):
        super().__init__()
        assert in_planes % groups == 0
        if groups != 1 or stride != 1 or dilation != 1 or kernel_size not in (1, 3) or padding != kernel_size // 2:
            raise ValueError("Dynamic_conv3d: groups=1, stride=1, dilation=1 and 'same' padding with k in {1,3} are implemented")
        self.in_planes, self.out_planes, self.kernel_size, self.K, self.compute_dtype = in_planes, out_planes, kernel_size, K, dtype
        self.attention = _Attention3d(in_planes, ratio, K, temperature)
        self.weight = nn.Parameter(torch.randn(K, out_planes, in_planes // groups, kernel_size, kernel_size, kernel_size),
                                   requires_grad=False)
        self.bias = nn.Parameter(torch.zeros(K, out_planes), requires_grad=False) if bias else None

    @torch.no_grad()
    def forward(self, x):
        x = _cuda_f32(x, "x")
        B, Ci, N, H, W = x.shape
        hidden = self.attention.fc1.weight.shape[0]
        y = torch.empty(B, self.out_planes, N, H, W, device=x.device)
        w = _cuda_f32(self.weight, "weight")
        b = None if self.bias is None else _cuda_f32(self.bias, "bias")
        f1 = _cuda_f32(self.attention.fc1.weight, "fc1").reshape(hidden, Ci)
        f2 = _cuda_f32(self.attention.fc2.weight, "fc2").reshape(self.K, hidden)
        code = _lib.F32 if self.compute_dtype == "fp32" else _lib.BF16
        ws = _dynconv_workspace(self, x.device, B, Ci, self.out_planes, N, H, W, self.kernel_size, self.K, code)
        with torch.cuda.device(x.device):
            _lib.check(_lib.load().tmd_dynamic_conv_ws(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(f1), _lib.ptr(f2), None, _lib.ptr(y),
                                                       B, Ci, self.out_planes, N, H, W, self.kernel_size, 3, self.K, hidden,
                                                       float(self.attention.temperature), code, _lib.ptr(ws), ws.numel(), _lib.stream_ptr()))
        return y


class _DynamicConvNd(nn.Module):
    """Dynamic_conv1d / Dynamic_conv2d (core/dynamic_conv.py:45-104, 140-190): same expert mixing as the 3-D layer; their
    attention carries a bias on fc2 (:18, :108)."""
    ndim = 2

    def __init__(self, in_planes, out_planes, kernel_size, ratio=0.25, stride=1, padding=0, dilation=1, groups=1, bias=True, K=4,
                 temperature=34, dtype="fp32"):
        super().__init__()
        assert in_planes % groups == 0 and temperature % 3 == 1
        if groups != 1 or stride != 1 or dilation != 1 or kernel_size not in (1, 3) or padding != kernel_size // 2:
            raise ValueError("Dynamic_conv: groups=1, stride=1, dilation=1 and 'same' padding with k in {1,3} are implemented")
        conv = nn.Conv1d if self.ndim == 1 else nn.Conv2d
        hidden = int(in_planes * ratio) + 1 if in_planes != 3 else K
        self.in_planes, self.out_planes, self.kernel_size, self.K, self.compute_dtype = in_planes, out_planes, kernel_size, K, dtype
        self.attention = nn.Module()
        self.attention.fc1 = conv(in_planes, hidden, 1, bias=False)
        self.attention.fc2 = conv(hidden, K, 1, bias=True)
        self.attention.temperature = temperature
        self.weight = nn.Parameter(torch.randn(K, out_planes, in_planes // groups, *([kernel_size] * self.ndim)), requires_grad=False)
        self.bias = nn.Parameter(torch.zeros(K, out_planes), requires_grad=False) if bias else None

    @torch.no_grad()
    def forward(self, x):
        x = _cuda_f32(x, "x")
        B, Ci = x.shape[:2]
        sp = [1] * (3 - self.ndim) + list(x.shape[2:])
        hidden = self.attention.fc1.weight.shape[0]
        y = torch.empty(B, self.out_planes, *x.shape[2:], device=x.device)
        w = _cuda_f32(self.weight, "weight")
        b = None if self.bias is None else _cuda_f32(self.bias, "bias")
        f1 = _cuda_f32(self.attention.fc1.weight, "fc1").reshape(hidden, Ci)
        f2 = _cuda_f32(self.attention.fc2.weight, "fc2").reshape(self.K, hidden)
        f2b = _cuda_f32(self.attention.fc2.bias, "fc2.bias")
        code = _lib.F32 if self.compute_dtype == "fp32" else _lib.BF16
        ws = _dynconv_workspace(self, x.device, B, Ci, self.out_planes, sp[0], sp[1], sp[2], self.kernel_size, self.K, code)
        with torch.cuda.device(x.device):
            _lib.check(_lib.load().tmd_dynamic_conv_ws(_lib.ptr(x), _lib.ptr(w), _lib.ptr(b), _lib.ptr(f1), _lib.ptr(f2), _lib.ptr(f2b), _lib.ptr(y),
                                                       B, Ci, self.out_planes, sp[0], sp[1], sp[2], self.kernel_size, self.ndim, self.K, hidden,
                                                       float(self.attention.temperature), code, _lib.ptr(ws), ws.numel(), _lib.stream_ptr()))
        return y


class Dynamic_conv1d(_DynamicConvNd):
    ndim = 1


class Dynamic_conv2d(_DynamicConvNd):
    ndim = 2


def layer_norm(x, weight=None, bias=None, eps=1e-5):
    """F.layer_norm over the last dim of x [..., C]."""
    x = _cuda_f32(x, "x")
    Cc = x.shape[-1]
    y = torch.empty_like(x)
    w = None if weight is None else _cuda_f32(weight, "weight")
    b = None if bias is None else _cuda_f32(bias, "bias")
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().tmd_layernorm(_lib.ptr(x), _lib.ptr(y), _lib.ptr(w), _lib.ptr(b), x.numel() // Cc, Cc, float(eps), _lib.stream_ptr()))
    return y


def geglu(proj):
    """proj [..., 2D] -> proj[..., :D] * gelu(proj[..., D:])."""
    proj = _cuda_f32(proj, "proj")
    D = proj.shape[-1] // 2
    out = torch.empty(*proj.shape[:-1], D, device=proj.device)
    with torch.cuda.device(proj.device):
        _lib.check(_lib.load().tmd_geglu(_lib.ptr(proj), _lib.ptr(out), proj.numel() // (2 * D), D, _lib.stream_ptr()))
    return out


class GEGLU(nn.Module):
    def __init__(self, dim_in, dim_out, dtype="fp32"):
        super().__init__()
        self.proj, self.compute_dtype = nn.Linear(dim_in, dim_out * 2), dtype

    @torch.no_grad()
    def forward(self, x):
        return geglu(_linear_tokens(x, self.proj.weight, self.proj.bias, self.compute_dtype))


class FeedForward(nn.Module):
    def __init__(self, dim, dim_out=None, mult=4, glu=False, dropout=0.0, dtype="fp32"):
        super().__init__()
        if not glu:
            raise NotImplementedError("the reference only builds the gated feed-forward (BasicTransformerBlock gated_ff=True)")
        inner = int(dim * mult)
        self.compute_dtype = dtype
        self.net = nn.Sequential(GEGLU(dim, inner, dtype), nn.Dropout(dropout), nn.Linear(inner, dim if dim_out is None else dim_out))

    @torch.no_grad()
    def forward(self, x):
        return _linear_tokens(self.net[0](x), self.net[2].weight, self.net[2].bias, self.compute_dtype)


class BasicTransformerBlock(nn.Module):
    def __init__(self, dim, n_heads, d_head, dropout=0.0, context_dim=None, gated_ff=True, checkpoint=True, disable_self_attn=False,
                 dtype="fp32"):
        super().__init__()
        self.disable_self_attn = disable_self_attn
        self.attn1 = CrossAttention(query_dim=dim, heads=n_heads, dim_head=d_head, dropout=dropout,
                                    context_dim=context_dim if disable_self_attn else None, dtype=dtype)
        self.ff = FeedForward(dim, dropout=dropout, glu=gated_ff, dtype=dtype)
        self.attn2 = CrossAttention(query_dim=dim, context_dim=context_dim, heads=n_heads, dim_head=d_head, dropout=dropout, dtype=dtype)
        self.norm1, self.norm2, self.norm3 = nn.LayerNorm(dim), nn.LayerNorm(dim), nn.LayerNorm(dim)

    @torch.no_grad()
    def forward(self, x, context=None):
        ln = lambda n, t: layer_norm(t, n.weight, n.bias, n.eps)
        x = self.attn1(ln(self.norm1, x), context=context if self.disable_self_attn else None) + x
        x = self.attn2(ln(self.norm2, x), context=context) + x
        return self.ff(ln(self.norm3, x)) + x


class SpatialTransformer(nn.Module):
    def __init__(self, in_channels, n_heads, d_head, depth=1, dropout=0.0, context_dim=None, disable_self_attn=False, use_linear=False,
                 use_checkpoint=True, dtype="fp32"):
        super().__init__()
        if context_dim is not None and not isinstance(context_dim, list):
            context_dim = [context_dim]
        if context_dim is None:
            context_dim = [None] * depth
        if use_linear:
            raise NotImplementedError("use_linear=True: the reference's proj_out has mismatched dimensions there (core/Attention.py:340)")
        inner = n_heads * d_head
        self.in_channels, self.compute_dtype = in_channels, dtype
        self.norm = nn.GroupNorm(num_groups=32, num_channels=in_channels, eps=1e-6, affine=True)
        self.proj_in = nn.Conv2d(in_channels, inner, 1)
        self.transformer_blocks = nn.ModuleList([BasicTransformerBlock(inner, n_heads, d_head, dropout=dropout, context_dim=context_dim[d],
                                                                       disable_self_attn=disable_self_attn, dtype=dtype) for d in range(depth)])
        self.proj_out = nn.Conv2d(inner, in_channels, 1)
        nn.init.zeros_(self.proj_out.weight)
        nn.init.zeros_(self.proj_out.bias)

    @torch.no_grad()
    def forward(self, x, context=None):
        if not isinstance(context, list):
            context = [context]
        B, Cc, H, W = x.shape
        inner = self.proj_in.weight.shape[0]
        h = group_norm(x, 32, self.norm.weight, self.norm.bias, eps=1e-6)
        h = _conv1x1(h, self.proj_in.weight.reshape(inner, Cc), self.proj_in.bias, self.compute_dtype)
        h = h.reshape(B, inner, H * W).transpose(1, 2).contiguous()
        for i, blk in enumerate(self.transformer_blocks):
            h = blk(h, context=context[i])
        h = h.transpose(1, 2).reshape(B, inner, H, W).contiguous()
        return _conv1x1(h, self.proj_out.weight.reshape(Cc, inner), self.proj_out.bias, self.compute_dtype) + x


class _PSFKernel(nn.Module):
    """Depthwise learned-PSF blur + every scale_factor-th output (core/kernelAdaption.py:117-180) as one kernel: the PSF is built
    from the parameter inside the kernel and only the kept outputs are computed."""
    mode = 0

    def __init__(self, kernel_size, bands, scale_factor):
        super().__init__()
        self.kernel_size, self.bands, self.scale_factor = kernel_size, bands, scale_factor
        self.psf = None

    def _run(self, Z, param, k):
        Z = _cuda_f32(Z, "Z")
        B, Cc, H, W = Z.shape
        s = self.scale_factor
        out = torch.empty(B, Cc, (H - k) // s + 1, (W - k) // s + 1, device=Z.device)
        psf = torch.empty(k, k, device=Z.device)
        p = _cuda_f32(param, "param")
        with torch.cuda.device(Z.device):
            _lib.check(_lib.load().tmd_psf_downsample(_lib.ptr(Z), _lib.ptr(out), _lib.ptr(p), _lib.ptr(psf), self.mode, B, Cc, H, W, k, s,
                                                      _lib.stream_ptr()))
        return out, psf


class EntireKernel(_PSFKernel):
    mode = 0

    def __init__(self, kernel_size, bands, scale_factor):
        super().__init__(kernel_size, bands, scale_factor)
        self.KernelAdaption = nn.Parameter(torch.randn(scale_factor, scale_factor), requires_grad=False)

    @torch.no_grad()
    def forward(self, Z):
        out, self.psf = self._run(Z, self.KernelAdaption, self.scale_factor)      # the PSF is scale_factor x scale_factor (:134)
        return out


class GaussianKernel(_PSFKernel):
    mode = 1

    def __init__(self, kernel_size, bands, scale_factor):
        super().__init__(kernel_size, bands, scale_factor)
        self.KernelParam = nn.Parameter(5 * torch.eye(2, 2), requires_grad=False)

    def re_init(self):
        self.KernelParam = nn.Parameter(5 * torch.eye(2, 2), requires_grad=False)

    @torch.no_grad()
    def forward(self, Z):
        out, psf = self._run(Z, self.KernelParam, self.kernel_size)
        self.KernelAdaption = psf[None, None]
        return out


GaussianKernel_new = GaussianKernel
