"""Torch-fp32 CPU restatement of the WavBEST denoiser (TEST INFRASTRUCTURE ONLY).

Follows /root/reference/GeneralModel/Hyper_unet_general.py op for op, but as a
flat functional program over a plain ``dict`` of tensors keyed by the
reference's state-dict names, so it can run where the reference is not mounted
(the GPU box).  Parity pin: ``tests/golden/make_golden.py`` checks it against
the imported reference; ``tests/test_oracle.py`` re-checks the committed vectors.

``faithful=True`` reproduces the reference's formulation exactly (weights scaled
per sample and run as a ``groups=B`` conv, Haar transform as dense matmuls,
nothing cached) and is what the CPU baseline times.  ``faithful=False`` uses
the algebraically identical shared-weight form ``conv3d(x*s, w)`` which is much
faster on CPU for B>1 and is used by the larger parity tests.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import torch
import torch.nn.functional as F

SENSORS = ("QB", "WV3", "GF2", "WV2", "WV4")  # Hyper_unet_general.py:587-598


# ----------------------------------------------------------------------------
# parameter inventory (Hyper_unet_general.py:523-564 and the block ctors)
# ----------------------------------------------------------------------------
def _conv(P, name, co, ci, k):
    P[name + ".weight"] = (co, ci, k, k, k)
    P[name + ".bias"] = (co,)


def _lin(P, name, co, ci):
    P[name + ".weight"] = (co, ci)
    P[name + ".bias"] = (co,)


def _resblock(P, pre, ci, co, E):  # ResBlockModulateBEST :220-235
    _conv(P, pre + ".conv20", co, ci, 3)
    _conv(P, pre + ".conv21", co, co, 3)
    _lin(P, pre + ".dense1.dense", ci, E)
    _lin(P, pre + ".dense2.dense", co, E)
    if ci != co:
        _conv(P, pre + ".res_conv", co, ci, 1)


def _wavelet(P, pre, c, E, hi=None):  # WaveletUPorDown :334-367
    _conv(P, pre + ".Conv_0", c, c, 3)
    _lin(P, pre + ".Dense_0", c, E)
    _conv(P, pre + ".Conv_1", c, c, 3)
    _conv(P, pre + ".Conv_2", c, c, 1)
    if hi is not None:
        P[pre + ".convH_0.0.weight"] = (3 * c, hi, 3, 3, 3)  # groups=3 -> ci/groups = hi
        P[pre + ".convH_0.0.bias"] = (3 * c,)
    _lin(P, pre + ".dense1.dense", c, E)
    _lin(P, pre + ".dense2.dense", c, E)


def param_shapes(channels=(32, 64, 128, 256), embed_dim=128, inter_dim=32) -> "OrderedDict[str, tuple]":
    """Name -> shape of every ``denoise_fn`` tensor, in the reference's registration order."""
    c0, c1, c2, c3 = channels
    E = embed_dim
    P: "OrderedDict[str, tuple]" = OrderedDict()
    _lin(P, "embed.0", E, inter_dim)
    _lin(P, "embed.2", E, E)
    _lin(P, "embed2.0", 4 * E, 768)
    _lin(P, "embed2.2", 4 * E, 4 * E)
    _lin(P, "embed2.4", E, 4 * E)
    for stem in ("conv1", "conv2"):  # AdaptionModulateBEST :158-166
        _conv(P, stem + ".conv20", c0, 1, 1)
        _conv(P, stem + ".conv21", c0, c0, 3)
        _lin(P, stem + ".dense2.dense", c0, E)
    for name, ci, co in (("down1", c0, c1), ("down2", c1, c2), ("down3", c2, c3),
                         ("down1_1", c0, c1), ("down2_1", c1, c2), ("down3_1", c2, c3)):
        _resblock(P, name + ".conv20", ci, co, E)
        _wavelet(P, name + ".down", co, E)
    _resblock(P, "middle1", c3, c3, E)
    for name, ci, co in (("up1", c3, c2), ("up2", c2, c1), ("up3", c1, c0)):
        _wavelet(P, name + ".up1", co, E, hi=ci)
        _resblock(P, name + ".conv20", 3 * ci, co, E)
    for i in range(4):  # FinalBlockModulateBEST :252-263
        _resblock(P, "final.conv2%d" % i, 3 * c0 if i == 0 else c0, c0, E)
    _conv(P, "final.conv24", 1, c0, 1)
    _lin(P, "final.dense2.dense", c0, E)
    return P


def seeded_weights(channels=(32, 64, 128, 256), seed=7, embed_dim=128, inter_dim=32):
    """Seeded fan-in-scaled weights (SURVEY.md 8(c) 'weight fixture'): N(0, 1/fan_in) for
    >=2-D tensors, N(0, 0.1^2) for 1-D, so every branch contributes O(1e-2..1) outputs."""
    g = torch.Generator().manual_seed(seed)
    sd = OrderedDict()
    for name, shape in param_shapes(channels, embed_dim, inter_dim).items():
        if len(shape) >= 2:
            fan_in = 1
            for d in shape[1:]:
                fan_in *= d
            sd[name] = torch.randn(shape, generator=g) / math.sqrt(fan_in)
        else:
            sd[name] = torch.randn(shape, generator=g) * 0.1
    return sd


def seeded_text_embeddings(seed=1234):
    """Stand-ins for the five cached CLIP pooled embeddings ([1,768] each)."""
    g = torch.Generator().manual_seed(seed)
    return {s: torch.randn(1, 768, generator=g) for s in SENSORS}


# ----------------------------------------------------------------------------
# operators
# ----------------------------------------------------------------------------
def swish(x):  # :111-113
    return x * torch.sigmoid(x)


def gamma_embedding(t, dim, max_period=10000):  # :80-97
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32) / half).to(t.device)
    args = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


def haar_matrices(n, dtype=torch.float32):
    """[n/2, n] low / high analysis matrices, as DWT_2D.get_matrix builds them from
    rec_lo=[s,s], rec_hi=[s,-s] (DWT_IDWT_layer.py:270-308)."""
    s = 1.0 / math.sqrt(2.0)
    lo = torch.zeros(n // 2, n, dtype=torch.float64)
    hi = torch.zeros(n // 2, n, dtype=torch.float64)
    for i in range(n // 2):
        lo[i, 2 * i], lo[i, 2 * i + 1] = s, s
        hi[i, 2 * i], hi[i, 2 * i + 1] = s, -s
    return lo.to(dtype), hi.to(dtype)


def dwt2d(x):
    """Haar analysis by the reference's six matmuls (DWT_IDWT_Functions.py:52-57)."""
    H, W = x.shape[-2:]
    # built on the host and moved to the input's device on every call, as DWT_2D.get_matrix does (DWT_IDWT_layer.py:310-319)
    l0, h0 = (m.to(x.device) for m in haar_matrices(H))
    l1, h1 = (m.to(x.device) for m in haar_matrices(W))
    l1, h1 = l1.t(), h1.t()
    L = torch.matmul(l0, x)
    Hh = torch.matmul(h0, x)
    return torch.matmul(L, l1), torch.matmul(L, h1), torch.matmul(Hh, l1), torch.matmul(Hh, h1)


def idwt2d(LL, LH, HL, HH):
    """Haar synthesis (DWT_IDWT_Functions.py:95-100; synthesis filters are the reversed
    dec_lo/dec_hi = the same [s,s], [s,-s] taps, DWT_IDWT_layer.py:352-356)."""
    h, w = LL.shape[-2:]
    l0, h0 = (m.to(LL.device) for m in haar_matrices(2 * h))
    l1, h1 = (m.to(LL.device) for m in haar_matrices(2 * w))
    l1, h1 = l1.t(), h1.t()
    L = torch.matmul(LL, l1.t()) + torch.matmul(LH, h1.t())
    Hh = torch.matmul(HL, l1.t()) + torch.matmul(HH, h1.t())
    return torch.matmul(l0.t(), L) + torch.matmul(h0.t(), Hh)


def dwt2d_closed(x):
    """Closed-form Haar butterflies -- what the CUDA kernels compute."""
    a, b = x[..., 0::2, 0::2], x[..., 0::2, 1::2]
    c, d = x[..., 1::2, 0::2], x[..., 1::2, 1::2]
    return (a + b + c + d) / 2, (a - b + c - d) / 2, (a + b - c - d) / 2, (a - b - c + d) / 2


def idwt2d_closed(LL, LH, HL, HH):
    out = LL.new_empty(*LL.shape[:-2], 2 * LL.shape[-2], 2 * LL.shape[-1])
    out[..., 0::2, 0::2] = (LL + LH + HL + HH) / 2
    out[..., 0::2, 1::2] = (LL - LH + HL - HH) / 2
    out[..., 1::2, 0::2] = (LL + LH - HL - HH) / 2
    out[..., 1::2, 1::2] = (LL - LH - HL + HH) / 2
    return out


def modulated_conv3d(x, w, s, padding, faithful=True):
    """:51-77.  s is [B,Ci,1,1] (Dense output).  No bias, no demodulation."""
    B = x.shape[0]
    co, ci = w.shape[:2]
    if not faithful:
        return F.conv3d(x * s.reshape(B, ci, 1, 1, 1), w, None, padding=padding)
    wb = w.unsqueeze(0) * s.unsqueeze(1).unsqueeze(5)          # [B,O,I,k,k,k]
    xg = x.reshape(1, -1, *x.shape[2:])
    y = F.conv3d(xg, wb.reshape(-1, ci, *w.shape[2:]), None, padding=padding, groups=B)
    return y.reshape(B, -1, *y.shape[2:])


class _Net:
    """Functional view over the state dict."""

    def __init__(self, sd, faithful=True):
        self.sd, self.faithful = sd, faithful

    def conv(self, name, x, padding, groups=1):
        return F.conv3d(x, self.sd[name + ".weight"], self.sd[name + ".bias"], padding=padding, groups=groups)

    def lin(self, name, x):
        return F.linear(x, self.sd[name + ".weight"], self.sd[name + ".bias"])

    def dense(self, name, x):  # Dense :100-108 -> [B,C,1,1]
        return self.lin(name + ".dense", x)[..., None, None]

    def modconv(self, name, x, s, padding):
        return modulated_conv3d(x, self.sd[name + ".weight"], s, padding, self.faithful)

    # --- blocks ---------------------------------------------------------
    def stem(self, pre, x, prompt):  # AdaptionModulateBEST.forward :168-173
        h = swish(self.conv(pre + ".conv20", x, 0))
        return self.modconv(pre + ".conv21", h, self.dense(pre + ".dense2", prompt), 1)

    def resblock(self, pre, x, embed, prompt, flag=False):  # :237-249
        h = x if flag else x + self.dense(pre + ".dense1", embed).unsqueeze(-1)
        h = self.conv(pre + ".conv20", swish(h), 1)
        h = self.modconv(pre + ".conv21", swish(h), self.dense(pre + ".dense2", prompt), 1)
        if (pre + ".res_conv.weight") in self.sd:
            return h + self.conv(pre + ".res_conv", x, 0)
        return h + x

    def wavelet(self, pre, x, temb, zemb, skipH=None, up=False, flag=False):  # :369-414
        B, C, N, H, W = x.shape
        h = self.conv(pre + ".Conv_0", swish(x), 1)
        x = self.conv(pre + ".Conv_2", x, 0)
        h = h.reshape(B, C * N, H, W)
        x = x.reshape(B, C * N, H, W)
        hH = None
        if up:
            D = C * N
            s = self.conv(pre + ".convH_0.0", torch.cat(skipH, dim=1) / 2.0, 1, groups=3) * 2.0
            s = s.reshape(B, 3 * D, H, W)
            h = idwt2d(2.0 * h, s[:, :D], s[:, D:2 * D], s[:, 2 * D:])
            x = idwt2d(2.0 * x, s[:, :D], s[:, D:2 * D], s[:, 2 * D:])
        else:
            h, hLH, hHL, hHH = dwt2d(h)
            x, _, _, _ = dwt2d(x)
            hH = tuple(t.reshape(B, C, N, H // 2, W // 2) for t in (hLH, hHL, hHH))
            h, x = h / 2.0, x / 2.0
        h = h.reshape(B, C, N, *h.shape[-2:])
        x = x.reshape(B, C, N, *x.shape[-2:])
        if not flag:
            h = h + self.lin(pre + ".Dense_0", temb)[:, :, None, None, None]
        h = self.modconv(pre + ".Conv_1", swish(h), self.dense(pre + ".dense1", zemb), 1)
        out = x + h
        return out if up else (out, hH)

    def down(self, pre, x, embed, prompt, flag=False):  # ResblockDownOneModulateBEST :192-196
        h = self.resblock(pre + ".conv20", x, embed, prompt, flag)
        return self.wavelet(pre + ".down", h, embed, prompt, flag=flag)

    def up(self, pre, x, embed, skipH, prompt):  # ResblockUpOneModulateBEST :214-217
        h = self.resblock(pre + ".conv20", x, embed, prompt)
        return self.wavelet(pre + ".up1", h, embed, prompt, skipH=skipH, up=True)

    def final(self, x, embed, prompt):  # FinalBlockModulateBEST :265-273
        h = x
        for i in range(4):
            h = self.resblock("final.conv2%d" % i, h, embed, prompt)
        return self.modconv("final.conv24", swish(h), self.dense("final.dense2", prompt), 0)


def condition_embeddings(sd, t_input, text_emb, B, inter_dim=32):
    """``prompt`` and ``embed`` vectors of WavBEST.forward :601-604."""
    net = _Net(sd)
    t = t_input.reshape(-1)
    p = text_emb.repeat(B, 1)
    p = swish(net.lin("embed2.4", swish(net.lin("embed2.2", swish(net.lin("embed2.0", p))))))
    e = swish(net.lin("embed.2", swish(net.lin("embed.0", gamma_embedding(t, inter_dim)))))
    return e, p


@torch.no_grad()
def wavbest_forward(sd, x_t, t_input, PAN, MS, text_emb, faithful=True, inter_dim=32, taps=None):
    """WavBEST.forward :600-636.  ``text_emb`` is the [1,768] embedding of the sensor
    prompt (what ``get_embeding(prompt)`` returns).  ``taps``: optional dict that receives
    named intermediates for block-level parity tests."""
    net = _Net(sd, faithful)
    B = x_t.shape[0]
    embed, prompt = condition_embeddings(sd, t_input, text_emb, B, inter_dim)
    cond = (PAN.repeat(1, MS.shape[1], 1, 1) - MS).unsqueeze(1)
    x = x_t.unsqueeze(1)

    h0_0 = net.stem("conv1", cond, prompt)
    h1_1, S1 = net.down("down1_1", h0_0, embed, prompt, flag=True)
    h2_1, S2 = net.down("down2_1", h1_1, embed, prompt, flag=True)
    h3_1, S3 = net.down("down3_1", h2_1, embed, prompt, flag=True)

    h0 = net.stem("conv2", x, prompt)
    h1, _ = net.down("down1", h0, embed, prompt)
    h2, _ = net.down("down2", h1, embed, prompt)
    h3, _ = net.down("down3", h2, embed, prompt)

    m = net.resblock("middle1", h3, embed, prompt)
    u1 = net.up("up1", torch.cat([m, h3_1, h3], 1), embed, S3, prompt)
    u2 = net.up("up2", torch.cat([u1, h2_1, h2], 1), embed, S2, prompt)
    u3 = net.up("up3", torch.cat([u2, h1_1, h1], 1), embed, S1, prompt)
    out = net.final(torch.cat([u3, h0_0, h0], 1), embed, prompt)
    if taps is not None:
        taps.update(h0_0=h0_0, h1_1=h1_1, h2_1=h2_1, h3_1=h3_1, h0=h0, h1=h1, h2=h2, h3=h3,
                    m=m, u1=u1, u2=u2, u3=u3, S1=S1, S2=S2, S3=S3, embed=embed, prompt=prompt)
    return out.squeeze(1)
