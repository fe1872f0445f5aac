"""Stand-alone operators with the reference's layer interfaces, backed by the CUDA library.

DWT_2D / IDWT_2D mirror DWT_IDWT/DWT_IDWT_layer.py:247-430 (forward only, Haar only);
modulated_conv3d mirrors GeneralModel/Hyper_unet_general.py:51-77.  CUDA tensors only.
"""
import ctypes as C

import torch

from . import _lib


def _cuda_f32(t, name):
    if not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor (tmdiff_b200 has no CPU path)" % name)
    return t.contiguous().float()


class DWT_2D(torch.nn.Module):
    """(N,C,H,W) -> LL, LH, HL, HH each (N,C,H/2,W/2)."""

    def __init__(self, wavename="haar"):
        super().__init__()
        if wavename != "haar":
            raise ValueError("only the 'haar' wavelet is on the sampling path")

    def forward(self, input, ll_only=False):
        assert input.dim() == 4                       # DWT_IDWT_layer.py:330
        x = _cuda_f32(input, "input")
        B, Cc, H, W = x.shape
        outs = [torch.empty(B, Cc, H // 2, W // 2, device=x.device) for _ in range(1 if ll_only else 4)]
        p = [_lib.ptr(o) for o in outs] + [None] * (4 - len(outs))
        with torch.cuda.device(x.device):
            _lib.check(_lib.load().tmd_dwt2d(_lib.ptr(x), p[0], p[1], p[2], p[3], B * Cc, H, W, _lib.stream_ptr()))
        return outs[0] if ll_only else tuple(outs)


class IDWT_2D(torch.nn.Module):
    """LL, LH, HL, HH (N,C,h,w) -> (N,C,2h,2w)."""

    def __init__(self, wavename="haar"):
        super().__init__()
        if wavename != "haar":
            raise ValueError("only the 'haar' wavelet is on the sampling path")

    def forward(self, LL, LH, HL, HH):
        assert LL.dim() == LH.dim() == HL.dim() == HH.dim() == 4   # DWT_IDWT_layer.py:425
        t = [_cuda_f32(v, "band") for v in (LL, LH, HL, HH)]
        B, Cc, h, w = t[0].shape
        out = torch.empty(B, Cc, 2 * h, 2 * w, device=t[0].device)
        with torch.cuda.device(out.device):
            _lib.check(_lib.load().tmd_idwt2d(*[_lib.ptr(v) for v in t], _lib.ptr(out), B * Cc, h, w, _lib.stream_ptr()))
        return out


def modulated_conv3d(x, w, s=None, padding=None, bias=None, stride=None, dilation=1, dtype="fp32"):
    """conv3d(x, w * s[b, None, :, None, None, None]) per sample, as one shared-weight implicit GEMM
    over x * s.  x [B,Ci,N,H,W]; w [Co,Ci,k,k,k] with k in {1,3} and 'same' padding; s [B,Ci(,1,1)]."""
    k = w.shape[-1]
    pad = padding if isinstance(padding, int) else (padding[0] if padding is not None else k // 2)
    if w.shape[2:] != (k, k, k) or k not in (1, 3) or pad != k // 2:
        raise ValueError("only 1x1x1 / 3x3x3 kernels with same-padding are on the sampling path")
    if stride not in (None, 1, (1, 1, 1)) or dilation not in (1, (1, 1, 1)):
        raise ValueError("stride and dilation must be 1")
    x, w = _cuda_f32(x, "x"), _cuda_f32(w, "w")
    B, Ci, N, H, W = x.shape
    Co = w.shape[0]
    if s is not None:
        s = _cuda_f32(s, "s").reshape(B, Ci)
    if bias is not None:
        bias = _cuda_f32(bias, "bias")
    y = torch.empty(B, Co, N, H, W, device=x.device)
    code = {"fp32": _lib.F32, "bf16": _lib.BF16}[dtype]
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().tmd_conv3d_mod(_lib.ptr(x), _lib.ptr(w), _lib.ptr(s), _lib.ptr(bias), _lib.ptr(y),
                                              B, Ci, Co, N, H, W, k, code, _lib.stream_ptr()))
    return y


def dynamic_thresholding(x0, ratio=0.995, max_val=1.0):
    """DPM_Solver.dynamic_thresholding_fn (core/dpm_solver_pytorch.py:430-439); returns a new tensor."""
    x = _cuda_f32(x0, "x0").clone()
    B = x.shape[0]
    n = x[0].numel()
    q_lo, q_w = quantile_position(n, ratio)
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().tmd_dynamic_threshold(_lib.ptr(x), B, n, q_lo, q_w, max_val, _lib.stream_ptr()))
    return x


def quantile_position(n, q):
    """(floor(rank), rank - floor(rank)) with rank = q*(n-1) rounded the way torch.quantile rounds it
    for an fp32 input (ATen Sorting.cpp: q is a tensor of the input dtype)."""
    rank = torch.tensor(q, dtype=torch.float32) * (n - 1)
    lo = torch.floor(rank)
    return int(lo.item()), float((rank - lo).item())


def quality_metrics(ref, img, ratio=0.25):
    """Per-image [SAM (degrees), ERGAS, Q] of img against ref ([B, C, H, W], C <= 8) as one fused device reduction
    (core/metrics.py SAM_numpy / ERGAS_numpy / UIQC_numpy semantics); returns a [B, 3] CUDA tensor."""
    a, b = _cuda_f32(ref, "ref"), _cuda_f32(img, "img")
    if a.shape != b.shape or a.dim() != 4:
        raise ValueError("ref and img must both be [B, C, H, W]")
    B, Cc, H, W = a.shape
    out = torch.empty(B, 3, device=a.device)
    with torch.cuda.device(a.device):
        _lib.check(_lib.load().tmd_quality_metrics(_lib.ptr(a), _lib.ptr(b), _lib.ptr(out), B, Cc, H, W, float(ratio), _lib.stream_ptr()))
    return out


def ancestral_update(x, eps, noise, c_recip, c_recipm1, coef1, coef2, sigma, clip=True, seed=0, step=0):
    """out = coef1 * clamp(c_recip*x - c_recipm1*eps, -1, 1) + coef2*x + sigma*z  (GeneralDiffusion.p_sample's arithmetic,
    GeneralModel/diffusion_general.py:134-138, 155-174, 196-201) in one kernel.  z = `noise`, or Philox normals drawn in the
    kernel from (seed, step) when noise is None."""
    x, eps = _cuda_f32(x, "x"), _cuda_f32(eps, "eps")
    if eps.shape != x.shape or x.numel() % 4:
        raise ValueError("ancestral_update: x and eps must have the same shape and a multiple of 4 elements")
    z = None if noise is None else _cuda_f32(noise, "noise")
    out = torch.empty_like(x)
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().tmd_ancestral_update(_lib.ptr(x), _lib.ptr(eps), _lib.ptr(z), _lib.ptr(out), float(c_recip), float(c_recipm1),
                                                    float(coef1), float(coef2), float(sigma), int(bool(clip)), int(seed), int(step),
                                                    x.numel(), _lib.stream_ptr()))
    return out
