"""On-device no-reference / full-reference quality metrics with the reference's names and return conventions
(SURVEY.md 8(f) rank 3): core/metrics.py:386-503 (SAM_torch is ops.quality_metrics' first column), utils/util.py:375-386.

Every image pair is reduced ONCE by the fused band-moment kernel (csrc/operators2.cu: per-band sums, squared sums and all
band-pair products, optionally behind the 3x3 high-pass of sobel_torch); what is left is arithmetic on [B, C, C] moment
tables, done with a handful of torch ops on those tiny tensors.  CUDA tensors only, no CPU fallback.
"""
import torch

from . import _lib
from .ops import _cuda_f32

_EPS32 = float(torch.finfo(torch.float32).eps)


def band_moments(a, b, highpass=False):
    """a [B, Ca, H, W], b [B, Cb, H, W] -> dict of per-image moment tables (sa, saa [B,Ca]; sb, sbb [B,Cb]; sab [B,Ca,Cb]; n)."""
    a, b = _cuda_f32(a, "a"), _cuda_f32(b, "b")
    if a.dim() != 4 or b.dim() != 4 or a.shape[0] != b.shape[0] or a.shape[2:] != b.shape[2:]:
        raise ValueError("band_moments: a [B, Ca, H, W] and b [B, Cb, H, W] expected")
    B, Ca, H, W = a.shape
    Cb = b.shape[1]
    out = torch.empty(B, 2 * Ca + 2 * Cb + Ca * Cb + 1, device=a.device)
    with torch.cuda.device(a.device):
        _lib.check(_lib.load().tmd_band_moments(_lib.ptr(a), _lib.ptr(b), _lib.ptr(out), B, Ca, Cb, H, W, int(bool(highpass)),
                                                _lib.stream_ptr()))
    o = 0
    res = {}
    for name, n in (("sa", Ca), ("saa", Ca), ("sb", Cb), ("sbb", Cb)):
        res[name] = out[:, o:o + n]
        o += n
    res["sab"] = out[:, o:o + Ca * Cb].reshape(B, Ca, Cb)
    res["n"] = out[:, -1]
    return res


def q_index_matrix(a, b, eps=1e-8):
    """QIndex_torch's per-image value (core/metrics.py:440-460) for every band pair: [B, Ca, Cb]."""
    m = band_moments(a, b)
    n = m["n"][:, None]
    Ea, Eb = m["sa"] / n, m["sb"] / n
    va, vb = m["saa"] / n - Ea * Ea, m["sbb"] / n - Eb * Eb
    cov = m["sab"] / n[:, :, None] - Ea[:, :, None] * Eb[:, None, :]
    return 4 * cov * Ea[:, :, None] * Eb[:, None, :] / ((va[:, :, None] + vb[:, None, :]) * (Ea[:, :, None] ** 2 + Eb[:, None, :] ** 2) + eps)


def QIndex_torch(a, b, eps=1e-8):
    """a, b one-channel images [N, H, W] -> mean Q over the N images."""
    return q_index_matrix(a[:, None], b[:, None], eps)[:, 0, 0].mean()


def D_lambda_torch(l_ms, ps):
    """Spectral distortion (core/metrics.py:463-481): mean over band pairs i != j of |Q(ps_i, ps_j) - Q(lms_i, lms_j)|,
    each Q already averaged over the batch like the reference."""
    L = ps.shape[1]
    d = (q_index_matrix(ps, ps).mean(0) - q_index_matrix(l_ms, l_ms).mean(0)).abs()
    off = d.sum() - d.diagonal().sum()
    return (off / L / (L - 1)).reshape(1)


def D_s_torch(l_ms, pan, l_pan, ps):
    """Spatial distortion (core/metrics.py:484-503): mean over bands of |Q(ps_i, pan) - Q(lms_i, l_pan)|."""
    L = ps.shape[1]
    d = (q_index_matrix(ps, pan[:, :1]).mean(0) - q_index_matrix(l_ms, l_pan[:, :1]).mean(0)).abs()
    return (d.sum() / L).reshape(1)


def qnr(l_ms, pan, l_pan, ps, alpha=1.0, beta=1.0):
    """(1 - D_lambda)^alpha (1 - D_s)^beta from the two indices above."""
    return (1 - D_lambda_torch(l_ms, ps)) ** alpha * (1 - D_s_torch(l_ms, pan, l_pan, ps)) ** beta


def _corr(m, eps):
    n = m["n"][:, None]
    i = torch.arange(m["sa"].shape[1], device=n.device)
    cov = m["sab"][:, i, i] - m["sa"] * m["sb"] / n
    vx, vy = m["saa"] - m["sa"] ** 2 / n, m["sbb"] - m["sb"] ** 2 / n
    return cov, vx.clamp_min(0), vy.clamp_min(0)


def SCC_torch(x, y):
    """Spatial correlation coefficient (core/metrics.py:425-437): correlation of the 3x3 high-passed bands, mean over (n, c)."""
    cov, vx, vy = _corr(band_moments(x, y, highpass=True), 0.0)
    return (cov / vx.sqrt() / vy.sqrt()).mean()


def cc(img1, img2):
    """Correlation coefficient per image (utils/util.py:375-386): [N], the mean over bands of the clamped band correlation."""
    cov, vx, vy = _corr(band_moments(img1, img2), _EPS32)
    return (cov / (_EPS32 + vx.sqrt() * vy.sqrt())).clamp(-1.0, 1.0).mean(-1)
