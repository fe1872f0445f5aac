"""numpy restatement of the reference's full-reference quality metrics (TEST INFRASTRUCTURE ONLY), used to check that
the CUDA path's fused images score within 0.5 % of the reference path's (north star).  Images are [H, W, C].
Pinned against the reference functions by tests/golden/make_golden_scene.py.  The reference has no `Q2n`: as SURVEY.md
8(d) notes, the per-band universal quality index (UIQC_numpy) stands in for it."""
import numpy as np


def sam(x_true, x_pred):
    """SAM_numpy (core/metrics.py:88-111): mean spectral angle in degrees, NaN pixels counted as 0."""
    dot = np.sum(x_true * x_pred, axis=2)
    res = np.arccos(dot / np.linalg.norm(x_pred, axis=2) / np.linalg.norm(x_true, axis=2))
    res[np.isnan(res)] = 0
    return float(np.mean(res) * 180 / np.pi)


def rmse(ms, ps):
    """RMSE_numpy (core/metrics.py:203-218): note the sum runs over every axis but the divisor is H*W only."""
    d = (ms - ps) ** 2
    return float(np.sqrt(np.sum(d) / (d.shape[0] * d.shape[1])))


def ergas(ms, ps, ratio=0.25):
    """ERGAS_numpy (core/metrics.py:221-241)."""
    d = ms.shape[2]
    summed = sum(rmse(ms[:, :, i], ps[:, :, i]) ** 2 / np.mean(ps[:, :, i]) ** 2 for i in range(d))
    return float(100 * ratio * np.sqrt(summed / d))


def q_index(a, b):
    """QIndex_numpy (core/metrics.py:262-284): universal image quality index of one band."""
    a, b = a.reshape(-1), b.reshape(-1)
    c = np.cov(a, b)
    m1, m2 = np.mean(a), np.mean(b)
    return float(4 * c[0, 1] * m1 * m2 / (c[0, 0] + c[1, 1]) / (m1 ** 2 + m2 ** 2))


def uiqc(ms, ps):
    """UIQC_numpy (core/metrics.py:244-259): mean per-band Q index."""
    return float(sum(q_index(ms[:, :, i], ps[:, :, i]) for i in range(ms.shape[2])) / ms.shape[2])
