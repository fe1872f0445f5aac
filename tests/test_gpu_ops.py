"""CUDA operators (through the C ABI) against the oracle and the reference's golden vectors."""
import pytest
import torch
import torch.nn.functional as F

from oracle import dpm_ref, wavbest_ref
from util import rel_l2

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops(lib_built):
    from tmdiff_b200 import ops
    return ops


def test_haar_golden(golden, ops):
    x = torch.from_numpy(golden["haar_x"]).cuda()
    got = ops.DWT_2D("haar")(x)
    for g, key in zip(got, ("haar_LL", "haar_LH", "haar_HL", "haar_HH")):
        assert torch.allclose(g.cpu(), torch.from_numpy(golden[key]), atol=1e-6), key
    rec = ops.IDWT_2D("haar")(*got)
    assert torch.allclose(rec.cpu(), torch.from_numpy(golden["haar_rec"]), atol=1e-6)
    assert torch.equal(ops.DWT_2D("haar")(x, ll_only=True), got[0])


@pytest.mark.parametrize("shape", [(1, 1, 2, 2), (3, 3, 24, 24), (2, 5, 6, 10), (2, 3, 8, 12), (2, 3, 6, 4), (1, 2048, 16, 16), (64, 512, 64, 64), (0, 4, 8, 8)])
def test_haar_shapes_and_round_trip(ops, shape):
    g = torch.Generator().manual_seed(5)
    x = torch.randn(shape, generator=g).cuda()
    bands = ops.DWT_2D("haar")(x)
    if x.numel():
        want = wavbest_ref.dwt2d_closed(x[:2].cpu())
        for b, w in zip(bands, want):
            assert torch.allclose(b[:2].cpu(), w, atol=1e-6)
    rec = ops.IDWT_2D("haar")(*bands)
    assert rec.shape == x.shape
    assert torch.allclose(rec, x, atol=1e-6)          # the reference's own round-trip check (DWT_IDWT_layer.py:650-667)
    # linearity / energy preservation of the orthonormal transform at full size
    if x.numel():
        e_in = float((x.double() ** 2).sum())
        e_out = sum(float((b.double() ** 2).sum()) for b in bands)
        assert abs(e_in - e_out) <= 1e-5 * e_in


def test_haar_rejects_odd_and_non_4d(ops):
    with pytest.raises(RuntimeError):
        ops.DWT_2D("haar")(torch.zeros(1, 1, 5, 8).cuda())
    with pytest.raises(AssertionError):
        ops.DWT_2D("haar")(torch.zeros(1, 8, 8).cuda())
    with pytest.raises(ValueError):
        ops.DWT_2D("db2")


CONV_CASES = [
    # B, Ci, Co, N, H, W, k
    (2, 32, 32, 4, 16, 16, 3),
    (1, 32, 64, 8, 16, 16, 3),
    (2, 64, 64, 4, 8, 8, 3),
    (1, 128, 256, 4, 8, 8, 3),
    (1, 768, 128, 4, 8, 8, 3),
    (2, 96, 32, 4, 16, 16, 1),
    (1, 256, 256, 8, 8, 8, 1),
    (1, 8, 16, 3, 4, 4, 3),
    (1, 16, 8, 1, 8, 8, 3),
    (3, 32, 32, 5, 24, 40, 3),
    (2, 64, 128, 8, 16, 16, 3),      # tcgen05 reuse schedule, 2 accumulators, two n-tiles
    (1, 128, 256, 4, 16, 24, 3),     # one accumulator per tile
    (2, 96, 32, 4, 32, 32, 3),       # 3 K-chunks, 4 accumulators
    (5, 64, 64, 4, 4, 4, 3),         # batch folded into the 128-row tile
    (3, 192, 48, 2, 8, 16, 1),
    (1, 32, 32, 8, 16, 24, 3),       # 8 accumulators per tile, merged-dz MMAs (N = 96)
    (2, 64, 48, 2, 32, 16, 3),       # 2 accumulators, merged (N = 144)
    (1, 32, 16, 6, 16, 8, 3),        # band count 6 -> 2 accumulators, three n-tiles
    (2, 64, 768, 1, 16, 8, 1),       # wider than one launch takes: three channel groups of 256 (stacked q / k / v projection)
    (1, 128, 320, 2, 8, 8, 3),       # 256 + 64
]


@pytest.mark.parametrize("case", CONV_CASES)
@pytest.mark.parametrize("dtype,tol", [("fp32", 5e-6), ("bf16", 1e-2)])  # fp32: summation-order noise over K <= 27*768 terms
def test_modulated_conv3d(ops, case, dtype, tol):
    B, Ci, Co, N, H, W, k = case
    g = torch.Generator().manual_seed(11)
    x = torch.randn(B, Ci, N, H, W, generator=g)
    w = torch.randn(Co, Ci, k, k, k, generator=g) / (Ci * k ** 3) ** 0.5
    s = torch.randn(B, Ci, 1, 1, generator=g)
    bias = torch.randn(Co, generator=g)
    want = wavbest_ref.modulated_conv3d(x, w, s, k // 2, faithful=True)             # the reference formulation
    got = ops.modulated_conv3d(x.cuda(), w.cuda(), s.cuda(), padding=k // 2, dtype=dtype)
    assert rel_l2(got, want) < tol
    want_b = F.conv3d(x, w, bias, padding=k // 2)
    got_b = ops.modulated_conv3d(x.cuda(), w.cuda(), None, padding=k // 2, bias=bias.cuda(), dtype=dtype)
    assert rel_l2(got_b, want_b) < tol


def test_modulated_conv3d_linearity_full_size(ops):
    """Size-independent property at a BASELINE-sized layer (B=16 of 64x64 4-band, 64->64): conv(a*x+y) = a*conv(x)+conv(y)."""
    g = torch.Generator().manual_seed(2)
    x = torch.randn(16, 64, 4, 64, 64, generator=g).cuda()
    y = torch.randn(16, 64, 4, 64, 64, generator=g).cuda()
    w = (torch.randn(64, 64, 3, 3, 3, generator=g) / 41.6).cuda()
    f = lambda t: ops.modulated_conv3d(t, w, None, padding=1, dtype="fp32")
    assert rel_l2(f(0.5 * x + y), 0.5 * f(x) + f(y)) < 1e-5


# the layer shapes of the benchmarked workload (BASELINE config 2: B = 64, 4 bands, 64x64): these run the code paths the
# headline number runs -- 512..2048 tiles over 148 persistent CTAs, NACC = 4 with pl_off = 1, several waves of accumulator
# double-buffering, the Co = 256 channel-group split, batch folded into M at the 8x8 level
BENCH_CONV_CASES = [
    (64, 32, 32, 4, 64, 64, 3),
    (64, 64, 64, 4, 64, 64, 3),
    (64, 96, 32, 4, 64, 64, 3),
    (64, 128, 128, 4, 32, 32, 3),
    (64, 256, 256, 4, 16, 16, 3),
    (64, 768, 128, 4, 8, 8, 3),
    (64, 256, 256, 4, 8, 8, 3),
    (64, 32, 32, 8, 64, 64, 3),      # 8 bands: two band tiles of 4 planes
    (64, 96, 32, 4, 64, 64, 1),
]


@pytest.mark.parametrize("case", BENCH_CONV_CASES)
def test_modulated_conv3d_benchmark_shapes(ops, case):
    """tcgen05 kernel at the benchmarked layer shapes against (i) the fp32 CUDA-core kernel on the whole tensor and (ii) the
    reference formulation (groups=B conv of per-sample modulated weights, Hyper_unet_general.py:51-77) on the CPU for the
    first, a middle and the last sample of the batch."""
    B, Ci, Co, N, H, W, k = case
    g = torch.Generator(device="cuda").manual_seed(23)
    x = torch.randn(B, Ci, N, H, W, generator=g, device="cuda")
    w = torch.randn(Co, Ci, k, k, k, generator=g, device="cuda") / (Ci * k ** 3) ** 0.5
    s = torch.randn(B, Ci, 1, 1, generator=g, device="cuda")
    got = ops.modulated_conv3d(x, w, s, padding=k // 2, dtype="bf16")
    ref32 = ops.modulated_conv3d(x, w, s, padding=k // 2, dtype="fp32")
    assert rel_l2(got, ref32) < 1e-2
    # per sample too: a tile-scheduling bug would hide in a whole-tensor norm
    per = ((got - ref32).flatten(1).norm(dim=1) / ref32.flatten(1).norm(dim=1)).max().item()
    assert per < 1e-2, per
    for b in (0, B // 2, B - 1):
        want = wavbest_ref.modulated_conv3d(x[b:b + 1].cpu(), w.cpu(), s[b:b + 1].cpu(), k // 2, faithful=True)
        assert rel_l2(ref32[b:b + 1], want) < 5e-6, b
        assert rel_l2(got[b:b + 1], want) < 1e-2, b


@pytest.mark.parametrize("shape", [(2, 8 * 128 * 128), (1, 8 * 512 * 512), (3, 51201), (2, 7), (64, 4 * 64 * 64)])
def test_dynamic_threshold_any_size(ops, shape):
    """No per-sample size limit (the reference's denoiser is fully convolutional and its self-test runs 512x512,
    Hyper_unet_general.py:664-671; dynamic_thresholding_fn has no limit, dpm_solver_pytorch.py:430-439); odd sizes take
    the scalar path."""
    g = torch.Generator().manual_seed(shape[1] % 1000)
    x = torch.randn(shape, generator=g) * 2.0
    x[0] *= 0.2
    want = dpm_ref.dynamic_threshold(x)
    got = ops.dynamic_thresholding(x.cuda()).cpu()
    assert torch.equal(got, want) or float((got - want).abs().max()) <= 1e-6, float((got - want).abs().max())


def test_dynamic_threshold_degenerate_values(ops):
    """All-equal samples (every radix level lands in one bin), exact ties at the two interpolated ranks, zeros and a huge
    outlier."""
    n = 4 * 64 * 64
    x = torch.full((4, n), 2.5)
    x[1] = 0.0
    x[2, ::2] = -2.5
    x[3] = torch.linspace(-3, 3, n)
    x[3, 7] = 1e30
    want = dpm_ref.dynamic_threshold(x)
    got = ops.dynamic_thresholding(x.cuda()).cpu()
    assert torch.allclose(got, want, rtol=0, atol=1e-6)


@pytest.mark.parametrize("scale", [0.3, 1.0, 3.0, 50.0])
@pytest.mark.parametrize("n", [4 * 32 * 32, 4 * 64 * 64, 8 * 64 * 64, 1000])
def test_dynamic_threshold(ops, scale, n):
    g = torch.Generator().manual_seed(int(scale * 10) + n)
    x = torch.randn(3, n, generator=g) * scale
    x[1] *= 0.01                       # a sample whose quantile stays below max_val
    want = dpm_ref.dynamic_threshold(x)
    got = ops.dynamic_thresholding(x.cuda())
    assert torch.allclose(got.cpu(), want, rtol=0, atol=1e-6), float((got.cpu() - want).abs().max())
    assert rel_l2(got, want) < 1e-6


def test_dynamic_threshold_boundary_counts(ops):
    """Exactly need_hi-1, need_hi, need_lo values above 1: exercises the three selection branches."""
    n = 4096
    lo, w = ops.quantile_position(n, 0.995)
    for k in (n - lo - 2, n - lo - 1, n - lo, n - lo + 5, n):
        x = torch.linspace(0.01, 0.99, n).repeat(2, 1)
        x[:, :k] = torch.linspace(1.5, 4.0, k)
        x[1] = -x[1]
        want = dpm_ref.dynamic_threshold(x)
        got = ops.dynamic_thresholding(x.cuda()).cpu()
        assert torch.allclose(got, want, rtol=0, atol=1e-6), k


@pytest.mark.gpu
def test_quality_metrics_match_reference(ops):
    """On-device SAM / ERGAS / Q against the reference's numpy functions (committed vectors) and the oracle restatement."""
    import numpy as np
    import os
    from oracle import metrics_ref
    vec = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "scene_vectors.npz"))
    a, b = vec["metric_a"], vec["metric_b"]                       # [H, W, C]
    to_nchw = lambda t: torch.from_numpy(t).permute(2, 0, 1)[None].float().cuda()
    got = ops.quality_metrics(to_nchw(a), to_nchw(b)).cpu().double().numpy()[0]
    assert np.allclose(got, vec["metric_values"], rtol=2e-5)
    g = torch.Generator().manual_seed(4)
    x = torch.rand(5, 8, 64, 64, generator=g) * 0.9 + 0.05
    y = (x + 0.03 * torch.randn(x.shape, generator=g)).clamp(0.01, 1)
    got = ops.quality_metrics(x.cuda(), y.cuda()).cpu().double().numpy()
    for i in range(5):
        hwc = lambda t: t[i].permute(1, 2, 0).double().numpy()
        want = [metrics_ref.sam(hwc(x), hwc(y)), metrics_ref.ergas(hwc(x), hwc(y)), metrics_ref.uiqc(hwc(x), hwc(y))]
        assert np.allclose(got[i], want, rtol=2e-5), i
    with pytest.raises(RuntimeError):
        ops.quality_metrics(torch.rand(1, 9, 8, 8).cuda(), torch.rand(1, 9, 8, 8).cuda())
