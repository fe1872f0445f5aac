"""tmd_dynamic_threshold_static: the allocation-free variant of tmd_dynamic_threshold (scratch in a file-scope
`static __device__` array of the library, tmdiff_b200/csrc/solver.cu).  It has no counterpart in the reference; the checker
here is the existing entry point, which tests/test_gpu_ops.py holds against the oracle.

Collected last on purpose (file name).
"""
import pytest
import torch

gpu = pytest.mark.gpu

MAX_BATCH = 64      # TMD_THRESHOLD_STATIC_MAX_BATCH, include/tmdiff_b200.h


def _call(name, x, ratio=0.995, max_val=1.0):
    """In place on a clone of x [B, ...] through the C-ABI entry point `name`; returns (rc, tensor)."""
    from tmdiff_b200 import _lib, ops
    y = x.clone()
    B, n = y.shape[0], y[0].numel()
    q_lo, q_w = ops.quantile_position(n, ratio)
    with torch.cuda.device(y.device):
        rc = getattr(_lib.load(), name)(_lib.ptr(y), B, n, q_lo, q_w, max_val, _lib.stream_ptr())
    return rc, y


@gpu
def test_static_workspace_threshold_matches_allocating_entry_point():
    from tmdiff_b200 import _lib
    assert torch.cuda.is_available()
    g = torch.Generator().manual_seed(20)
    # two calls in a row on the shared array: a full batch first, then fewer, larger, odd-sized samples (the scalar path),
    # whose histograms must not see what the first call left behind; then the first input again
    first = (torch.randn(MAX_BATCH, 4, 64, 64, generator=g) * 2.0).cuda()
    first[1] *= 0.01                                    # a sample whose quantile stays below max_val
    second = (torch.randn(3, 8 * 128 * 128 + 3, generator=g) * 5.0).cuda()
    second[2] = 0.75                                    # every radix level lands in one bin
    for x in (first, second, first):
        rc, want = _call("tmd_dynamic_threshold", x)
        _lib.check(rc)
        rc, got = _call("tmd_dynamic_threshold_static", x)
        _lib.check(rc)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        assert not torch.equal(got, x)                  # it did something
        assert torch.allclose(got, want, rtol=0, atol=1e-6), float((got - want).abs().max())


@gpu
def test_static_workspace_threshold_rejects_sizes_beyond_the_limit():
    from tmdiff_b200 import _lib
    x = torch.randn(MAX_BATCH + 1, 256).cuda()
    rc, y = _call("tmd_dynamic_threshold_static", x)
    assert rc != 0
    msg = _lib.load().tmd_last_error().decode()
    assert "TMD_THRESHOLD_STATIC_MAX_BATCH" in msg and str(MAX_BATCH) in msg, msg
    torch.cuda.synchronize()
    assert torch.equal(y, x)                            # rejected before any launch
    # a good call still works afterwards
    rc, got = _call("tmd_dynamic_threshold_static", x[:MAX_BATCH])
    _lib.check(rc)
    rc, want = _call("tmd_dynamic_threshold", x[:MAX_BATCH])
    _lib.check(rc)
    assert torch.allclose(got, want, rtol=0, atol=1e-6)
