"""Golden vectors for the remaining stand-alone operators and the torch metrics, produced by the UNMODIFIED reference on the
CPU: BasicTransformerBlock / SpatialTransformer (core/Attention.py), Dynamic_conv1d / Dynamic_conv2d (core/dynamic_conv.py),
EntireKernel / GaussianKernel (core/kernelAdaption.py), QIndex_torch / D_lambda_torch / D_s_torch / SCC_torch
(core/metrics.py) and cc (utils/util.py).  Stores inputs, every state-dict tensor and the outputs.

    python tests/golden/make_golden_operators2.py      (build container only: needs /root/reference)
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402,F401  (sys.path for the reference + shims)
from make_golden_operators import seeded  # noqa: E402


def main():
    import types
    for name in ("skimage", "skimage.metrics", "sewar", "sewar.full_ref", "h5py", "cv2", "torchvision", "torchvision.utils"):   # module-top imports of core/metrics.py, unused by the functions called here
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    sys.modules["skimage.metrics"].peak_signal_noise_ratio = sys.modules["skimage.metrics"].structural_similarity = None
    if not hasattr(sys.modules["torchvision.utils"], "make_grid"):
        sys.modules["torchvision.utils"].make_grid = None
    import scipy.ndimage
    sys.modules.setdefault("scipy.ndimage.filters", scipy.ndimage)
    mg._load_reference()
    import core.Attention as A
    import core.dynamic_conv as D
    import core.kernelAdaption as KA
    import core.metrics as M
    g = torch.Generator().manual_seed(23)
    out = {}

    def put(tag, module, inputs, result):
        for k, v in module.state_dict().items():
            out["%s/sd/%s" % (tag, k)] = v.numpy()
        for k, v in inputs.items():
            out["%s/in/%s" % (tag, k)] = v.numpy()
        out["%s/out" % tag] = result.detach().numpy()

    with torch.no_grad():
        m = seeded(A.BasicTransformerBlock(64, 4, 16, context_dim=96, checkpoint=False), 6)
        for n in (m.norm1, m.norm2, m.norm3):
            n.weight.add_(1.0)
        x, ctx = torch.randn(2, 48, 64, generator=g), torch.randn(2, 77, 96, generator=g)
        put("btb", m, {"x": x, "context": ctx}, m(x, context=ctx))

        m = seeded(A.SpatialTransformer(64, 4, 16, depth=2, context_dim=[96, 96], use_checkpoint=False), 7)
        for blk in m.transformer_blocks:
            for n in (blk.norm1, blk.norm2, blk.norm3):
                n.weight.add_(1.0)
        x, ctx, ctx2 = torch.randn(2, 64, 8, 6, generator=g), torch.randn(2, 77, 96, generator=g), torch.randn(2, 77, 96, generator=g)
        put("st", m, {"x": x, "context": ctx, "context2": ctx2}, m(x, context=[ctx, ctx2]))     # proj_out is re-seeded (not the zero init): a real residual

        m = seeded(D.Dynamic_conv2d(16, 32, 3, padding=1, K=4, temperature=34), 8)
        m.attention.fc2.weight.mul_(40.0)
        x = torch.randn(3, 16, 12, 10, generator=g)
        put("dyn2d", m, {"x": x}, m(x))

        m = seeded(D.Dynamic_conv1d(16, 32, 3, padding=1, K=4, temperature=34), 9)
        m.attention.fc2.weight.mul_(40.0)
        x = torch.randn(3, 16, 40, generator=g)
        put("dyn1d", m, {"x": x}, m(x))

        m = KA.EntireKernel(kernel_size=4, bands=4, scale_factor=4)
        m.KernelAdaption.copy_(torch.randn(4, 4, generator=g))
        x = torch.rand(2, 4, 32, 32, generator=g)
        put("psf_entire", m, {"x": x}, m(x))
        out["psf_entire/psf"] = m.psf.numpy()

        m = KA.GaussianKernel(kernel_size=5, bands=4, scale_factor=4)
        m.KernelParam.copy_(torch.tensor([[0.9, 0.3], [0.25, 0.6]]))
        x = torch.rand(2, 4, 33, 29, generator=g)
        put("psf_gauss", m, {"x": x}, m(x))
        out["psf_gauss/psf"] = m.KernelAdaption[0, 0].numpy()

        # metrics: a 4-band fused image, its low-resolution MS, PAN and low-resolution PAN
        sys.path.insert(0, "/root/reference")
        import utils.util as U
        ps = torch.rand(3, 4, 64, 64, generator=g)
        ps = 0.7 * ps + 0.3 * ps.mean(1, keepdim=True)
        lms = torch.nn.functional.avg_pool2d(ps, 4) + 0.02 * torch.randn(3, 4, 16, 16, generator=g)
        pan = ps.mean(1, keepdim=True) + 0.05 * torch.randn(3, 1, 64, 64, generator=g)
        lpan = torch.nn.functional.avg_pool2d(pan, 4)
        gt = ps + 0.05 * torch.randn(3, 4, 64, 64, generator=g)
        for k, v in (("ps", ps), ("lms", lms), ("pan", pan), ("lpan", lpan), ("gt", gt)):
            out["metrics/in/" + k] = v.numpy()
        out["metrics/QIndex"] = M.QIndex_torch(ps[:, 0], gt[:, 0]).numpy()
        out["metrics/D_lambda"] = M.D_lambda_torch(lms, ps).numpy()
        out["metrics/D_s"] = M.D_s_torch(lms, pan, lpan, ps).numpy()
        # sobel_torch's weight is [1,1,3,3]: the reference function takes one band at a time
        out["metrics/SCC"] = np.stack([M.SCC_torch(gt[:, c:c + 1], ps[:, c:c + 1]).numpy() for c in range(4)])
        out["metrics/cc"] = U.cc(ps, gt).numpy()
    np.savez_compressed(os.path.join(HERE, "operator_vectors2.npz"), **out)
    print({k: (v.shape, float(np.abs(v).mean())) for k, v in out.items() if k.endswith("out") or k.startswith("metrics/") and "/in/" not in k})


if __name__ == "__main__":
    main()
