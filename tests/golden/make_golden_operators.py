"""Golden vectors for the stand-alone operators, produced by the UNMODIFIED reference classes
on the CPU: AdaptiveGroupNorm / AttnBlockpp (GeneralModel/Hyper_unet_general.py), SpatialSelfAttention / CrossAttention
(core/Attention.py), Dynamic_conv3d (core/dynamic_conv.py).  Stores inputs, every state-dict tensor and the outputs.

    python tests/golden/make_golden_operators.py      (build container only: needs /root/reference)
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402,F401  (sys.path for the reference + shims)


def seeded(module, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in module.named_parameters():
            if p.dim() >= 2:
                fan_in = p[0].numel() if "NIN" not in name and name != "weight" else (p.shape[0] if "NIN" in name else p[0, 0].numel())
                p.copy_(torch.randn(p.shape, generator=g) / fan_in ** 0.5)
            else:
                p.copy_(torch.randn(p.shape, generator=g) * 0.1 + (1.0 if name.endswith("norm.weight") or "GroupNorm_0.weight" in name else 0.0))
    return module


def main():
    H = mg._load_reference()
    import core.Attention as A
    import core.dynamic_conv as D
    g = torch.Generator().manual_seed(11)
    out = {}

    def put(tag, module, inputs, result):
        for k, v in module.state_dict().items():
            out["%s/sd/%s" % (tag, k)] = v.numpy()
        for k, v in inputs.items():
            out["%s/in/%s" % (tag, k)] = v.numpy()
        out["%s/out" % tag] = result.detach().numpy()

    with torch.no_grad():
        m = seeded(H.AdaptiveGroupNorm(8, 32, 128), 1)
        x, st = torch.randn(2, 32, 4, 8, 8, generator=g) * 2 + 0.5, torch.randn(2, 128, generator=g)
        put("adagn", m, {"x": x, "style": st}, m(x, st))
        out["adagn/out_swish"] = H.Swish()(m(x, st)).numpy()

        m = seeded(H.AttnBlockpp(64, skip_rescale=True, init_scale=1.0), 2)
        x = torch.randn(2, 16, 4, 8, 8, generator=g)
        put("attnpp", m, {"x": x}, m(x))

        m = seeded(A.SpatialSelfAttention(64), 3)
        x = torch.randn(2, 64, 8, 12, generator=g)
        put("ssa", m, {"x": x}, m(x))

        m = seeded(A.CrossAttention(64, context_dim=128, heads=8, dim_head=64), 4)
        x, ctx = torch.randn(2, 96, 64, generator=g), torch.randn(2, 77, 128, generator=g)
        put("xattn", m, {"x": x, "context": ctx}, m(x, context=ctx))

        m = seeded(D.Dynamic_conv3d(16, 32, 3, padding=1, K=4, temperature=34), 5)
        with torch.no_grad():
            m.attention.fc2.weight.mul_(40.0)          # spread the expert softmax away from uniform
        x = torch.randn(3, 16, 4, 8, 8, generator=g)
        put("dyn3d", m, {"x": x}, m(x))
    np.savez_compressed(os.path.join(HERE, "operator_vectors.npz"), **out)
    print({k: v.shape for k, v in out.items() if k.endswith("out")})


if __name__ == "__main__":
    main()
