"""Generate the committed golden vectors by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
It (1) imports the reference with the fixtures of SURVEY.md 8(c) (pywt / ml_collections
shims, a stub FrozenCLIPEmbedder returning seeded [1,768] vectors), (2) checks the oracle
restatement (oracle/) against it, printing the errors, and (3) writes small .npz vectors
that tests/ compare against on machines where the reference is not mounted.
"""
import json
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [os.path.join(HERE, "_shims"), "/root/reference", ROOT]

from oracle import dpm_ref, wavbest_ref  # noqa: E402
from oracle.inputs import make_inputs  # noqa: E402

EMB = wavbest_ref.seeded_text_embeddings()
PROMPT_OF_TEXT = {}


class _Emb(torch.Tensor):  # CPU host: defuse the hard-coded .to("cuda") at Hyper_unet_general.py:602
    def to(self, *a, **k):
        return self.as_subclass(torch.Tensor)


def _load_reference():
    import core.clip as clipmod

    class FakeCLIP(torch.nn.Module):
        def __init__(self, *a, **k):
            super().__init__()

        def encode(self, text):
            return PROMPT_OF_TEXT[text].as_subclass(_Emb)

    clipmod.FrozenCLIPEmbedder = FakeCLIP
    import GeneralModel.Hyper_unet_general as H
    H.FrozenCLIPEmbedder = FakeCLIP
    # map prompt text -> seeded embedding (get_prompt is a pure function of the sensor name)
    for s in wavbest_ref.SENSORS:
        PROMPT_OF_TEXT[H.WavBEST.get_prompt(None, s)] = EMB[s].clone()
    # the reference's "WV2" text equals no other text, fine; GF2/WV2 differ by resolution digits
    return H


def rel(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def main():
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count())
    H = _load_reference()
    from GeneralModel.diffusion_general import GeneralDiffusion
    from core.dpm_solver_pytorch import DPM_Solver, NoiseScheduleVP, model_wrapper
    from DWT_IDWT.DWT_IDWT_layer import DWT_2D, IDWT_2D

    report = {}
    out = {}

    # ---- Haar ----------------------------------------------------------
    x = torch.randn(2, 6, 16, 24)
    ref = DWT_2D("haar")(x)
    mine = wavbest_ref.dwt2d(x)
    closed = wavbest_ref.dwt2d_closed(x)
    report["dwt_matmul_vs_ref_maxabs"] = max(float((a - b).abs().max()) for a, b in zip(mine, ref))
    report["dwt_closed_vs_ref_maxabs"] = max(float((a - b).abs().max()) for a, b in zip(closed, ref))
    rec_ref = IDWT_2D("haar")(*ref)
    report["idwt_matmul_vs_ref_maxabs"] = float((wavbest_ref.idwt2d(*ref) - rec_ref).abs().max())
    report["idwt_closed_vs_ref_maxabs"] = float((wavbest_ref.idwt2d_closed(*ref) - rec_ref).abs().max())
    out.update(haar_x=x.numpy(), haar_LL=ref[0].numpy(), haar_LH=ref[1].numpy(), haar_HL=ref[2].numpy(),
               haar_HH=ref[3].numpy(), haar_rec=rec_ref.numpy())

    # ---- parameter inventory ------------------------------------------
    chans = [32, 64, 128, 256]
    net = H.WavBEST(channels=chans).eval()
    ref_shapes = {k: tuple(v.shape) for k, v in net.state_dict().items()}
    mine_shapes = dict(wavbest_ref.param_shapes(chans))
    assert ref_shapes == mine_shapes, (set(ref_shapes) ^ set(mine_shapes))
    report["n_tensors"] = len(ref_shapes)
    report["n_params"] = int(sum(np.prod(s) for s in ref_shapes.values()))
    with open(os.path.join(HERE, "param_shapes.json"), "w") as f:
        json.dump({k: list(v) for k, v in wavbest_ref.param_shapes(chans).items()}, f, indent=0)
    sd = wavbest_ref.seeded_weights(chans, seed=7)
    net.load_state_dict(sd, strict=True)

    # ---- one denoiser evaluation ---------------------------------------
    for tag, (B, N, HW, sensor) in {"wv3_b1": (1, 8, 64, "WV3"), "qb_b2": (2, 4, 32, "QB")}.items():
        pan, ms, x_T = make_inputs(B, N, HW, HW)
        t_in = torch.tensor([999.0, 476.19][:B])
        with torch.no_grad():
            y_ref = net(x_T, t_in, PAN=pan, MS=ms, prompt=sensor)
        taps = {}
        y_f = wavbest_ref.wavbest_forward(sd, x_T, t_in, pan, ms, EMB[sensor], faithful=True, taps=taps)
        y_s = wavbest_ref.wavbest_forward(sd, x_T, t_in, pan, ms, EMB[sensor], faithful=False)
        report["fwd_%s_faithful_rel" % tag] = rel(y_f, y_ref)
        report["fwd_%s_shared_rel" % tag] = rel(y_s, y_ref)
        report["fwd_%s_absmean" % tag] = float(y_ref.abs().mean())
        out["fwd_%s_t" % tag] = t_in.numpy()
        out["fwd_%s_out" % tag] = y_ref.numpy()
        if tag == "qb_b2":
            for k in ("h0_0", "h3_1", "h0", "h3", "m", "u1", "u3", "embed", "prompt"):
                out["fwd_qb_b2_" + k] = taps[k].numpy()

    # ---- schedule scalars ---------------------------------------------
    G = GeneralDiffusion(net, loss_type="l1")
    G.set_new_noise_schedule({"schedule": "cosine", "n_timestep": 1000}, "cpu")
    betas_mine = dpm_ref.cosine_betas(1000)
    report["betas_bit_exact"] = bool(torch.equal(betas_mine, G.betas))
    ns_ref = NoiseScheduleVP("discrete", betas=G.betas)
    ns = dpm_ref.ScheduleVP(betas_mine)
    tt = torch.cat([torch.linspace(1e-3, 1.0, 257), torch.tensor([1e-3, 0.5, 1.0, 0.0005, 0.37])])
    for name, a, b in (("log_alpha", ns.log_alpha(tt), ns_ref.marginal_log_mean_coeff(tt)),
                       ("sigma", ns.sigma(tt), ns_ref.marginal_std(tt)),
                       ("lambda", ns.lam(tt), ns_ref.marginal_lambda(tt)),
                       ("inv_lambda", ns.inv_lam(ns.lam(tt)), ns_ref.inverse_lambda(ns_ref.marginal_lambda(tt)))):
        report["sched_%s_bit_exact" % name] = bool(torch.equal(a, b))
        report["sched_%s_maxabs" % name] = float((a - b).abs().max())
    out["betas"] = G.betas.numpy()
    out["sched_t"] = tt.numpy()
    out["sched_log_alpha"] = ns_ref.marginal_log_mean_coeff(tt).numpy()
    out["sched_lambda"] = ns_ref.marginal_lambda(tt).numpy()

    # ---- full trajectories --------------------------------------------
    for tag, (B, N, HW, sensor, steps) in {"cfg1_wv3_10": (1, 8, 64, "WV3", 10),
                                           "qb32_20": (2, 4, 32, "QB", 20)}.items():
        pan, ms, x_T = make_inputs(B, N, HW, HW)
        # The reference's x_start branch multiplies a [B] alpha into [B,N,H,W] without
        # expand_dims (dpm_solver_pytorch.py:304-306), so it only runs at B=1 -- which is what
        # its loaders use (data/__init__.py:18-19).  A batch is therefore B independent B=1 runs.
        calls, finals = [], []
        t0 = time.time()
        for b in range(B):
            cb = []

            class Spy(torch.nn.Module):
                def forward(self, x, t, **kw):
                    y = net(x, t, **kw)
                    cb.append((t.clone(), x.clone(), y.clone()))
                    return y

            fn = model_wrapper(Spy(), ns_ref, model_type="x_start",
                               model_kwargs={"PAN": pan[b:b + 1], "MS": ms[b:b + 1], "prompt": sensor})
            sol = DPM_Solver(fn, ns_ref, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding")
            finals.append(sol.sample(x_T[b:b + 1], steps=steps, order=3, skip_type="logSNR", method="singlestep",
                                     denoise_to_zero=True))
            calls.append(cb)
        calls = [tuple(torch.cat([calls[b][i][j] for b in range(B)]) for j in range(3)) for i in range(steps + 1)]
        x_ref = torch.cat(finals)
        report["traj_%s_ref_seconds" % tag] = time.time() - t0
        trace = []
        smp = dpm_ref.Sampler(lambda x, t: wavbest_ref.wavbest_forward(sd, x, t, pan, ms, EMB[sensor], faithful=True),
                              ns, trace=trace)
        x_mine = smp.sample(x_T, steps=steps, order=3)
        assert len(trace) == len(calls) == steps + 1
        report["traj_%s_nfe" % tag] = len(calls)
        report["traj_%s_final_rel" % tag] = rel(x_mine, x_ref)
        report["traj_%s_t_input_maxabs" % tag] = max(float((a["t_input"] - c[0]).abs().max()) for a, c in zip(trace, calls))
        report["traj_%s_out_rel_per_nfe" % tag] = [rel(a["out"], c[2]) for a, c in zip(trace, calls)]
        report["traj_%s_x_rel_per_nfe" % tag] = [rel(a["x"], c[1]) for a, c in zip(trace, calls)]
        out["traj_%s_t_input" % tag] = np.stack([c[0].numpy() for c in calls])
        out["traj_%s_final" % tag] = x_ref.numpy()
        keep = range(len(calls)) if HW == 32 else (0, steps // 2, steps)
        for i in keep:
            out["traj_%s_x_%d" % (tag, i)] = calls[i][1].numpy()
            out["traj_%s_out_%d" % (tag, i)] = calls[i][2].numpy()

    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **out)
    with open(os.path.join(HERE, "oracle_vs_reference.json"), "w") as f:
        json.dump(report, f, indent=1)
    print(json.dumps(report, indent=1))


if __name__ == "__main__":
    main()
