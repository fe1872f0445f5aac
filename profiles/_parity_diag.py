"""Per-evaluation parity of one patch of the benchmarked workload against the oracle: free-running trajectory (raw output, eps
from the trajectory's own x) and the x divergence.  Usage: python profiles/_parity_diag.py BANDS SENSOR [PATCH] [DTYPE]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import dpm_ref, wavbest_ref
from oracle.inputs import make_inputs
from tmdiff_b200 import GeneralDiffusion, WavBEST

bands, sensor = int(sys.argv[1]), sys.argv[2]
p = int(sys.argv[3]) if len(sys.argv) > 3 else 0
dtype = sys.argv[4] if len(sys.argv) > 4 else "bf16"
pan, ms, x_T = make_inputs(64, bands, 64, 64)
pan, ms, x_T = pan[p:p + 1], ms[p:p + 1], x_T[p:p + 1]
sd, emb = wavbest_ref.seeded_weights(), wavbest_ref.seeded_text_embeddings()
m = WavBEST(channels=[32, 64, 128, 256], dtype=dtype, text_embeddings=emb)
m.load_state_dict(sd, strict=True)
m = m.cuda()
G = GeneralDiffusion(m)
G.set_new_noise_schedule({"schedule": "cosine", "n_timestep": 1000}, "cuda")
img, (trace, trace_x) = G.sample_by_dpmsolver({"PAN": pan.cuda(), "MS": ms.cuda(), "Res": x_T}, sensor, steps=20, x_T=x_T.cuda(), return_trace="x")
trace, trace_x = trace.cpu(), trace_x.cpu()
ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))
tr = []
net = lambda x, t: wavbest_ref.wavbest_forward(sd, x, t, pan, ms, emb[sensor], faithful=False)
final = dpm_ref.Sampler(net, ns, trace=tr).sample(x_T, steps=20, order=3)
rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
print("lib", os.environ.get("TMDIFF_B200_LIB", "default"), "bands", bands, "patch", p, dtype)
print(" i   t_input    alpha     sigma   |out|rms  |eps|rms   rel_out   rel_eps   rel_x   rel_out_tf")
for i, rec in enumerate(tr):
    t = torch.tensor([rec["t"]])
    a, s = ns.alpha(t), ns.sigma(t)
    eps_c = (trace_x[i] - a * trace[i]) / s
    # teacher-forced: the CUDA network on the oracle's x
    out_tf = m(rec["x"].cuda(), rec["t_input"].cuda(), PAN=pan.cuda(), MS=ms.cuda(), prompt=sensor).cpu()
    print("%2d %9.3f %9.2e %9.2e %9.2e %9.2e %9.2e %9.2e %9.2e %9.2e" % (
        i, float(rec["t_input"][0]), float(a), float(s), float(rec["out"].pow(2).mean().sqrt()), float(rec["eps"].pow(2).mean().sqrt()),
        rel(trace[i], rec["out"]), rel(eps_c, rec["eps"]), rel(trace_x[i], rec["x"]), rel(out_tf, rec["out"])))
print("final rel", rel(img.cpu() - ms, final))
