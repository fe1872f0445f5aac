"""WavBEST on the GPU (through the drop-in module and the C ABI) against the reference's golden
vectors and the oracle."""
import pytest
import torch

from oracle import dpm_ref, wavbest_ref
from oracle.inputs import make_inputs
from util import psnr, rel_l2

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-5     # north_star: relative L2 <= 1e-5 in fp32 mode
BF16_TOL = 2e-2     # north_star: relative L2 <= 2e-2 in bf16 mode


def make_model(dtype, channels=(32, 64, 128, 256), seed=7):
    from tmdiff_b200 import WavBEST
    m = WavBEST(channels=list(channels), dtype=dtype, text_embeddings=wavbest_ref.seeded_text_embeddings())
    m.load_state_dict(wavbest_ref.seeded_weights(channels, seed=seed), strict=True)
    return m.cuda()


@pytest.fixture(scope="module")
def model_fp32(lib_built):
    return make_model("fp32")


@pytest.fixture(scope="module")
def model_bf16(lib_built):
    return make_model("bf16")


@pytest.mark.parametrize("tag,B,N,HW,sensor", [("qb_b2", 2, 4, 32, "QB"), ("wv3_b1", 1, 8, 64, "WV3")])
def test_forward_fp32_golden(golden, model_fp32, tag, B, N, HW, sensor):
    pan, ms, x_T = make_inputs(B, N, HW, HW)
    t = torch.from_numpy(golden["fwd_%s_t" % tag])
    y = model_fp32(x_T.cuda(), t.cuda(), PAN=pan.cuda(), MS=ms.cuda(), prompt=sensor)
    assert rel_l2(y, torch.from_numpy(golden["fwd_%s_out" % tag])) < FP32_TOL
    # [B,1] integer time, as p_mean_variance passes it (diffusion_general.py:160-164)
    ti = torch.tensor([[7]] * B)
    y2 = model_fp32(x_T.cuda(), ti.cuda(), pan.cuda(), ms.cuda(), sensor)
    want = wavbest_ref.wavbest_forward(wavbest_ref.seeded_weights(), x_T, ti.float().reshape(-1), pan, ms,
                                       wavbest_ref.seeded_text_embeddings()[sensor], faithful=False)
    assert rel_l2(y2, want) < FP32_TOL


@pytest.mark.parametrize("tag,B,N,HW,sensor", [("qb_b2", 2, 4, 32, "QB"), ("wv3_b1", 1, 8, 64, "WV3")])
def test_forward_bf16_golden(golden, model_bf16, tag, B, N, HW, sensor):
    pan, ms, x_T = make_inputs(B, N, HW, HW)
    t = torch.from_numpy(golden["fwd_%s_t" % tag])
    y = model_bf16(x_T.cuda(), t.cuda(), PAN=pan.cuda(), MS=ms.cuda(), prompt=sensor)
    assert rel_l2(y, torch.from_numpy(golden["fwd_%s_out" % tag])) < BF16_TOL


def test_small_channels_and_odd_band_count(lib_built):
    """channels (8,16,32,64), 3 bands, 16x24 patch, mixed per-sample prompts, vs the oracle."""
    ch = (8, 16, 32, 64)
    m = make_model("fp32", ch, seed=3)
    sd = wavbest_ref.seeded_weights(ch, seed=3)
    emb = wavbest_ref.seeded_text_embeddings()
    g = torch.Generator().manual_seed(9)
    B, N, H, W = 3, 3, 16, 24
    pan, ms, x = torch.rand(B, 1, H, W, generator=g), torch.rand(B, N, H, W, generator=g), torch.randn(B, N, H, W, generator=g)
    t = torch.tensor([999.0, 500.5, 0.0])
    prompts = ["QB", "WV3", "GF2"]
    y = m(x.cuda(), t.cuda(), PAN=pan.cuda(), MS=ms.cuda(), prompt=prompts)
    for b in range(B):
        want = wavbest_ref.wavbest_forward(sd, x[b:b + 1], t[b:b + 1], pan[b:b + 1], ms[b:b + 1], emb[prompts[b]], faithful=True)
        assert rel_l2(y[b:b + 1], want) < FP32_TOL, b


def test_condition_cache_is_refreshed(model_fp32):
    pan, ms, x_T = make_inputs(1, 4, 32, 32)
    pan, ms, x = pan.cuda(), ms.cuda(), x_T.cuda()
    t = torch.tensor([300.0]).cuda()
    y1 = model_fp32(x, t, PAN=pan, MS=ms, prompt="QB")
    y1b = model_fp32(x, t, PAN=pan, MS=ms, prompt="QB")
    assert torch.equal(y1, y1b)                     # deterministic, cached condition branch
    ms2 = ms.clone()
    ms2.mul_(0.5)
    y2 = model_fp32(x, t, PAN=pan, MS=ms2, prompt="QB")
    assert not torch.allclose(y1, y2)
    ms.mul_(0.5)                                    # in-place edit of the cached tensor is noticed
    assert torch.equal(model_fp32(x, t, PAN=pan, MS=ms, prompt="QB"), y2)
    y3 = model_fp32(x, t, PAN=pan, MS=ms, prompt="GF2")
    assert not torch.allclose(y2, y3)


def test_errors(model_fp32):
    pan, ms, x_T = make_inputs(1, 4, 32, 32)
    with pytest.raises(ValueError):
        model_fp32(x_T.cuda(), torch.zeros(1).cuda(), PAN=pan.cuda(), MS=ms.cuda(), prompt="LANDSAT")
    with pytest.raises(ValueError):
        model_fp32(x_T.cuda(), torch.zeros(1).cuda())
    with pytest.raises(RuntimeError):
        model_fp32(x_T.cuda(), torch.zeros(1).cuda(), PAN=pan, MS=ms, prompt="QB")     # CPU tensors: no fallback
    with pytest.raises(RuntimeError):
        bad = torch.zeros(1, 4, 36, 32).cuda()                                          # not a multiple of 8
        model_fp32(bad, torch.zeros(1).cuda(), PAN=bad[:, :1], MS=bad, prompt="QB")


# ------------------------------------------------------------------------------------------------
# full trajectories
# ------------------------------------------------------------------------------------------------
def _diffusion(model):
    from tmdiff_b200 import GeneralDiffusion
    G = GeneralDiffusion(model)
    G.set_new_noise_schedule({"schedule": "cosine", "n_timestep": 1000}, "cuda")
    return G


def test_trajectory_config1_fp32(golden, model_fp32):
    """BASELINE config 1: WV3 64x64, B=1, 10 steps (11 NFE), against the reference run."""
    pan, ms, x_T = make_inputs(1, 8, 64, 64)
    G = _diffusion(model_fp32)
    img, trace = G.sample_by_dpmsolver({"PAN": pan.cuda(), "MS": ms.cuda(), "Res": x_T}, "WV3", steps=10, x_T=x_T.cuda(),
                                       return_trace=True)
    assert trace.shape[0] == 11
    for i in (0, 5, 10):
        assert rel_l2(trace[i], torch.from_numpy(golden["traj_cfg1_wv3_10_out_%d" % i])) < FP32_TOL, i
    final = torch.from_numpy(golden["traj_cfg1_wv3_10_final"])
    assert rel_l2(img.cpu() - ms, final) < FP32_TOL
    assert psnr(img.cpu(), final + ms) >= 45.0


@pytest.mark.parametrize("dtype,tol", [("fp32", FP32_TOL), ("bf16", BF16_TOL)])
def test_trajectory_qb_20_steps(golden, model_fp32, model_bf16, dtype, tol):
    """21 NFE on two 4-band 32x32 patches: raw network output and eps at every evaluation."""
    model = model_fp32 if dtype == "fp32" else model_bf16
    pan, ms, x_T = make_inputs(2, 4, 32, 32)
    G = _diffusion(model)
    img, trace = G.sample_by_dpmsolver({"PAN": pan.cuda(), "MS": ms.cuda(), "Res": x_T}, "QB", steps=20, x_T=x_T.cuda(),
                                       return_trace=True)
    ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))
    worst_out = worst_eps = 0.0
    for i in range(21):
        ref_out = torch.from_numpy(golden["traj_qb32_20_out_%d" % i])
        ref_x = torch.from_numpy(golden["traj_qb32_20_x_%d" % i])
        worst_out = max(worst_out, rel_l2(trace[i], ref_out))
        # eps as the reference forms it from (x, out) at this evaluation (dpm_solver_pytorch.py:304-306)
        t = (torch.from_numpy(golden["traj_qb32_20_t_input"][i][:1]) / 1000.0 + 1.0 / 1000)
        a, s = ns.alpha(t), ns.sigma(t)
        eps_ref = (ref_x - a * ref_out) / s
        eps_got = (ref_x - a * trace[i].cpu()) / s
        worst_eps = max(worst_eps, rel_l2(eps_got, eps_ref))
    assert worst_out < tol, worst_out
    assert worst_eps < tol, worst_eps
    final = torch.from_numpy(golden["traj_qb32_20_final"])
    assert psnr(img.cpu(), final + ms) >= 45.0
    if dtype == "fp32":
        assert rel_l2(img.cpu() - ms, final) < tol


def test_batch_is_independent_patches(model_fp32):
    """A batch equals its patches run one at a time; also covers the multi-GPU shard semantics."""
    pan, ms, x_T = make_inputs(3, 4, 32, 32)
    G = _diffusion(model_fp32)
    full = G.sample_by_dpmsolver({"PAN": pan.cuda(), "MS": ms.cuda(), "Res": x_T}, "QB", steps=10, x_T=x_T.cuda())
    one = G.sample_by_dpmsolver({"PAN": pan[1:2].cuda(), "MS": ms[1:2].cuda(), "Res": x_T[1:2]}, "QB", steps=10, x_T=x_T[1:2].cuda())
    assert rel_l2(full[1:2], one) < 1e-6


def test_reference_call_surface(model_fp32):
    """model_wrapper -> model_fn(x, t) and p_sample, as the reference drives the module."""
    from tmdiff_b200 import NoiseScheduleVP, model_wrapper
    pan, ms, x_T = make_inputs(1, 4, 32, 32)
    G = _diffusion(model_fp32)
    ns = NoiseScheduleVP("discrete", betas=G.betas)
    fn = model_wrapper(model_fp32, ns, model_type="x_start", model_kwargs={"PAN": pan.cuda(), "MS": ms.cuda(), "prompt": "QB"})
    t = torch.tensor([0.5]).cuda()
    eps = fn(x_T.cuda(), t)
    sd, emb = wavbest_ref.seeded_weights(), wavbest_ref.seeded_text_embeddings()["QB"]
    o_ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))
    out = wavbest_ref.wavbest_forward(sd, x_T, (t.cpu() - 1 / 1000) * 1000, pan, ms, emb, faithful=False)
    want = (x_T - o_ns.alpha(t.cpu()) * out) / o_ns.sigma(t.cpu())
    assert rel_l2(eps, want) < FP32_TOL
    x1 = G.p_sample(x_T.cuda(), 0, condition_x={"PAN": pan.cuda(), "MS": ms.cuda()}, prompt="QB")
    assert x1.shape == x_T.shape and torch.isfinite(x1).all()


def test_ancestral_step_fused(model_fp32):
    """p_sample's fused update == the reference's torch expressions (p_mean_variance + noise * exp(0.5 log_var), :155-201) for
    the same noise tensor, bit for bit; the in-kernel Philox draw is standard normal, differs per step and is reproducible."""
    pan, ms, x_T = [t.cuda() for t in make_inputs(2, 4, 32, 32)]
    G = _diffusion(model_fp32)
    cond = {"PAN": pan, "MS": ms}
    for t in (999, 500, 1, 0):
        z = torch.randn(x_T.shape, generator=torch.Generator().manual_seed(t)).cuda()
        mean, log_var = G.p_mean_variance(x=x_T, t=t, clip_denoised=True, x_in=cond, prompt="QB")
        want = mean + (z if t > 0 else torch.zeros_like(z)) * (0.5 * log_var).exp()
        got = G.p_sample(x_T, t, condition_x=cond, prompt="QB", noise=z)
        assert rel_l2(got, want) < 1e-6
    from tmdiff_b200 import ops
    x = torch.zeros(64, 4, 64, 64).cuda()
    a = ops.ancestral_update(x, x, None, 1.0, 0.0, 0.0, 0.0, 1.0, True, seed=7, step=3)      # out = z
    b = ops.ancestral_update(x, x, None, 1.0, 0.0, 0.0, 0.0, 1.0, True, seed=7, step=4)
    assert torch.equal(a, ops.ancestral_update(x, x, None, 1.0, 0.0, 0.0, 0.0, 1.0, True, seed=7, step=3))
    assert abs(float(a.mean())) < 5e-3 and abs(float(a.var()) - 1) < 1e-2 and abs(float((a * b).mean())) < 5e-3
    assert abs(float((a ** 4).mean()) - 3) < 0.1                                              # kurtosis of a normal
    torch.manual_seed(0)
    r1 = G.p_sample_loop_steps(cond, x_T, [5, 4, 3, 2, 1, 0], prompt="QB")
    torch.manual_seed(0)
    r2 = G.p_sample_loop_steps(cond, x_T, [5, 4, 3, 2, 1, 0], prompt="QB")
    assert torch.equal(r1, r2) and torch.isfinite(r1).all()


@pytest.mark.parametrize("tag", ["ms3_noise_12", "ms2_cfg_8", "ms3_xstart_9"])
@pytest.mark.parametrize("dtype,tol", [("fp32", FP32_TOL), ("bf16", BF16_TOL)])
def test_sampler_variants(golden_samplers, model_fp32, model_bf16, tag, dtype, tol):
    """sample_by_dpmsolver_noise / _guidance call shapes (multistep orders 2/3, model_type 'noise', classifier-free
    guidance as one 2B batch) and multistep over the x_start wrapper, against the reference's own runs."""
    from conftest import SAMPLER_CASES
    from tmdiff_b200 import DPM_Solver, NoiseScheduleVP, model_wrapper
    N, HW, sensor, steps, order, mtype, g, dtz = SAMPLER_CASES[tag]
    model = model_fp32 if dtype == "fp32" else model_bf16
    pan, ms, x_T = [t.cuda() for t in make_inputs(1, N, HW, HW)]
    G = _diffusion(model)
    ns = NoiseScheduleVP("discrete", betas=G.betas)
    if g is None:
        fn = model_wrapper(model, ns, model_type=mtype, model_kwargs={"PAN": pan, "MS": ms, "prompt": sensor})
    else:
        fn = model_wrapper(model, ns, model_type=mtype, model_kwargs={"MS": torch.cat([torch.zeros_like(ms), ms]), "prompt": sensor},
                           guidance_type="classifier-free", condition=pan, unconditional_condition=torch.zeros_like(pan),
                           guidance_scale=g)
    sol = DPM_Solver(fn, ns, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding")
    got, trace = sol.sample(x_T, steps=steps, order=order, skip_type="logSNR", method="multistep", denoise_to_zero=dtz,
                            return_trace=True)
    n_eval = golden_samplers[tag + "_t_input"].shape[0]
    assert trace.shape[0] == n_eval and trace.shape[1] == (2 if g is not None else 1)
    worst = max(rel_l2(trace[i], torch.from_numpy(golden_samplers["%s_out_%d" % (tag, i)])) for i in range(n_eval))
    assert worst < tol, worst
    final = torch.from_numpy(golden_samplers[tag + "_final"])
    assert psnr(got.cpu() + ms.cpu(), final + ms.cpu()) >= 45.0
    if dtype == "fp32":
        assert rel_l2(got, final) < tol


def test_sampler_entry_points(model_fp32):
    """The GeneralDiffusion methods themselves: shapes, finiteness, and guidance == 1 equals the conditional run."""
    pan, ms, x_T = [t.cuda() for t in make_inputs(2, 4, 32, 32)]
    G = _diffusion(model_fp32)
    x_in = {"PAN": pan, "MS": ms, "Res": x_T}
    a = G.sample_by_dpmsolver_noise(x_in, "QB", steps=6, x_T=x_T)
    b = G.sample_by_dpmsolver_guidance(x_in, "QB", 1.0, steps=6, x_T=x_T)
    c = G.sample_by_dpmsolver_guidance(x_in, "QB", 3.0, steps=6, x_T=x_T)
    r = G.sample_by_regression(x_in, "QB")
    for t in (a, b, c, r):
        assert t.shape == ms.shape and bool(torch.isfinite(t).all())
    # guidance_scale == 1 -> noise_pred_fn(x, t, cond=condition) only (dpm_solver_pytorch.py:336-337)
    from tmdiff_b200 import DPM_Solver, NoiseScheduleVP, model_wrapper
    ns = NoiseScheduleVP("discrete", betas=G.betas)
    fn = model_wrapper(model_fp32, ns, model_type="noise", model_kwargs={"PAN": pan, "MS": ms, "prompt": "QB"})
    want = DPM_Solver(fn, ns, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding").sample(
        x_T, steps=6, order=2, skip_type="logSNR", method="multistep") + ms
    assert rel_l2(b, want) < 1e-6
    assert rel_l2(c, want) > 1e-4


def test_mixed_sensor_batch(model_fp32):
    """BASELINE config 3: per-sample prompts inside one band bucket equal the single-prompt runs of each patch."""
    pan, ms, x_T = [t.cuda() for t in make_inputs(3, 4, 32, 32)]
    G = _diffusion(model_fp32)
    prompts = ["QB", "GF2", "WV2"]
    joint = G.sample_by_dpmsolver({"PAN": pan, "MS": ms, "Res": x_T}, prompts, steps=6, x_T=x_T)
    for i, p in enumerate(prompts):
        one = G.sample_by_dpmsolver({"PAN": pan[i:i + 1], "MS": ms[i:i + 1], "Res": x_T[i:i + 1]}, p, steps=6, x_T=x_T[i:i + 1])
        assert rel_l2(joint[i:i + 1], one) < 1e-6, p
    same = G.sample_by_dpmsolver({"PAN": pan, "MS": ms, "Res": x_T}, "QB", steps=6, x_T=x_T)
    assert rel_l2(joint[1:2], same[1:2]) > 1e-4          # the text modulation really differs between sensors


# ------------------------------------------------------------------------------------------------
# the benchmarked workload itself (BASELINE config 2 / the 8-band twin): B = 64, 64x64, 20 steps
# ------------------------------------------------------------------------------------------------
def _oracle_trace(p, pan, ms, x_T, sensor, steps):
    sd, emb = wavbest_ref.seeded_weights(), wavbest_ref.seeded_text_embeddings()[sensor]
    ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))
    tr = []
    net = lambda x, t: wavbest_ref.wavbest_forward(sd, x, t, pan[p:p + 1], ms[p:p + 1], emb, faithful=False)
    final = dpm_ref.Sampler(net, ns, trace=tr).sample(x_T[p:p + 1], steps=steps, order=3)
    return tr, final, ns


@pytest.mark.parametrize("bands,sensor", [(4, "QB"), (8, "WV3")])
def test_benchmarked_shape_parity(lib_built, bands, sensor):
    """What bench.py times -- 64 patches of 64x64, 20 DPM-Solver++ steps, bf16, CUDA-graph replay -- against the oracle run at
    B = 1 on patches {0, 31, 63} (batch independence: test_batch_is_independent_patches): raw network output and eps at every
    evaluation, eps formed from the x the CUDA trajectory itself fed to the network (core/dpm_solver_pytorch.py:296-306,
    :447-456), final image PSNR; the fp32 mode of the same batch is the second witness at 1e-5."""
    from tmdiff_b200 import WavBEST
    B, HW, steps = 64, 64, 20
    pan, ms, x_T = make_inputs(B, bands, HW, HW)                 # seed 3407: rank 0's batch in bench.py
    x_in = {"PAN": pan.cuda(), "MS": ms.cuda(), "Res": x_T}
    results = {}
    for dtype in ("bf16", "fp32"):
        m = WavBEST(channels=[32, 64, 128, 256], dtype=dtype, text_embeddings=wavbest_ref.seeded_text_embeddings(), max_batch=B)
        m.load_state_dict(wavbest_ref.seeded_weights(), strict=True)
        m = m.cuda()
        G = _diffusion(m)
        img_graph = G.sample_by_dpmsolver(x_in, sensor, steps=steps, x_T=x_T.cuda())          # captures the graph
        m.reset_condition_cache()
        img_replay = G.sample_by_dpmsolver(x_in, sensor, steps=steps, x_T=x_T.cuda())         # replays it
        assert torch.equal(img_graph, img_replay)
        m.reset_condition_cache()
        img, (trace, trace_x) = G.sample_by_dpmsolver(x_in, sensor, steps=steps, x_T=x_T.cuda(), return_trace="x")
        assert torch.equal(img, img_graph)                        # the graph replays exactly the eager launch list
        assert trace.shape == (steps + 1, B, bands, HW, HW) and trace_x.shape == trace.shape
        results[dtype] = (img.cpu(), trace.cpu(), trace_x.cpu())
        m.release()
    for p in (0, 31, 63):
        tr, final, ns = _oracle_trace(p, pan, ms, x_T, sensor, steps)
        assert len(tr) == steps + 1
        for dtype, tol in (("bf16", BF16_TOL), ("fp32", FP32_TOL)):
            img, trace, trace_x = results[dtype]
            worst_out = worst_eps = 0.0
            for i, rec in enumerate(tr):
                t = torch.tensor([rec["t"]])
                a, s = ns.alpha(t), ns.sigma(t)
                eps_cuda = (trace_x[i, p] - a * trace[i, p]) / s          # from the CUDA trajectory's own x
                worst_out = max(worst_out, rel_l2(trace[i, p], rec["out"][0]))
                worst_eps = max(worst_eps, rel_l2(eps_cuda, rec["eps"][0]))
            assert worst_out < tol, (dtype, p, worst_out)
            assert worst_eps < tol, (dtype, p, worst_eps)
            assert psnr(img[p], final[0] + ms[p]) >= 45.0, (dtype, p)
    # every patch of the batch: the bf16 image against the fp32 one
    for p in range(B):
        assert psnr(results["bf16"][0][p], results["fp32"][0][p]) >= 45.0, p


def test_large_patch_8band_128(lib_built):
    """The denoiser is fully convolutional (the reference's self-test runs 512x512, Hyper_unet_general.py:664-671) and
    dynamic_thresholding_fn has no size limit: an 8-band 128x128 patch (131072 values per sample) samples."""
    pan, ms, x_T = make_inputs(1, 8, 128, 128)
    m = make_model("fp32")
    G = _diffusion(m)
    img, trace = G.sample_by_dpmsolver({"PAN": pan.cuda(), "MS": ms.cuda(), "Res": x_T}, "WV3", steps=4, x_T=x_T.cuda(), return_trace=True)
    tr, final, _ = _oracle_trace(0, pan, ms, x_T, "WV3", 4)
    for i, rec in enumerate(tr):
        assert rel_l2(trace[i], rec["out"]) < FP32_TOL, i
    assert rel_l2(img.cpu() - ms, final) < FP32_TOL
    m.release()


def test_condition_cache_survives_address_reuse(model_fp32):
    """Non-contiguous / fp16 PAN and MS are converted copies inside the module; when the caller frees the originals the
    caching allocator hands the next batch the same addresses with version 0 -- the cache must not hit."""
    pan, ms, x_T = make_inputs(1, 4, 32, 32)
    t = torch.tensor([300.0]).cuda()
    x = x_T.cuda()
    outs = []
    for scale in (1.0, 0.5):
        pan_h, ms_h = (pan * scale).cuda().half(), (ms * scale).cuda().half()
        outs.append(model_fp32(x, t, PAN=pan_h, MS=ms_h, prompt="QB"))
        want = model_fp32(x, t, PAN=pan_h.float().clone(), MS=ms_h.float().clone(), prompt="QB")
        assert torch.equal(outs[-1], want)
        del pan_h, ms_h, want
    assert not torch.allclose(outs[0], outs[1])
    # weights: a second load_state_dict is noticed, and refresh_weights() covers `.data` writes
    sd = wavbest_ref.seeded_weights()
    y0 = model_fp32(x, t, PAN=pan.cuda(), MS=ms.cuda(), prompt="QB")
    p = model_fp32.get_parameter("final.conv24.weight")
    p.data.mul_(2.0)
    model_fp32.refresh_weights()
    y1 = model_fp32(x, t, PAN=pan.cuda(), MS=ms.cuda(), prompt="QB")
    assert rel_l2(y1, 2.0 * y0) < 1e-6
    model_fp32.load_state_dict(sd, strict=True)
    assert torch.equal(model_fp32(x, t, PAN=pan.cuda(), MS=ms.cuda(), prompt="QB"), y0)


def test_sample_validates_x_against_condition(model_fp32):
    """DPM_Solver.sample raises (instead of reading out of bounds) when x does not match the bound condition."""
    from tmdiff_b200 import DPM_Solver, NoiseScheduleVP, model_wrapper
    pan, ms, x_T = [t.cuda() for t in make_inputs(2, 4, 32, 32)]
    G = _diffusion(model_fp32)
    ns = NoiseScheduleVP("discrete", betas=G.betas)
    fn = model_wrapper(model_fp32, ns, model_type="x_start", model_kwargs={"PAN": pan, "MS": ms, "prompt": "QB"})
    sol = DPM_Solver(fn, ns, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding")
    for bad in (x_T[:1], x_T[:, :2], torch.cat([x_T, x_T]), x_T[..., :16]):
        with pytest.raises(ValueError):
            sol.sample(bad, steps=3, order=3, skip_type="logSNR", method="singlestep")


# ------------------------------------------------------------------------------------------------
# ancestral sampler against the reference's own p_mean_variance / p_sample / p_sample_loop runs
# (tests/golden/make_golden_ancestral.py; GeneralModel/diffusion_general.py:155-225, 338-339)
# ------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def golden_ancestral():
    import numpy as np
    import os
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ancestral_vectors.npz"))


def test_ancestral_steps_match_reference(golden_ancestral, model_fp32):
    """p_mean_variance and p_sample(noise = the reference's draw) at t in {999, 500, 1, 0}."""
    g = golden_ancestral
    pan, ms, x_T = [t.cuda() for t in make_inputs(2, 4, 32, 32)]
    G = _diffusion(model_fp32)
    cond = {"PAN": pan, "MS": ms}
    for t in (999, 500, 1, 0):
        x = torch.from_numpy(g["step_%d_x" % t]).cuda()
        mean, log_var = G.p_mean_variance(x=x, t=t, clip_denoised=True, x_in=cond, prompt="QB")
        assert rel_l2(mean, torch.from_numpy(g["step_%d_mean" % t])) < FP32_TOL, t
        assert float(log_var) == pytest.approx(float(g["step_%d_log_var" % t]), rel=1e-6)
        out = G.p_sample(x, t, condition_x=cond, prompt="QB", noise=torch.from_numpy(g["step_%d_z" % t]).cuda())
        assert rel_l2(out, torch.from_numpy(g["step_%d_out" % t])) < FP32_TOL, t


def test_ancestral_loop_matches_reference(golden_ancestral, model_fp32):
    """The last six steps of the 1000-step loop, and the reference's own p_sample_loop / super_resolution on a 6-step
    schedule (continous = True / False; super_resolution passes its prompt into `continous`), replaying the recorded draws."""
    from tmdiff_b200 import GeneralDiffusion
    g = golden_ancestral
    pan, ms, x_T = [t.cuda() for t in make_inputs(2, 4, 32, 32)]
    G = _diffusion(model_fp32)
    cond = {"PAN": pan, "MS": ms}
    zs = iter([torch.from_numpy(g["tail_z_%d" % j]) for j in range(5)])
    out = G.p_sample_loop_steps(cond, torch.from_numpy(g["tail_x5"]).cuda(), [5, 4, 3, 2, 1, 0], prompt="QB", noise_source=lambda s: next(zs))
    assert rel_l2(out, torch.from_numpy(g["tail_out"])) < FP32_TOL
    G6 = GeneralDiffusion(model_fp32)
    G6.set_new_noise_schedule({"schedule": "cosine", "n_timestep": 6}, "cuda")
    assert torch.equal(G6.betas.cpu(), torch.from_numpy(g["loop6_betas"]))
    cond1 = {"PAN": pan[:1], "MS": ms[:1], "Res": x_T[:1]}
    replay = lambda: (lambda it: (lambda s: next(it)))(iter([torch.from_numpy(g["loop6_noise_%d" % j]) for j in range(6)]))
    cont = G6.p_sample_loop(cond1, continous=True, prompt="QB", noise_source=replay())
    want = torch.from_numpy(g["loop6_continuous"])
    assert cont.shape == want.shape
    assert rel_l2(cont, want) < FP32_TOL
    last = G6.p_sample_loop(cond1, continous=False, prompt="QB", noise_source=replay())
    assert last.shape == tuple(g["loop6_last"].shape) and rel_l2(last, torch.from_numpy(g["loop6_last"])) < FP32_TOL
    # super_resolution(x_in, continous, prompt, guidance) == p_sample_loop(x_in, prompt): the whole series (:338-339)
    torch.manual_seed(0)
    sr = G6.super_resolution(cond1, False, "QB", 1.0)
    assert sr.shape == want.shape


def test_mixed_sensor_two_band_buckets(lib_built):
    """BASELINE config 3 in miniature: a joint batch of 8-band (WV3) and 4-band (QB, GF2) patches with per-sample prompts through
    parallel.sample_mixed -- one engine per band count, chunked buckets -- equals the oracle run patch by patch."""
    from tmdiff_b200 import parallel
    sd, emb = wavbest_ref.seeded_weights(), wavbest_ref.seeded_text_embeddings()
    m = make_model("fp32")
    G = _diffusion(m)
    p8, m8, x8 = make_inputs(3, 8, 32, 32, seed=11)
    p4, m4, x4 = make_inputs(4, 4, 32, 32, seed=12)
    order = [("WV3", 8, 0), ("QB", 4, 0), ("GF2", 4, 1), ("WV3", 8, 1), ("QB", 4, 2), ("WV3", 8, 2), ("GF2", 4, 3)]
    patches, noises = [], []
    for sensor, n, i in order:
        pan, ms, x = (p8, m8, x8) if n == 8 else (p4, m4, x4)
        patches.append((pan[i].cuda(), ms[i].cuda(), sensor))
        noises.append(x[i])
    got = parallel.sample_mixed(G, patches, steps=6, x_T=noises, batch=2)       # chunks of 2: buckets of 3 and 4 split unevenly
    assert len(m._engines) == 2                                                 # one engine per band count, both alive
    ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))
    for k, (sensor, n, i) in enumerate(order):
        pan, ms, x = (p8, m8, x8) if n == 8 else (p4, m4, x4)
        net = lambda xx, t: wavbest_ref.wavbest_forward(sd, xx, t, pan[i:i + 1], ms[i:i + 1], emb[sensor], faithful=False)
        want = dpm_ref.Sampler(net, ns).sample(x[i:i + 1], steps=6, order=3)[0] + ms[i]
        assert got[k].shape == want.shape
        assert rel_l2(got[k].cpu() - ms[i], want - ms[i]) < FP32_TOL, (k, sensor)
    m.release()


def test_arena_redzones_intact(lib_built, monkeypatch):
    """The library's own bounds check of its workspace (compute-sanitizer is closed on this pool, profiles/sanitizer_r02.md):
    with TMD_ARENA_REDZONE=1 every activation tensor is followed by a canary-filled red zone; none may be touched by a forward
    pass or a full sampling run -- bf16 and fp32, 4 and 8 bands, a patch that is not a multiple of the 16 x 8 tile."""
    monkeypatch.setenv("TMD_ARENA_REDZONE", "1")
    for dtype in ("bf16", "fp32"):
        for B, N, H, W, sensor in ((3, 4, 32, 32, "QB"), (2, 8, 64, 64, "WV3"), (2, 4, 24, 40, "GF2")):
            m = make_model(dtype)
            G = _diffusion(m)
            g = torch.Generator().manual_seed(B + N)
            pan, ms, x = torch.rand(B, 1, H, W, generator=g), torch.rand(B, N, H, W, generator=g), torch.randn(B, N, H, W, generator=g)
            img = G.sample_by_dpmsolver({"PAN": pan.cuda(), "MS": ms.cuda(), "Res": x}, sensor, steps=4, x_T=x.cuda())
            assert bool(torch.isfinite(img).all())
            assert m.check_redzones() > 40                      # every named buffer of the plan has one
            m.release()
