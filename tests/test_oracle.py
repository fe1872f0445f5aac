"""The oracle (oracle/) against the golden vectors produced by the unmodified reference
(tests/golden/make_golden.py).  CPU only."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import dpm_ref, wavbest_ref
from oracle.inputs import make_inputs
from util import rel_l2

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def sd():
    return wavbest_ref.seeded_weights()


def test_param_inventory_matches_reference():
    ref = json.load(open(os.path.join(HERE, "golden", "param_shapes.json")))
    mine = {k: list(v) for k, v in wavbest_ref.param_shapes().items()}
    assert mine == ref
    assert len(mine) == 272 and sum(int(np.prod(v)) for v in mine.values()) == 30928513


def test_haar_against_reference(golden):
    x = torch.from_numpy(golden["haar_x"])
    for fn in (wavbest_ref.dwt2d, wavbest_ref.dwt2d_closed):
        for got, key in zip(fn(x), ("haar_LL", "haar_LH", "haar_HL", "haar_HH")):
            assert torch.allclose(got, torch.from_numpy(golden[key]), atol=1e-6)
    bands = [torch.from_numpy(golden[k]) for k in ("haar_LL", "haar_LH", "haar_HL", "haar_HH")]
    for fn in (wavbest_ref.idwt2d, wavbest_ref.idwt2d_closed):
        assert torch.allclose(fn(*bands), torch.from_numpy(golden["haar_rec"]), atol=1e-6)
    assert torch.allclose(wavbest_ref.idwt2d_closed(*wavbest_ref.dwt2d_closed(x)), x, atol=1e-6)  # round trip (DWT_IDWT_layer.py:650-667)


@pytest.mark.parametrize("tag,B,N,HW,sensor", [("qb_b2", 2, 4, 32, "QB"), ("wv3_b1", 1, 8, 64, "WV3")])
def test_denoiser_against_reference(golden, sd, tag, B, N, HW, sensor):
    pan, ms, x_T = make_inputs(B, N, HW, HW)
    t = torch.from_numpy(golden["fwd_%s_t" % tag])
    emb = wavbest_ref.seeded_text_embeddings()[sensor]
    ref = torch.from_numpy(golden["fwd_%s_out" % tag])
    taps = {}
    y = wavbest_ref.wavbest_forward(sd, x_T, t, pan, ms, emb, faithful=True, taps=taps)
    assert rel_l2(y, ref) < 1e-6
    assert rel_l2(wavbest_ref.wavbest_forward(sd, x_T, t, pan, ms, emb, faithful=False), ref) < 2e-6
    if tag == "qb_b2":
        for k in ("h0_0", "h3_1", "h0", "h3", "m", "u1", "u3", "embed", "prompt"):
            assert rel_l2(taps[k], torch.from_numpy(golden["fwd_qb_b2_" + k])) < 1e-6, k


def test_schedule_bit_exact(golden):
    betas = dpm_ref.cosine_betas(1000)
    assert np.array_equal(betas.numpy(), golden["betas"])
    ns = dpm_ref.ScheduleVP(betas)
    t = torch.from_numpy(golden["sched_t"])
    assert np.array_equal(ns.log_alpha(t).numpy(), golden["sched_log_alpha"])
    assert np.array_equal(ns.lam(t).numpy(), golden["sched_lambda"])


def test_quantile_threshold_matches_torch():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(3, 4, 16, 16, generator=g) * 1.5
    y = dpm_ref.dynamic_threshold(x)
    s = torch.quantile(x.abs().reshape(3, -1), 0.995, dim=1).clamp_min(1.0)
    assert torch.equal(y, torch.clamp(x, -s[:, None, None, None], s[:, None, None, None]) / s[:, None, None, None])


def test_trajectory_against_reference(golden, sd):
    """20-step (21 NFE) singlestep DPM-Solver++ on two 4-band 32x32 patches."""
    B, N, HW, steps = 2, 4, 32, 20
    pan, ms, x_T = make_inputs(B, N, HW, HW)
    emb = wavbest_ref.seeded_text_embeddings()["QB"]
    ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))
    trace = []
    smp = dpm_ref.Sampler(lambda x, t: wavbest_ref.wavbest_forward(sd, x, t, pan, ms, emb, faithful=False), ns, trace=trace)
    x = smp.sample(x_T, steps=steps, order=3)
    assert len(trace) == steps + 1
    t_ref = golden["traj_qb32_20_t_input"]
    for i, tr in enumerate(trace):
        assert np.array_equal(tr["t_input"].numpy(), t_ref[i])
        assert rel_l2(tr["out"], torch.from_numpy(golden["traj_qb32_20_out_%d" % i])) < 5e-6, i
    assert rel_l2(x, torch.from_numpy(golden["traj_qb32_20_final"])) < 5e-6


@pytest.mark.parametrize("tag", ["ms3_noise_12", "ms2_cfg_8", "ms3_xstart_9"])
def test_sampler_variants_against_reference(golden_samplers, sd, tag):
    """Multistep DPM-Solver++ (orders 2/3), model_type 'noise' and classifier-free guidance: the oracle restatement
    against the reference's own DPM_Solver / model_wrapper run (committed vectors)."""
    from conftest import SAMPLER_CASES
    N, HW, sensor, steps, order, mtype, g, dtz = SAMPLER_CASES[tag]
    pan, ms, x_T = make_inputs(1, N, HW, HW)
    emb = wavbest_ref.seeded_text_embeddings()[sensor]
    if g is not None:
        pan, ms = torch.cat([torch.zeros_like(pan), pan]), torch.cat([torch.zeros_like(ms), ms])
    trace = []
    smp = dpm_ref.Sampler(lambda x, t: wavbest_ref.wavbest_forward(sd, x, t, pan, ms, emb, faithful=True),
                          dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000)), trace=trace, model_type=mtype, guidance_scale=g)
    got = smp.sample_multistep(x_T, steps=steps, order=order, denoise_to_zero=dtz)
    assert len(trace) == golden_samplers[tag + "_t_input"].shape[0]
    for i, tr in enumerate(trace):
        assert np.array_equal(tr["t_input"].numpy(), golden_samplers[tag + "_t_input"][i])
        assert rel_l2(tr["out"], torch.from_numpy(golden_samplers["%s_out_%d" % (tag, i)])) < 2e-6
    assert rel_l2(got, torch.from_numpy(golden_samplers[tag + "_final"])) < 2e-6
