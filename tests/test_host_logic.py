"""Host-side logic of the product (no GPU): schedule scalars, sampler program, module
state-dict layout, the C-ABI library's exported symbols, multi-rank sharding."""
import ctypes
import json
import os
import re

import numpy as np
import pytest
import torch

from oracle import dpm_ref
from tmdiff_b200 import _lib, diffusion, dpm_solver, networks, ops, parallel
from util import run_plan_on_cpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def test_betas_and_schedule_match_reference_vectors(golden):
    betas = torch.tensor(diffusion.make_beta_schedule("cosine", 1000), dtype=torch.float32)
    assert np.array_equal(betas.numpy(), golden["betas"])
    ns = dpm_solver.NoiseScheduleVP("discrete", betas=betas)
    t = torch.from_numpy(golden["sched_t"])
    assert np.array_equal(ns.marginal_log_mean_coeff(t).numpy(), golden["sched_log_alpha"])
    assert np.array_equal(ns.marginal_lambda(t).numpy(), golden["sched_lambda"])
    with pytest.raises(ValueError):
        dpm_solver.NoiseScheduleVP("cosine")


class _FakeModel:
    pass


def _solver():
    betas = torch.tensor(diffusion.make_beta_schedule("cosine", 1000), dtype=torch.float32)
    ns = dpm_solver.NoiseScheduleVP("discrete", betas=betas)
    fn = dpm_solver.model_wrapper(_FakeModel(), ns, model_type="x_start", model_kwargs={})
    return dpm_solver.DPM_Solver(fn, ns, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding"), ns


@pytest.mark.parametrize("steps,orders", [(10, [3, 3, 3, 1]), (20, [3] * 6 + [2]), (30, [3] * 9 + [2, 1]), (11, [3, 3, 3, 2])])
def test_singlestep_orders(steps, orders):
    assert dpm_solver.singlestep_orders(steps, 3) == orders
    assert sum(orders) == steps


@pytest.mark.parametrize("tag,steps", [("cfg1_wv3_10", 10), ("qb32_20", 20)])
def test_plan_t_inputs_bit_exact_with_reference(golden, tag, steps):
    sol, _ = _solver()
    plan = sol.build_plan(steps=steps, order=3, skip_type="logSNR", method="singlestep", denoise_to_zero=True)
    assert plan.n_eval == steps + 1
    ref = golden["traj_%s_t_input" % tag][:, 0]
    assert np.array_equal(np.array(plan.t_inputs(), dtype=np.float32), ref)


def test_plan_reproduces_oracle_trajectory_on_a_toy_model():
    """Execute the straight-line program with torch on the CPU against the oracle's sampler for a
    cheap analytic 'network' -- checks every coefficient and slot assignment bit for bit."""
    sol, ns = _solver()
    net = lambda x, t_in: torch.tanh(x * 0.3) * (1.0 - t_in.reshape(-1, 1, 1, 1) / 2000.0)
    g = torch.Generator().manual_seed(1)
    xT = torch.randn(1, 4, 8, 8, generator=g)
    for steps in (10, 20, 11):
        want = dpm_ref.Sampler(net, dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))).sample(xT, steps=steps, order=3)
        plan = sol.build_plan(steps=steps, order=3, skip_type="logSNR", method="singlestep", denoise_to_zero=True)
        got = run_plan_on_cpu(plan, net, xT)
        assert torch.equal(got, want), steps


@pytest.mark.parametrize("steps,order,mtype,g,dtz", [(12, 3, "noise", None, True), (8, 2, "noise", 2.0, False),
                                                     (9, 3, "x_start", None, True), (50, 3, "noise", None, True),
                                                     (5, 3, "x_start", 1.5, True), (3, 1, "noise", None, False)])
def test_multistep_plan_reproduces_oracle_on_a_toy_model(steps, order, mtype, g, dtz):
    """The multistep / noise-type / classifier-free-guidance programs (sample_by_dpmsolver_noise, _guidance) against
    the oracle's restatement of DPM_Solver.sample(method='multistep'), bit for bit."""
    betas = torch.tensor(diffusion.make_beta_schedule("cosine", 1000), dtype=torch.float32)
    ns = dpm_solver.NoiseScheduleVP("discrete", betas=betas)
    fn = dpm_solver.model_wrapper(_FakeModel(), ns, model_type=mtype, model_kwargs={})
    sol = dpm_solver.DPM_Solver(fn, ns, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding")

    def net(x, t_in):
        y = torch.tanh(x * 0.3) * (1.0 - t_in.reshape(-1, 1, 1, 1) / 2000.0)
        if g is not None:                       # the two halves of a guidance batch see different conditions
            y = torch.cat([y[: x.shape[0] // 2] * 0.5, y[x.shape[0] // 2:]])
        return y

    xT = torch.randn(2, 4, 8, 8, generator=torch.Generator().manual_seed(2))
    want = dpm_ref.Sampler(net, dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000)), model_type=mtype,
                           guidance_scale=g).sample_multistep(xT, steps=steps, order=order, denoise_to_zero=dtz)
    plan = sol.build_plan(steps=steps, order=order, skip_type="logSNR", method="multistep", denoise_to_zero=dtz)
    assert plan.n_eval == steps + (1 if dtz else 0)
    assert all(0 <= v < _lib.NUM_SLOTS for st in plan.steps if st[0] == "update" for v in (st[1], st[2], st[3]))
    got = run_plan_on_cpu(plan, net, xT, model_type=mtype, guidance_scale=g)
    assert torch.equal(got, want)
    cp = plan.to_c(4 * 8 * 8, model_type=mtype, cfg=g is not None, guidance_scale=g or 1.0)
    assert cp.n_steps == len(plan.steps) and cp.model_type == (1 if mtype == "noise" else 0) and cp.cfg == (g is not None)


def test_model_wrapper_rejects_what_the_reference_samplers_never_use():
    betas = torch.tensor(diffusion.make_beta_schedule("cosine", 1000), dtype=torch.float32)
    ns = dpm_solver.NoiseScheduleVP("discrete", betas=betas)
    with pytest.raises(NotImplementedError):
        dpm_solver.model_wrapper(_FakeModel(), ns, model_type="v")
    with pytest.raises(NotImplementedError):
        dpm_solver.model_wrapper(_FakeModel(), ns, model_type="noise", guidance_type="classifier")
    with pytest.raises(AssertionError):
        dpm_solver.model_wrapper(_FakeModel(), ns, model_type="bogus")
    fn = dpm_solver.model_wrapper(_FakeModel(), ns, model_type="noise")
    sol = dpm_solver.DPM_Solver(fn, ns, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding")
    with pytest.raises(NotImplementedError):
        sol.build_plan(steps=10, order=3, method="adaptive")
    with pytest.raises(AssertionError):
        sol.build_plan(steps=2, order=3, method="multistep")        # `assert steps >= order` (:1210)


def test_quantile_position_matches_torch_quantile():
    g = torch.Generator().manual_seed(3)
    for n in (4 * 32 * 32, 4 * 64 * 64, 8 * 64 * 64, 1000):
        v = torch.rand(n, generator=g)
        lo, w = ops.quantile_position(n, 0.995)
        sv = torch.sort(v).values
        want = torch.quantile(v, 0.995)
        got = torch.lerp(sv[lo], sv[min(lo + 1, n - 1)], torch.tensor(w))
        assert torch.equal(got, want), n


def test_state_dict_layout_matches_reference():
    ref = json.load(open(os.path.join(HERE, "golden", "param_shapes.json")))
    m = networks.WavBEST(channels=[32, 64, 128, 256])
    sd = m.state_dict()
    assert {k: list(v.shape) for k, v in sd.items()} == ref
    assert networks.WavBEST().channels == [16, 32, 64, 128]            # reference default (:526-527)
    m.set_text_embedding("QB", torch.zeros(1, 768))
    assert m.get_embeding("QB").shape == (1, 768) and m.get_embeding("WV3") is None and m.get_embeding("nope") is None
    with pytest.raises(ValueError):
        m.set_text_embedding("LANDSAT", torch.zeros(1, 768))


def test_no_cpu_fallback():
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    m = networks.WavBEST(channels=[8, 16, 32, 64], text_embeddings={"QB": torch.zeros(1, 768)})
    x = torch.zeros(1, 4, 8, 8)
    with pytest.raises(RuntimeError):
        m(x, torch.zeros(1), PAN=torch.zeros(1, 1, 8, 8), MS=x, prompt="QB")
    with pytest.raises(RuntimeError):
        ops.DWT_2D("haar")(torch.zeros(1, 1, 8, 8))


def test_library_exports_every_declared_symbol(lib_built):
    header = open(os.path.join(ROOT, "include", "tmdiff_b200.h")).read()
    declared = set(re.findall(r"\b(tmd_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    lib = ctypes.CDLL(lib_built)
    for name in declared:
        assert hasattr(lib, name), name
    assert _lib.load().tmd_version() >= 1
    # argument validation happens before any device work
    h = ctypes.c_void_p()
    cfg = _lib.Config((ctypes.c_int32 * 4)(32, 64, 128, 256), 128, 32, 1, 4, 60, 64, 0)
    assert _lib.load().tmd_create(ctypes.byref(cfg), ctypes.byref(h)) != 0
    assert _lib.load().tmd_last_error()


def test_shard_range_partitions():
    for n in (0, 1, 7, 64, 513):
        for w in (1, 2, 3, 8):
            spans = [parallel.shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(e - s for s, e in spans) - min(e - s for s, e in spans) <= 1


def _worker(rank, world, port, n_total, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = torch.arange(n_total * 6, dtype=torch.float32).reshape(n_total, 2, 3)
    s, e = parallel.shard_range(n_total, rank, world)
    got = parallel.gather_tiles(full[s:e] * 2.0, n_total)       # "sampling" = x2 on the local shard
    q.put((rank, bool(torch.equal(got, full * 2.0))))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_total", [5, 8])
def test_two_rank_gather_gloo(n_total):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000 + n_total
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
    assert sorted(res) == [(0, True), (1, True)]


def test_bucket_by_bands_and_sample_mixed_order():
    g = torch.Generator().manual_seed(0)
    patches = []
    for i, (n, s) in enumerate([(8, "WV3"), (4, "QB"), (4, "GF2"), (8, "WV3"), (4, "QB")]):
        patches.append((torch.rand(1, 8, 8, generator=g), torch.rand(n, 8, 8, generator=g), s))
    b = parallel.bucket_by_bands(patches)
    assert sorted(b) == [4, 8] and b[4][0] == [1, 2, 4] and b[8][0] == [0, 3] and b[4][3] == ["QB", "GF2", "QB"]
    assert b[8][2].shape == (2, 8, 8, 8) and b[4][1].shape == (3, 1, 8, 8)

    class Fake:
        def sample_by_dpmsolver(self, x_in, prompt, steps=20, x_T=None):
            assert len(prompt) == x_in["MS"].shape[0]
            return x_in["MS"] + x_in["PAN"] + x_T

    noise = [torch.randn(p[1].shape, generator=g) for p in patches]
    out = parallel.sample_mixed(Fake(), patches, steps=3, x_T=noise)
    for o, (pan, ms, _), z in zip(out, patches, noise):
        assert torch.equal(o, ms + pan + z)


def test_ancestral_schedule_buffers_match_reference_run():
    """set_new_noise_schedule's buffers (GeneralModel/diffusion_general.py:86-132) against values the reference itself
    produced (tests/golden/make_golden_ancestral.py): the 6-step cosine betas and posterior_log_variance_clipped[t]."""
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ancestral_vectors.npz"))
    G = diffusion.GeneralDiffusion(None)
    G.set_new_noise_schedule({"schedule": "cosine", "n_timestep": 6}, "cpu")
    assert torch.equal(G.betas, torch.from_numpy(g["loop6_betas"]))
    G.set_new_noise_schedule({"schedule": "cosine", "n_timestep": 1000}, "cpu")
    for t in (999, 500, 1, 0):
        assert float(G.posterior_log_variance_clipped[t]) == float(g["step_%d_log_var" % t])
