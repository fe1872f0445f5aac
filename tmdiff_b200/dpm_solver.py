"""Host side of the DPM-Solver++ sampler: the reference's `NoiseScheduleVP` / `model_wrapper` /
`DPM_Solver` call surface (core/dpm_solver_pytorch.py:6-181, :184-348, :351-1289) for the
configuration family the reference's `sample_by_dpmsolver` uses, compiled into a straight-line
device program (`tmd_plan`) that `tmd_sample` executes without returning to Python.

All schedule scalars are evaluated here once, on the CPU, with fp32 torch ops in the reference's
order -- the reference re-derives them on the GPU at every step through `interpolate_fn`
(~390 tiny kernels per network evaluation).  They are bit-identical to the
reference's values (tests/test_host_logic.py), which matters because x0 = (x - sigma*eps)/alpha
divides by alpha_T ~ 5e-5.
"""
import ctypes as C

import torch

from . import _lib
from .ops import quantile_position

_CPU = torch.device("cpu")


def _interp(x, xp, yp):
    """Piecewise-linear y(x) on ascending keypoints with outer-segment extrapolation; final
    expression as in interpolate_fn (:1296-1335) so the fp32 rounding is the same."""
    xs, xp, yp = x.reshape(-1), xp.reshape(-1), yp.reshape(-1)
    K = xp.numel()
    idx = torch.searchsorted(xp, xs, right=False)
    lo = torch.where(idx == 0, torch.zeros_like(idx), torch.where(idx == K, torch.full_like(idx, K - 2), idx - 1))
    x0, x1, y0, y1 = xp[lo], xp[lo + 1], yp[lo], yp[lo + 1]
    return y0 + (xs - x0) * (y1 - y0) / (x1 - x0)


class NoiseScheduleVP:
    """`NoiseScheduleVP('discrete', betas=...)` or `alphas_cumprod=...` (:6-181).  The continuous
    'linear' / 'cosine' schedules are not on the reference's sampling path and are rejected."""

    def __init__(self, schedule="discrete", betas=None, alphas_cumprod=None, dtype=torch.float32):
        if schedule != "discrete":
            raise ValueError("Unsupported noise schedule {}: the TMDiff sampling path uses 'discrete'".format(schedule))
        if betas is not None:
            log_alphas = 0.5 * torch.log(1 - torch.as_tensor(betas).detach().to(_CPU, torch.float32)).cumsum(dim=0)
        else:
            assert alphas_cumprod is not None
            log_alphas = 0.5 * torch.log(torch.as_tensor(alphas_cumprod).detach().to(_CPU, torch.float32))
        self.schedule = schedule
        self.total_N = len(log_alphas)
        self.T = 1.0
        self.t_array = torch.linspace(0.0, 1.0, self.total_N + 1)[1:].reshape((1, -1)).to(dtype=dtype)
        self.log_alpha_array = log_alphas.reshape((1, -1)).to(dtype=dtype)

    def marginal_log_mean_coeff(self, t):
        t = torch.as_tensor(t)
        return _interp(t.detach().to(_CPU), self.t_array, self.log_alpha_array).to(t.device)

    def marginal_alpha(self, t):
        return torch.exp(self.marginal_log_mean_coeff(t))

    def marginal_std(self, t):
        return torch.sqrt(1.0 - torch.exp(2.0 * self.marginal_log_mean_coeff(t)))

    def marginal_lambda(self, t):
        la = self.marginal_log_mean_coeff(t)
        return la - 0.5 * torch.log(1.0 - torch.exp(2.0 * la))

    def inverse_lambda(self, lamb):
        lamb = torch.as_tensor(lamb)
        l_cpu = lamb.detach().to(_CPU)
        la = -0.5 * torch.logaddexp(torch.zeros((1,)), -2.0 * l_cpu)
        t = _interp(la, torch.flip(self.log_alpha_array, [1]), torch.flip(self.t_array, [1]))
        return t.to(lamb.device)


def model_wrapper(model, noise_schedule, model_type="noise", model_kwargs={}, guidance_type="uncond", condition=None,
                  unconditional_condition=None, guidance_scale=1.0, classifier_fn=None, classifier_kwargs={}):
    """:184-348.  Returns `model_fn(x, t_continuous)` (the noise prediction); the returned callable also
    carries what `DPM_Solver.sample` needs to run the whole trajectory on the device.

    Implemented: model_type 'x_start' (sample_by_dpmsolver) and 'noise' (sample_by_dpmsolver_noise /
    _guidance, diffusion_general.py:257-335); guidance_type 'uncond' and 'classifier-free' (:336-344; the
    condition is the PAN tensor, passed to the denoiser as its third positional argument, :299-302)."""
    assert model_type in ["noise", "x_start", "v", "score"]
    assert guidance_type in ["uncond", "classifier", "classifier-free"]
    if model_type not in ("x_start", "noise"):
        raise NotImplementedError("tmdiff_b200 implements model_type 'x_start' and 'noise' (the reference's samplers)")
    if guidance_type == "classifier":
        raise NotImplementedError("classifier guidance needs a backward pass; the reference's samplers do not use it")
    cfg_active = guidance_type == "classifier-free" and not (guidance_scale == 1.0 or unconditional_condition is None)

    def noise_pred_fn(x, t_continuous, cond=None):
        t_input = (t_continuous - 1.0 / noise_schedule.total_N) * 1000.0           # :291-292
        output = model(x, t_input, **model_kwargs) if cond is None else model(x, t_input, cond, **model_kwargs)
        if model_type == "noise":
            return output
        B = x.shape[0]
        # the reference broadcasts a [B] alpha over W (:304-306) and therefore only runs at B=1
        alpha_t = noise_schedule.marginal_alpha(t_continuous).reshape(B, 1, 1, 1)
        sigma_t = noise_schedule.marginal_std(t_continuous).reshape(B, 1, 1, 1)
        return (x - alpha_t * output) / sigma_t

    def model_fn(x, t_continuous):
        if guidance_type == "uncond":
            return noise_pred_fn(x, t_continuous)
        if not cfg_active:
            return noise_pred_fn(x, t_continuous, cond=condition)
        x_in = torch.cat([x] * 2)
        t_in = torch.cat([t_continuous] * 2)
        c_in = torch.cat([unconditional_condition, condition])
        noise_uncond, noise = noise_pred_fn(x_in, t_in, cond=c_in).chunk(2)
        return noise_uncond + guidance_scale * (noise - noise_uncond)

    model_fn.model = model
    model_fn.noise_schedule = noise_schedule
    model_fn.model_kwargs = dict(model_kwargs)
    model_fn.model_type = model_type
    model_fn.guidance_type = guidance_type
    model_fn.condition = condition
    model_fn.unconditional_condition = unconditional_condition
    model_fn.guidance_scale = float(guidance_scale)
    model_fn.cfg_active = cfg_active
    return model_fn


def singlestep_orders(steps, order):
    """get_orders_and_timesteps_for_singlestep_solver (:537-555), order list only."""
    if order == 3:
        K = steps // 3 + 1
        if steps % 3 == 0:
            return [3] * (K - 2) + [2, 1]
        if steps % 3 == 1:
            return [3] * (K - 1) + [1]
        return [3] * (K - 1) + [2]
    if order == 2:
        if steps % 2 == 0:
            return [2] * (steps // 2)
        return [2] * (steps // 2) + [1]
    if order == 1:
        return [1] * steps
    raise ValueError("'order' must be '1' or '2' or '3'.")


class SamplePlan:
    """A straight-line program over tensor slots: evaluations of the denoiser (+ x0 conversion and
    dynamic thresholding) and dpmsolver++ updates."""
    X, M0, M1, M2, XS = 0, 1, 2, 3, 4

    def __init__(self):
        self.steps = []
        self.result_slot = 0

    def eval(self, ns, src, dst, t):
        t = t.reshape(1)
        t_input = (t - 1.0 / ns.total_N) * 1000.0
        self.steps.append(("eval", src, dst, t_input.item(), ns.marginal_alpha(t).item(), ns.marginal_std(t).item()))

    def update(self, dst, x, m, c1, c2, m1=-1, c3=0.0, form=0, m2=-1, k=()):
        f = lambda v: float(v.item()) if torch.is_tensor(v) else float(v)
        k = [f(v) for v in k] + [0.0] * (5 - len(k))
        self.steps.append(("update", dst, x, m, m1, f(c1), f(c2), f(c3), int(form), m2, k))

    @property
    def n_eval(self):
        return sum(1 for s in self.steps if s[0] == "eval")

    def t_inputs(self):
        return [s[3] for s in self.steps if s[0] == "eval"]

    def to_c(self, n_per_sample, ratio=0.995, max_val=1.0, model_type="x_start", cfg=False, guidance_scale=1.0):
        arr = (_lib.Step * len(self.steps))()
        for i, s in enumerate(self.steps):
            if s[0] == "eval":
                arr[i].kind = 0
                arr[i].eval = _lib.Eval(s[1], s[2], s[3], s[4], s[5])
            else:
                arr[i].kind = 1
                arr[i].update = _lib.Update(*s[1:10], (C.c_float * 5)(*s[10]))
        q_lo, q_w = quantile_position(n_per_sample, ratio)
        plan = _lib.Plan(C.cast(arr, C.POINTER(_lib.Step)), len(self.steps), self.result_slot, q_w, q_lo, max_val,
                         _lib.MODEL_NOISE if model_type == "noise" else _lib.MODEL_X_START, 1 if cfg else 0,
                         float(guidance_scale))
        plan._keepalive = arr
        return plan


class DPM_Solver:
    """`DPM_Solver(model_fn, noise_schedule, algorithm_type="dpmsolver++",
    correcting_x0_fn="dynamic_thresholding")` (:351-428) with `.sample(...)` (:1081-1289) for
    method='singlestep' (orders 1-3, skip_type 'logSNR' / 'time_uniform')."""

    def __init__(self, model_fn, noise_schedule, algorithm_type="dpmsolver++", correcting_x0_fn=None, correcting_xt_fn=None,
                 thresholding_max_val=1.0, dynamic_thresholding_ratio=0.995):
        assert algorithm_type in ["dpmsolver", "dpmsolver++"]
        if algorithm_type != "dpmsolver++" or correcting_x0_fn != "dynamic_thresholding" or correcting_xt_fn is not None:
            raise NotImplementedError("tmdiff_b200 implements algorithm_type='dpmsolver++' with "
                                      "correcting_x0_fn='dynamic_thresholding' (the reference's sample_by_dpmsolver)")
        if not hasattr(model_fn, "model"):
            raise TypeError("model_fn must come from tmdiff_b200.dpm_solver.model_wrapper")
        self.model_fn_ = model_fn
        self.noise_schedule = noise_schedule
        self.algorithm_type = algorithm_type
        self.dynamic_thresholding_ratio = dynamic_thresholding_ratio
        self.thresholding_max_val = thresholding_max_val

    # ---- host-side schedule logic ------------------------------------------------------------
    def get_time_steps(self, skip_type, t_T, t_0, N, device=_CPU):
        ns = self.noise_schedule
        if skip_type == "logSNR":                                                   # :482-486
            lambda_T = ns.marginal_lambda(torch.tensor(t_T))
            lambda_0 = ns.marginal_lambda(torch.tensor(t_0))
            return ns.inverse_lambda(torch.linspace(lambda_T.item(), lambda_0.item(), N + 1))
        if skip_type == "time_uniform":
            return torch.linspace(t_T, t_0, N + 1)
        if skip_type == "time_quadratic":
            return torch.linspace(t_T ** 0.5, t_0 ** 0.5, N + 1).pow(2)
        raise ValueError("Unsupported skip_type {}, need to be 'logSNR' or 'time_uniform' or 'time_quadratic'".format(skip_type))

    def get_orders_and_timesteps_for_singlestep_solver(self, steps, order, skip_type, t_T, t_0, device=_CPU):
        orders = singlestep_orders(steps, order)
        if skip_type == "logSNR":
            outer = self.get_time_steps(skip_type, t_T, t_0, len(orders))
        else:
            outer = self.get_time_steps(skip_type, t_T, t_0, steps)[torch.cumsum(torch.tensor([0] + orders), 0)]
        return outer, orders

    def _plan_multistep(self, P, steps, order, skip_type, t_T, t_0, lower_order_final):
        """`sample`'s multistep branch (:1208-1252) + multistep_dpm_solver_{second,third}_update (:817-927) and
        dpm_solver_first_update (:563-608), dpmsolver++ / solver_type 'dpmsolver'.  The rotating
        model_prev_list becomes a rotation of three slots."""
        ns = self.noise_schedule
        assert steps >= order
        ts = self.get_time_steps(skip_type, t_T, t_0, steps)
        assert ts.shape[0] - 1 == steps
        X = P.X
        free = [P.M0, P.M1, P.M2]
        prev = []   # [(slot, t)] oldest first

        def eval_into_free(t):
            slot = free.pop(0)
            P.eval(ns, X, slot, t)
            prev.append((slot, t))

        def update(t, k):
            (m0, t0) = prev[-1]
            lam_0, lam_t = ns.marginal_lambda(t0), ns.marginal_lambda(t)
            sigma_0, sigma_t = ns.marginal_std(t0), ns.marginal_std(t)
            alpha_t = torch.exp(ns.marginal_log_mean_coeff(t))
            h = lam_t - lam_0
            phi_1 = torch.expm1(-h)
            if k == 1:
                P.update(X, X, m0, sigma_t / sigma_0, alpha_t * phi_1)
                return
            (m1, t1) = prev[-2]
            h_0 = lam_0 - ns.marginal_lambda(t1)
            r0 = h_0 / h
            if k == 2:
                P.update(X, X, m0, sigma_t / sigma_0, alpha_t * phi_1, m1, 0.5 * (alpha_t * phi_1), form=1, k=[1.0 / r0])
                return
            assert len(prev) == 3                                  # model_prev_2, model_prev_1, model_prev_0 = model_prev_list
            (m2, t2) = prev[-3]
            h_1 = ns.marginal_lambda(t1) - ns.marginal_lambda(t2)
            r1 = h_1 / h
            phi_2 = phi_1 / h + 1.0
            phi_3 = phi_2 / h - 0.5
            P.update(X, X, m0, sigma_t / sigma_0, alpha_t * phi_1, m1, alpha_t * phi_2, form=2, m2=m2,
                     k=[1.0 / r0, 1.0 / r1, r0 / (r0 + r1), 1.0 / (r0 + r1), alpha_t * phi_3])

        eval_into_free(ts[0])
        for step in range(1, order):
            update(ts[step], step)
            eval_into_free(ts[step])
        for step in range(order, steps + 1):
            step_order = min(order, steps + 1 - step) if (lower_order_final and steps < 10) else order
            update(ts[step], step_order)
            if len(prev) == order:
                free.append(prev.pop(0)[0])
            if step < steps:                                       # "We do not need to evaluate the final model value." (:1247)
                eval_into_free(ts[step])
        return free

    def build_plan(self, steps=20, order=3, skip_type="logSNR", method="singlestep", denoise_to_zero=False, t_start=None,
                   t_end=None, lower_order_final=True):
        """The scalar program of `sample` (:1253-1281) + the singlestep updates (:563-815, dpmsolver++ /
        solver_type 'dpmsolver' branches)."""
        ns = self.noise_schedule
        t_0 = 1.0 / ns.total_N if t_end is None else t_end
        t_T = ns.T if t_start is None else t_start
        assert t_0 > 0 and t_T > 0
        P = SamplePlan()
        X, M0, M1, M2, XS = P.X, P.M0, P.M1, P.M2, P.XS
        if method == "multistep":
            free = self._plan_multistep(P, steps, order, skip_type, t_T, t_0, lower_order_final)
            P.result_slot = X
            if denoise_to_zero:
                P.eval(ns, X, free[0] if free else XS, torch.ones((1,)) * t_0)
                P.result_slot = P.steps[-1][2]
            return P
        if method == "singlestep":
            outer, orders = self.get_orders_and_timesteps_for_singlestep_solver(steps, order, skip_type, t_T, t_0)
        elif method == "singlestep_fixed":
            orders = [order] * (steps // order)
            outer = self.get_time_steps(skip_type, t_T, t_0, len(orders))
        else:
            raise NotImplementedError("method %r: tmdiff_b200 implements 'singlestep', 'singlestep_fixed' and 'multistep' "
                                      "(the adaptive solver is not used by the reference's samplers)" % (method,))
        for k, o in enumerate(orders):
            s, t = outer[k], outer[k + 1]
            inner = self.get_time_steps(skip_type, s.item(), t.item(), o)
            li = ns.marginal_lambda(inner)
            h_in = li[-1] - li[0]
            r1 = None if o <= 1 else (li[1] - li[0]) / h_in
            r2 = None if o <= 2 else (li[2] - li[0]) / h_in
            lambda_s, lambda_t = ns.marginal_lambda(s), ns.marginal_lambda(t)
            h = lambda_t - lambda_s
            sigma_s, sigma_t = ns.marginal_std(s), ns.marginal_std(t)
            alpha_t = torch.exp(ns.marginal_log_mean_coeff(t))
            phi_1 = torch.expm1(-h)
            P.eval(ns, X, M0, s)
            if o == 1:                                                             # :589-592
                P.update(X, X, M0, sigma_t / sigma_s, alpha_t * phi_1)
            elif o == 2:                                                           # :648-658
                s1 = ns.inverse_lambda(lambda_s + r1 * h)
                sigma_s1, alpha_s1 = ns.marginal_std(s1), torch.exp(ns.marginal_log_mean_coeff(s1))
                phi_11 = torch.expm1(-r1 * h)
                P.update(XS, X, M0, sigma_s1 / sigma_s, alpha_s1 * phi_11)
                P.eval(ns, XS, M1, s1)
                P.update(X, X, M0, sigma_t / sigma_s, alpha_t * phi_1, M1, -((0.5 / r1) * (alpha_t * phi_1)))
            else:                                                                  # :733-760
                s1 = ns.inverse_lambda(lambda_s + r1 * h)
                s2 = ns.inverse_lambda(lambda_s + r2 * h)
                sigma_s1, sigma_s2 = ns.marginal_std(s1), ns.marginal_std(s2)
                alpha_s1 = torch.exp(ns.marginal_log_mean_coeff(s1))
                alpha_s2 = torch.exp(ns.marginal_log_mean_coeff(s2))
                phi_11, phi_12 = torch.expm1(-r1 * h), torch.expm1(-r2 * h)
                phi_22 = torch.expm1(-r2 * h) / (r2 * h) + 1.0
                phi_2 = phi_1 / h + 1.0
                P.update(XS, X, M0, sigma_s1 / sigma_s, alpha_s1 * phi_11)
                P.eval(ns, XS, M1, s1)
                P.update(XS, X, M0, sigma_s2 / sigma_s, alpha_s2 * phi_12, M1, r2 / r1 * (alpha_s2 * phi_22))
                P.eval(ns, XS, M2, s2)
                P.update(X, X, M0, sigma_t / sigma_s, alpha_t * phi_1, M2, (1.0 / r2) * (alpha_t * phi_2))
        P.result_slot = X
        if denoise_to_zero:                                                         # :1279-1281
            P.eval(ns, X, M0, torch.ones((1,)) * t_0)
            P.result_slot = M0
        return P

    # ---- device execution --------------------------------------------------------------------
    @torch.no_grad()
    def sample(self, x, steps=20, t_start=None, t_end=None, order=2, skip_type="time_uniform", method="multistep",
               lower_order_final=True, denoise_to_zero=False, solver_type="dpmsolver", atol=0.0078, rtol=0.05,
               return_intermediate=False, return_trace=False):
        if solver_type != "dpmsolver":
            raise NotImplementedError("solver_type must be 'dpmsolver'")
        if return_intermediate:
            raise NotImplementedError("return_intermediate: use return_trace for the per-evaluation network outputs")
        plan = self.build_plan(steps, order, skip_type, method, denoise_to_zero, t_start, t_end, lower_order_final)
        fn = self.model_fn_
        model, kw = fn.model, fn.model_kwargs
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor (no CPU path)")
        xT = x.contiguous().float()
        if xT.dim() != 4:
            raise ValueError("x must be [B, bands, H, W]")
        B = xT.shape[0]
        MS = kw["MS"]
        cfg = bool(getattr(fn, "cfg_active", False))
        if fn.guidance_type == "uncond":
            PAN = kw["PAN"]
        elif cfg:                                   # c_in = cat([unconditional_condition, condition]) (:340)
            PAN = torch.cat([fn.unconditional_condition, fn.condition])
            if MS.shape[0] != 2 * B:
                raise ValueError("classifier-free guidance: model_kwargs['MS'] must hold 2B patches [unconditional | conditional] "
                                 "(diffusion_general.py:308)")
        else:                                       # guidance_scale == 1: conditional half only (:336-337)
            PAN = fn.condition
            if MS.shape[0] == 2 * B:
                MS = MS[B:]
        if tuple(MS.shape) != (MS.shape[0],) + tuple(xT.shape[1:]) or MS.shape[0] != (2 * B if cfg else B):
            raise ValueError("x %s does not match the condition MS %s%s" % (tuple(xT.shape), tuple(MS.shape),
                                                                            " (classifier-free guidance: MS holds 2B patches)" if cfg else ""))
        h = model.bind_condition(PAN, MS, kw["prompt"])
        n_per = xT[0].numel()
        cplan = plan.to_c(n_per, self.dynamic_thresholding_ratio, self.thresholding_max_val, fn.model_type, cfg, fn.guidance_scale)
        out = torch.empty_like(xT)
        nB = 2 * B if cfg else B
        trace = torch.empty((plan.n_eval, nB) + tuple(xT.shape[1:]), device=xT.device) if return_trace else None
        # return_trace == "x": also the x every evaluation was given (eps parity from the trajectory's own x)
        trace_x = torch.empty((plan.n_eval,) + tuple(xT.shape), device=xT.device) if return_trace == "x" else None
        with torch.cuda.device(xT.device):
            _lib.check(_lib.load().tmd_sample(h, _lib.ptr(xT), B, xT.shape[1], xT.shape[2], xT.shape[3], C.byref(cplan), _lib.ptr(out),
                                              _lib.ptr(trace), _lib.ptr(trace_x), _lib.stream_ptr()))
        out = out.to(x.dtype)
        if trace_x is not None:
            return out, (trace, trace_x)
        return (out, trace) if return_trace else out
