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
    """:184-348.  Returns `model_fn(x, t_continuous)`; the returned callable also carries what
    `DPM_Solver.sample` needs to run the whole trajectory on the device."""
    if model_type != "x_start" or guidance_type != "uncond":
        raise NotImplementedError("tmdiff_b200 implements the reference's sample_by_dpmsolver configuration: "
                                  "model_type='x_start', guidance_type='uncond'")

    def model_fn(x, t_continuous):
        t_input = (t_continuous - 1.0 / noise_schedule.total_N) * 1000.0           # :291-292
        output = model(x, t_input, **model_kwargs)
        B = x.shape[0]
        alpha_t = noise_schedule.marginal_alpha(t_continuous).reshape(B, 1, 1, 1)   # the reference broadcasts [B] over W (:304-306) and only runs at B=1
        sigma_t = noise_schedule.marginal_std(t_continuous).reshape(B, 1, 1, 1)
        return (x - alpha_t * output) / sigma_t

    model_fn.model = model
    model_fn.noise_schedule = noise_schedule
    model_fn.model_kwargs = dict(model_kwargs)
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

    def update(self, dst, x, m, c1, c2, m1=-1, c3=0.0):
        f = lambda v: float(v.item()) if torch.is_tensor(v) else float(v)
        self.steps.append(("update", dst, x, m, m1, f(c1), f(c2), f(c3)))

    @property
    def n_eval(self):
        return sum(1 for s in self.steps if s[0] == "eval")

    def t_inputs(self):
        return [s[3] for s in self.steps if s[0] == "eval"]

    def to_c(self, n_per_sample, ratio=0.995, max_val=1.0):
        arr = (_lib.Step * len(self.steps))()
        for i, s in enumerate(self.steps):
            if s[0] == "eval":
                arr[i].kind = 0
                arr[i].eval = _lib.Eval(s[1], s[2], s[3], s[4], s[5])
            else:
                arr[i].kind = 1
                arr[i].update = _lib.Update(*s[1:])
        q_lo, q_w = quantile_position(n_per_sample, ratio)
        plan = _lib.Plan(C.cast(arr, C.POINTER(_lib.Step)), len(self.steps), self.result_slot, q_w, q_lo, max_val)
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

    def build_plan(self, steps=20, order=3, skip_type="logSNR", method="singlestep", denoise_to_zero=False, t_start=None,
                   t_end=None):
        """The scalar program of `sample` (:1253-1281) + the singlestep updates (:563-815, dpmsolver++ /
        solver_type 'dpmsolver' branches)."""
        ns = self.noise_schedule
        t_0 = 1.0 / ns.total_N if t_end is None else t_end
        t_T = ns.T if t_start is None else t_start
        assert t_0 > 0 and t_T > 0
        if method == "singlestep":
            outer, orders = self.get_orders_and_timesteps_for_singlestep_solver(steps, order, skip_type, t_T, t_0)
        elif method == "singlestep_fixed":
            orders = [order] * (steps // order)
            outer = self.get_time_steps(skip_type, t_T, t_0, len(orders))
        else:
            raise NotImplementedError("method %r: tmdiff_b200 implements the singlestep solvers" % (method,))
        P = SamplePlan()
        X, M0, M1, M2, XS = P.X, P.M0, P.M1, P.M2, P.XS
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
        plan = self.build_plan(steps, order, skip_type, method, denoise_to_zero, t_start, t_end)
        model = self.model_fn_.model
        kw = self.model_fn_.model_kwargs
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor (no CPU path)")
        h = model.bind_condition(kw["PAN"], kw["MS"], kw["prompt"])
        xT = x.contiguous().float()
        B = xT.shape[0]
        n_per = xT[0].numel()
        cplan = plan.to_c(n_per, self.dynamic_thresholding_ratio, self.thresholding_max_val)
        out = torch.empty_like(xT)
        trace = torch.empty((plan.n_eval,) + tuple(xT.shape), device=xT.device) if return_trace else None
        with torch.cuda.device(xT.device):
            _lib.check(_lib.load().tmd_sample(h, _lib.ptr(xT), C.byref(cplan), _lib.ptr(out), _lib.ptr(trace), _lib.stream_ptr()))
        out = out.to(x.dtype)
        return (out, trace) if return_trace else out
