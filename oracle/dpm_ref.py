"""Torch-fp32 CPU restatement of the DPM-Solver++ sampling path (TEST INFRASTRUCTURE ONLY).

Restates /root/reference/core/dpm_solver_pytorch.py for the configuration the
reference's ``sample_by_dpmsolver`` uses (GeneralModel/diffusion_general.py:227-255):
discrete VP schedule from betas, ``model_type="x_start"``, ``dpmsolver++`` with
dynamic thresholding, singlestep order<=3 with logSNR spacing, plus the multistep
variant used by ``sample_by_dpmsolver_noise``.  All scalar arithmetic is done with
fp32 torch tensors in the reference's order so that the trajectory can be
compared tightly.  Parity pin: tests/golden/make_golden.py.
"""
from __future__ import annotations

import math

import numpy as np
import torch


def cosine_betas(n_timestep=1000, cosine_s=8e-3):
    """make_beta_schedule('cosine') (GeneralModel/diffusion_general.py:29-63): float64
    torch -> clamp(max=0.999) -> fp32 buffer (set_new_noise_schedule :86-101)."""
    t = torch.arange(n_timestep + 1, dtype=torch.float64) / n_timestep + cosine_s
    a = torch.cos(t / (1 + cosine_s) * math.pi / 2).pow(2)
    a = a / a[0]
    betas = (1 - a[1:] / a[:-1]).clamp(max=0.999)
    return torch.tensor(betas.numpy(), dtype=torch.float32)


def piecewise_linear(x, xp, yp):
    """interpolate_fn (:1296-1335) for C=1: y at x on keypoints (xp ascending), linear
    extrapolation from the outermost segment.  Same final expression, so same rounding."""
    xs = x.reshape(-1)
    xp1, yp1 = xp.reshape(-1), yp.reshape(-1)
    K = xp1.numel()
    # index the query would take in sort([x, xp...]) ; ties: x sorts first (stable sort, x at slot 0)
    idx = torch.searchsorted(xp1, xs, right=False)
    lo = torch.where(idx == 0, torch.zeros_like(idx), torch.where(idx == K, torch.full_like(idx, K - 2), idx - 1))
    x0, x1, y0, y1 = xp1[lo], xp1[lo + 1], yp1[lo], yp1[lo + 1]
    return (y0 + (xs - x0) * (y1 - y0) / (x1 - x0)).reshape(-1, 1)


class ScheduleVP:
    """NoiseScheduleVP('discrete', betas=...) (:6-181)."""

    def __init__(self, betas):
        la = 0.5 * torch.log(1 - betas).cumsum(dim=0)
        self.total_N = la.numel()
        self.T = 1.0
        self.device = betas.device              # CPU for the oracle; a CUDA device for bench.py's eager-GPU baseline arm
        self.t_array = torch.linspace(0.0, 1.0, self.total_N + 1)[1:].reshape(1, -1).to(betas.device)
        self.log_alpha_array = la.reshape(1, -1)

    def log_alpha(self, t):
        return piecewise_linear(t.reshape(-1, 1), self.t_array, self.log_alpha_array).reshape(-1)

    def alpha(self, t):
        return torch.exp(self.log_alpha(t))

    def sigma(self, t):
        return torch.sqrt(1.0 - torch.exp(2.0 * self.log_alpha(t)))

    def lam(self, t):
        la = self.log_alpha(t)
        return la - 0.5 * torch.log(1.0 - torch.exp(2.0 * la))

    def inv_lam(self, lamb):
        la = -0.5 * torch.logaddexp(torch.zeros(1, device=lamb.device), -2.0 * lamb)
        t = piecewise_linear(la.reshape(-1, 1), torch.flip(self.log_alpha_array, [1]), torch.flip(self.t_array, [1]))
        return t.reshape(-1)


def singlestep_orders(steps, order):  # :537-548
    if order == 3:
        K = steps // 3 + 1
        r = steps % 3
        return [3] * (K - 2) + [2, 1] if r == 0 else [3] * (K - 1) + ([1] if r == 1 else [2])
    if order == 2:
        return [2] * (steps // 2) + ([1] if steps % 2 else [])
    return [1] * steps


def dynamic_threshold(x0, ratio=0.995, max_val=1.0):  # :430-439
    s = torch.quantile(torch.abs(x0).reshape(x0.shape[0], -1), ratio, dim=1)
    s = torch.maximum(s, max_val * torch.ones_like(s))[(...,) + (None,) * (x0.dim() - 1)]
    return torch.clamp(x0, -s, s) / s


class Sampler:
    """DPM_Solver(model_fn, ns, 'dpmsolver++', 'dynamic_thresholding') over a raw x0-predicting
    network ``net(x, t_input) -> out`` wrapped the way model_wrapper does (:184-348)."""

    def __init__(self, net, ns: ScheduleVP, trace=None, model_type="x_start", guidance_scale=None):
        """model_type 'x_start' | 'noise' (:296-306).  guidance_scale not None: classifier-free guidance
        (:336-344) -- `net` is then called on cat([x]*2) and must return [uncond | cond] halves."""
        self.net, self.ns, self.trace = net, ns, trace
        self.model_type, self.guidance_scale = model_type, guidance_scale

    def _noise_pred(self, x, tB):
        ns = self.ns
        t_input = (tB - 1.0 / ns.total_N) * 1000.0                  # :291-292
        out = self.net(x, t_input)
        if self.model_type == "noise":
            return out, out, t_input
        a, s = ns.alpha(tB), ns.sigma(tB)
        return (x - a[:, None, None, None] * out) / s[:, None, None, None], out, t_input   # x_start -> noise  :304-306

    def data_prediction(self, x, t):
        ns = self.ns
        tB = t.expand(x.shape[0])                                   # :418
        if self.guidance_scale is None:
            eps, out, t_input = self._noise_pred(x, tB)
        else:                                                       # :338-344
            e2, out, t_input = self._noise_pred(torch.cat([x] * 2), torch.cat([tB] * 2))
            e_u, e_c = e2.chunk(2)
            eps = e_u + self.guidance_scale * (e_c - e_u)
        a1, s1 = ns.alpha(t), ns.sigma(t)
        x0 = dynamic_threshold((x - s1 * eps) / a1)                 # :447-456
        if self.trace is not None:
            self.trace.append(dict(t=float(t), t_input=t_input.clone(), x=x.clone(), out=out.clone(),
                                   eps=eps.clone(), x0=x0.clone()))
        return x0

    def time_steps_logsnr(self, t_T, t_0, N):  # :482-486
        dev = self.ns.device
        lT, l0 = self.ns.lam(torch.tensor(t_T, device=dev)), self.ns.lam(torch.tensor(t_0, device=dev))
        return self.ns.inv_lam(torch.linspace(lT.item(), l0.item(), N + 1).to(dev))

    def time_steps(self, skip_type, t_T, t_0, N):
        if skip_type == "logSNR":
            return self.time_steps_logsnr(t_T, t_0, N)
        if skip_type == "time_uniform":
            return torch.linspace(t_T, t_0, N + 1).to(self.ns.device)
        raise ValueError(skip_type)

    def _update(self, x, s, t, order, r1, r2):
        """first/second/third singlestep updates, dpmsolver++ branch, solver_type 'dpmsolver'
        (:563-608, :610-691, :693-815)."""
        ns, f = self.ns, self.data_prediction
        h = ns.lam(t) - ns.lam(s)
        sig_s, sig_t, a_t = ns.sigma(s), ns.sigma(t), torch.exp(ns.log_alpha(t))
        phi_1 = torch.expm1(-h)
        if order == 1:
            return sig_t / sig_s * x - a_t * phi_1 * f(x, s)
        if order == 2:
            r1 = 0.5 if r1 is None else r1
            s1 = ns.inv_lam(ns.lam(s) + r1 * h)
            sig_s1, a_s1 = ns.sigma(s1), torch.exp(ns.log_alpha(s1))
            phi_11 = torch.expm1(-r1 * h)
            m_s = f(x, s)
            x_s1 = (sig_s1 / sig_s) * x - (a_s1 * phi_11) * m_s
            m_s1 = f(x_s1, s1)
            return (sig_t / sig_s) * x - (a_t * phi_1) * m_s - (0.5 / r1) * (a_t * phi_1) * (m_s1 - m_s)
        r1 = 1.0 / 3.0 if r1 is None else r1
        r2 = 2.0 / 3.0 if r2 is None else r2
        s1 = ns.inv_lam(ns.lam(s) + r1 * h)
        s2 = ns.inv_lam(ns.lam(s) + r2 * h)
        sig_s1, sig_s2 = ns.sigma(s1), ns.sigma(s2)
        a_s1, a_s2 = torch.exp(ns.log_alpha(s1)), torch.exp(ns.log_alpha(s2))
        phi_11, phi_12 = torch.expm1(-r1 * h), torch.expm1(-r2 * h)
        phi_22 = torch.expm1(-r2 * h) / (r2 * h) + 1.0
        phi_2 = phi_1 / h + 1.0
        m_s = f(x, s)
        x_s1 = (sig_s1 / sig_s) * x - (a_s1 * phi_11) * m_s
        m_s1 = f(x_s1, s1)
        x_s2 = (sig_s2 / sig_s) * x - (a_s2 * phi_12) * m_s + r2 / r1 * (a_s2 * phi_22) * (m_s1 - m_s)
        m_s2 = f(x_s2, s2)
        return (sig_t / sig_s) * x - (a_t * phi_1) * m_s + (1.0 / r2) * (a_t * phi_2) * (m_s2 - m_s)

    def _multistep_update(self, x, m_list, t_list, t, order):
        """multistep_dpm_solver_update (:929-956 dispatch; :563-608, :817-927), dpmsolver++ / 'dpmsolver'."""
        ns = self.ns
        t0 = t_list[-1]
        lam_0, lam_t = ns.lam(t0), ns.lam(t)
        sig_0, sig_t, a_t = ns.sigma(t0), ns.sigma(t), torch.exp(ns.log_alpha(t))
        h = lam_t - lam_0
        phi_1 = torch.expm1(-h)
        if order == 1:
            return sig_t / sig_0 * x - a_t * phi_1 * m_list[-1]
        m1, m0 = m_list[-2], m_list[-1]
        h_0 = lam_0 - ns.lam(t_list[-2])
        r0 = h_0 / h
        D1_0 = (1.0 / r0) * (m0 - m1)
        if order == 2:
            return (sig_t / sig_0) * x - (a_t * phi_1) * m0 - 0.5 * (a_t * phi_1) * D1_0
        m2 = m_list[-3]
        h_1 = ns.lam(t_list[-2]) - ns.lam(t_list[-3])
        r1 = h_1 / h
        D1_1 = (1.0 / r1) * (m1 - m2)
        D1 = D1_0 + (r0 / (r0 + r1)) * (D1_0 - D1_1)
        D2 = (1.0 / (r0 + r1)) * (D1_0 - D1_1)
        phi_2 = phi_1 / h + 1.0
        phi_3 = phi_2 / h - 0.5
        return (sig_t / sig_0) * x - (a_t * phi_1) * m0 + (a_t * phi_2) * D1 - (a_t * phi_3) * D2

    @torch.no_grad()
    def sample_multistep(self, x, steps=50, order=3, skip_type="logSNR", denoise_to_zero=False, lower_order_final=True):
        """DPM_Solver.sample(method='multistep') (:1208-1252)."""
        ns = self.ns
        t_0, t_T = 1.0 / ns.total_N, ns.T
        ts = self.time_steps(skip_type, t_T, t_0, steps)
        t_list, m_list = [ts[0]], [self.data_prediction(x, ts[0])]
        for step in range(1, order):
            x = self._multistep_update(x, m_list, t_list, ts[step], step)
            t_list.append(ts[step])
            m_list.append(self.data_prediction(x, ts[step]))
        for step in range(order, steps + 1):
            step_order = min(order, steps + 1 - step) if (lower_order_final and steps < 10) else order
            x = self._multistep_update(x, m_list, t_list, ts[step], step_order)
            t_list = t_list[1:] + [ts[step]]
            if step < steps:
                m_list = m_list[1:] + [self.data_prediction(x, ts[step])]
            else:
                m_list = m_list[1:] + [m_list[-1]]
        if denoise_to_zero:
            x = self.data_prediction(x, torch.ones(1, device=x.device) * t_0)
        return x

    @torch.no_grad()
    def sample(self, x, steps=20, order=3, skip_type="logSNR", denoise_to_zero=True):
        """DPM_Solver.sample(method='singlestep') (:1253-1281)."""
        ns = self.ns
        t_0, t_T = 1.0 / ns.total_N, ns.T
        orders = singlestep_orders(steps, order)
        if skip_type == "logSNR":
            outer = self.time_steps(skip_type, t_T, t_0, len(orders))
        else:
            outer = self.time_steps(skip_type, t_T, t_0, steps)[torch.cumsum(torch.tensor([0] + orders), 0)]
        for k, o in enumerate(orders):
            s, t = outer[k], outer[k + 1]
            inner = self.time_steps(skip_type, s.item(), t.item(), o)
            li = ns.lam(inner)
            h = li[-1] - li[0]
            r1 = None if o <= 1 else (li[1] - li[0]) / h
            r2 = None if o <= 2 else (li[2] - li[0]) / h
            x = self._update(x, s, t, o, r1, r2)
        if denoise_to_zero:
            x = self.data_prediction(x, torch.ones(1, device=x.device) * t_0)
        return x
