"""`GeneralDiffusion` sampling surface (GeneralModel/diffusion_general.py:66-339) over the CUDA path.

Sampling methods only; the training loss (`p_losses_dynamic`, :349-370) needs a backward pass and
is out of scope for this hot path.
"""
import math

import numpy as np
import torch
from torch import nn

from .dpm_solver import DPM_Solver, NoiseScheduleVP, model_wrapper


def make_beta_schedule(schedule, n_timestep):
    """:29-63 (float64 numpy, clipped at 0.999 for the cosine schedule)."""
    if schedule == "linear":
        scale = 1000 / n_timestep
        return np.linspace(scale * 1e-6, scale * 1e-2, n_timestep, dtype=np.float64)
    if schedule == "cosine":
        bar = lambda t: math.cos((t + 0.008) / 1.008 * math.pi / 2) ** 2
        return np.array([min(1 - bar((i + 1) / n_timestep) / bar(i / n_timestep), 0.999) for i in range(n_timestep)])
    raise NotImplementedError(schedule)


def res2img(img_, img_lr_up):
    """utils/util.py:135-137"""
    return img_ + img_lr_up


class GeneralDiffusion(nn.Module):
    def __init__(self, denoise_fn, loss_type="l1"):
        super().__init__()
        self.denoise_fn = denoise_fn
        self.loss_type = loss_type

    def set_loss(self, device):
        pass  # training only

    def set_new_noise_schedule(self, schedule_opt, device):
        """:86-132 -- only the buffers the sampling methods read."""
        betas = make_beta_schedule(schedule_opt["schedule"], schedule_opt["n_timestep"])
        to_torch = lambda a: torch.tensor(a, dtype=torch.float32, device=device)
        alphas = 1.0 - betas
        ac = np.cumprod(alphas, axis=0)
        ac_prev = np.append(1.0, ac[:-1])
        self.sqrt_alphas_cumprod_prev = np.sqrt(np.append(1.0, ac))
        self.num_timesteps = int(betas.shape[0])
        self.register_buffer("betas", to_torch(betas))
        self.register_buffer("alphas_cumprod", to_torch(ac))
        self.register_buffer("sqrt_recip_alphas_cumprod_1", to_torch(np.sqrt(1.0 / ac)))
        self.register_buffer("sqrt_recipm1_alphas_cumprod_1", to_torch(np.sqrt(1.0 / ac - 1)))
        pv = betas * (1.0 - ac_prev) / (1.0 - ac)
        self.register_buffer("posterior_variance", to_torch(pv))
        self.register_buffer("posterior_log_variance_clipped", to_torch(np.log(np.maximum(pv, 1e-20))))
        self.register_buffer("posterior_mean_coef1", to_torch(betas * np.sqrt(ac_prev) / (1.0 - ac)))
        self.register_buffer("posterior_mean_coef2", to_torch((1.0 - ac_prev) * np.sqrt(alphas) / (1.0 - ac)))
        # fp32 host copies of the per-step scalars for the fused ancestral step (no device read-back per step)
        self._host_schedule = {k: getattr(self, k).detach().cpu().numpy() for k in
                               ("sqrt_recip_alphas_cumprod_1", "sqrt_recipm1_alphas_cumprod_1", "posterior_mean_coef1",
                                "posterior_mean_coef2", "posterior_log_variance_clipped")}

    # ---- ancestral DDPM step (:155-225): one denoiser call + element-wise torch ops -------------
    def predict_start_from_noise(self, x_t, t, noise):                     # :376-378
        return self.sqrt_recip_alphas_cumprod_1[t] * x_t - self.sqrt_recipm1_alphas_cumprod_1[t] * noise

    def q_posterior(self, x_start, x_t, t):                                # :134-138
        mean = self.posterior_mean_coef1[t] * x_start + self.posterior_mean_coef2[t] * x_t
        return mean, self.posterior_log_variance_clipped[t]

    @torch.no_grad()
    def p_mean_variance(self, x, t, clip_denoised=True, x_in=None, prompt="QB", guidance=1.0):
        B = x.shape[0]
        time_in = torch.full((B, 1), t + 1, device=x.device)
        # the reference treats the network output as eps here (:163-164)
        x_recon = self.predict_start_from_noise(x, t=t, noise=self.denoise_fn(x, time_in, x_in["PAN"], x_in["MS"], prompt))
        if clip_denoised:
            x_recon = x_recon.clamp(-1.0, 1.0)
        return self.q_posterior(x_start=x_recon, x_t=x, t=t)

    @torch.no_grad()
    def p_sample(self, x, t, clip_denoised=True, condition_x=None, prompt="QB", guidance=1.0, noise=None, seed=None):
        """:196-201.  One denoiser evaluation, then predict_start_from_noise + clamp + q_posterior + the noise term as ONE
        kernel (tmd_ancestral_update) with host-side schedule scalars.  noise: the step's z (default torch.randn_like, i.e.
        the reference's draw under the same torch seed); seed: draw z inside the kernel instead (Philox, no noise tensor)."""
        from . import ops
        B = x.shape[0]
        time_in = torch.full((B, 1), t + 1, device=x.device)
        eps = self.denoise_fn(x, time_in, condition_x["PAN"], condition_x["MS"], prompt)
        h = self._host_schedule
        sigma = float(np.exp(np.float32(0.5) * h["posterior_log_variance_clipped"][t])) if t > 0 else 0.0
        if t > 0 and noise is None and seed is None:
            noise = torch.randn_like(x)
        return ops.ancestral_update(x, eps, noise if t > 0 else None, float(h["sqrt_recip_alphas_cumprod_1"][t]),
                                    float(h["sqrt_recipm1_alphas_cumprod_1"][t]), float(h["posterior_mean_coef1"][t]),
                                    float(h["posterior_mean_coef2"][t]), sigma, clip_denoised, 0 if seed is None else int(seed), t)

    @torch.no_grad()
    def p_sample_loop(self, x_in, continous=False, prompt="QB", guidance=1.0, seed=None, noise_source=None):
        """:203-225.  seed=None: x_T and every step's noise come from torch's generator exactly where the reference draws
        them; seed=int: the per-step noise is drawn inside the update kernel (nothing but x and eps is read).
        noise_source (optional): callable(shape) -> tensor that replaces those draws, in the reference's order (x_T first,
        then one per step with t > 0) -- how a recorded reference run is replayed."""
        device = self.betas.device
        sample_inter = 1 | (self.num_timesteps // 10)
        shape = tuple(x_in["Res"].shape)
        img = torch.randn(shape, device=device) if noise_source is None else noise_source(shape).to(device)
        ret = res2img(img, x_in["MS"])
        for i in reversed(range(self.num_timesteps)):
            z = noise_source(shape).to(device) if (noise_source is not None and i > 0) else None
            img = self.p_sample(img, i, condition_x=x_in, prompt=prompt, guidance=guidance, seed=seed, noise=z)
            if i % sample_inter == 0:
                ret = torch.cat([ret, res2img(img, x_in["MS"])], dim=0)
        return ret if continous else ret[-1]

    @torch.no_grad()
    def p_sample_loop_steps(self, x_in, img, steps, prompt="QB", seed=None, noise_source=None):
        """The body of p_sample_loop over an explicit list of timesteps from a given x_t (tests, truncated schedules)."""
        for i in steps:
            z = noise_source(tuple(img.shape)).to(img.device) if (noise_source is not None and i > 0) else None
            img = self.p_sample(img, i, condition_x=x_in, prompt=prompt, seed=seed, noise=z)
        return img

    @torch.no_grad()
    def super_resolution(self, x_in, continous, prompt, guidance):
        return self.p_sample_loop(x_in, prompt)    # sic: the reference passes prompt into `continous` (:338-339)

    # ---- DPM-Solver++ (:227-255), the hot path --------------------------------------------------
    @torch.no_grad()
    def sample_by_dpmsolver(self, x_in, prompt, steps=30, x_T=None, return_trace=False):
        """steps defaults to the reference's hard-coded 30; BASELINE configs use 10 / 20."""
        device = self.betas.device
        if x_T is None:
            x_T = torch.randn(x_in["Res"].shape, device=device)
        ns = NoiseScheduleVP(schedule="discrete", betas=self.betas)
        fn = model_wrapper(self.denoise_fn, ns, model_type="x_start",
                           model_kwargs={"PAN": x_in["PAN"], "MS": x_in["MS"], "prompt": prompt})
        solver = DPM_Solver(fn, ns, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding")
        res = solver.sample(x_T, steps=steps, order=3, skip_type="logSNR", method="singlestep", denoise_to_zero=True,
                            return_trace=return_trace)
        if return_trace:
            return res2img(res[0], x_in["MS"]), res[1]
        return res2img(res, x_in["MS"])

    @torch.no_grad()
    def sample_by_dpmsolver_noise(self, x_in, prompt, steps=50, x_T=None, return_trace=False):
        """:257-286 -- model_type 'noise', multistep order 3, logSNR, denoise_to_zero.  The reference also puts
        x_in['wav'] into model_kwargs, which `WavBEST.forward` (:600) does not accept; it is ignored here."""
        device = self.betas.device
        if x_T is None:
            x_T = torch.randn(x_in["Res"].shape, device=device)
        ns = NoiseScheduleVP(schedule="discrete", betas=self.betas)
        fn = model_wrapper(self.denoise_fn, ns, model_type="noise",
                           model_kwargs={"PAN": x_in["PAN"], "MS": x_in["MS"], "prompt": prompt})
        solver = DPM_Solver(fn, ns, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding")
        res = solver.sample(x_T, steps=steps, order=3, skip_type="logSNR", method="multistep", denoise_to_zero=True,
                            return_trace=return_trace)
        if return_trace:
            return res2img(res[0], x_in["MS"]), res[1]
        return res2img(res, x_in["MS"])

    @torch.no_grad()
    def sample_by_regression(self, x_in, prompt):
        """:288-297 -- one denoiser evaluation at time_in = 1001 on noise (the reference's extra `wav` positional
        argument is dropped, as above)."""
        device = self.betas.device
        x_T = torch.randn(x_in["Res"].shape, device=device)
        B = x_T.shape[0]
        time_in = torch.tensor([1000 + 1] * B, device=device).view(B, -1)
        return res2img(self.denoise_fn(x_T, time_in, x_in["PAN"], x_in["MS"], prompt), x_in["MS"])

    @torch.no_grad()
    def sample_by_dpmsolver_guidance(self, x_in, prompt, guidance, steps=50, x_T=None, return_trace=False):
        """:299-335 -- classifier-free guidance: condition = PAN, unconditional = zeros (PAN and MS), model_type
        'noise', multistep order 2, logSNR, no denoise_to_zero.  Both halves run as one 2B batch per evaluation."""
        device = self.betas.device
        if x_T is None:
            x_T = torch.randn(x_in["Res"].shape, device=device)
        condition = x_in["PAN"]
        ns = NoiseScheduleVP(schedule="discrete", betas=self.betas)
        fn = model_wrapper(self.denoise_fn, ns, model_type="noise",
                           model_kwargs={"MS": torch.cat([torch.zeros_like(x_in["MS"]), x_in["MS"]]), "prompt": prompt},
                           guidance_type="classifier-free", condition=condition,
                           unconditional_condition=torch.zeros_like(condition), guidance_scale=guidance)
        solver = DPM_Solver(fn, ns, algorithm_type="dpmsolver++", correcting_x0_fn="dynamic_thresholding")
        res = solver.sample(x_T, steps=steps, order=2, skip_type="logSNR", method="multistep", return_trace=return_trace)
        if return_trace:
            return res2img(res[0], x_in["MS"]), res[1]
        return res2img(res, x_in["MS"])

    def forward(self, x, *args, **kwargs):
        raise NotImplementedError("training loss (p_losses_dynamic) is outside the sampling hot path")
