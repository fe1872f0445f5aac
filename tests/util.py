import torch


def rel_l2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def psnr(a, b, peak=1.0):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    mse = float(((a - b) ** 2).mean())
    return float("inf") if mse == 0 else 10.0 * torch.log10(torch.tensor(peak * peak / mse)).item()


def run_plan_on_cpu(plan, net, xT, model_type="x_start", guidance_scale=None):
    """Interpret a tmdiff_b200 SamplePlan with fp32 torch ops on the CPU, rounding where the CUDA kernels
    round (csrc/solver.cu).  `net(x, t_input)`; with guidance_scale it is called on cat([x]*2)."""
    from oracle import dpm_ref
    T = torch.tensor
    slots = {0: xT.clone()}
    for s in plan.steps:
        if s[0] == "eval":
            _, src, dst, t_in, a, sg = s
            x, a, sg = slots[src], T(a), T(sg)
            conv = (lambda xx, o: o) if model_type == "noise" else (lambda xx, o: (xx - a * o) / sg)
            if guidance_scale is None:
                eps = conv(x, net(x, T([t_in]).expand(x.shape[0])))
            else:
                x2 = torch.cat([x] * 2)
                e_u, e_c = conv(x2, net(x2, T([t_in]).expand(x2.shape[0]))).chunk(2)
                eps = e_u + T(float(guidance_scale)) * (e_c - e_u)
            slots[dst] = dpm_ref.dynamic_threshold((x - sg * eps) / a)
        else:
            _, dst, x, m, m1, c1, c2, c3, form, m2, k = s
            r = T(c1) * slots[x] - T(c2) * slots[m]
            if form == 0:
                if m1 >= 0:
                    r = r + T(c3) * (slots[m1] - slots[m])
            else:
                d10 = T(k[0]) * (slots[m] - slots[m1])
                if form == 1:
                    r = r - T(c3) * d10
                else:
                    d11 = T(k[1]) * (slots[m1] - slots[m2])
                    dd = d10 - d11
                    r = (r + T(c3) * (d10 + T(k[2]) * dd)) - T(k[4]) * (T(k[3]) * dd)
            slots[dst] = r
    return slots[plan.result_slot]
