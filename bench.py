#!/usr/bin/env python
"""Headline benchmark of the TMDiff sampling hot path on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one pass of the hot path over one batch of synthetic patches: the step-invariant
condition branch once, then the whole DPM-Solver++ trajectory (20 solver steps = 21 network
evaluations, order 3, logSNR spacing, dynamic thresholding, denoise_to_zero), then res2img.
The workload at every N is BASELINE config 2 per GPU: 64 four-band 64x64 patches, bf16
(weak scaling; N=8 is config 3's 512 patches).  N>1 is launched by torchrun, one rank per GPU;
patches are independent, so the only collective is the final all-gather of output tiles.

One JSON line on stdout (rank 0):
  value      patches/s with PAN/MS/x_T already resident in HBM
  e2e        the same through GeneralDiffusion.sample_by_dpmsolver with HOST (pinned) buffers:
             H2D of PAN/MS/x_T and D2H of the fused image inside the timed region
  roofline   the dominant kernel family (3x3x3 implicit-GEMM conv): algorithmic FLOPs of the
             launches / their summed CUDA-event time, against the measured bf16 peak
  cpu_baseline  the oracle's faithful restatement of the reference (torch fp32 on the host
             cores) timed on one patch of the same workload
`--impl reference` times that CPU path alone (rank 0 only), one patch per step.
"""
import argparse
import gc
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CHANNELS = (32, 64, 128, 256)          # config/general_finetune.json:63-66
METRIC = "patches/s (64x64 PAN patches through the full DPM-Solver++ trajectory)"
# conv FLOPs per 4-band 64x64 denoiser evaluation (2*MACs), split into the part
# that depends on (x_t, t) and the step-invariant part (condition branch + convH_0)
GFLOP_VARIANT_PER_BAND = 50.71 / 4
GFLOP_INVARIANT_PER_BAND = 35.49 / 4


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="patches per GPU per step")
    ap.add_argument("--bands", type=int, default=4)
    ap.add_argument("--patch", type=int, default=64)
    ap.add_argument("--solver-steps", type=int, default=20)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--sensor", default="QB")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-threads", type=int, default=0)
    ap.add_argument("--launch-csv", default=None, help="write the per-launch CUDA-event times of the profiled step here")
    ap.add_argument("--workload", default="config2", choices=["config2", "mixed512", "scene2048"],
                    help="config2 (default, what the driver runs): --batch patches per GPU, weak scaling; mixed512: BASELINE config 3, one "
                         "512-patch WV3+QB+GF2 batch over the GPUs (strong scaling); scene2048: BASELINE config 4, a 2048x2048 PAN / 8-band "
                         "512x512 MS scene, 1024 tiles over the GPUs (strong scaling)")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle check of the timed batch")
    ap.add_argument("--no-eager-baseline", action="store_true", help="skip the PyTorch-eager-on-GPU (cuDNN) baseline arm")
    ap.add_argument("--eager-budget-s", type=float, default=45.0, help="time budget per eager-baseline variant")
    return ap.parse_args()


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def kernel_source_hash():
    """sha256 (16 hex digits) of the conv kernel's sources: an ncu capture describes one build of them."""
    import hashlib
    h = hashlib.sha256()
    for f in ("conv_umma.cu", "conv_op.h"):
        h.update(open(os.path.join(ROOT, "tmdiff_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def conv_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per conv launch from the committed ncu capture
    (profiles/capture.sh -> profiles/traffic_from_ncu.py).  The capture is stamped with the hash of the kernel sources it
    was taken from; None (null) when no capture is committed or the kernel has changed since."""
    p = os.path.join(ROOT, "profiles", "conv_traffic.json")
    if not os.path.exists(p):
        return None, "no capture committed"
    d = json.load(open(p))
    if d.get("kernel_source_sha16") != kernel_source_hash():
        return None, "stale: capture %s is of another build of conv_umma.cu" % d.get("tag")
    return d.get("traffic_bytes_per_launch"), "ncu capture %s of this build (sha %s)" % (d.get("tag"), d.get("kernel_source_sha16"))


# ------------------------------------------------------------------------------------------------
# clocks during the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU every 100 ms through NVML (the same
    counters as the profiling recipe's nvidia-smi line)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self._stop_evt, self.max_mhz = index, [], set(), threading.Event(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                 "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop_evt.wait(0.1)

    def finish(self):
        self._stop_evt.set()
        self.join(2.0)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle's faithful restatement of the reference, on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_sample_seconds(bands, patch, solver_steps, sensor, threads):
    """Seconds for ONE patch through the reference algorithm on the CPU (B=1, as the reference's
    loaders run it, data/__init__.py:18-19)."""
    import torch
    from oracle import dpm_ref, wavbest_ref
    from oracle.inputs import make_inputs
    torch.set_num_threads(threads)
    sd, emb = wavbest_ref.seeded_weights(CHANNELS), wavbest_ref.seeded_text_embeddings()
    pan, ms, x_T = make_inputs(1, bands, patch, patch)
    ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))
    net = lambda x, t: wavbest_ref.wavbest_forward(sd, x, t, pan, ms, emb[sensor], faithful=True)
    t0 = time.perf_counter()
    with torch.no_grad():
        out = dpm_ref.Sampler(net, ns).sample(x_T, steps=solver_steps, order=3) + ms
    dt = time.perf_counter() - t0
    assert bool(torch.isfinite(out).all())
    return dt


def run_reference(args):
    """`--impl reference`: the reference's CPU implementation of the path (the oracle port -- the
    reference is Python and /root/reference does not exist on the GPU box), all host threads,
    one patch of the workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    threads = args.cpu_threads or os.cpu_count()
    for _ in range(args.warmup):
        cpu_sample_seconds(args.bands, args.patch, max(1, args.solver_steps // 10), args.sensor, threads)  # short warm-up passes
    times = [cpu_sample_seconds(args.bands, args.patch, args.solver_steps, args.sensor, threads) for _ in range(args.steps)]
    total = sum(times)
    v = args.steps / total
    sample = "1 patch (%d-band %dx%d, %d solver steps = %d NFE) per step, B=1 as the reference runs it" % (
        args.bands, args.patch, args.patch, args.solver_steps, args.solver_steps + 1)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "patches/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "fp32", "data": "synthetic (seeded PAN/MS/x_T, seeded weights and text embeddings)",
        "config": workload_config(args, per_step_patches=1),
        "cpu_baseline": {"value": v, "unit": "patches/s", "cores": threads, "kind": "port", "sample": sample,
                         "torch_threads": torch.get_num_threads()},
        "e2e": {"value": v, "unit": "patches/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def workload_config(args, per_step_patches):
    return {"workload": "BASELINE config 2: %d-band %dx%d patches, batch %d per GPU, %d DPM-Solver++ steps (%d NFE), %s"
                        % (args.bands, args.patch, args.patch, args.batch, args.solver_steps, args.solver_steps + 1, args.dtype),
            "patches_per_step_per_gpu": per_step_patches, "bands": args.bands, "patch": args.patch,
            "solver": "dpmsolver++ singlestep order 3, logSNR, dynamic thresholding 0.995, denoise_to_zero",
            "solver_steps": args.solver_steps, "channels": list(CHANNELS), "sensor_prompt": args.sensor,
            "l2": "no flush: each step streams >2 GB of activations per GPU (L2 is 126 MB) and every step re-reads its inputs from HBM",
            "parallelism": "patches sharded across GPUs, one all-gather of output tiles per step"}


# ------------------------------------------------------------------------------------------------
# parity of the timed batch (the oracle as checker)
# ------------------------------------------------------------------------------------------------
def parity_of_timed_batch(G, model, x_in, xT_d, pan_h, ms_h, xT_h, args, img_timed):
    """The tensors that were just timed, checked: (i) the CUDA-graph replay equals the eager launch list bit for bit,
    (ii) patches {0, B/2-1, B-1} against the oracle run at B = 1 on the host -- raw network output and eps (from the CUDA
    trajectory's own x, dpm_solver_pytorch.py:296-306) at every evaluation, PSNR of the final image."""
    import torch
    from oracle import dpm_ref, wavbest_ref
    B = xT_d.shape[0]
    model.reset_condition_cache()
    img, (trace, trace_x) = G.sample_by_dpmsolver(x_in, args.sensor, steps=args.solver_steps, x_T=xT_d, return_trace="x")
    graph_equal = bool(torch.equal(img, img_timed))
    img, trace, trace_x = img.cpu(), trace.cpu(), trace_x.cpu()
    sd, emb = wavbest_ref.seeded_weights(CHANNELS), wavbest_ref.seeded_text_embeddings()[args.sensor]
    ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000))
    rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
    worst_out = worst_eps = 0.0
    psnr_min = float("inf")
    patches = sorted({0, max(B // 2 - 1, 0), B - 1})
    for p in patches:
        tr = []
        net = lambda x, t: wavbest_ref.wavbest_forward(sd, x, t, pan_h[p:p + 1], ms_h[p:p + 1], emb, faithful=False)
        final = dpm_ref.Sampler(net, ns, trace=tr).sample(xT_h[p:p + 1], steps=args.solver_steps, order=3) + ms_h[p:p + 1]
        for i, rec in enumerate(tr):
            t = torch.tensor([rec["t"]])
            a, sg = ns.alpha(t), ns.sigma(t)
            worst_out = max(worst_out, rel(trace[i, p], rec["out"][0]))
            worst_eps = max(worst_eps, rel((trace_x[i, p] - a * trace[i, p]) / sg, rec["eps"][0]))
        mse = float(((img[p].double() - final[0].double()) ** 2).mean())
        psnr_min = min(psnr_min, float("inf") if mse == 0 else 10.0 * torch.log10(torch.tensor(1.0 / mse)).item())
    tol = 1e-5 if args.dtype == "fp32" else 2e-2
    return {"rel_l2_out_max": worst_out, "rel_l2_eps_max": worst_eps, "psnr_min": psnr_min, "patches_checked": patches,
            "nfe_checked": args.solver_steps + 1, "graph_replay_equals_eager": graph_equal, "tolerance": tol,
            "ok": bool(worst_out < tol and worst_eps < tol and psnr_min >= 45.0 and graph_equal),
            "against": "oracle (torch fp32 on the host, B=1 per patch, bit-exact restatement of the reference)"}


# ------------------------------------------------------------------------------------------------
# PyTorch-eager-on-GPU baseline (library kernels on the same B200)
# ------------------------------------------------------------------------------------------------
def gpu_eager_baseline(args, dev, pan_d, ms_d, xT_d):
    """The bar SURVEY.md 0.1 sets: the reference algorithm in eager PyTorch on this same GPU -- the oracle's faithful
    restatement (per-sample modulated weights as a groups=B conv3d, Haar transforms as dense matmuls, nothing cached, every
    schedule scalar computed on the device) moved to CUDA, i.e. cuDNN / cuBLAS / ATen kernels, `cudnn.benchmark = True` as the
    reference sets it (general_sharpening_joint_random_batch_finetune.py:31).  fp32 (TF32 convolutions are torch's default)
    and torch.autocast(bf16); B = the workload's batch and B = 1 (how the reference's loaders run it).  Not our code path:
    nothing of tmdiff_b200 is called here."""
    import torch
    from oracle import dpm_ref, wavbest_ref
    torch.backends.cudnn.benchmark = True
    sd = {k: v.to(dev) for k, v in wavbest_ref.seeded_weights(CHANNELS).items()}
    emb = wavbest_ref.seeded_text_embeddings()[args.sensor].to(dev)
    ns = dpm_ref.ScheduleVP(dpm_ref.cosine_betas(1000).to(dev))
    out = {}

    def run(B, autocast, steps):
        pan, ms, xT = pan_d[:B], ms_d[:B], xT_d[:B]
        net = lambda x, t: wavbest_ref.wavbest_forward(sd, x, t, pan, ms, emb, faithful=True).float()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
            img = dpm_ref.Sampler(net, ns).sample(xT, steps=steps, order=3) + ms
        e1.record()
        torch.cuda.synchronize()
        assert bool(torch.isfinite(img).all())
        return e0.elapsed_time(e1) * 1e-3

    for B in (args.batch, 1):
        for name, autocast in (("fp32_tf32conv", False), ("autocast_bf16", True)):
            key = "B%d_%s" % (B, name)
            try:
                run(B, autocast, 2)                               # warm-up: cuDNN autotuning of every layer shape, 3 NFE
                t_probe = run(B, autocast, 2)
                est = t_probe / 3 * (args.solver_steps + 1)
                if est <= args.eager_budget_s:
                    sec, extrapolated = run(B, autocast, args.solver_steps), False
                else:                                             # too slow to time whole: 3 evaluations scaled to the trajectory
                    sec, extrapolated = est, True
                out[key] = {"value": B / sec, "unit": "patches/s", "seconds_per_batch": sec, "batch": B, "extrapolated_from_3_nfe": extrapolated}
            except Exception as exc:                              # e.g. out of memory in the per-sample weight copies
                out[key] = {"error": "%s: %s" % (type(exc).__name__, str(exc)[:200])}
            torch.cuda.empty_cache()
    try:                                                          # kernels per evaluation, from the profiler's CUDA activity list
        from torch.profiler import ProfilerActivity, profile
        t = torch.full((1,), 500.0, device=dev)
        with torch.no_grad(), profile(activities=[ProfilerActivity.CUDA]) as prof:
            wavbest_ref.wavbest_forward(sd, xT_d[:1], t, pan_d[:1], ms_d[:1], emb, faithful=True)
            torch.cuda.synchronize()
        out["gpu_kernels_per_evaluation"] = sum(1 for ev in prof.events() if ev.device_type == torch.autograd.DeviceType.CUDA)
    except Exception:
        out["gpu_kernels_per_evaluation"] = None
    out["what"] = ("oracle.wavbest_ref.wavbest_forward(faithful=True) + oracle.dpm_ref.Sampler on cuda: eager PyTorch "
                   "(cuDNN conv3d, cuBLAS matmul Haar, ATen elementwise), cudnn.benchmark=True")
    torch.backends.cudnn.benchmark = False
    return out


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from oracle import wavbest_ref                       # seeded weights/embeddings only (inputs, not compute)
    from oracle.inputs import make_inputs
    from tmdiff_b200 import GeneralDiffusion, WavBEST
    from tmdiff_b200.parallel import gather_tiles

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; tmdiff_b200 has no CPU path (use --impl reference for the CPU arm)")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched with torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    # libraries may write to the C-level stdout (NCCL prints its version line there under NCCL_DEBUG=VERSION): keep fd 1 clean for
    # the ONE JSON line by pointing it at stderr until that line is printed
    sys.stdout.flush()
    _real_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    B, N, P = args.batch, args.bands, args.patch
    model = WavBEST(channels=list(CHANNELS), dtype=args.dtype, text_embeddings=wavbest_ref.seeded_text_embeddings(),
                    max_batch=B)
    model.load_state_dict(wavbest_ref.seeded_weights(CHANNELS), strict=True)
    model = model.to(dev)
    G = GeneralDiffusion(model)
    G.set_new_noise_schedule({"schedule": "cosine", "n_timestep": 1000}, dev)

    pan_h, ms_h, xT_h = [t.pin_memory() for t in make_inputs(B, N, P, P, seed=3407 + rank)]
    out_h = torch.empty_like(ms_h).pin_memory()
    pan_d, ms_d, xT_d = pan_h.to(dev), ms_h.to(dev), xT_h.to(dev)
    total = world * B

    def step_resident():
        model.reset_condition_cache()                    # a new batch of patches every step: nothing carried over
        img = G.sample_by_dpmsolver({"PAN": pan_d, "MS": ms_d, "Res": xT_d}, args.sensor, steps=args.solver_steps, x_T=xT_d)
        return gather_tiles(img, total) if world > 1 else img

    def step_e2e():
        pan, ms, xT = pan_h.to(dev, non_blocking=True), ms_h.to(dev, non_blocking=True), xT_h.to(dev, non_blocking=True)
        model.reset_condition_cache()
        img = G.sample_by_dpmsolver({"PAN": pan, "MS": ms, "Res": xT}, args.sensor, steps=args.solver_steps, x_T=xT)
        out_h.copy_(img, non_blocking=True)              # this rank's fused tiles back to the host
        if world > 1:
            gather_tiles(img, total)
        torch.cuda.current_stream().synchronize()        # the caller holds the host result before the next batch

    def timed(fn, k):
        """k steps bracketed by barrier + synchronize; device time from CUDA events, max over ranks.  The cyclic garbage
        collector is held off for the k steps (as timeit does): a generation-2 pass of this process takes ~100 ms of host
        time, which the end-to-end loop -- one host synchronisation per step -- cannot hide behind queued GPU work."""
        gc.collect()
        gc.disable()
        try:
            return _timed(fn, k)
        finally:
            gc.enable()

    def _timed(fn, k):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(args.warmup, 3)):
        img_timed = step_resident()
    torch.cuda.synchronize()
    img_timed = img_timed[rank * B:(rank + 1) * B].clone() if world > 1 else img_timed.clone()

    sampler = ClockSampler(local)
    sampler.start()
    l0 = model.launch_count()
    t_ms = timed(step_resident, args.steps)
    launches = model.launch_count() - l0
    clocks = sampler.finish()

    step_e2e()
    step_e2e()
    e2e_ms = timed(step_e2e, args.steps)

    # per-kernel-kind device time: the same step again with every launch bracketed by CUDA events on
    # the launching stream (kept out of the timed regions above so the event records cost them nothing)
    prof = None
    if rank == 0:
        model.profile_begin()
        model.reset_condition_cache()                    # rank-local: no collective inside the profiled step
        G.sample_by_dpmsolver({"PAN": pan_d, "MS": ms_d, "Res": xT_d}, args.sensor, steps=args.solver_steps, x_T=xT_d)
        prof = model.profile_end(args.launch_csv)
    if world > 1:
        dist.barrier()

    if rank == 0:
        peaks = measured_peaks()
        value = total * args.steps / (t_ms * 1e-3)
        e2e = total * args.steps / (e2e_ms * 1e-3)
        c3 = prof["conv3x3x3"]
        ach = c3["flops"] / (c3["ms"] * 1e-3) / 1e12 if c3["ms"] > 0 else 0.0
        ach_exec = c3["flops_executed"] / (c3["ms"] * 1e-3) / 1e12 if c3["ms"] > 0 else 0.0
        prof_total = sum(k["ms"] for k in prof.values())
        peak = peaks["bf16_tflops_sustained"]
        traffic, traffic_note = conv_traffic()
        mem_ms = sum(prof[k]["ms"] for k in ("elementwise", "solver", "conv1x1x1"))
        mem_bytes = sum(prof[k]["bytes"] for k in ("elementwise", "solver", "conv1x1x1"))
        mem_gbs = mem_bytes / (mem_ms * 1e-3) / 1e9 if mem_ms > 0 else 0.0
        gflop_patch = N * (GFLOP_INVARIANT_PER_BAND + (args.solver_steps + 1) * GFLOP_VARIANT_PER_BAND)
        line = {
            "metric": METRIC, "value": value, "unit": "patches/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": t_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": args.dtype, "data": "synthetic (seeded PAN/MS/x_T, seeded weights and text embeddings)",
            "config": workload_config(args, B),
            "e2e": {"value": e2e, "unit": "patches/s", "h2d_bytes_per_step": int(world * (pan_h.numel() + ms_h.numel() + xT_h.numel()) * 4),
                    "d2h_bytes_per_step": int(world * out_h.numel() * 4), "ms_per_step": e2e_ms / args.steps,
                    "api": "tmdiff_b200.GeneralDiffusion.sample_by_dpmsolver on pinned host tensors"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"kernel": "conv3d 3x3x3 implicit GEMM (all %d launches of one step)" % c3["launches"],
                         "bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                         "achieved_executed": ach_exec, "frac_executed": ach_exec / peak,
                         "executed_note": "achieved counts the zero-padding band planes as the reference executes them (SURVEY.md 8d); "
                                          "achieved_executed only the MMAs the kernel issues ((3N-2)/(3N) of the 3x3x3 taps)",
                         "traffic": traffic, "traffic_note": traffic_note, "peak_source": "%s bf16_tflops_sustained (kernel timed inside a long step)" % peaks["source"],
                         "share_of_step": c3["ms"] / prof_total if prof_total else None,
                         "how": "CUDA events around every launch on the launching stream, one extra step after the timed region"},
            "roofline_hbm": {"kernel": "memory-bound launches of one step (stems, slice activations, 32->1 projection, 1x1x1 convs, "
                                       "x0 conversion + dynamic threshold + fused solver update)",
                             "bound": "hbm", "achieved": mem_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": mem_gbs / peaks["hbm_gbs"],
                             "share_of_step": mem_ms / prof_total if prof_total else None},
            "launch_gap_ms": t_ms / args.steps - prof_total,
            "launch_gap_note": "ms_per_step minus the summed CUDA-event time of every launch of one step: prologue / tail / "
                               "dependency gaps between the launches of the graph",
            "end_to_end_tensor_frac": value / world * gflop_patch * 1e9 / (peak * 1e12),
            "kernel_kinds": {k: {"ms": round(v["ms"], 3), "launches": v["launches"],
                                 "tflops": round(v["flops"] / (v["ms"] * 1e-3) / 1e12, 2) if v["ms"] > 0 else 0.0,
                                 "gbs": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9, 1) if v["ms"] > 0 else 0.0}
                             for k, v in prof.items()},
        }
        if not args.no_parity:
            line["parity"] = parity_of_timed_batch(G, model, {"PAN": pan_d, "MS": ms_d, "Res": xT_d}, xT_d, pan_h, ms_h, xT_h, args, img_timed)
        if not args.no_eager_baseline and world == 1:
            model.release()                                # the eager arm gets the whole GPU
            line["gpu_eager_baseline"] = gpu_eager_baseline(args, dev, pan_d, ms_d, xT_d)
        if not args.no_cpu_baseline and world == 1:        # the reported CPU baseline is an N = 1 item (rank 0 of a multi-rank run has peers waiting)
            threads = args.cpu_threads or os.cpu_count()
            sec = cpu_sample_seconds(N, P, args.solver_steps, args.sensor, threads)
            line["cpu_baseline"] = {"value": 1.0 / sec, "unit": "patches/s", "cores": threads, "kind": "port",
                                    "sample": "1 patch of the workload (%d-band %dx%d, %d NFE), B=1, oracle faithful=True, %.1f s"
                                              % (N, P, P, args.solver_steps + 1, sec)}
        sys.stdout.flush()
        os.dup2(_real_stdout, 1)
        print(json.dumps(line))
        sys.stdout.flush()
        sys.stdout.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# BASELINE configs 3 and 4: a fixed job split over the GPUs (strong scaling)
# ------------------------------------------------------------------------------------------------
def run_job(args):
    """--workload mixed512 / scene2048.  The job is the same at every N (strong scaling): rank r takes its contiguous share of
    every band bucket / of the tile list, there is no per-step communication, one all-gather per bucket (or of the tiles) at the
    end.  `value` with the inputs resident in HBM, `e2e` from pinned host buffers with the result read back.  `result_sha16` is
    a hash of the bytes of the result on rank 0: equal across N means the sharded job is bit-identical to the single-GPU one."""
    import hashlib
    import torch
    import torch.distributed as dist
    from oracle import wavbest_ref
    from oracle.inputs import make_inputs
    from tmdiff_b200 import GeneralDiffusion, WavBEST, ops
    from tmdiff_b200.parallel import sample_buckets
    from tmdiff_b200.tiling import sample_scene

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; tmdiff_b200 has no CPU path")
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    if world == 1 and args.gpus > 1:
        raise SystemExit("bench.py --gpus %d must be launched with torchrun --nproc-per-node %d" % (args.gpus, args.gpus))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    sys.stdout.flush()
    _real_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    chunk = args.batch
    model = WavBEST(channels=list(CHANNELS), dtype=args.dtype, text_embeddings=wavbest_ref.seeded_text_embeddings(), max_batch=chunk)
    model.load_state_dict(wavbest_ref.seeded_weights(CHANNELS), strict=True)
    model = model.to(dev)
    G = GeneralDiffusion(model)
    G.set_new_noise_schedule({"schedule": "cosine", "n_timestep": 1000}, dev)
    P = args.patch
    gen = torch.Generator().manual_seed(3407)

    if args.workload == "mixed512":
        # 192 WV3 (8 bands) + 160 QB + 160 GF2 (4 bands): the two 4-band sensors share a bucket with per-sample prompts
        spec = {8: [("WV3", 192)], 4: [("QB", 160), ("GF2", 160)]}
        host = {}
        for n, parts in spec.items():
            b = sum(c for _, c in parts)
            pan, ms, xT = make_inputs(b, n, P, P, seed=3407 + n)
            prompts = [s for s, c in parts for _ in range(c)]
            perm = torch.randperm(b, generator=gen).tolist()            # sensors interleaved inside the bucket
            host[n] = (pan.pin_memory(), ms.pin_memory(), [prompts[i] for i in perm], xT.pin_memory())
        resident = {n: (p.to(dev), m.to(dev), pr, x.to(dev)) for n, (p, m, pr, x) in host.items()}
        n_patches = sum(v[1].shape[0] for v in host.values())
        gflop = sum(v[1].shape[0] * n * (GFLOP_INVARIANT_PER_BAND + (args.solver_steps + 1) * GFLOP_VARIANT_PER_BAND) for n, v in host.items())
        out_host = {n: torch.empty_like(v[1]).pin_memory() for n, v in host.items()}
        h2d = sum((v[0].numel() + v[1].numel() + v[3].numel()) * 4 for v in host.values())
        d2h = sum(v.numel() * 4 for v in out_host.values())

        def step(buckets, read_back):
            model.reset_condition_cache()
            fused = sample_buckets(G, buckets, steps=args.solver_steps, batch=chunk)
            if read_back:
                for n, t in fused.items():
                    out_host[n].copy_(t, non_blocking=True)
                torch.cuda.current_stream().synchronize()
            return fused
        step_resident = lambda: step(resident, False)
        step_e2e = lambda: step(host, True)
        digest = lambda r: hashlib.sha256(b"".join(r[n].cpu().numpy().tobytes() for n in sorted(r))).hexdigest()[:16]
        what = "BASELINE config 3: one joint batch of 192 WV3 (8-band) + 160 QB + 160 GF2 (4-band) %dx%d patches, per-sample prompts" % (P, P)
        extra = lambda r: {}
    else:
        Hs = Ws = 2048
        C8, scale = 8, 2047.0
        pan_h = (torch.rand(1, Hs, Ws, generator=gen) * scale).pin_memory()                     # sensor DN
        lrms_h = (torch.rand(C8, Hs // 4, Ws // 4, generator=gen) * scale).pin_memory()
        T = (Hs // P) * (Ws // P)
        xT_h = torch.randn(T, C8, P, P, generator=gen).pin_memory()
        pan_d, lrms_d, xT_d = pan_h.to(dev), lrms_h.to(dev), xT_h.to(dev)
        out_host = torch.empty(C8, Hs, Ws).pin_memory()
        n_patches = T
        gflop = T * C8 * (GFLOP_INVARIANT_PER_BAND + (args.solver_steps + 1) * GFLOP_VARIANT_PER_BAND)
        h2d, d2h = (pan_h.numel() + lrms_h.numel() + xT_h.numel()) * 4, out_host.numel() * 4

        def step(pan, lrms, xT, read_back):
            model.reset_condition_cache()
            if read_back:
                pan, lrms = pan.to(dev, non_blocking=True), lrms.to(dev, non_blocking=True)
            scene = sample_scene(G, pan, lrms, "WV3", tile=P, steps=args.solver_steps, batch=chunk, x_T=xT, img_scale=scale, clamp=True)
            if read_back:
                out_host.copy_(scene, non_blocking=True)
                torch.cuda.current_stream().synchronize()
            return scene
        step_resident = lambda: step(pan_d, lrms_d, xT_d, False)
        step_e2e = lambda: step(pan_h, lrms_h, xT_h, True)
        digest = lambda r: hashlib.sha256(r.cpu().numpy().tobytes()).hexdigest()[:16]
        what = ("BASELINE config 4: PAN %dx%d / 8-band MS %dx%d scene in sensor DN, %d tiles of %dx%d in unpatch_16 order, DN scaling + "
                "bilinear x4 + tile cut and stitch + clamp + scale back on the device" % (Hs, Ws, Hs // 4, Ws // 4, T, P, P))

        def extra(r):        # on-device quality of the fused scene against a seeded synthetic HR scene (SAM / ERGAS / Q, core/metrics.py)
            hr = (r / scale + 0.02 * torch.randn(r.shape, generator=torch.Generator(device=dev).manual_seed(9), device=dev)).clamp(0.01, 1.0)
            q = ops.quality_metrics(hr[None], (r / scale).clamp(0.01, 1.0)[None])[0].tolist()
            return {"quality_vs_synthetic_hr": {"SAM_deg": q[0], "ERGAS": q[1], "Q": q[2]}}

    def timed(fn, k):
        gc.collect()
        gc.disable()
        try:
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(k):
                r = fn()
            e1.record()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            return float(ms.item()), r
        finally:
            gc.enable()

    for _ in range(max(args.warmup, 3)):
        step_resident()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = model.launch_count()
    t_ms, result = timed(step_resident, args.steps)
    launches = model.launch_count() - l0
    clocks = sampler.finish()
    step_e2e()
    e2e_ms, _ = timed(step_e2e, args.steps)
    if rank == 0:
        peaks = measured_peaks()
        peak = peaks["bf16_tflops_sustained"]
        value = n_patches * args.steps / (t_ms * 1e-3)
        ach = gflop * 1e9 * args.steps / (t_ms * 1e-3) / 1e12 / world
        line = {
            "metric": METRIC, "value": value, "unit": "patches/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": t_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.dtype,
            "data": "synthetic (seeded PAN/MS/x_T, seeded weights and text embeddings)",
            "config": {"workload": what, "patches_per_step": n_patches, "chunk": chunk, "solver_steps": args.solver_steps, "channels": list(CHANNELS),
                       "l2": "no flush: every chunk streams >2 GB of activations per GPU (L2 is 126 MB)",
                       "parallelism": "each band bucket / the tile list split contiguously over the GPUs, one all-gather per bucket at the end"},
            "e2e": {"value": n_patches * args.steps / (e2e_ms * 1e-3), "unit": "patches/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": e2e_ms / args.steps,
                    "api": "tmdiff_b200.parallel.sample_buckets" if args.workload == "mixed512" else "tmdiff_b200.tiling.sample_scene"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"kernel": "whole step: minimal algorithmic conv FLOPs (SURVEY.md 8d) / step time, per GPU", "bound": "tensor", "achieved": ach,
                         "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": None,
                         "peak_source": "%s bf16_tflops_sustained" % peaks["source"]},
            "result_sha16": digest(result),
        }
        line.update(extra(result))
        sys.stdout.flush()
        os.dup2(_real_stdout, 1)
        print(json.dumps(line))
        sys.stdout.flush()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload != "config2":
        run_job(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
