#!/usr/bin/env python
"""bench.py -- the reference's headline metric on B200: PC samples/s of the CIFAR-10-shape NCSN++ VE-SDE
predictor-corrector sampler (ReverseDiffusion + Langevin, 1000 steps; BASELINE.json configs[2]).

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run)
    python bench.py --impl reference ...                      (CPU arm: the oracle port of the reference's path)

A "step" is one PC step of the whole batch = Langevin corrector (1 score-net forward + norm pass + update) followed
by the reverse-diffusion predictor (1 forward + update): 2 score-network forwards.  Every PC step costs the same, so
`value` = batch / (measured seconds per PC step x 1000 steps); K steps are timed after W warm-up steps, on the
device, max over ranks.  `e2e` runs the public API call a user makes -- `sampling.get_pc_sampler(...)(model)` --
with the prior drawn on the host (pinned) and copied in, and the samples copied back out, inside the timed region.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_STEPS = 1000                     # sde.N of configs/ve/cifar10_ncsnpp_continuous.py
CONFIG = "ve/cifar10_ncsnpp_continuous"
TOTAL_BATCH = 1024                 # BASELINE.json configs[2]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch-per-gpu", type=int, default=0, help="default: 1024 at N=1, 1024/N otherwise ... see --scaling")
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                    help="strong (default, BASELINE.json configs[2]): 1024 images TOTAL, 1024/N per GPU; weak: every GPU "
                         "samples 1024 images (reported as the secondary `weak` object of the strong line anyway)")
    ap.add_argument("--no-weak", action="store_true", help="skip the secondary weak-scaling measurement at N > 1")
    ap.add_argument("--no-ref-gpu", action="store_true", help="skip the reference-ops-on-this-GPU leg (aux.reference_gpu)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--e2e-steps", type=int, default=0, help="PC steps of the end-to-end API run (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-aux", action="store_true", help="skip the upfirdn2d / ns_step side measurements")
    ap.add_argument("--cpu-batch", type=int, default=8)
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], 0, [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [v.strip() for v in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax = max(smax, float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax or None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "tc_burst": d["bf16_tflops"], "tc_sustained": d["bf16_tflops_sustained"],
                "source": "measured"}
    return {"hbm": 6650.0, "tc_burst": 1590.0, "tc_sustained": 1400.0, "source": "fallback"}


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference's path (oracle/score_net_ref.py + oracle/sampling_ref.py), torch CPU fp32
# ------------------------------------------------------------------------------------------------------------------
def cpu_pc_steps(batch, n_pc_steps, threads):
    """Times `n_pc_steps` PC steps (Langevin + reverse diffusion, 2 forwards each) of the C3 workload on the host.
    Returns (seconds per PC step, batch)."""
    import torch
    from oracle import sampling_ref, score_net_ref, weights
    from b_pinn_kalman_filter_b200 import configs
    from b_pinn_kalman_filter_b200.models import utils as mutils
    import b_pinn_kalman_filter_b200.models  # noqa: F401
    from b_pinn_kalman_filter_b200.models.engine import parameter_spec
    torch.set_num_threads(threads)
    cfg = configs.get_config(CONFIG)
    spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
    sd = weights.trained_like_state_dict(spec, cfg, seed=42)
    net = score_net_ref.RefScoreNet(cfg, sd)
    sde = sampling_ref.RefSDE.from_config(cfg)
    shape = (batch, 3, 32, 32)
    gen = torch.Generator().manual_seed(0)

    def noises():
        while True:
            yield torch.randn(shape, generator=gen)

    with torch.no_grad():
        sampling_ref.pc_sample(sde, net, shape, "reverse_diffusion", "langevin", 0.16, 1, True, noises(), eps=1e-5,
                               n_iters=1)                                       # warm-up
        t0 = time.perf_counter()
        sampling_ref.pc_sample(sde, net, shape, "reverse_diffusion", "langevin", 0.16, 1, True, noises(), eps=1e-5,
                               n_iters=n_pc_steps)
        dt = time.perf_counter() - t0
    return dt / n_pc_steps, batch


def cpu_baseline_obj(batch, n_pc_steps):
    threads = os.cpu_count() or 1
    s_per_step, b = cpu_pc_steps(batch, n_pc_steps, threads)
    return {"value": b / (s_per_step * N_STEPS), "unit": "samples/s", "cores": threads, "kind": "port",
            "sample": "%d PC steps (2 score-net forwards each) at batch %d of the same 1000-step workload, torch %s CPU "
                      "fp32 oracle port; per-step time x 1000" % (n_pc_steps, b, __import__("torch").__version__),
            "s_per_pc_step": s_per_step}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, a.steps)
    warm = a.warmup
    t0 = time.perf_counter()
    n_cpu = min(max(steps, 1), 20)
    s_per_step, b = cpu_pc_steps(a.cpu_batch, n_cpu, os.cpu_count() or 1)
    val = b / (s_per_step * N_STEPS)
    line = {"impl": "reference", "metric": "PC samples/s (NCSN++ 32x32, 1000 steps)", "value": val, "unit": "samples/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": warm, "ms_per_step": s_per_step * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C3 cifar10 NCSN++ VE PC sampler (reverse_diffusion + langevin, snr 0.16), N=1000",
                       "batch": b, "timed_pc_steps": n_cpu},
            "cpu_baseline": {"value": val, "unit": "samples/s", "cores": os.cpu_count() or 1, "kind": "port",
                             "sample": "%d PC steps at batch %d on the host, per-step time x 1000" % (n_cpu, b)},
            "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    emit(line)


# ------------------------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------------------------
def aux_measurements(dev, pk):
    """The metric's second half (upfirdn2d HBM GB/s at the C5 shape) and the ns_step rollout (configs[1]); each is
    one kernel timed alone with CUDA events on the launching stream, inputs (>= 268 MB) larger than L2."""
    import torch
    from b_pinn_kalman_filter_b200 import op
    from b_pinn_kalman_filter_b200.op import ns_step
    out = {}

    # each kernel is timed ALONE (the roofline's burst HBM figure is measured the same way): let the board come back
    # from the power-capped clocks of the sampling run before the first one
    torch.cuda.synchronize()
    time.sleep(2.0)

    def timeit(fn, iters=20):
        for _ in range(5):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    k1 = torch.tensor([1., 3., 3., 1.], device=dev)
    k = torch.outer(k1, k1)
    k = k / k.sum()
    x = torch.randn(8, 128, 256, 256, device=dev)
    from b_pinn_kalman_filter_b200.op.upfirdn2d import _launch
    xp = x.view(-1, 256, 256)
    for tag, fn, nbytes in [
        ("down2 [8x128,256,256]->[.,128,128] f32 planes", lambda: _launch(xp, k, (1, 1), (2, 2), (1, 1, 1, 1)),
         4 * (x.numel() + x.numel() // 4)),
    ]:
        ms = timeit(fn)
        out["upfirdn2d"] = {"shape": tag, "ms": ms, "achieved": nbytes / ms / 1e6, "peak": pk["hbm"], "unit": "GB/s",
                            "frac": nbytes / ms / 1e6 / pk["hbm"], "algorithmic_bytes": nbytes}
    xs = torch.randn(8, 128, 128, 128, device=dev).view(-1, 128, 128)
    ms = timeit(lambda: _launch(xs, k * 4, (2, 2), (1, 1), (2, 1, 2, 1)))
    nb = 4 * (xs.numel() * 5)
    out["upfirdn2d_up2"] = {"shape": "up2 [8x128,128,128]->[.,256,256] f32 planes", "ms": ms,
                            "achieved": nb / ms / 1e6, "peak": pk["hbm"], "unit": "GB/s", "frac": nb / ms / 1e6 / pk["hbm"]}
    del x, xp, xs
    # ns_step: B=256, 192x192 (pinn_kalman/simulator.py:35-61): velocity -> pressure -> density
    B, S = 256, 192
    g = torch.Generator(device="cpu").manual_seed(0)
    vel = ((torch.rand(B, 2, S, S, generator=g) * 0.75 + 0.05) * torch.where(torch.rand(B, 2, S, S, generator=g) > 0.5, 1.0, -1.0)).to(dev)
    pres = (torch.rand(B, 1, S, S, generator=g) * 0.1).to(dev)
    dens = (torch.rand(B, 1, S, S, generator=g) * 0.5 + 0.25).to(dev)
    dt, dx = 0.0025, 1 / 200

    def step():
        v = ns_step.update_velocity(vel, pres, dt, dx)
        p = ns_step.update_pressure(pres, v, dt, dx)
        ns_step.update_density(dens, v, dt, dx)
        return p

    ms = timeit(step)
    cells = B * S * S
    out["ns_step"] = {"shape": "B=256 192x192 fp32, velocity+pressure+density", "ms_per_step": ms,
                      "cell_updates_per_s": cells / ms * 1e3, "achieved": 52 * cells / ms / 1e6, "peak": pk["hbm"],
                      "unit": "GB/s", "frac": 52 * cells / ms / 1e6 / pk["hbm"],
                      "algorithmic_bytes_per_cell": 52,
                      "achieved_at_32B_per_cell": 32 * cells / ms / 1e6, "frac_at_32B_per_cell": 32 * cells / ms / 1e6 / pk["hbm"],
                      "bound": "fp32 instruction issue, not HBM: ~490 SASS instructions per cell per step (cuobjdump: "
                               "ns_velocity4 ~250, ns_cip4 160, ns_pressure4 82; 57 of the 160 are FP math) = >= 130 us at 4 "
                               "warp-instructions/clk/SM for 9.44 M cells"}
    # the kernel to beat: the reference's own extension (oracle/_ref, op/ns_step_kernel.cu rebuilt for sm_100) at the same
    # shape -- its update_velocity is only CORRECT at B = 1; timed anyway, it moves the same bytes
    try:
        from oracle import build_ref_ext
        ext = build_ref_ext.load_ext("ns_step_forward")
        if ext is None:
            raise RuntimeError("oracle/_ref/ns_step_forward not built")

        def ref_step():
            v = ext.update_velocity(vel, pres, dt, dx)
            p = ext.update_pressure(pres, v, dt, dx)
            ext.update_density(dens, v, dt, dx)
            return p

        rms = timeit(ref_step, iters=5)
        out["ns_step"]["reference_extension"] = {"ms_per_step": rms, "speedup": rms / ms,
                                                 "kernel": "op/ns_step_kernel.cu:96-234 + op/ns_step.cpp:45-102, 9 launches"}
    except Exception as e:  # noqa: BLE001
        out["ns_step"]["reference_extension"] = {"unavailable": repr(e)[:160]}
    del vel, pres, dens
    # CPU legs of the side metrics: the C restatement of ns_step (-fopenmp, all host cores) on a bounded
    # sample, and the reference's upfirdn2d_native arithmetic (F.pad + F.conv2d, op/upfirdn2d.py:159-200) at the C5 shape
    try:
        import numpy as np
        from oracle import ns_step_ref
        d0, v0, p0 = ns_step_ref.synthetic_fields(16, 192, seed=0)
        ns_step_ref.update_velocity(v0, p0, dt, dx)
        t0 = time.perf_counter()
        for _ in range(3):
            v1 = ns_step_ref.update_velocity(v0, p0, dt, dx)
            ns_step_ref.update_pressure(p0, v1, dt, dx)
            ns_step_ref.update_density(d0, v1, dt, dx)
        cs = (time.perf_counter() - t0) / 3
        out["ns_step"]["cpu_baseline"] = {"value": 16 * 192 * 192 / cs, "unit": "cell updates/s", "cores": os.cpu_count(),
                                          "kind": "port", "sample": "3 steps at B=16 192x192, oracle/ns_step_ref.c -fopenmp",
                                          "gpu_over_cpu": (cells / ms * 1e3) / (16 * 192 * 192 / cs)}
        import torch.nn.functional as F
        torch.set_num_threads(os.cpu_count() or 1)
        xc = torch.randn(128, 1, 256, 256)
        kc = torch.outer(torch.tensor([1., 3., 3., 1.]), torch.tensor([1., 3., 3., 1.]))
        kc = (kc / kc.sum()).flip(0, 1).view(1, 1, 4, 4)
        fcpu = lambda: F.conv2d(F.pad(xc, [1, 1, 1, 1]), kc)[:, :, ::2, ::2]
        fcpu()
        t0 = time.perf_counter()
        for _ in range(3):
            fcpu()
        cs = (time.perf_counter() - t0) / 3
        nbc = 4 * (xc.numel() + xc.numel() // 4)
        out["upfirdn2d"]["cpu_baseline"] = {"value": nbc / cs / 1e9, "unit": "GB/s", "cores": os.cpu_count(), "kind": "port",
                                            "sample": "down2 of 128 planes 256x256 (1/8 of the C5 shape), the arithmetic of "
                                                      "upfirdn2d_native: pad + conv2d + decimate, torch CPU"}
    except Exception as e:  # noqa: BLE001
        out["cpu_legs_error"] = repr(e)[:160]
    out["other_configs"] = other_config_measurements(dev, timeit)
    return out


def other_config_measurements(dev, timeit):
    """BASELINE.json's other configs through the same engine at their per-GPU batch of the 8-GPU split:
    C5 CelebA-HQ-256 NCSN++ forward (B=8), C1 DDPM 64x64 forward (B=4), C4 DPS right-hand side = forward + input
    gradient (B=16).  Device-resident timings, random-init weights; reported beside the headline, not as bench lines."""
    import torch
    from b_pinn_kalman_filter_b200 import _lib, configs
    from b_pinn_kalman_filter_b200.models import utils as mutils
    res = {}
    for tag, name, B, fwd_per_sample in (("C5_celebahq256_ncsnpp_forward", "ve/celebahq_256_ncsnpp_continuous", 8, 4000),
                                         ("C1_nc_ddpm_forward", "vp/nc_ddpmpp", 4, 2000)):
        cfg = configs.get_config(name)
        cfg.device = dev
        m = mutils.ScoreModel(cfg, device=dev, precision="bf16", init=False)
        m.init_weights(42, zero_init_scale=1.0)
        S, C = cfg.data.image_size, cfg.data.num_channels
        y = m(torch.randn(B, C, S, S, device=dev), torch.rand(B, device=dev) * 10 + 0.1)
        plan = m.engine.plan(B)
        ms = timeit(lambda: m.engine._run_ops(plan, "all", plan.ops), iters=5)     # as model(x, t) runs it: graph replay
        res[tag] = {"batch": B, "ms_per_forward": ms, "tflops": plan.flops() / ms / 1e9, "launches": len(plan.ops),
                    "samples_per_s_per_gpu": B / (ms * 1e-3 * fwd_per_sample), "finite": bool(torch.isfinite(y).all())}
        # the same config END TO END through the public API (get_pc_sampler: host prior -> H2D, PC steps, samples -> D2H)
        # over n_run steps of its N-step schedule; per-step cost is schedule-independent, so samples/s = B / (t / n_run * N)
        from b_pinn_kalman_filter_b200 import sampling, sde_lib
        n_run = 50
        mm = cfg.model
        if cfg.training.sde == "vesde":
            sde_r, eps = sde_lib.VESDE(mm.sigma_min, mm.sigma_max, n_run), 1e-5
        else:
            sde_r, eps = sde_lib.VPSDE(mm.beta_min, mm.beta_max, n_run), 1e-3
        shape = (B, C, S, S)
        host_prior = torch.randn(shape).pin_memory()
        sde_r.prior_sampling = lambda shp: host_prior
        fn = sampling.get_pc_sampler(sde_r, shape, sampling.get_predictor(cfg.sampling.predictor),
                                     sampling.get_corrector(cfg.sampling.corrector), lambda v: v, cfg.sampling.snr,
                                     n_steps=cfg.sampling.n_steps_each, continuous=cfg.training.continuous, denoise=True,
                                     eps=eps, device=dev)
        fn(m)                                             # warm-up: plans, graph capture
        host_out = torch.empty(shape).pin_memory()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        xs, nfe = fn(m)
        host_out.copy_(xs)
        torch.cuda.synchronize()
        dt_run = time.perf_counter() - t0
        res[tag]["e2e"] = {"api": "sampling.get_pc_sampler(%s + %s)(model), pinned-host prior -> H2D, %d of %d PC steps, "
                                  "samples -> D2H" % (cfg.sampling.predictor, cfg.sampling.corrector, n_run, mm.num_scales),
                           "seconds": dt_run, "ms_per_pc_step": dt_run / n_run * 1e3,
                           "samples_per_s_per_gpu": B / (dt_run / n_run * mm.num_scales), "nfe": nfe,
                           "finite": bool(torch.isfinite(host_out).all())}
        del m, plan
        torch.cuda.empty_cache()
    # training step (losses.get_step_fn with train=True): C3 network, bf16, 128 images per GPU (= 1024 over
    # 8 GPUs), dropout 0.1, Adam + EMA; forward + loss + input-gradient chain + parameter gradients + fused optimizer
    try:
        from b_pinn_kalman_filter_b200 import losses, sde_lib
        from b_pinn_kalman_filter_b200.models.ema import ExponentialMovingAverage
        cfg = configs.get_config(CONFIG)
        cfg.device = dev
        mt = mutils.ScoreModel(cfg, device=dev, precision="bf16", init=False)
        mt.init_weights(42, zero_init_scale=1.0)
        sde_t = sde_lib.VESDE(cfg.model.sigma_min, cfg.model.sigma_max, cfg.model.num_scales)
        state = dict(optimizer=losses.get_optimizer(cfg, mt.parameters()), model=mt,
                     ema=ExponentialMovingAverage(mt.parameters(), decay=cfg.model.ema_rate), step=0)
        fn = losses.get_step_fn(sde_t, train=True, optimize_fn=losses.optimization_manager(cfg), reduce_mean=False,
                                continuous=True, likelihood_weighting=False)
        Bt = 128
        xb = torch.rand(Bt, 3, 32, 32, device=dev)
        l0 = float(fn(state, xb))
        ms = timeit(lambda: fn(state, xb), iters=5)
        tr = mt.trainer
        res["C3_training_step"] = {"batch": Bt, "ms_per_step": ms, "images_per_s": Bt / ms * 1e3, "loss_first": l0,
                                   "loss_last": float(tr.loss[0]), "launches": len(tr.plan.ops),
                                   "parameters": tr.n, "finite": bool(torch.isfinite(tr.flat_p).all()),
                                   "note": "weight gradients: b200_wgrad_tc (split-K over pixels, MN-major "
                                           "operands, no im2col); gradient collection and weight re-packing: one "
                                           "b200_copy_batch launch each"}
        del mt, state, tr
        torch.cuda.empty_cache()
    except Exception as e:  # noqa: BLE001
        res["C3_training_step"] = {"unavailable": repr(e)[:200]}
    cfg = configs.get_config("inverse/nc_ddpmpp_inpaint_dps")
    cfg.device = dev
    m = mutils.ScoreModel(cfg, device=dev, precision="bf16", init=False)
    m.init_weights(42, zero_init_scale=1.0)
    B = 16
    x, lab, v = torch.randn(B, 1, 64, 64, device=dev), torch.rand(B, device=dev) * 900 + 5, torch.randn(B, 1, 64, 64, device=dev)

    def rhs():
        m.engine.forward_for_grad(x, lab)
        return m.engine.input_grad(v)

    ms = timeit(rhs, iters=5)
    plan = m.engine.plan(B, grad=True)
    res["C4_dps_rhs_forward_plus_input_grad"] = {"batch": B, "ms": ms, "rhs_evals_per_s": B / ms * 1e3,
                                                 "launches": len(plan.ops), "finite": bool(torch.isfinite(rhs()).all())}
    del m, plan
    # the precision modes of the C3 score network (north_star: "TF32/bf16 conv modes"), one forward at 128 images:
    # bf16 = the headline's mode; tf32 / fp32 = fp32 activations with the convolutions on tcgen05 kind::tf32 (one product /
    # the fp32-faithful three-term split); fp32_simt = the CUDA-core strict mode the split replaced
    modes = {}
    for precision in ("bf16", "tf32", "fp32", "fp32_simt"):
        cfg = configs.get_config(CONFIG)
        cfg.device = dev
        m = mutils.ScoreModel(cfg, device=dev, precision=precision, init=False)
        m.init_weights(42, zero_init_scale=1.0)
        plan = m.engine.plan(128)
        plan.x_in.normal_()
        plan.labels.fill_(1.0)
        ms = timeit(lambda: plan.run(_lib.stream()), iters=3)
        modes[precision] = {"ms_per_forward": ms, "tflops": plan.flops() / ms / 1e9,
                            "convs_tcgen05": m.engine.tc_convs + m.engine.tc32_convs, "convs_cuda_cores": m.engine.simt_convs,
                            "finite": bool(torch.isfinite(plan.out).all())}
        del m, plan
    res["C3_precision_modes_forward_b128"] = modes
    return res


def reference_gpu_leg(dev, our_fwd_ms, B, precision):
    """The "kernel to beat" (SURVEY 2b / BASELINE.md section 4): what the reference's own GPU path costs on THIS B200.
    (i) its score-network forward = the same torch ops (F.conv2d / group_norm / einsum / silu ...) on cuDNN / cuBLAS
        sm_100 kernels, run through the oracle's functional restatement of NCSNpp.forward (models/ncsnpp.py:232-381) on
        the device, in fp32, TF32 and bf16 autocast, at the bench's batch;
    (ii) its own CUDA extensions (oracle/_ref, built from the unmodified op/*.cu for sm_100) at the aux shapes.
    Outside the timed `value`; oracle/ is used here as the thing measured AGAINST, never as the product."""
    import torch
    from oracle import build_ref_ext, score_net_ref, weights
    from b_pinn_kalman_filter_b200 import configs, op
    from b_pinn_kalman_filter_b200.models import utils as mutils
    from b_pinn_kalman_filter_b200.models.engine import parameter_spec
    out = {}

    def timeit(fn, iters):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters

    try:
        cfg = configs.get_config(CONFIG)
        spec, _ = parameter_spec(cfg, mutils.get_model(cfg.model.name))
        sd = weights.trained_like_state_dict(spec, cfg, seed=42)
        net = score_net_ref.RefScoreNet(cfg, sd, device=dev)
        x = torch.rand(B, 3, 32, 32, device=dev)
        lab = torch.exp(torch.rand(B, device=dev) * 8.5 - 4.6)
        saved = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.benchmark)
        torch.backends.cudnn.benchmark = True
        modes = {}
        try:
            for tag, tf32, autocast in (("fp32", False, False), ("tf32", True, False), ("bf16_autocast", True, True)):
                torch.backends.cudnn.allow_tf32 = tf32
                torch.backends.cuda.matmul.allow_tf32 = tf32
                with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                    ms = timeit(lambda: net(x, lab), 3 if not tf32 else 5)
                modes[tag] = {"ms_per_forward": ms, "speedup_ours_vs_reference": ms / our_fwd_ms}
        finally:
            torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.benchmark = saved
        out["score_net_forward"] = {"batch": B, "ours_ms_per_forward": our_fwd_ms, "ours_precision": precision,
                                    "reference_ops": "torch %s F.conv2d / group_norm / einsum on cuDNN + cuBLAS (sm_100), "
                                                     "eager, as the reference runs them" % torch.__version__, **modes}
        del net, x
        torch.cuda.empty_cache()
    except Exception as e:  # noqa: BLE001 -- a side measurement must never take the bench line down
        out["score_net_forward"] = {"unavailable": repr(e)[:200]}

    try:
        ext = build_ref_ext.load_ext("upfirdn2d")
        if ext is None:
            raise RuntimeError("oracle/_ref/upfirdn2d not built")
        k1 = torch.tensor([1., 3., 3., 1.], device=dev)
        k = torch.outer(k1, k1)
        k = k / k.sum()
        xs = torch.randn(8 * 128, 256, 256, 1, device=dev)
        from b_pinn_kalman_filter_b200.op.upfirdn2d import _launch
        nb = 4 * (xs.numel() + xs.numel() // 4)
        ref_ms = timeit(lambda: ext.upfirdn2d(xs, k, 1, 1, 2, 2, 1, 1, 1, 1), 10)
        xp = xs.view(-1, 256, 256)
        our_ms = timeit(lambda: _launch(xp, k, (1, 1), (2, 2), (1, 1, 1, 1)), 10)
        out["upfirdn2d_down2"] = {"shape": "[8x128,256,256] -> [.,128,128] fp32", "reference_kernel_ms": ref_ms,
                                  "reference_kernel_GBps": nb / ref_ms / 1e6, "ours_ms": our_ms,
                                  "ours_GBps": nb / our_ms / 1e6, "speedup": ref_ms / our_ms,
                                  "reference_kernel": "op/upfirdn2d_kernel.cu:107-207 rebuilt for sm_100 (oracle/_ref)"}
        del xs, xp
        ext = build_ref_ext.load_ext("fused")
        if ext is None:
            raise RuntimeError("oracle/_ref/fused not built")
        xa = torch.randn(128, 128, 32, 32, device=dev)
        bias = torch.randn(128, device=dev)
        empty = xa.new_empty(0)
        ref_ms = timeit(lambda: ext.fused_bias_act(xa, bias, empty, 3, 0, 0.2, 2 ** 0.5), 20)
        our_ms = timeit(lambda: op.fused_leaky_relu(xa, bias, 0.2, 2 ** 0.5), 20)
        nb = 8 * xa.numel()
        out["fused_bias_act"] = {"shape": "[128,128,32,32] fp32", "reference_kernel_ms": ref_ms,
                                 "reference_kernel_GBps": nb / ref_ms / 1e6, "ours_ms": our_ms,
                                 "ours_GBps": nb / our_ms / 1e6, "speedup": ref_ms / our_ms,
                                 "note": "ours includes the python wrapper's output allocation",
                                 "reference_kernel": "op/fused_bias_act_kernel.cu:18-49 rebuilt for sm_100 (oracle/_ref)"}
    except Exception as e:  # noqa: BLE001
        out["reference_extensions"] = {"unavailable": repr(e)[:200]}
    return out


def run_b200(a):
    import torch
    import torch.distributed as dist
    from b_pinn_kalman_filter_b200 import _lib, configs, sampling, sde_lib
    from b_pinn_kalman_filter_b200.models import utils as mutils
    import b_pinn_kalman_filter_b200.models  # noqa: F401

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL writes its debug output (the version banner at NCCL_DEBUG >= VERSION) to STDOUT unless told otherwise;
        # stdout carries exactly one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()
    pk = peaks()

    B = a.batch_per_gpu or (TOTAL_BATCH if a.scaling == "weak" else TOTAL_BATCH // world)
    cfg = configs.get_config(CONFIG)
    cfg.device = dev
    model = mutils.ScoreModel(cfg, device=dev, precision=a.precision, init=False)
    model.init_weights(42, zero_init_scale=1.0)     # random-init weights of the named architecture (see init_weights)
    sde = sde_lib.VESDE(cfg.model.sigma_min, cfg.model.sigma_max, cfg.model.num_scales)
    shape = (B, 3, 32, 32)
    pg = dist.group.WORLD if world > 1 else None
    loop = sampling.FusedPCLoop(model, sde, shape, sampling.ReverseDiffusionPredictor, sampling.LangevinCorrector,
                                cfg.sampling.snr, cfg.sampling.n_steps_each, False, True, 1e-5, process_group=pg,
                                rank_offset=rank * B, total_batch=B * world, seed=1234)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing: W warm-up + K timed PC steps -------------------------------------------------
    x0 = torch.randn(shape, device=dev) * sde.sigma_max
    loop.reset(x0)
    loop.run(max(a.warmup, 3))        # >= 3: one eager step + CUDA-graph capture + replays (single-GPU runs)
    barrier()
    clocks = ClockSampler(local)
    clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    loop.run(a.steps)
    launches = a.steps * loop.launches_per_step
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    finite = bool(torch.isfinite(loop.x).all())
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / a.steps
    # nvidia-smi samples every 100 ms: a timed region shorter than a few periods (small per-GPU batches at N = 8) would
    # leave no sample, so the SAME loop keeps running, untimed, until the sampler has seen ~0.6 s of this load (the step
    # count comes from the max-over-ranks time: every rank of a sharded loop must run the same number of steps)
    extra = 0
    if ms_per_step * a.steps < 600.0:
        extra = int(math.ceil((600.0 - ms_per_step * a.steps) / ms_per_step))
        loop.run(extra)
        barrier()
    clk = clocks.stop()
    clk["sampled_over"] = "the timed region" if not extra else \
        "the timed region + %d more untimed steps of the same loop (region shorter than the sampling period)" % extra
    total_B = B * world
    value = total_B / (ms_per_step * 1e-3 * N_STEPS)

    # ---- secondary: weak scaling (1024 images on EVERY GPU) beside the strong headline ---------------------------
    weak = None
    if world > 1 and a.scaling == "strong" and not a.no_weak and not a.batch_per_gpu:
        loop_w = sampling.FusedPCLoop(model, sde, (TOTAL_BATCH, 3, 32, 32), sampling.ReverseDiffusionPredictor,
                                      sampling.LangevinCorrector, cfg.sampling.snr, cfg.sampling.n_steps_each, False, True,
                                      1e-5, process_group=pg, rank_offset=rank * TOTAL_BATCH,
                                      total_batch=TOTAL_BATCH * world, seed=1234)
        loop_w.reset(torch.randn((TOTAL_BATCH, 3, 32, 32), device=dev) * sde.sigma_max)
        loop_w.run(max(a.warmup, 3))
        barrier()
        w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0.record()
        loop_w.run(a.steps)
        w1.record()
        barrier()
        tw = torch.tensor([w0.elapsed_time(w1)], device=dev, dtype=torch.float64)
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        wms = float(tw.item()) / a.steps
        weak = {"value": TOTAL_BATCH * world / (wms * 1e-3 * N_STEPS), "unit": "samples/s", "ms_per_step": wms,
                "batch_per_gpu": TOTAL_BATCH, "global_batch": TOTAL_BATCH * world, "scaling": "weak",
                "graph_replay": loop_w.graph is not None}
        del loop_w

    # ---- per-kernel breakdown of one score-net forward (CUDA events around every launch) -----------------------
    plan = loop.plan
    torch.cuda.synchronize()
    table = plan.run_timed(_lib.stream())
    fwd_ms = sum(v[1] for v in table.values())
    conv_names = [n for n in table if "conv" in n or "gemm" in n]
    top = max(table, key=lambda n: table[n][1])
    dom = max(conv_names, key=lambda n: table[n][1]) if conv_names else top
    n_l, d_ms, d_fl = table[dom]
    if dom.startswith("b200_conv2d_tc"):
        # one kernel template (conv_tc_kernel<...>) behind three entry points: plain, GroupNorm-fused input, stride 2
        fam = [n for n in table if n.startswith("b200_conv2d_tc")]
        n_l, d_ms, d_fl = (sum(table[n][i] for n in fam) for i in range(3))
        dom = "b200_conv2d_tc"
    achieved = d_fl / (d_ms * 1e-3) / 1e12 if d_ms > 0 else 0.0
    # DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture of this command
    # (profiles/summarize_ncu.py full ...); null when no capture is committed
    traffic, traffic_src = None, None
    tpath = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "conv_tc_traffic.json")
    if dom.startswith("b200_conv2d_tc") and os.path.exists(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        traffic, traffic_src = tj.get("dram_bytes_per_launch"), "profiles/conv_tc_traffic.json (%s, %d launches)" % (
            tj.get("source", "ncu --set full"), tj.get("launches", 0))
    roof = {"bound": "tensor", "kernel": dom, "achieved": achieved, "peak": pk["tc_sustained"], "unit": "TFLOP/s",
            "frac": achieved / pk["tc_sustained"], "frac_of_burst_peak": achieved / pk["tc_burst"], "traffic": traffic, "traffic_unit": "bytes per launch (DRAM read + write)",
            "traffic_source": traffic_src, "peak_source": pk["source"] + " (sustained bf16)",
            "launches_per_forward": n_l, "avg_launch_ms": d_ms / max(n_l, 1),
            "flops_per_forward": d_fl, "share_of_forward": d_ms / fwd_ms if fwd_ms else None}
    breakdown = {n: {"launches": v[0], "ms": round(v[1], 4), "gflop": round(v[2] / 1e9, 3)} for n, v in
                 sorted(table.items(), key=lambda kv: -kv[1][1])}

    # ---- end to end through the public API with host buffers ---------------------------------------------------
    e2e_steps = a.e2e_steps or N_STEPS
    est = ms_per_step * 1e-3 * e2e_steps
    if not a.e2e_steps and est > 150.0:
        e2e_steps = max(a.steps, int(150.0 / (ms_per_step * 1e-3)))
    sde_e = sde_lib.VESDE(cfg.model.sigma_min, cfg.model.sigma_max, e2e_steps)
    host_prior = torch.empty(shape, dtype=torch.float32).pin_memory()
    host_prior.normal_(generator=torch.Generator().manual_seed(rank)).mul_(sde.sigma_max)
    host_out = torch.empty(shape, dtype=torch.float32).pin_memory()
    sde_e.prior_sampling = lambda shp: host_prior            # the API draws the prior on the host (sde_lib.py:290)
    fn = sampling.get_pc_sampler(sde_e, shape, sampling.ReverseDiffusionPredictor, sampling.LangevinCorrector,
                                 lambda v: v, cfg.sampling.snr, n_steps=cfg.sampling.n_steps_each, continuous=True,
                                 denoise=True, eps=1e-5, device=dev, process_group=pg, rank_offset=rank * B,
                                 total_batch=total_B)
    barrier()
    t0 = time.perf_counter()
    xs, nfe = fn(model)
    host_out.copy_(xs, non_blocking=True)
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    e2e_val = total_B / (e2e_s / e2e_steps * N_STEPS)
    e2e_finite = bool(torch.isfinite(host_out).all())
    nbytes = host_prior.numel() * 4

    line = {"metric": "PC samples/s (NCSN++ 32x32, 1000 steps)", "value": value, "unit": "samples/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": a.scaling, "vs_baseline": None, "dtype": a.precision, "data": "synthetic",
            "config": {"workload": "C3: cifar10 NCSN++ (62.8M params) VE-SDE PC sampler, reverse_diffusion + langevin, "
                                   "snr 0.16, N=1000, 3x32x32", "batch_per_gpu": B, "global_batch": total_B,
                       "pc_steps_per_sample": N_STEPS, "forwards_per_pc_step": loop.forwards_per_step,
                       "value_definition": "global_batch / (measured s per PC step x 1000 PC steps)",
                       "l2": "inputs larger than L2, no flush: %.0f MB of activations written and re-read per forward "
                             "+ 126 MB of weights vs the 126 MB L2" % (1.25 * B),
                       "weights": "random init seed 42 (init_scale=0 layers drawn at scale 1), Philox noise on device",
                       "parallelism": "dp%d (batch shards; the two Langevin norm sums cross the GPUs inside the reducing "
                                      "kernel over NVLink peer memory%s)" % (world, "" if loop.xchg is not None or world == 1
                                                                             else " -- UNAVAILABLE here: NCCL all-reduce"),
                       "graph_replay": loop.graph is not None},
            "score_evals_per_s": total_B * loop.forwards_per_step / (ms_per_step * 1e-3),
            "model_tflops": plan.flops() * loop.forwards_per_step * world / (ms_per_step * 1e-3) / 1e12,
            "gpu_launches": launches, "finite": finite and e2e_finite,
            "clocks": clk, "roofline": roof, "forward_breakdown": breakdown, "forward_ms_sum_of_kernels": fwd_ms,
            "e2e": {"value": e2e_val, "unit": "samples/s", "h2d_bytes_per_step": nbytes / e2e_steps,
                    "d2h_bytes_per_step": nbytes / e2e_steps, "h2d_bytes_total": nbytes, "d2h_bytes_total": nbytes,
                    "pc_steps_run": e2e_steps, "seconds": e2e_s, "nfe": nfe,
                    "api": "sampling.get_pc_sampler(VESDE(N=%d), ...)(model): pinned-host prior -> H2D, %d PC steps, "
                           "samples -> D2H; scaled to 1000 steps" % (e2e_steps, e2e_steps)}}
    if weak is not None:
        line["weak"] = weak
    if rank == 0:
        if not a.no_aux:
            line["aux"] = aux_measurements(dev, pk)
            if not a.no_ref_gpu and world == 1:
                line["aux"]["reference_gpu"] = reference_gpu_leg(dev, fwd_ms, B, a.precision)
        if not a.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline_obj(a.cpu_batch, 20)      # >= 20 PC steps
        elif not a.no_cpu_baseline:
            line["cpu_baseline"] = {"value": None, "unit": "samples/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": "timed at N=1 only"}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_JSON_FD = None


def emit(line):
    """The ONE JSON line, on the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    a = parse()
    # stdout carries exactly one JSON line: libraries that print there (NCCL's version banner does, whatever
    # NCCL_DEBUG_FILE says) are sent to stderr for the life of the process
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
