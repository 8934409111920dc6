"""Score-network forward of the other BASELINE configs through the engine: python profiles/prof_configs.py [name ...]
Prints, per config, the per-entry-point time of one forward (CUDA events around every launch) and the implied rates."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from b_pinn_kalman_filter_b200 import _lib, configs  # noqa: E402
from b_pinn_kalman_filter_b200.models import utils as mutils  # noqa: E402
import b_pinn_kalman_filter_b200.models  # noqa: E402,F401

CASES = {
    # name: (config, per-GPU batch, forwards per generated sample)
    "C5": ("ve/celebahq_256_ncsnpp_continuous", 8, 4000),
    "C1": ("vp/nc_ddpmpp", 4, 2000),
    "C4net": ("inverse/nc_ddpmpp_inpaint_dps", 16, 0),
    "C1b": ("vp/cifar10_ncsnpp_continuous", 4, 1000),
}
dev = torch.device("cuda:0")
for name in [n for n in (sys.argv[1:] or list(CASES)) if n in CASES]:
    cfgname, B, fps = CASES[name]
    cfg = configs.get_config(cfgname)
    cfg.device = dev
    t0 = time.time()
    model = mutils.ScoreModel(cfg, device=dev, precision="bf16", init=False)
    model.init_weights(42, zero_init_scale=1.0)
    S, C = cfg.data.image_size, cfg.data.num_channels
    x = torch.randn(B, C, S, S, device=dev)
    lab = torch.rand(B, device=dev) * 10 + 0.1
    y = model(x, lab)
    torch.cuda.synchronize()
    plan = model.engine.plan(B)
    for _ in range(2):
        plan.run(_lib.stream())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        plan.run(_lib.stream())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    tab = plan.run_timed(_lib.stream())
    fl = plan.flops()
    print("== %s  %s  B=%d  %dx%dx%d  finite=%s  build %.1fs" % (name, cfgname, B, C, S, S, bool(torch.isfinite(y).all()),
                                                                time.time() - t0))
    print("   forward %.3f ms  (%d launches)  %.1f GFLOP/sample  %.1f TFLOP/s  %.1f forwards/s/GPU%s"
          % (ms, len(plan.ops), fl / B / 1e9, fl / ms / 1e9, B / ms * 1e3,
             "  => %.4f samples/s/GPU" % (B / ms * 1e3 / fps) if fps else ""))
    for k, (n, t, f) in sorted(tab.items(), key=lambda kv: -kv[1][1]):
        print("   %-28s x%-4d %8.3f ms %s" % (k, n, t, ("%7.1f TFLOP/s" % (f / t / 1e9)) if f else ""))
    del model, plan
    torch.cuda.empty_cache()

# ---- C4: one DPS right-hand side = forward (activations kept) + input gradient, DDPM 64x64x1 --------------------
if not sys.argv[1:] or "C4" in sys.argv[1:]:
    cfg = configs.get_config("inverse/nc_ddpmpp_inpaint_dps")
    cfg.device = dev
    model = mutils.ScoreModel(cfg, device=dev, precision="bf16", init=False)
    model.init_weights(42, zero_init_scale=1.0)
    for B in (2, 16):
        x = torch.randn(B, 1, 64, 64, device=dev)
        lab = torch.rand(B, device=dev) * 900 + 5
        v = torch.randn(B, 1, 64, 64, device=dev)
        eng = model.engine
        for _ in range(2):
            eng.forward_for_grad(x, lab)
            g = eng.input_grad(v)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            eng.forward_for_grad(x, lab)
            g = eng.input_grad(v)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        plan = eng.plan(B, grad=True)
        print("== C4 DPS rhs  B=%d  fwd+input-grad %.3f ms  (%d + %d launches)  %.1f rhs-evals/s  finite=%s"
              % (B, ms, plan.n_fwd, len(plan.ops) - plan.n_fwd, B / ms * 1e3, bool(torch.isfinite(g).all())))
        eng.forward_for_grad(x, lab)
        tab = plan.run_timed(_lib.stream())          # whole list (forward + backward) with per-launch events
        for k, (n, t, f) in sorted(tab.items(), key=lambda kv: -kv[1][1])[:12]:
            print("   %-28s x%-4d %8.3f ms %s" % (k, n, t, ("%7.1f TFLOP/s" % (f / t / 1e9)) if f else ""))
