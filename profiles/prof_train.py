"""Where a C3 training step spends its time (B = 128, bf16): python profiles/prof_train.py [batch]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200 import _lib, configs, losses, sde_lib
from b_pinn_kalman_filter_b200.models import backward, utils as mutils
from b_pinn_kalman_filter_b200.models.ema import ExponentialMovingAverage

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dev = torch.device("cuda:0")
cfg = configs.get_config("ve/cifar10_ncsnpp_continuous")
cfg.device = dev
m = mutils.ScoreModel(cfg, device=dev, precision="bf16", init=False)
m.init_weights(42, zero_init_scale=1.0)
sde = sde_lib.VESDE(cfg.model.sigma_min, cfg.model.sigma_max, cfg.model.num_scales)
state = dict(optimizer=losses.get_optimizer(cfg, m.parameters()), model=m,
             ema=ExponentialMovingAverage(m.parameters(), decay=cfg.model.ema_rate), step=0)
fn = losses.get_step_fn(sde, train=True, optimize_fn=losses.optimization_manager(cfg), reduce_mean=False, continuous=True)
x = torch.rand(B, 3, 32, 32, device=dev)
for _ in range(3):
    fn(state, x)
tr = m.trainer
p = tr.plan


def timed(label, f, n=3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n):
        f()
    torch.cuda.synchronize()
    print("%-34s %8.2f ms" % (label, (time.perf_counter() - t0) / n * 1e3))


st = _lib.stream()
timed("whole step_fn", lambda: fn(state, x))
timed("forward launches (%d)" % p.n_fwd, lambda: tr._run(p.ops[:p.n_fwd], st))
timed("backward launches (%d)" % (len(p.ops) - p.n_fwd), lambda: tr._run(p.ops[p.n_fwd:], st))
timed("collect (packed grads -> flat)", tr._collect)
timed("refresh_packed", tr.eng.refresh_packed)
timed("refresh dgrad weights", lambda: [backward.refresh_dgrad_weights(pl) for pl in tr.eng.plans.values()])
timed("grad_sumsq + adam_ema kernels", lambda: (_lib.lib().b200_grad_sumsq(_lib.ptr(tr.sumsq), _lib.ptr(tr.flat_g), tr.n, st),
                                                _lib.lib().b200_adam_ema(_lib.ptr(tr.flat_p), _lib.ptr(tr.flat_m), _lib.ptr(tr.flat_v), _lib.ptr(tr.flat_ema), _lib.ptr(tr.flat_g), _lib.ptr(tr.hyper), _lib.ptr(tr.sumsq), 1.0, tr.n, st)))
tab = {}
evs = []
for fnc, args, name in p.ops:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fnc(*args, st); e1.record(); evs.append((name, e0, e1))
torch.cuda.synchronize()
for name, e0, e1 in evs:
    a = tab.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += e0.elapsed_time(e1)
print("--- device time per entry point (events around every launch)")
for k, v in sorted(tab.items(), key=lambda kv: -kv[1][1])[:14]:
    print("%-34s n=%4d %8.2f ms" % (k, v[0], v[1]))
print("sum %.2f ms" % sum(v[1] for v in tab.values()))
# the convolution launches by role (forward / backward-data / weight gradient) and, for the weight gradients, by shape
role = {}
for (name, e0, e1), tag in zip(evs, p.tags):
    if not name.startswith("b200_conv2d"):
        continue
    r = "wgrad" if tag.startswith("wgrad") else ("dgrad" if tag.startswith("dgrad") else "forward / other")
    a = role.setdefault(r, [0, 0.0]); a[0] += 1; a[1] += e0.elapsed_time(e1)
    if r == "wgrad":
        a = role.setdefault("  " + tag, [0, 0.0]); a[0] += 1; a[1] += e0.elapsed_time(e1)
print("--- convolution launches by role")
for k, v in sorted(role.items(), key=lambda kv: (kv[0].startswith("  "), -kv[1][1]))[:24]:
    print("%-44s n=%4d %8.2f ms" % (k, v[0], v[1]))
