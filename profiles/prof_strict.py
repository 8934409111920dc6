"""One C3 forward per precision mode: python profiles/prof_strict.py [B]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200 import _lib, configs
from b_pinn_kalman_filter_b200.models import utils as mutils
import b_pinn_kalman_filter_b200.models  # noqa: F401
B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dev = torch.device("cuda", 0)
for precision in ("bf16", "tf32", "fp32", "fp32_simt"):
    cfg = configs.get_config("ve/cifar10_ncsnpp_continuous"); cfg.device = dev
    m = mutils.ScoreModel(cfg, device=dev, precision=precision, init=False)
    m.init_weights(42, zero_init_scale=1.0)
    plan = m.engine.plan(B); plan.x_in.normal_(); plan.labels.fill_(1.0)
    st = _lib.stream()
    for _ in range(2): plan.run(st)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    n = 3
    for _ in range(n): plan.run(st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    t = plan.run_timed(st)
    top = sorted(t.items(), key=lambda kv: -kv[1][1])[:5]
    print("%-9s B=%d: %.2f ms per forward, %.1f TFLOP/s; tc32 convs %d, simt convs %d; top: %s" % (
        precision, B, ms, plan.flops() / ms / 1e9, m.engine.tc32_convs, m.engine.simt_convs,
        ", ".join("%s %.2f" % (k.replace("b200_", ""), v[1]) for k, v in top)))
    del m, plan
