"""Per-shape timing of one score-network forward (CUDA events around every launch).  Profiling aid:
    python profiles/shape_breakdown.py [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from b_pinn_kalman_filter_b200 import _lib, configs  # noqa: E402
from b_pinn_kalman_filter_b200.models import utils as mutils  # noqa: E402
import b_pinn_kalman_filter_b200.models  # noqa: E402,F401

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
name = sys.argv[2] if len(sys.argv) > 2 else "ve/cifar10_ncsnpp_continuous"
cfg = configs.get_config(name)
dev = torch.device("cuda:0")
m = mutils.ScoreModel(cfg, device=dev, precision="bf16", init=False)
m.init_weights(42, zero_init_scale=1.0)
plan = m.engine.plan(B)
for _ in range(3):
    plan.run(_lib.stream())
torch.cuda.synchronize()
acc = {}
for _ in range(3):
    for k, v in plan.run_timed(_lib.stream(), by_tag=True).items():
        a = acc.setdefault(k, [0, 0.0, 0])
        a[0] += v[0]; a[1] += v[1]; a[2] += v[2]
tot = sum(v[1] for v in acc.values()) / 3
print("forward: %.3f ms (sum of kernels), batch %d" % (tot, B))
for k, v in sorted(acc.items(), key=lambda kv: -kv[1][1]):
    ms = v[1] / 3
    tf = v[2] / 3 / (ms * 1e-3) / 1e12 if ms > 0 else 0
    print("%-52s n=%3d %8.3f ms %5.1f%% %8.1f TFLOP/s" % (k, v[0] // 3, ms, 100 * ms / tot, tf))
