"""Per-shape breakdown of one C3 score-network forward: python profiles/shape_breakdown.py [B] [config]
CUDA events around every launch of the plan (eager, so each figure carries the launch gap), and beside it the whole
forward replayed as a CUDA graph -- the number the PC loop actually pays."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200 import _lib, configs
from b_pinn_kalman_filter_b200.models import utils as mutils
import b_pinn_kalman_filter_b200.models  # noqa: F401

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
name = sys.argv[2] if len(sys.argv) > 2 else "ve/cifar10_ncsnpp_continuous"
dev = torch.device("cuda", 0)
cfg = configs.get_config(name)
cfg.device = dev
model = mutils.ScoreModel(cfg, device=dev, precision="bf16", init=False)
model.init_weights(42, zero_init_scale=1.0)
plan = model.engine.plan(B)
plan.x_in.normal_()
plan.labels.fill_(1.0)
st = _lib.stream()
for _ in range(3):
    plan.run(st)
torch.cuda.synchronize()
acc = {}
REP = 5
for _ in range(REP):
    t = plan.run_timed(st, by_tag=True)
    for k, v in t.items():
        a = acc.setdefault(k, [v[0], 0.0, v[2]])
        a[1] += v[1] / REP
total = sum(v[1] for v in acc.values())
# the same launches as one CUDA graph
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    sp = _lib.stream()
    plan.run(sp)
    s.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        plan.run(_lib.stream())
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    for _ in range(20):
        g.replay()
    e1.record(s)
    s.synchronize()
graph_ms = e0.elapsed_time(e1) / 20
print("forward: %.3f ms sum of per-launch events, %.3f ms as a CUDA graph, batch %d, %d launches, %.1f TFLOP/s in graph"
      % (total, graph_ms, B, len(plan.ops), plan.flops() / graph_ms / 1e9))
for k, (n, ms, fl) in sorted(acc.items(), key=lambda kv: -kv[1][1]):
    print("%-52s n=%3d %8.3f ms %5.1f%% %8.1f TFLOP/s  %6.1f us/launch" % (k, n, ms, 100 * ms / total, fl / ms / 1e9 if ms else 0,
                                                                         1e3 * ms / n))
if os.environ.get("DUMP_TAGS"):
    with open(os.environ["DUMP_TAGS"], "w") as f:
        for (nm, fl), tag in zip(plan.meta, plan.tags):
            f.write("%s %s\t%d\n" % (nm, tag, fl))
