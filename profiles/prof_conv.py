"""One conv_tc shape in isolation (for ncu): python profiles/prof_conv.py B H C1 C2 Cout k [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from b_pinn_kalman_filter_b200 import _lib  # noqa: E402
from b_pinn_kalman_filter_b200.models import tc  # noqa: E402,F401

B, H, C1, C2, Cout, k = (int(v) for v in sys.argv[1:7])
iters = int(sys.argv[7]) if len(sys.argv) > 7 else 5
dev = torch.device("cuda:0")
x1 = torch.randn(B, H, H, C1, device=dev).to(torch.bfloat16)
x2 = torch.randn(B, H, H, C2, device=dev).to(torch.bfloat16) if C2 else None
w = (torch.randn(Cout, k, k, C1 + C2, device=dev) * 0.02).to(torch.bfloat16)
bias = torch.randn(Cout, device=dev)
out = torch.empty(B, H, H, Cout, device=dev, dtype=torch.bfloat16)
lib = _lib.lib()
T = lib.b200_conv2d_tc_stat_slots(B, H, H)
stats = torch.empty(B, T, Cout, 2, device=dev)
res = torch.randn(B, H, H, Cout, device=dev).to(torch.bfloat16) if os.environ.get("RES", "0") == "1" else None


def run():
    _lib.check(lib.b200_conv2d_tc(_lib.ptr(x1), C1, _lib.ptr(x2), C2, _lib.ptr(w), _lib.ptr(bias), None, 0, _lib.ptr(res), 1.0,
                                  _lib.ptr(out), 0, None, _lib.ptr(stats) if os.environ.get("STATS", "1") == "1" else None, B, H, H, Cout, k, _lib.stream()), "conv")


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    run()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
fl = 2.0 * B * H * H * Cout * (C1 + C2) * k * k
print("conv %s: %.3f ms  %.1f TFLOP/s" % (sys.argv[1:7], ms, fl / ms / 1e9))
