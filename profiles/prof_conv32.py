"""Strict-mode conv in isolation: python profiles/prof_conv32.py B H C1 C2 Cout k  -> tc32 (3 terms, 1 term) vs conv_simt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200 import _lib
from b_pinn_kalman_filter_b200.models import engine, tc  # noqa: F401
B, H, C1, C2, Cout, k = (int(v) for v in sys.argv[1:7])
dev = torch.device("cuda:0")
lib = _lib.lib(); P = _lib.ptr
x1 = torch.randn(B, H, H, C1, device=dev); x2 = torch.randn(B, H, H, C2, device=dev) if C2 else None
w = torch.randn(Cout, k, k, C1 + C2, device=dev) / (k * k * (C1 + C2)) ** 0.5
hi, lo = torch.empty_like(w), torch.empty_like(w)
bias = torch.randn(Cout, device=dev); out = torch.empty(B, H, H, Cout, device=dev)
lib.b200_split_tf32(P(w), P(hi), P(lo), w.numel(), _lib.stream())
fl = 2.0 * B * H * H * Cout * k * k * (C1 + C2)
def t(fn, iters=10):
    for _ in range(2): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
for terms in (3, 1):
    ms = t(lambda: _lib.check(lib.b200_conv2d_tc32(P(x1), C1, P(x2), C2, P(hi), P(lo), P(bias), None, 0, None, 1.0, P(out), B, H, H, Cout, k, terms, _lib.stream()), "tc32"))
    print("tc32 terms=%d %s: %.3f ms  %.1f TFLOP/s" % (terms, sys.argv[1:7], ms, fl / ms / 1e9))
ms = t(lambda: _lib.check(lib.b200_conv2d_simt(P(x1), C1, P(x2), C2, P(w), P(bias), None, 0, None, 1.0, P(out), 0, None, 0, B, H, H, H, H, Cout, k, k, 1, k // 2, k // 2, _lib.stream()), "simt"), 3)
print("conv_simt fp32 %s: %.3f ms  %.1f TFLOP/s" % (sys.argv[1:7], ms, fl / ms / 1e9))
