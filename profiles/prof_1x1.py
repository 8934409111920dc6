"""1x1 convolutions with K = 256 as the C3 network runs them at 16x16 (q/k/v projection without statistics, NIN_3 with
residual and statistics): python profiles/prof_1x1.py [B]   (REPS=1 for an ncu capture)"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b_pinn_kalman_filter_b200 import _lib
from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
lib = _lib.lib(); dev = torch.device("cuda:0"); P = _lib.ptr
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
REPS = int(os.environ.get("REPS", "20"))
for (H, C, Cout, with_res, with_stats) in [(16, 256, 768, 0, 0), (16, 256, 256, 1, 1)]:
    x = torch.randn(B, H, H, C, device=dev).to(torch.bfloat16)
    w = (torch.randn(Cout, C, device=dev) / C ** 0.5).to(torch.bfloat16)
    res = torch.randn(B, H, H, Cout, device=dev).to(torch.bfloat16) if with_res else None
    bias = torch.randn(Cout, device=dev); out = torch.empty(B, H, H, Cout, device=dev, dtype=torch.bfloat16)
    T = lib.b200_conv2d_tc_stat_slots(B, H, H); stats = torch.empty(B, T, Cout, 2, device=dev)
    fn = lambda: _lib.check(lib.b200_conv2d_tc(P(x), C, None, 0, P(w), P(bias), None, 0, P(res) if with_res else None, 0.7071,
                                               P(out), 0, None, P(stats) if with_stats else None, B, H, H, Cout, 1, _lib.stream()), "conv")
    for _ in range(3 if REPS > 1 else 1): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(REPS): fn()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / REPS * 1e3
    nb = (x.numel() + out.numel() * (2 if with_res else 1)) * 2
    print("1x1 %dx%d %d->%d%s%s B=%d: %.1f us  %.0f GB/s (floor %.1f us at 6555 GB/s)" % (H, H, C, Cout, " +res" if with_res else "", " +stats" if with_stats else "", B, us, nb / us * 1e-3, nb / 6555e3), flush=True)
