"""upfirdn2d channels-last bf16 at the C3 / C5 network shapes: python profiles/prof_upfirdn_nhwc.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200 import _lib
from b_pinn_kalman_filter_b200.models import engine  # noqa: F401
dev = torch.device("cuda:0")
lib, P, st = _lib.lib(), _lib.ptr, _lib.stream()
k1 = torch.tensor([1., 3., 3., 1.], device=dev)
k = torch.outer(k1, k1); k = k / k.sum()
for (B, H, C, up, down, pad) in [(1024, 32, 128, 1, 2, (1, 1)), (1024, 16, 256, 1, 2, (1, 1)), (1024, 16, 256, 2, 1, (2, 1)),
                                 (1024, 8, 256, 2, 1, (2, 1)), (1024, 16, 256, 1, 1, (2, 2)), (8, 256, 128, 1, 2, (1, 1)),
                                 (8, 128, 128, 2, 1, (2, 1))]:
    Ho = (H * up + pad[0] + pad[1] - 4) // down + 1
    x = torch.randn(B, H, H, C, device=dev).to(torch.bfloat16)
    o = torch.empty(B, Ho, Ho, C, device=dev, dtype=torch.bfloat16)
    kk = k * up * up
    f = lambda: lib.b200_upfirdn2d_nhwc(P(x), P(kk), P(o), 1, B, H, H, C, 4, 4, up, down, pad[0], pad[1], st)
    REPS = int(os.environ.get('PROF_REPS', '20'))
    for _ in range(3 if REPS > 1 else 0): f()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(REPS): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / REPS
    nb = 2 * (x.numel() + o.numel())
    print("nhwc bf16 B=%d %dx%d C=%d up%d down%d: %.3f ms  %.0f GB/s (%.0f%% of 6553.6)" % (B, H, H, C, up, down, ms, nb / ms / 1e6, nb / ms / 1e6 / 65.536))
