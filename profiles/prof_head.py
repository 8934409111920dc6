"""Network head conv (3x3 C -> 3, NCHW fp32 output, out_mode 1) in a CUDA graph: python profiles/prof_head.py
B200_CONV_HEAD_HALO=0 selects the one-box-per-tap path for comparison."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b_pinn_kalman_filter_b200 import _lib
from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
lib = _lib.lib(); dev = torch.device("cuda:0"); P = _lib.ptr
def graph_us(fn, n=20):
    s2 = torch.cuda.Stream()
    with torch.cuda.stream(s2):
        fn(); s2.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s2):
            for _ in range(n): fn()
        for _ in range(3): g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s2)
        for _ in range(5): g.replay()
        e1.record(s2); s2.synchronize()
    return e0.elapsed_time(e1) / 5 / n * 1e3
for (B, H, C) in [(1024, 32, 128), (128, 32, 128), (8, 256, 128)]:
    x = torch.randn(B, H, H, C, device=dev).to(torch.bfloat16)
    w = (torch.randn(3, 3, 3, C, device=dev) / (9 * C) ** 0.5).to(torch.bfloat16)
    bias = torch.randn(3, device=dev); mul = torch.rand(B, device=dev) + 0.5
    out = torch.empty(B, 3, H, H, device=dev)
    fn = lambda: _lib.check(lib.b200_conv2d_tc(P(x), C, None, 0, P(w), P(bias), None, 0, None, 1.0, P(out), 1, P(mul), None, B, H, H, 3, 3, _lib.stream()), "head")
    us = graph_us(fn)
    nb = x.numel() * 2 + out.numel() * 4
    print("head 3x3 %dx%d %d->3 B=%d: %.1f us  %.0f GB/s (floor %.1f us)  head_halo=%s" % (H, H, C, B, us, nb / us * 1e-3, nb / 6555e3, os.environ.get("B200_CONV_HEAD_HALO", "1")), flush=True)
