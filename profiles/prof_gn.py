"""GroupNorm(+SiLU) apply in isolation: python profiles/prof_gn.py B HW C [iters]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200 import _lib
from b_pinn_kalman_filter_b200.models import engine  # noqa: F401  (registers signatures)
from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
B, HW, C = (int(v) for v in sys.argv[1:4])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 20
dev = torch.device("cuda:0")
lib = _lib.lib()
x = torch.randn(B, HW, C, device=dev).to(torch.bfloat16)
out = torch.empty_like(x)
G = min(C // 4, 32)
ns = int(os.environ.get('NS', '0')) or lib.b200_gn_nsplit(B, HW)
H = int(HW ** 0.5)
T = lib.b200_conv2d_tc_stat_slots(B, H, H)
st = torch.rand(B, T, C, 2, device=dev) + 1.0
st[..., 1] = st[..., 0] ** 2 / (HW / T) * 1.5 + 1.0
part = torch.empty(B * ns * G * 2, device=dev)
gamma = torch.ones(C, device=dev); beta = torch.zeros(C, device=dev)
def t(fn):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
P = _lib.ptr
s = _lib.stream()
nbytes = x.numel() * 4
ms = t(lambda: lib.b200_gn_apply_stats(P(x), C, P(st), T, None, 0, None, 0, 1, B, HW, G, ns, P(gamma), P(beta), 1e-6, 1, P(out), s))
print("gn_apply_stats B=%d HW=%d C=%d ns=%d T=%d: %.3f ms  %.0f GB/s" % (B, HW, C, ns, T, ms, nbytes / ms / 1e6))
ms = t(lambda: lib.b200_gn_partial(P(x), C, None, 0, 1, B, HW, G, ns, P(part), s))
print("gn_partial: %.3f ms  %.0f GB/s" % (ms, x.numel() * 2 / ms / 1e6))
ms = t(lambda: lib.b200_gn_apply(P(x), C, None, 0, 1, B, HW, G, ns, P(part), P(gamma), P(beta), 1e-6, 1, P(out), s))
print("gn_apply (partials): %.3f ms  %.0f GB/s" % (ms, nbytes / ms / 1e6))
ms = t(lambda: out.copy_(x))
print("torch copy: %.3f ms  %.0f GB/s" % (ms, nbytes / ms / 1e6))
# the same launch 40 times inside a CUDA graph (what the PC loop pays: no host launch cost), for each NS in NS_LIST
def graph_us(fn, n=40):
    s2 = torch.cuda.Stream()
    with torch.cuda.stream(s2):
        st2 = _lib.stream()
        fn(st2); s2.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s2):
            for _ in range(n): fn(_lib.stream())
        for _ in range(3): g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s2)
        for _ in range(5): g.replay()
        e1.record(s2); s2.synchronize()
    return e0.elapsed_time(e1) / 5 / n * 1e3
for nsv in [int(v) for v in os.environ.get("NS_LIST", str(ns)).split(",")]:
    us = graph_us(lambda st_: lib.b200_gn_apply_stats(P(x), C, P(st), T, None, 0, None, 0, 1, B, HW, G, nsv, P(gamma), P(beta), 1e-6, 1, P(out), st_))
    print("graph: gn_apply_stats B=%d HW=%d C=%d ns=%d: %.2f us  %.0f GB/s" % (B, HW, C, nsv, us, nbytes / us / 1e3))
if HW <= 64:     # thread-per-group kernel (default) against the block-per-image kernel
    lib.b200_gn_small_enable(0)
    us = graph_us(lambda st_: lib.b200_gn_apply_stats(P(x), C, P(st), T, None, 0, None, 0, 1, B, HW, G, ns, P(gamma), P(beta), 1e-6, 1, P(out), st_))
    lib.b200_gn_small_enable(1)
    print("graph: block-per-image kernel        ns=%d: %.2f us  %.0f GB/s" % (ns, us, nbytes / us / 1e3))
