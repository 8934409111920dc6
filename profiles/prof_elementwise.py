"""One pass of each memory-bound kernel at its BASELINE shape (for ncu --set full): python profiles/prof_elementwise.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200 import _lib, op, sampling
from b_pinn_kalman_filter_b200.op.upfirdn2d import _launch
from b_pinn_kalman_filter_b200.models import engine, tc  # noqa: F401
dev = torch.device("cuda:0")
lib, P, st = _lib.lib(), _lib.ptr, _lib.stream()
k1 = torch.tensor([1., 3., 3., 1.], device=dev)
k = torch.outer(k1, k1); k = k / k.sum()
for rep in range(2):
    # upfirdn2d fp32 planes, C5's largest shapes (the metric's second half)
    x = torch.randn(8 * 128, 256, 256, device=dev)
    _launch(x, k, (1, 1), (2, 2), (1, 1, 1, 1))
    xs = torch.randn(8 * 128, 128, 128, device=dev)
    _launch(xs, k * 4, (2, 2), (1, 1), (2, 1, 2, 1))
    # upfirdn2d channels-last bf16 inside the networks (C3: [1024, 32, 32, 128] -> 16x16)
    xn = torch.randn(1024, 32, 32, 128, device=dev).to(torch.bfloat16)
    on = torch.empty(1024, 16, 16, 128, device=dev, dtype=torch.bfloat16)
    lib.b200_upfirdn2d_nhwc(P(xn), P(k), P(on), 1, 1024, 32, 32, 128, 4, 4, 1, 2, 1, 1, st)
    # GroupNorm + SiLU apply from conv statistics, C3 top level
    B, HW, C, G = 1024, 1024, 128, 32
    xg = torch.randn(B, HW, C, device=dev).to(torch.bfloat16); og = torch.empty_like(xg)
    T = lib.b200_conv2d_tc_stat_slots(B, 32, 32)
    sts = torch.rand(B, T, C, 2, device=dev) + 1.0
    sts[..., 1] = sts[..., 0] ** 2 / (HW / T) * 1.5 + 1.0
    ga, be = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    lib.b200_gn_apply_stats(P(xg), C, P(sts), T, None, 0, None, 0, 1, B, HW, G, lib.b200_gn_nsplit(B, HW), P(ga), P(be), 1e-6, 1, P(og), st)
    # fused SDE update with in-kernel Philox noise, C3 state [1024, 3, 32, 32]
    xx = torch.randn(1024, 3, 32, 32, device=dev); sc = torch.randn_like(xx)
    a = torch.ones(1024, device=dev); b = torch.full((1024,), 0.1, device=dev); c = torch.full((1024,), 0.3, device=dev)
    sampling.fused_update(xx, sc, a, b, c)
    # fused bias + leaky ReLU at an NCSN++ activation shape
    xf = torch.randn(8, 128, 256, 256, device=dev); bf = torch.randn(128, device=dev)
    op.fused_leaky_relu(xf, bf)
    torch.cuda.synchronize()
print("ok")
