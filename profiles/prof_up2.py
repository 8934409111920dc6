"""upfirdn2d fp32 planes, the two factor-2 shapes of BASELINE's metric, alone (for ncu): python profiles/prof_up2.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200.op.upfirdn2d import _launch
dev = torch.device("cuda:0")
k1 = torch.tensor([1., 3., 3., 1.], device=dev); k = torch.outer(k1, k1); k = k / k.sum()
xd = torch.randn(8 * 128, 256, 256, device=dev)
xu = torch.randn(8 * 128, 128, 128, device=dev)
for _ in range(5):
    _launch(xd, k, (1, 1), (2, 2), (1, 1, 1, 1))
    _launch(xu, k * 4, (2, 2), (1, 1), (2, 1, 2, 1))
torch.cuda.synchronize()
