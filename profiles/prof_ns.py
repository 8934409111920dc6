"""One launch of each ns_step kernel at the C2 size (for ncu): python profiles/prof_ns.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200.op import ns_step
dev = torch.device("cuda:0")
B, S = 256, 192
vel = (torch.rand(B, 2, S, S, device=dev) * 0.75 + 0.05) * torch.where(torch.rand(B, 2, S, S, device=dev) > 0.5, 1.0, -1.0)
pres = torch.rand(B, 1, S, S, device=dev) * 0.1
dens = torch.rand(B, 1, S, S, device=dev) * 0.5 + 0.25
for _ in range(2):
    v = ns_step.update_velocity(vel, pres, 0.0025, 0.005)
    p = ns_step.update_pressure(pres, v, 0.0025, 0.005)
    d = ns_step.update_density(dens, v, 0.0025, 0.005)
torch.cuda.synchronize()
print("ok", float(v.abs().mean()), float(p.abs().mean()), float(d.abs().mean()))
