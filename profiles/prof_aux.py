"""Side measurements alone (upfirdn2d planes, ns_step): python profiles/prof_aux.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
dev = torch.device("cuda:0")
print(json.dumps(bench.aux_measurements(dev, bench.peaks()), indent=1))

# ns_step kernels one by one through the C-ABI (preallocated outputs: no wrapper / allocator time)
from b_pinn_kalman_filter_b200 import _lib
L = _lib.lib()
B, S = 256, 192
vel = (torch.rand(B, 2, S, S, device=dev) * 0.75 + 0.05) * torch.where(torch.rand(B, 2, S, S, device=dev) > 0.5, 1.0, -1.0)
pres = torch.rand(B, 1, S, S, device=dev) * 0.1
dens = torch.rand(B, 1, S, S, device=dev) * 0.5 + 0.25
vo, po, do_ = torch.empty_like(vel), torch.empty_like(pres), torch.empty_like(dens)
st = _lib.stream()
P = _lib.ptr
calls = {
    "velocity (20 B/cell)": (lambda: L.b200_ns_update_velocity(P(vo), P(vel), P(pres), None, B, S, S, 0.0025, 0.005, 0, st), 20),
    "pressure (16 B/cell)": (lambda: L.b200_ns_update_pressure(P(po), P(pres), P(vel), B, S, S, 0.0025, 0.005, st), 16),
    "density  (16 B/cell)": (lambda: L.b200_ns_update_density(P(do_), P(dens), P(vel), B, S, S, 0.0025, 0.005, st), 16),
}
for name, (fn, bpc) in calls.items():
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print("ns %-22s %.4f ms  %.0f GB/s" % (name, ms, bpc * B * S * S / ms / 1e6))
