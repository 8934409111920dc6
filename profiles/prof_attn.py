"""Attention kernel timing: b200_attention_tc at the C3 shape (L = 256, C = 256), CUDA events over repeated launches."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b_pinn_kalman_filter_b200 import _lib
from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
lib = _lib.lib()
dev = torch.device("cuda:0")
for B in (128, 1024):
    L = C = 256
    qkv = torch.randn(B, L, 3 * C, device=dev).to(torch.bfloat16)
    out = torch.empty(B, L, C, device=dev, dtype=torch.bfloat16)
    run = lambda: _lib.check(lib.b200_attention_tc(_lib.ptr(qkv), _lib.ptr(out), B, L, C, C ** -0.5, _lib.stream()), "attn")
    REPS = int(os.environ.get('REPS', '50'))
    for _ in range(5 if REPS > 1 else 0): run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = REPS
    e0.record()
    for _ in range(n): run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    fl = 4.0 * B * L * L * C
    q, k, v = (t.float() for t in qkv[:4].split(C, dim=-1))
    ref = torch.softmax(q @ k.transpose(1, 2) * C ** -0.5, dim=-1) @ v
    rel = ((out[:4].float() - ref).norm() / ref.norm()).item()
    print(f"B={B:5d}  {ms*1e3:8.1f} us/launch  {fl/ms*1e-9:7.1f} TFLOP/s  rel-L2 vs fp32 {rel:.2e}  HBM floor {(B*L*4*C*2)/6545e3:.1f} us", flush=True)
