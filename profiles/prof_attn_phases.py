"""Phase timeline of CTA 0 of the streaming attention kernel (clock64 stamps).  Needs a trace build:
  B200_NVCC_EXTRA=-DB200_ATTN_TRACE python -m b_pinn_kalman_filter_b200.build"""
import os, sys, ctypes, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from b_pinn_kalman_filter_b200 import _lib
from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
lib = _lib.lib()
dev = torch.device("cuda:0")
B, L, C = 1024, 256, 256
qkv = torch.randn(B, L, 3 * C, device=dev).to(torch.bfloat16)
out = torch.empty(B, L, C, device=dev, dtype=torch.bfloat16)
dbg = torch.zeros(32 * 16, device=dev, dtype=torch.int64)
run = lambda: _lib.check(lib.b200_attention_tc(_lib.ptr(qkv), _lib.ptr(out), B, L, C, C ** -0.5, _lib.stream()), "attn")
for _ in range(3): run()
lib.b200_attention_tc_trace.argtypes = [ctypes.c_void_p]; lib.b200_attention_tc_trace.restype = ctypes.c_int
assert lib.b200_attention_tc_trace(dbg.data_ptr()) == 0
run(); torch.cuda.synchronize()
lib.b200_attention_tc_trace(None)
d = dbg.cpu().view(16, 32)
t0 = int(d[0, 16])   # producer start
names = {0: "iss QK wait-free", 1: "iss QK start", 2: "iss QK issued", 3: "iss PV wait-P", 4: "iss PV start", 5: "iss PV issued",
         8: "simt wait S", 9: "simt S ready", 10: "simt max done", 11: "simt P free", 12: "simt P done", 13: "simt O ready",
         14: "simt stored", 17: "prod V begin", 18: "prod V end"}
for n in range(8):
    print(f"tile {n}: " + "  ".join(f"{names[k]}={int(d[n, k]) - t0}" for k in sorted(names) if int(d[n, k]) != 0))
