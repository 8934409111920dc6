"""Does running a chain of layers chunk-by-chunk (so that a chunk's activations stay in the 126 MB L2 between the
producing and the consuming kernel) beat running each layer over the whole batch?
    python profiles/prof_l2_chunk.py [H C nblocks]
Chain = `nblocks` BigGAN-style residual blocks at H x H, C channels: GN+SiLU -> conv3x3 -> GN+SiLU -> conv3x3(+x)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from b_pinn_kalman_filter_b200 import _lib
from b_pinn_kalman_filter_b200.models import engine  # noqa: F401
from b_pinn_kalman_filter_b200.models import tc  # noqa: F401

H = int(sys.argv[1]) if len(sys.argv) > 1 else 32
C = int(sys.argv[2]) if len(sys.argv) > 2 else 128
NB = int(sys.argv[3]) if len(sys.argv) > 3 else 4
dev = torch.device("cuda:0")
lib = _lib.lib()
P = _lib.ptr
s = _lib.stream()
HW = H * H
G = min(C // 4, 32)
BT = 1036
T = lib.b200_conv2d_tc_stat_slots(BT, H, H)
assert T == lib.b200_conv2d_tc_stat_slots(74, H, H)
es = 2
bufs = [torch.randn(BT, HW, C, device=dev).to(torch.bfloat16) for _ in range(5)]
stats = [torch.rand(BT, T, C, 2, device=dev) + 1.0 for _ in range(3)]
for st in stats:
    st[..., 1] = st[..., 0] ** 2 / (HW / T) * 1.5 + 1.0
w = (torch.randn(C, 3, 3, C, device=dev) * 0.03).to(torch.bfloat16)
bias = torch.zeros(C, device=dev)
gamma = torch.ones(C, device=dev); beta = torch.zeros(C, device=dev)
SB = HW * C * es            # bytes per sample of an activation
SS = T * C * 2 * 4          # bytes per sample of a statistics buffer


def chain(b0, B):
    X, A, Hh, Bn, O = (t.data_ptr() + b0 * SB for t in bufs)
    sX, sH, sO = (t.data_ptr() + b0 * SS for t in stats)
    ns = lib.b200_gn_nsplit(B, HW)
    for _ in range(NB):
        rc = lib.b200_gn_apply_stats(X, C, sX, T, None, 0, None, 0, 1, B, HW, G, ns, P(gamma), P(beta), 1e-6, 1, A, s)
        rc |= lib.b200_conv2d_tc(A, C, None, 0, P(w), P(bias), None, 0, None, 1.0, Hh, 0, None, sH, B, H, H, C, 3, s)
        rc |= lib.b200_gn_apply_stats(Hh, C, sH, T, None, 0, None, 0, 1, B, HW, G, ns, P(gamma), P(beta), 1e-6, 1, Bn, s)
        rc |= lib.b200_conv2d_tc(Bn, C, None, 0, P(w), P(bias), None, 0, X, 0.70710678, O, 0, None, sO, B, H, H, C, 3, s)
        assert rc == 0, rc
        X, O = O, X
        sX, sO = sO, sX


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def full():
    chain(0, BT)


def chunked(bc):
    def run():
        for b0 in range(0, BT, bc):
            chain(b0, min(bc, BT - b0))
    return run


fl = NB * 2 * 2 * BT * HW * C * C * 9
ms = timeit(full)
print("H=%d C=%d blocks=%d B=%d  full-batch layers: %.3f ms  (%.0f TFLOP/s conv-equivalent)" % (H, C, NB, BT, ms, fl / ms / 1e9))
for bc in (37, 74, 111, 148, 222, 296, 518):
    ms = timeit(chunked(bc))
    print("  chunks of %4d (%.1f MB / activation): %.3f ms  (%.0f TFLOP/s)" % (bc, bc * SB / 1e6, ms, fl / ms / 1e9))
