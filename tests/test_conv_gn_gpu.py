"""GroupNorm + SiLU fused into the consumer convolution's operand pipeline (b200_gn_coef + b200_conv2d_tc_gn,
csrc/conv_tc.cu GNF variants) against the two-launch form it replaces (b200_gn_apply_stats + b200_conv2d_tc) and a
plain PyTorch fp32 reference of `conv(silu(group_norm(x)))` (models/layerspp.py:242-246, 264-266).

The fused path evaluates exactly the arithmetic of gn_apply_kernel<bf16> (same fp32 statistics fold, same fma, same
tanh.approx SiLU, one rounding to bf16) on exactly the same operands, so the two must agree BIT FOR BIT."""
import pytest
import torch

import cases

pytestmark = pytest.mark.gpu

SHAPES = [  # B, H, W, C1, C2, Cout      (3x3, stride 1: the shapes the halo + pair variants take)
    (2, 32, 32, 128, 0, 128),        # BN = 128 halo + pair
    (3, 32, 32, 256, 128, 128),      # two sources (skip concat), K = 9 * 384
    (4, 16, 16, 256, 0, 256),        # BN = 256 halo + pair (too few tiles at this B: runs as BN = 128)
    (160, 16, 16, 256, 256, 256),    # BN = 256 halo + pair, more tiles than SM pairs: persistent loop, both stages
    (83, 32, 32, 128, 0, 128),       # many tiles, BN = 128
    (2, 64, 64, 128, 128, 128),      # 64 x 64 level (DDPM / C5 shapes)
    (1, 16, 16, 128, 0, 128),        # a single pair of tiles
]


def _inputs(dev, B, H, W, C1, C2, Cout, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    rn = lambda *s: torch.randn(*s, generator=g)
    # sources with per-channel offsets / scales so that the normalisation matters
    x1 = (rn(B, H, W, C1) * (0.5 + rn(C1).abs()) + rn(C1)).to(dev, torch.bfloat16)
    x2 = (rn(B, H, W, C2) * (0.5 + rn(C2).abs()) + 2 * rn(C2)).to(dev, torch.bfloat16) if C2 else None
    Ct = C1 + C2
    w = (rn(Cout, 3, 3, Ct) / (9 * Ct) ** 0.5).to(dev, torch.bfloat16)
    bias = rn(Cout).to(dev)
    temb = rn(B, Cout).to(dev)
    res = rn(B, H, W, Cout).to(dev, torch.bfloat16)
    gamma = (1 + 0.2 * rn(Ct)).to(dev)
    beta = (0.3 * rn(Ct)).to(dev)
    return x1, x2, w, bias, temb, res, gamma, beta


def _stats_of(x):
    """[B][1 slot][C][2] sum / sum of squares of a bf16 tensor: what a producing conv's epilogue leaves behind."""
    B, H, W, C = x.shape
    f = x.float().reshape(B, H * W, C)
    return torch.stack([f.sum(1), (f * f).sum(1)], -1).reshape(B, 1, C, 2).contiguous()


@pytest.mark.parametrize("shape", SHAPES)
def test_fused_gn_conv_equals_apply_then_conv(cuda, shape):
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import engine, tc  # noqa: F401  (register the signatures)
    B, H, W, C1, C2, Cout = shape
    lib = _lib.lib()
    assert lib.b200_conv2d_tc_gn_supported(C1, C2, B, H, W, Cout, 3) == 1
    x1, x2, w, bias, temb, res, gamma, beta = _inputs(cuda, B, H, W, C1, C2, Cout, seed=sum(shape))
    Ct = C1 + C2
    G = min(Ct // 4, 32)
    s1 = _stats_of(x1)
    s2 = _stats_of(x2) if C2 else None
    st = _lib.stream()
    # --- two launches: normalised tensor written, then convolved
    h = torch.empty(B, H, W, Ct, device=cuda, dtype=torch.bfloat16)
    ns = lib.b200_gn_nsplit(B, H * W)
    _lib.check(lib.b200_gn_apply_stats(_lib.ptr(x1), C1, _lib.ptr(s1), 1, _lib.ptr(x2), C2, _lib.ptr(s2), 1 if C2 else 0,
                                       1, B, H * W, G, ns, _lib.ptr(gamma), _lib.ptr(beta), 1e-6, 1, _lib.ptr(h), st),
               "b200_gn_apply_stats")
    T = lib.b200_conv2d_tc_stat_slots(B, H, W)
    out_a = torch.full((B, H, W, Cout), float("nan"), device=cuda, dtype=torch.bfloat16)
    stats_a = torch.zeros(B, T, Cout, 2, device=cuda)
    _lib.check(lib.b200_conv2d_tc(_lib.ptr(h), Ct, None, 0, _lib.ptr(w), _lib.ptr(bias), _lib.ptr(temb), Cout,
                                  _lib.ptr(res), 0.70710678, _lib.ptr(out_a), 0, None, _lib.ptr(stats_a), B, H, W, Cout, 3,
                                  st), "b200_conv2d_tc")
    # --- fused: coefficient fold + GroupNorm inside the conv
    coef = torch.empty(B, Ct, 2, device=cuda)
    _lib.check(lib.b200_gn_coef(_lib.ptr(s1), 1, C1, _lib.ptr(s2), 1 if C2 else 0, C2, B, H * W, G, _lib.ptr(gamma),
                                _lib.ptr(beta), 1e-6, _lib.ptr(coef), st), "b200_gn_coef")
    out_b = torch.full((B, H, W, Cout), float("nan"), device=cuda, dtype=torch.bfloat16)
    stats_b = torch.zeros(B, T, Cout, 2, device=cuda)
    _lib.check(lib.b200_conv2d_tc_gn(_lib.ptr(x1), C1, _lib.ptr(x2), C2, _lib.ptr(coef), _lib.ptr(w), _lib.ptr(bias),
                                     _lib.ptr(temb), Cout, _lib.ptr(res), 0.70710678, _lib.ptr(out_b), _lib.ptr(stats_b),
                                     B, H, W, Cout, 3, st), "b200_conv2d_tc_gn")
    torch.cuda.synchronize()
    assert torch.isfinite(out_b.float()).all()
    assert torch.equal(out_a, out_b), (out_a.float() - out_b.float()).abs().max().item()
    assert torch.equal(stats_a, stats_b)
    # --- and both against torch fp32 on the same operands
    xx = torch.cat([x1, x2], -1) if C2 else x1
    hn = torch.nn.functional.group_norm(xx.float().permute(0, 3, 1, 2), G, gamma, beta, eps=1e-6)
    hn = torch.nn.functional.silu(hn).to(torch.bfloat16).float()
    ref = torch.nn.functional.conv2d(hn, w.float().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    ref = (ref + bias + temb[:, None, None, :] + res.float()) * 0.70710678
    rel = ((out_b.float() - ref).norm() / ref.norm()).item()
    assert rel < 6e-3, rel


def test_fused_gn_conv_refuses_other_shapes(cuda):
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
    lib = _lib.lib()
    assert lib.b200_conv2d_tc_gn_supported(256, 0, 64, 8, 8, 256, 3) == 0       # 8 x 8: batch-spanning tiles, no halo
    assert lib.b200_conv2d_tc_gn_supported(128, 0, 4, 32, 32, 128, 1) == 0      # 1 x 1
    assert lib.b200_conv2d_tc_gn_supported(8, 0, 4, 32, 32, 128, 3) == 0        # image stem


@pytest.mark.parametrize("B", [2, 40])
def test_network_with_and_without_gn_fusion_is_bit_identical(cuda, B):
    """The whole C3 network: the plan with GroupNorm folded into its consumer convs vs the plan with standalone
    GroupNorm launches -- identical bits, far fewer GroupNorm launches (95 -> 56 in C3: the 39 GroupNorms in front of a
    3x3 convolution at 32x32 / 16x16; the 8x8 / 4x4 levels, attention and FIR-resampled blocks keep theirs)."""
    from b_pinn_kalman_filter_b200.models import utils as mutils
    cfg, sd, x, labels, out = cases.score_case("cifar10_ncsnpp_ve")
    g = torch.Generator().manual_seed(B)
    xb = (torch.rand(B, 3, 32, 32, generator=g) * 1.5 - 0.25).to(cuda)
    lb = torch.exp(torch.rand(B, generator=g) * 8.5 - 4.6).to(cuda)
    outs, counts = [], []
    for fuse in (False, "all"):
        m = mutils.ScoreModel(cfg, device=cuda, precision="bf16", init=False)
        m.load_state_dict(sd)
        m.engine.fuse_gn_conv = fuse
        outs.append(m(xb, lb))
        names = [n for _, _, n in m.engine.plan(B).ops]
        # one b200_gn_coef launch per GroupNorm-fused convolution (b200_conv2d_tc_gn, or b200_conv2d_tc_sc with gn_coef)
        counts.append((sum(n.startswith("b200_gn_apply") for n in names), names.count("b200_gn_coef")))
    assert torch.equal(outs[0], outs[1]), (outs[0] - outs[1]).abs().max().item()
    (gn0, f0), (gn1, f1) = counts
    assert f0 == 0 and f1 >= 30, counts
    assert gn1 == gn0 - f1, counts              # every fused conv removed its GroupNorm launch
    if B == 2:      # and the fused plan still reproduces the reference's golden output
        assert cases.rel_l2(m(x.to(cuda), labels.to(cuda)).cpu(), out) < 2e-2


@pytest.mark.parametrize("B,HW,C1,C2,G,T1,T2", [(3, 1024, 256, 0, 32, 8, 0), (300, 256, 256, 256, 32, 2, 2),
                                               (2, 1024, 128, 0, 32, 16, 0), (2, 64, 256, 128, 32, 1, 1)])
def test_gn_coef_thread_per_group_kernel_is_bit_identical(cuda, B, HW, C1, C2, G, T1, T2):
    """b200_gn_coef takes a thread-per-(image, 8 channels) kernel for <= 16 statistics slots: the same sums in the
    same order as the block kernel, so the coefficients must be the same bits."""
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
    lib = _lib.lib()
    g = torch.Generator(device="cpu").manual_seed(HW + C1 + T1)
    C = C1 + C2
    sts = [(torch.rand(B, T, c, 2, generator=g) * 50 + 1).to(cuda) for c, T in ((C1, T1), (C2, T2)) if c]
    for s in sts:
        s[..., 1] = s[..., 0] ** 2 / (HW / s.shape[1]) * 1.5 + 1.0          # keeps the variance positive
    gamma = torch.randn(C, generator=g).to(cuda)
    beta = torch.randn(C, generator=g).to(cuda)
    outs = []
    for small in (1, 0):
        lib.b200_gn_small_enable(small)
        try:
            coef = torch.full((B, C, 2), float("nan"), device=cuda)
            _lib.check(lib.b200_gn_coef(_lib.ptr(sts[0]), T1, C1, _lib.ptr(sts[1]) if C2 else None, T2, C2, B, HW, G,
                                        _lib.ptr(gamma), _lib.ptr(beta), 1e-6, _lib.ptr(coef), _lib.stream()), "gn_coef")
            torch.cuda.synchronize()
            outs.append(coef)
        finally:
            lib.b200_gn_small_enable(1)
    assert torch.isfinite(outs[0]).all()
    assert torch.equal(outs[0], outs[1])
