"""Strict-mode convolution on the tensor cores (csrc/conv_tc32.cu: tcgen05.mma kind::tf32, three-term split with fp32
accumulation) through the C ABI, against the exact convolution (torch, fp64) of the same fp32 operands.

The bar for terms = 3 is "as good as an fp32 FMA loop": its error against fp64 must be within 2x of what torch's own fp32
convolution (cuDNN, TF32 off) leaves on the same operands, or below 1e-6 of the output scale.  terms = 1 is plain TF32."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [  # B, H, W, C1, C2, Cout, k
    (2, 32, 32, 128, 0, 128, 3),
    (3, 16, 16, 256, 128, 256, 3),      # two sources (skip concat), two N tiles
    (9, 8, 8, 256, 0, 256, 3),          # tiles span 2 images, odd batch: a half-empty tail tile
    (5, 4, 4, 256, 256, 256, 3),        # tiles span 8 images
    (4, 32, 32, 128, 0, 256, 1),        # 1x1 (NIN / attention projections)
    (2, 64, 64, 128, 32, 128, 3),       # 64 x 64 level, a 32-channel second source
    (40, 16, 16, 128, 0, 128, 3),       # more tiles than SMs: persistent loop, both accumulators, all stages
]


def _sig():
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401  (registers the signatures)
    return _lib


def _run(lib, _lib, x1, x2, w, bias, temb, res, scale, k, terms):
    B, H, W, C1 = x1.shape
    C2 = x2.shape[-1] if x2 is not None else 0
    Cout = w.shape[0]
    wp = w.permute(0, 2, 3, 1).contiguous()                   # [Cout][kh][kw][Cin]
    hi, lo = torch.empty_like(wp), torch.empty_like(wp)
    st = _lib.stream()
    _lib.check(lib.b200_split_tf32(_lib.ptr(wp), _lib.ptr(hi), _lib.ptr(lo), wp.numel(), st), "b200_split_tf32")
    assert torch.equal(hi + lo, wp) and bool(((hi.view(torch.int32) & 0x1fff) == 0).all())
    out = torch.full((B, H, W, Cout), float("nan"), device=x1.device)
    assert lib.b200_conv2d_tc32_supported(C1, C2, B, H, W, Cout, k) == 1
    _lib.check(lib.b200_conv2d_tc32(_lib.ptr(x1), C1, _lib.ptr(x2), C2, _lib.ptr(hi), _lib.ptr(lo), _lib.ptr(bias),
                                    _lib.ptr(temb), Cout if temb is not None else 0, _lib.ptr(res), scale, _lib.ptr(out),
                                    B, H, W, Cout, k, terms, st), "b200_conv2d_tc32")
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("shape", SHAPES)
def test_tc32_matches_exact_convolution(cuda, shape):
    _lib = _sig()
    lib = _lib.lib()
    B, H, W, C1, C2, Cout, k = shape
    g = torch.Generator().manual_seed(sum(shape))
    rn = lambda *s: torch.randn(*s, generator=g)
    Ct = C1 + C2
    x1 = (rn(B, H, W, C1) * (0.5 + rn(C1).abs())).to(cuda)
    x2 = (rn(B, H, W, C2) + 0.3).to(cuda) if C2 else None
    w = (rn(Cout, Ct, k, k) / (k * k * Ct) ** 0.5).to(cuda)
    bias, temb, res = rn(Cout).to(cuda), rn(B, Cout).to(cuda), rn(B, H, W, Cout).to(cuda)
    scale = 0.70710678
    xx = torch.cat([x1, x2], -1) if C2 else x1

    def ref(dt):
        y = torch.nn.functional.conv2d(xx.to(dt).permute(0, 3, 1, 2), w.to(dt), padding=k // 2).permute(0, 2, 3, 1)
        return (y + bias.to(dt) + temb.to(dt)[:, None, None, :] + res.to(dt)) * scale
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        exact = ref(torch.float64)
        e_torch = (ref(torch.float32).double() - exact).abs().max().item()
    finally:
        torch.backends.cudnn.allow_tf32 = old
    top = exact.abs().max().item()
    y3 = _run(lib, _lib, x1, x2, w, bias, temb, res, scale, k, 3)
    assert torch.isfinite(y3).all()
    e3 = (y3.double() - exact).abs().max().item()
    assert e3 <= max(2.0 * e_torch, 1e-6 * top), (e3, e_torch, top)
    y1 = _run(lib, _lib, x1, x2, w, bias, temb, res, scale, k, 1)
    e1 = (y1.double() - exact).abs().max().item()
    assert e3 < e1 < 4e-3 * top, (e1, e3, top)               # plain TF32: ~2^-11 per operand, and worse than the split


def test_tc32_refuses_unsupported_shapes(cuda):
    _lib = _sig()
    lib = _lib.lib()
    assert lib.b200_conv2d_tc32_supported(3, 0, 4, 32, 32, 128, 3) == 0        # image stem
    assert lib.b200_conv2d_tc32_supported(128, 0, 4, 32, 32, 3, 3) == 0        # head
    assert lib.b200_conv2d_tc32_supported(128, 0, 4, 32, 32, 128, 5) == 0
    assert lib.b200_conv2d_tc32_supported(128, 0, 4, 24, 24, 128, 3) == 0      # 24 does not tile 128 pixels


@pytest.mark.parametrize("name", ["cifar10_ncsnpp_ve", "nc_ddpm_vp"])
def test_strict_network_on_tensor_cores(cuda, name):
    """The full-size networks in strict mode against the reference's fp32 golden output: `fp32` (three-term tf32 split on
    the tensor cores) must hold the same 1e-5 rel-L2 bar as the CUDA-core strict mode it replaces; `tf32` (one product)
    is the reference's cudnn.allow_tf32 precision."""
    import cases
    from b_pinn_kalman_filter_b200.models import utils as mutils
    cfg, sd, x, labels, out = cases.score_case(name)
    errs = {}
    for precision, tol in (("fp32", 1e-5), ("fp32_simt", 1e-5), ("tf32", 5e-3)):
        m = mutils.ScoreModel(cfg, device=cuda, precision=precision, init=False)
        m.load_state_dict(sd)
        y = m(x.to(cuda), labels.to(cuda)).cpu()
        errs[precision] = cases.rel_l2(y, out)
        assert errs[precision] < tol, (precision, errs)
        if precision == "fp32_simt":
            assert m.engine.tc32_convs == 0
        else:
            assert m.engine.tc32_convs > 40 and m.engine.tc32_convs > 5 * m.engine.simt_convs, (
                m.engine.tc32_convs, m.engine.simt_convs)
    assert errs["fp32"] < 3 * errs["fp32_simt"] + 1e-6, errs
    assert errs["tf32"] > errs["fp32"], errs


@pytest.mark.parametrize("B,L,C", [(3, 256, 256), (2, 128, 384)])
def test_strict_attention_on_tensor_cores(cuda, B, L, C):
    """q k^T -> softmax -> p v as two batched tf32-split GEMMs (b200_attn32_prep + b200_gemm_tc32_batched +
    b200_softmax_rows_f32) against the exact attention of the same fp32 operands (models/layerspp.py:82-86)."""
    _lib = _sig()
    lib = _lib.lib()
    g = torch.Generator().manual_seed(B + L + C)
    qkv = (torch.randn(B, L, 3 * C, generator=g) * 1.3).to(cuda)
    q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]

    def ref(dt):
        w = torch.einsum("blc,bmc->blm", q.to(dt), k.to(dt)) * (int(C) ** (-0.5))
        return torch.einsum("blm,bmc->blc", torch.softmax(w, -1), v.to(dt))
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        exact = ref(torch.float64)
        e_torch = (ref(torch.float32).double() - exact).abs().max().item()
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old
    kh, kl, vh, vl = (torch.empty(B, L, C, device=cuda) for _ in range(4))
    sc = torch.empty(B, L, L, device=cuda)
    out = torch.full((B, L, C), float("nan"), device=cuda)
    st = _lib.stream()
    P = _lib.ptr
    _lib.check(lib.b200_attn32_prep(P(qkv), P(kh), P(kl), P(vh), P(vl), B, L, C, st), "prep")
    assert torch.equal(kh + kl, k) and torch.equal((vh + vl).view(B, C, L), v.transpose(1, 2))
    errs = {}
    for terms in (3, 1):
        _lib.check(lib.b200_gemm_tc32_batched(P(qkv), 3 * C, P(kh), P(kl), P(sc), B, L, L, C, int(C) ** (-0.5), terms, st), "qk")
        _lib.check(lib.b200_softmax_rows_f32(P(sc), B * L, L, st), "softmax")
        _lib.check(lib.b200_gemm_tc32_batched(P(sc), L, P(vh), P(vl), P(out), B, L, C, L, 1.0, terms, st), "pv")
        torch.cuda.synchronize()
        assert torch.isfinite(out).all()
        errs[terms] = (out.double() - exact).abs().max().item()
    top = exact.abs().max().item()
    assert errs[3] <= max(2.0 * e_torch, 1e-6 * top), (errs, e_torch, top)
    assert errs[3] < errs[1] < 1e-2 * top, (errs, top)
