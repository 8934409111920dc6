"""tcgen05 implicit-GEMM convolution (csrc/conv_tc.cu) through the C ABI vs a plain PyTorch fp32 reference of the same
op on the same bf16-rounded operands.  Tolerance: bf16 output rounding (2^-8 relative) + fp32 accumulation order."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [  # B, H, W, C1, C2, Cout, k
    (4, 32, 32, 128, 0, 128, 3),
    (3, 16, 16, 256, 256, 256, 3),
    (5, 8, 8, 256, 0, 256, 3),       # B not a multiple of the 2-image tile
    (9, 4, 4, 256, 256, 256, 1),     # 8-image tiles, ragged batch
    (2, 32, 32, 128, 0, 768, 1),     # fused q,k,v projection: three N tiles
    (2, 64, 64, 128, 128, 128, 3),   # DDPM 64x64 level, two-row tiles
    (1, 128, 128, 64, 0, 128, 3),    # one tile per image row
    (1, 16, 256, 64, 0, 128, 3),     # two tiles per image row
    (160, 16, 16, 256, 128, 256, 3),  # more tiles than SMs: persistent loop + both TMEM buffers
    (83, 32, 32, 128, 128, 128, 3),   # Cout = 128 with many tiles: paired M tiles per CTA (MT = 2)
    (1190, 8, 8, 128, 0, 128, 1),     # MT = 2 with an odd number of 2-image M tiles (last group half empty)
    (297, 8, 8, 256, 0, 256, 3),      # CTA-pair variant (cta_group::2), odd number of M tiles: the last pair is half empty
    (1186, 4, 4, 256, 256, 256, 3),   # CTA-pair variant, 8-image tiles, ragged batch, two sources
    (152, 16, 16, 256, 0, 512, 3),    # CTA-pair variant, two 256-wide N tiles
]


def _run(dev, B, H, W, C1, C2, Cout, k, with_bias=True, with_temb=True, with_res=True, scale=0.70710678):
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401  registers the signatures
    g = torch.Generator(device="cpu").manual_seed(B * 1000 + H + C1 + Cout + k)
    rn = lambda *s: torch.randn(*s, generator=g)
    x1 = rn(B, H, W, C1).to(dev, torch.bfloat16)
    x2 = rn(B, H, W, C2).to(dev, torch.bfloat16) if C2 else None
    Ct = C1 + C2
    w = (rn(Cout, k, k, Ct) / (k * k * Ct) ** 0.5).to(dev, torch.bfloat16)
    bias = rn(Cout).to(dev) if with_bias else None
    temb = rn(B, Cout + 64).to(dev) if with_temb else None   # a column slice of a wider [B, sum(out_ch)] table
    res = rn(B, H, W, Cout).to(dev, torch.bfloat16) if with_res else None
    out = torch.full((B, H, W, Cout), float("nan"), device=dev, dtype=torch.bfloat16)
    lib = _lib.lib()
    assert lib.b200_conv2d_tc_supported(C1, C2, B, H, W, Cout, k, 1, 0) == 1
    T = lib.b200_conv2d_tc_stat_slots(B, H, W)
    stats = torch.full((B, T, Cout, 2), float("nan"), device=dev)
    stats_ptr = _lib.ptr(stats)
    rc = lib.b200_conv2d_tc(_lib.ptr(x1), C1, _lib.ptr(x2), C2, _lib.ptr(w), _lib.ptr(bias),
                            (temb.data_ptr() + 4 * 32) if with_temb else None, Cout + 64, _lib.ptr(res), scale,
                            _lib.ptr(out), 0, None, stats_ptr, B, H, W, Cout, k, _lib.stream())
    _lib.check(rc, "b200_conv2d_tc")
    torch.cuda.synchronize()
    xx = torch.cat([x1, x2], -1) if C2 else x1
    ref = torch.nn.functional.conv2d(xx.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), padding=k // 2)
    ref = ref.permute(0, 2, 3, 1)
    if with_bias:
        ref = ref + bias
    if with_temb:
        ref = ref + temb[:, None, None, 32:32 + Cout]
    if with_res:
        ref = ref + res.float()
    ref = ref * scale
    # statistics by-product: per-channel sum / sum of squares of the bf16 output, folded over the slots
    st = stats.double().sum(1)
    o = out.double().reshape(B, H * W, Cout)
    assert torch.allclose(st[..., 0], o.sum(1), rtol=1e-4, atol=1e-2)
    assert torch.allclose(st[..., 1], (o * o).sum(1), rtol=1e-4, atol=1e-2)
    return out.float(), ref


@pytest.mark.parametrize("shape", SHAPES)
def test_conv_tc_matches_torch_fp32(cuda, shape):
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    out, ref = _run(cuda, *shape)
    assert torch.isfinite(out).all()
    err = (out - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err <= 2 ** -7 * scale, (err, scale)
    rel = ((out - ref).norm() / ref.norm()).item()
    assert rel < 4e-3, rel


def test_conv_tc_plain_epilogue(cuda):
    out, ref = _run(cuda, 2, 16, 16, 128, 0, 256, 3, with_bias=False, with_temb=False, with_res=False, scale=1.0)
    assert ((out - ref).norm() / ref.norm()).item() < 4e-3


def test_conv_tc_unsupported_shapes_are_refused(cuda):
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
    lib = _lib.lib()
    assert lib.b200_conv2d_tc_supported(8, 0, 4, 32, 32, 128, 3, 1, 0) == 0      # Cin = 8 (image stem)
    assert lib.b200_conv2d_tc_supported(128, 0, 4, 32, 32, 8, 3, 1, 0) == 0      # Cout = 8 (image head)
    assert lib.b200_conv2d_tc_supported(128, 0, 4, 33, 33, 128, 3, 2, 0) == 0    # stride 2
    assert lib.b200_conv2d_tc_supported(128, 0, 4, 24, 24, 128, 3, 1, 0) == 0    # 128 % W != 0


@pytest.mark.parametrize("B", [1, 3, 150, 75, 520])     # 2 ... 1040 tiles: 1 to 8 per persistent CTA, odd and even
def test_attention_tc_matches_torch_fp32(cuda, B):
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
    L, C = 256, 256
    g = torch.Generator(device="cpu").manual_seed(B)
    qkv = torch.randn(B, L, 3 * C, generator=g)
    qkv[:, :, :C] *= 2.0          # sharper softmax rows
    qkv = qkv.to(cuda, torch.bfloat16)
    out = torch.full((B, L, C), float("nan"), device=cuda, dtype=torch.bfloat16)
    lib = _lib.lib()
    assert lib.b200_attention_tc_supported(L, C) == 1 and lib.b200_attention_tc_supported(64, 64) == 0
    _lib.check(lib.b200_attention_tc(_lib.ptr(qkv), _lib.ptr(out), B, L, C, C ** -0.5, _lib.stream()), "attention_tc")
    torch.cuda.synchronize()
    q, k, v = (t.float() for t in qkv.split(C, dim=-1))
    ref = torch.softmax(q @ k.transpose(1, 2) * C ** -0.5, dim=-1) @ v
    assert torch.isfinite(out).all()
    rel = ((out.float() - ref).norm() / ref.norm()).item()
    assert rel < 8e-3, rel          # P and the output are rounded to bf16 (2^-9 relative each)
    assert (out.float() - ref).abs().max().item() < 2e-2 * ref.abs().max().item()


@pytest.mark.parametrize("B,H,C,Cout,k", [(3, 32, 128, 3, 3), (2, 64, 128, 1, 3), (5, 8, 256, 3, 1)])
def test_conv_tc_head_nchw(cuda, B, H, C, Cout, k):
    """Network head: conv -> NCHW fp32 scaled per sample (out_mode 1)."""
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
    g = torch.Generator(device="cpu").manual_seed(B + C)
    x = torch.randn(B, H, H, C, generator=g).to(cuda, torch.bfloat16)
    w = (torch.randn(Cout, k, k, C, generator=g) / (k * k * C) ** 0.5).to(cuda, torch.bfloat16)
    bias = torch.randn(Cout, generator=g).to(cuda)
    mul = (torch.rand(B, generator=g) + 0.5).to(cuda)
    out = torch.full((B, Cout, H, H), float("nan"), device=cuda)
    lib = _lib.lib()
    assert lib.b200_conv2d_tc_supported(C, 0, B, H, H, Cout, k, 1, 1) == 1
    _lib.check(lib.b200_conv2d_tc(_lib.ptr(x), C, None, 0, _lib.ptr(w), _lib.ptr(bias), None, 0, None, 1.0, _lib.ptr(out),
                                  1, _lib.ptr(mul), None, B, H, H, Cout, k, _lib.stream()), "conv2d_tc head")
    torch.cuda.synchronize()
    ref = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), bias, padding=k // 2)
    ref = ref * mul[:, None, None, None]
    assert (out - ref).abs().max().item() <= 1e-4 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("B,H,C,k,res", [(3, 32, 128, 3, True), (2, 64, 128, 3, False), (5, 8, 256, 3, True), (2, 256, 128, 3, True)])
def test_conv_tc_narrow_pyramid(cuda, B, H, C, k, res):
    """Image-pyramid conv of progressive='output_skip' (out_mode 2): 3 channels stored 8 wide, + residual pyramid."""
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
    g = torch.Generator(device="cpu").manual_seed(B + C + H)
    x = torch.randn(B, H, H, C, generator=g).to(cuda, torch.bfloat16)
    w = torch.zeros(8, k, k, C)
    w[:3] = torch.randn(3, k, k, C, generator=g) / (k * k * C) ** 0.5
    w = w.to(cuda, torch.bfloat16)
    bias = torch.zeros(8)
    bias[:3] = torch.randn(3, generator=g)
    bias = bias.to(cuda)
    r = torch.zeros(B, H, H, 8)
    r[..., :3] = torch.randn(B, H, H, 3, generator=g)
    r = r.to(cuda, torch.bfloat16)
    out = torch.full((B, H, H, 8), float("nan"), device=cuda, dtype=torch.bfloat16)
    lib = _lib.lib()
    assert lib.b200_conv2d_tc_supported(C, 0, B, H, H, 8, k, 1, 2) == 1
    _lib.check(lib.b200_conv2d_tc(_lib.ptr(x), C, None, 0, _lib.ptr(w), _lib.ptr(bias), None, 0, _lib.ptr(r) if res else None,
                                  0.5, _lib.ptr(out), 2, None, None, B, H, H, 8, k, _lib.stream()), "conv2d_tc narrow")
    torch.cuda.synchronize()
    ref = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w.float().permute(0, 3, 1, 2), bias, padding=k // 2)
    ref = ref.permute(0, 2, 3, 1)
    if res:
        ref = ref + r.float()
    ref = ref * 0.5
    assert torch.isfinite(out.float()).all()
    assert (out.float() - ref).abs().max().item() <= 1e-2 * max(1.0, ref.abs().max().item())   # bf16 output rounding
    assert out[..., 3:].float().abs().max().item() == 0.0


@pytest.mark.parametrize("B,Hin,C,Cout,res", [(3, 33, 64, 128, True), (5, 17, 128, 256, True), (9, 9, 256, 256, False),
                                              (2, 64, 128, 128, False)])
def test_conv_tc_stride2(cuda, B, Hin, C, Cout, res):
    """3x3 stride-2 VALID conv with zero fill beyond the input (FIR conv-downsample / DDPM downsample)."""
    from b_pinn_kalman_filter_b200 import _lib
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401
    Ho = Hin // 2
    g = torch.Generator(device="cpu").manual_seed(Hin + C)
    x = torch.randn(B, Hin, Hin, C, generator=g).to(cuda, torch.bfloat16)
    w = (torch.randn(Cout, 3, 3, C, generator=g) / (9 * C) ** 0.5).to(cuda, torch.bfloat16)
    bias = torch.randn(Cout, generator=g).to(cuda)
    r = torch.randn(B, Ho, Ho, Cout, generator=g).to(cuda, torch.bfloat16) if res else None
    out = torch.full((B, Ho, Ho, Cout), float("nan"), device=cuda, dtype=torch.bfloat16)
    lib = _lib.lib()
    assert lib.b200_conv2d_tc_s2_supported(C, B, Ho, Ho, Cout) == 1
    _lib.check(lib.b200_conv2d_tc_s2(_lib.ptr(x), C, _lib.ptr(w), _lib.ptr(bias), _lib.ptr(r), 0.5, _lib.ptr(out), None, B, Hin,
                                     Hin, Ho, Ho, Cout, _lib.stream()), "conv2d_tc_s2")
    torch.cuda.synchronize()
    xin = x.float().permute(0, 3, 1, 2)
    pad = 2 * Ho + 1 - Hin                     # even inputs: one zero row/column at the bottom/right (F.pad(0,1,0,1))
    xin = torch.nn.functional.pad(xin, (0, pad, 0, pad))
    ref = torch.nn.functional.conv2d(xin, w.float().permute(0, 3, 1, 2), bias, stride=2).permute(0, 2, 3, 1)
    if res:
        ref = ref + r.float()
    ref = ref * 0.5
    assert torch.isfinite(out).all()
    assert ((out.float() - ref).norm() / ref.norm()).item() < 4e-3
