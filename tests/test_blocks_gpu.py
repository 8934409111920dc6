"""Score-network building-block kernels through the C ABI vs plain PyTorch fp32 references (floating-point kernels)
and the numpy oracle (upfirdn2d)."""
import numpy as np
import pytest
import torch

from oracle import upfirdn2d_ref

pytestmark = pytest.mark.gpu


def _lib():
    from b_pinn_kalman_filter_b200 import _lib
    import b_pinn_kalman_filter_b200.models.engine  # noqa: F401  registers the entry points
    return _lib


@pytest.mark.parametrize("B,K,N", [(3, 512, 700), (40, 512, 1000), (1024, 256, 512), (130, 64, 96)])
@pytest.mark.parametrize("acts", [(0, 0), (1, 0), (1, 1)])
def test_linear(cuda, B, K, N, acts):
    L = _lib()
    g = torch.Generator().manual_seed(B + K)
    x = torch.randn(B, K, generator=g).to(cuda)
    W = (torch.randn(N, K, generator=g) / K ** 0.5).to(cuda)
    b = torch.randn(N, generator=g).to(cuda)
    out = torch.empty(B, N, device=cuda)
    L.check(L.lib().b200_linear_f32(L.ptr(x), L.ptr(W), L.ptr(b), L.ptr(out), B, K, N, acts[0], acts[1], L.stream()), "linear")
    xi = torch.nn.functional.silu(x.double()) if acts[0] else x.double()
    ref = xi @ W.double().t() + b.double()
    ref = torch.nn.functional.silu(ref) if acts[1] else ref
    assert (out.double() - ref).abs().max().item() <= 2e-5 * ref.abs().max().item()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,HW,C1,C2,G", [(2, 1024, 128, 0, 32), (3, 256, 256, 256, 32), (300, 16, 256, 0, 32),
                                          (2, 4096, 128, 128, 32), (5, 64, 32, 0, 8), (1, 65536, 128, 0, 32)])
def test_groupnorm_silu(cuda, dtype, B, HW, C1, C2, G):
    L = _lib()
    lib = L.lib()
    g = torch.Generator().manual_seed(HW + C1)
    x1 = (torch.randn(B, HW, C1, generator=g) * 2 + 0.5).to(cuda, dtype)
    x2 = (torch.randn(B, HW, C2, generator=g) - 1.0).to(cuda, dtype) if C2 else None
    C = C1 + C2
    gamma = (1 + 0.1 * torch.randn(C, generator=g)).to(cuda)
    beta = (0.1 * torch.randn(C, generator=g)).to(cuda)
    dt = 0 if dtype == torch.float32 else 1
    ns = lib.b200_gn_nsplit(B, HW)
    part = torch.empty(B * ns * G * 2, device=cuda)
    out = torch.empty(B, HW, C, device=cuda, dtype=dtype)
    for act in (0, 1):
        L.check(lib.b200_gn_partial(L.ptr(x1), C1, L.ptr(x2), C2, dt, B, HW, G, ns, L.ptr(part), L.stream()), "gn_partial")
        L.check(lib.b200_gn_apply(L.ptr(x1), C1, L.ptr(x2), C2, dt, B, HW, G, ns, L.ptr(part), L.ptr(gamma), L.ptr(beta),
                                  1e-6, act, L.ptr(out), L.stream()), "gn_apply")
        xx = torch.cat([x1, x2], -1) if C2 else x1
        ref = torch.nn.functional.group_norm(xx.double().permute(0, 2, 1), G, gamma.double(), beta.double(), eps=1e-6)
        ref = ref.permute(0, 2, 1)
        ref = torch.nn.functional.silu(ref) if act else ref
        tol = 1e-5 if dtype == torch.float32 else 2 ** -7
        assert (out.double() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,HW,C1,C2,G,T1,T2", [(5, 16, 256, 0, 32, 1, 0), (3, 64, 256, 256, 32, 1, 1), (2, 64, 128, 0, 32, 2, 0),
                                               (3, 49, 256, 128, 32, 1, 2), (2, 16, 512, 0, 32, 4, 0),
                                               (300, 16, 256, 0, 32, 1, 0)])
@pytest.mark.parametrize("act", [0, 1])
def test_groupnorm_from_statistics_small_images(cuda, dtype, B, HW, C1, C2, G, T1, T2, act):
    """b200_gn_apply_stats on images of <= 64 pixels takes the thread-per-group kernel: its output must be
    BIT-identical with the block-per-image kernel's (same sums in the same order) and match torch's group_norm on
    statistics that are exact for the input."""
    L = _lib()
    from b_pinn_kalman_filter_b200.models import tc  # noqa: F401  (registers the signatures)
    lib = L.lib()
    g = torch.Generator(device="cpu").manual_seed(HW + C1 + C2 + T1)
    C = C1 + C2
    xs = [(torch.randn(B, HW, c, generator=g) * 1.5 + 0.3).to(cuda, dtype) for c in (C1, C2) if c]

    def stats(x, T):                       # [B][T][c][2]: the image's (sum, sum of squares) spread over T slots
        xf = x.float()
        parts = xf.tensor_split(T, dim=1)
        return torch.stack([torch.stack([p.sum(1), (p * p).sum(1)], -1) for p in parts], 1).contiguous()
    sts = [stats(x, T) for x, T in zip(xs, (T1, T2))]
    gamma = torch.randn(C, generator=g).to(cuda)
    beta = torch.randn(C, generator=g).to(cuda)
    outs = []
    for small in (1, 0):
        lib.b200_gn_small_enable(small)
        try:
            out = torch.full((B, HW, C), float("nan"), device=cuda, dtype=dtype)
            x2, s2 = (xs[1], sts[1]) if C2 else (None, None)
            L.check(lib.b200_gn_apply_stats(L.ptr(xs[0]), C1, L.ptr(sts[0]), T1, L.ptr(x2) if C2 else None, C2,
                                            L.ptr(s2) if C2 else None, T2, 0 if dtype == torch.float32 else 1, B, HW, G,
                                            1, L.ptr(gamma), L.ptr(beta), 1e-6, act, L.ptr(out), L.stream()), "gn_apply_stats")
            torch.cuda.synchronize()
            outs.append(out)
        finally:
            lib.b200_gn_small_enable(1)
    assert torch.isfinite(outs[0].float()).all()
    assert torch.equal(outs[0], outs[1])
    xx = torch.cat(xs, -1)
    ref = torch.nn.functional.group_norm(xx.double().permute(0, 2, 1), G, gamma.double(), beta.double(), eps=1e-6)
    ref = ref.permute(0, 2, 1)
    ref = torch.nn.functional.silu(ref) if act else ref
    tol = 2e-5 if dtype == torch.float32 else 2 ** -7
    assert (outs[0].double() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,H,W,C,up,down,pad", [(2, 32, 32, 128, 1, 2, (1, 1)), (2, 4, 4, 256, 2, 1, (2, 1)),
                                                (3, 16, 16, 256, 2, 1, (2, 1)), (2, 32, 32, 8, 1, 1, (2, 2)),
                                                (2, 8, 8, 256, 1, 1, (2, 2)), (1, 64, 64, 128, 1, 2, (1, 1)),
                                                # more up-2 shapes: one block row, ragged extents, C = 64
                                                (2, 8, 8, 128, 2, 1, (2, 1)), (2, 20, 20, 64, 2, 1, (2, 1)),
                                                (1, 128, 128, 64, 2, 1, (2, 1)),
                                                # non-square, C / 8 = 48 (not a power of two), odd extents
                                                (2, 10, 6, 384, 1, 2, (1, 1)), (2, 6, 10, 384, 2, 1, (2, 1)),
                                                (2, 7, 9, 64, 1, 1, (2, 2)), (3, 5, 5, 8, 1, 2, (1, 1)),
                                                # other paddings take the generic kernel
                                                (2, 8, 8, 64, 1, 2, (2, 2)), (2, 8, 8, 64, 2, 1, (1, 1))])
def test_upfirdn2d_nhwc(cuda, dtype, B, H, W, C, up, down, pad):
    L = _lib()
    rng = np.random.default_rng(H + C)
    x = rng.standard_normal((B, C, H, W)).astype(np.float32)
    k = upfirdn2d_ref.setup_kernel((1, 3, 3, 1)) * up * up
    xt = torch.from_numpy(x).to(cuda, dtype)
    ref = upfirdn2d_ref.upfirdn2d(xt.float().cpu().numpy(), k, up, down, pad)     # on the dtype-rounded input
    Ho, Wo = ref.shape[2:]
    xin = xt.permute(0, 2, 3, 1).contiguous()
    out = torch.empty(B, Ho, Wo, C, device=cuda, dtype=dtype)
    kt = torch.from_numpy(k).to(cuda)
    L.check(L.lib().b200_upfirdn2d_nhwc(L.ptr(xin), L.ptr(kt), L.ptr(out), 0 if dtype == torch.float32 else 1, B, H, W,
                                        C, 4, 4, up, down, pad[0], pad[1], L.stream()), "upfirdn2d_nhwc")
    got = out.float().permute(0, 3, 1, 2).cpu().numpy()
    tol = 1e-5 if dtype == torch.float32 else 2 ** -7
    assert np.abs(got - ref).max() <= tol * max(1.0, np.abs(ref).max())
