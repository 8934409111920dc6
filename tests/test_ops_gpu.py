"""GPU parity of the native operators (through the C ABI) against the CPU oracle and the reference-made goldens."""
import numpy as np
import pytest
import torch

from conftest import golden_path
from oracle import fused_act_ref, ns_step_ref, upfirdn2d_ref

pytestmark = pytest.mark.gpu


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


# ---------------------------------------------------------------- upfirdn2d
def test_upfirdn2d_golden(cuda):
    from b_pinn_kalman_filter_b200 import op
    g = np.load(golden_path("ops.npz"))
    for tag in sorted({k.split("_")[1] for k in g.files if k.startswith("ufd_")}):
        up, down, p0, p1 = (int(v) for v in g["ufd_%s_cfg" % tag])
        y = op.upfirdn2d(_t(g["ufd_%s_x" % tag], cuda), _t(g["ufd_%s_k" % tag], cuda), up, down, (p0, p1))
        ref = g["ufd_%s_y" % tag]
        assert tuple(y.shape) == ref.shape, tag
        np.testing.assert_allclose(y.cpu().numpy(), ref, rtol=1e-5, atol=1e-6, err_msg=tag)


NCSNPP_SHAPES = [  # (N, C, H, W, up, down, pad): the call shapes of SURVEY.md 8a-N1 (C3) at a small batch
    (2, 128, 32, 32, 1, 2, (1, 1)), (2, 256, 16, 16, 1, 2, (1, 1)), (2, 256, 8, 8, 1, 2, (1, 1)),
    (2, 256, 4, 4, 2, 1, (2, 1)), (2, 256, 8, 8, 2, 1, (2, 1)), (2, 256, 16, 16, 2, 1, (2, 1)),
    (2, 3, 32, 32, 1, 1, (2, 2)), (2, 128, 16, 16, 1, 1, (2, 2)), (2, 256, 8, 8, 1, 1, (2, 2)),
    (1, 5, 37, 19, 1, 2, (1, 1)), (1, 3, 64, 64, 2, 1, (2, 1)), (1, 2, 130, 70, 1, 1, (2, 2)),
]


@pytest.mark.parametrize("shape", NCSNPP_SHAPES)
def test_upfirdn2d_vs_oracle(cuda, shape):
    from b_pinn_kalman_filter_b200 import op
    N, C, H, W, up, down, pad = shape
    rng = np.random.default_rng(H * 131 + W)
    x = rng.standard_normal((N, C, H, W)).astype(np.float32)
    k = upfirdn2d_ref.setup_kernel((1, 3, 3, 1)) * (up ** 2)
    ref = upfirdn2d_ref.upfirdn2d(x, k, up, down, pad)
    y = op.upfirdn2d(_t(x, cuda), _t(k, cuda), up, down, pad).cpu().numpy()
    assert y.shape == ref.shape
    np.testing.assert_allclose(y, ref, rtol=1e-5, atol=2e-6)


def test_upfirdn2d_generic_filters(cuda):
    from b_pinn_kalman_filter_b200 import op
    rng = np.random.default_rng(3)
    for (kh, kw, up, down, pad) in [(3, 5, 3, 2, (1, 2)), (1, 1, 1, 1, (0, 0)), (6, 6, 2, 1, (3, 2)), (4, 4, 1, 1, (-1, 0)),
                                    (7, 2, 1, 3, (4, 1))]:
        x = rng.standard_normal((2, 3, 11, 13)).astype(np.float32)
        k = rng.standard_normal((kh, kw)).astype(np.float32)
        ref = upfirdn2d_ref.upfirdn2d(x, k, up, down, pad)
        y = op.upfirdn2d(_t(x, cuda), _t(k, cuda), up, down, pad).cpu().numpy()
        assert y.shape == ref.shape
        np.testing.assert_allclose(y, ref, rtol=1e-5, atol=1e-5)


def test_upfirdn2d_linearity_full_size(cuda):
    """Size-independent property at the C5 micro-benchmark shape: upfirdn2d is linear, and down2(const) == const."""
    from b_pinn_kalman_filter_b200 import op
    k = torch.tensor(upfirdn2d_ref.setup_kernel((1, 3, 3, 1)), device=cuda)
    a = torch.randn(8, 128, 256, 256, device=cuda)
    b = torch.randn(8, 128, 256, 256, device=cuda)
    ya, yb = op.upfirdn2d(a, k, 1, 2, (1, 1)), op.upfirdn2d(b, k, 1, 2, (1, 1))
    yab = op.upfirdn2d(a * 0.5 + b * 2.0, k, 1, 2, (1, 1))
    assert tuple(ya.shape) == (8, 128, 128, 128)
    torch.testing.assert_close(yab, ya * 0.5 + yb * 2.0, rtol=1e-4, atol=1e-5)
    ones = op.upfirdn2d(torch.ones(1, 4, 256, 256, device=cuda), k, 1, 2, (1, 1))
    torch.testing.assert_close(ones[:, :, 1:-1, 1:-1], torch.ones(1, 4, 126, 126, device=cuda))
    up = op.upfirdn2d(torch.ones(1, 4, 64, 64, device=cuda), k * 4, 2, 1, (2, 1))
    torch.testing.assert_close(up[:, :, 2:-2, 2:-2], torch.ones(1, 4, 124, 124, device=cuda))


def test_upfirdn2d_backward_is_adjoint(cuda):
    """<upfirdn(x), g> == <x, upfirdn^T(g)> and gradcheck-style double backward runs (op/upfirdn2d.py:19-85)."""
    from b_pinn_kalman_filter_b200 import op
    k = torch.tensor(upfirdn2d_ref.setup_kernel((1, 3, 3, 1)), device=cuda)
    for up, down, pad in [(1, 2, (1, 1)), (2, 1, (2, 1)), (1, 1, (2, 2))]:
        x = torch.randn(2, 3, 12, 12, device=cuda, requires_grad=True)
        y = op.upfirdn2d(x, k * up * up, up, down, pad)
        g = torch.randn_like(y).requires_grad_()
        (gx,) = torch.autograd.grad(y, x, g, create_graph=True)
        lhs = float((y.detach().double() * g.double()).sum())
        rhs = float((x.detach().double() * gx.detach().double()).sum())
        assert abs(lhs - rhs) <= 1e-4 * max(1.0, abs(lhs))
        # double backward (op/upfirdn2d.py:62-85): d<gx, w>/dg is the forward op applied to w
        w = torch.randn_like(x)
        (gg,) = torch.autograd.grad(gx, g, w)
        torch.testing.assert_close(gg, op.upfirdn2d(w, k * up * up, up, down, pad), rtol=1e-5, atol=1e-5)


def test_upfirdn2d_rejects_cpu_tensor():
    from b_pinn_kalman_filter_b200 import op
    with pytest.raises(RuntimeError):
        op.upfirdn2d(torch.zeros(1, 1, 4, 4), torch.ones(2, 2))


# ---------------------------------------------------------------- fused bias + leaky relu
def test_fused_leaky_relu_golden_and_oracle(cuda):
    from b_pinn_kalman_filter_b200 import op
    g = np.load(golden_path("ops.npz"))
    y = op.fused_leaky_relu(_t(g["lrelu_x"], cuda), _t(g["lrelu_b"], cuda))
    np.testing.assert_allclose(y.cpu().numpy(), g["lrelu_y"], rtol=1e-6, atol=1e-7)
    rng = np.random.default_rng(0)
    for shape in [(3, 7, 5, 5), (2, 128, 32, 32), (4, 6), (1, 3, 2, 2, 2)]:
        x = rng.standard_normal(shape).astype(np.float32)
        b = rng.standard_normal(shape[1]).astype(np.float32)
        for slope, scale in [(0.2, 2 ** 0.5), (0.01, 1.0)]:
            ref = fused_act_ref.fused_leaky_relu(x, b, slope, scale)
            out = op.fused_leaky_relu(_t(x, cuda), _t(b, cuda), slope, scale).cpu().numpy()
            np.testing.assert_array_equal(out, ref)


def test_fused_leaky_relu_backward(cuda):
    from b_pinn_kalman_filter_b200 import op
    x = torch.randn(2, 5, 4, 4, device=cuda, requires_grad=True)
    b = torch.randn(5, device=cuda, requires_grad=True)
    y = op.fused_leaky_relu(x, b, 0.2, 2 ** 0.5)
    gy = torch.randn_like(y)
    gx, gb = torch.autograd.grad(y, (x, b), gy)
    xr = x.detach().clone().requires_grad_()
    br = b.detach().clone().requires_grad_()
    yr = torch.nn.functional.leaky_relu(xr + br.view(1, -1, 1, 1), 0.2) * 2 ** 0.5
    gxr, gbr = torch.autograd.grad(yr, (xr, br), gy)
    torch.testing.assert_close(gx, gxr, rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(gb, gbr, rtol=1e-5, atol=1e-5)
    m = op.FusedLeakyReLU(5).to(cuda)
    assert m(x).shape == x.shape


# ---------------------------------------------------------------- ns_step
DT, DX = 0.0025, 1 / 200


@pytest.mark.parametrize("B,S", [(1, 16), (3, 24), (2, 64), (5, 33)])
def test_ns_step_vs_oracle(cuda, B, S):
    from b_pinn_kalman_filter_b200.op import ns_step
    dens, vel, pres = ns_step_ref.synthetic_fields(B, S, seed=B * 100 + S)
    d, v, p = _t(dens, cuda), _t(vel, cuda), _t(pres, cuda)
    # tolerance: 1e-5 relative to the field scale (north_star); the stencil is fp32 with double-promoted literals
    for got, ref in [
        (ns_step.update_density(d, v, DT, DX), ns_step_ref.update_density(dens, vel, DT, DX)),
        (ns_step.update_velocity(v, p, DT, DX), ns_step_ref.update_velocity(vel, pres, DT, DX)),
        (ns_step.update_pressure(p, v, DT, DX), ns_step_ref.update_pressure(pres, vel, DT, DX)),
    ]:
        got = got.cpu().numpy()
        assert got.shape == ref.shape
        err = np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30)
        assert err <= 1e-5, err


def test_ns_step_bug_compat_matches_oracle(cuda):
    from b_pinn_kalman_filter_b200.op import ns_step
    dens, vel, pres = ns_step_ref.synthetic_fields(4, 20, seed=9)
    ref = ns_step_ref.update_velocity(vel, pres, DT, DX, bug_compat=True)
    got = ns_step.update_velocity(_t(vel, cuda), _t(pres, cuda), DT, DX, reference_bug_compat=True).cpu().numpy()
    err = np.abs(got - ref).max() / np.abs(ref).max()
    assert err <= 1e-5, err
    # and it differs from the per-sample semantics for B > 1 (that is the reference's defect)
    assert np.abs(ref - ns_step_ref.update_velocity(vel, pres, DT, DX)).max() > 1e-3


def test_ns_rollout_vs_oracle(cuda):
    """20 steps of simulator.step's loop (pinn_kalman/simulator.py:54-61): velocity -> pressure -> density."""
    from b_pinn_kalman_filter_b200.op import ns_step
    dens, vel, pres = ns_step_ref.synthetic_fields(2, 32, seed=4)
    d, v, p = _t(dens, cuda), _t(vel, cuda), _t(pres, cuda)
    for _ in range(20):
        vel = ns_step_ref.update_velocity(vel, pres, DT, DX)
        pres = ns_step_ref.update_pressure(pres, vel, DT, DX)
        dens = ns_step_ref.update_density(dens, vel, DT, DX)
        v = ns_step.update_velocity(v, p, DT, DX)
        p = ns_step.update_pressure(p, v, DT, DX)
        d = ns_step.update_density(d, v, DT, DX)
    for got, ref in [(d, dens), (v, vel), (p, pres)]:
        err = np.abs(got.cpu().numpy() - ref).max() / np.abs(ref).max()
        assert err <= 1e-4, err    # 20 compounding steps


def test_ns_step_full_size_properties(cuda):
    """BASELINE config C2 size (B=256, 192x192): batch independence and uniform-field fixed point."""
    from b_pinn_kalman_filter_b200.op import ns_step
    dens, vel, pres = ns_step_ref.synthetic_fields(4, 192, seed=2)
    d = _t(np.tile(dens, (64, 1, 1, 1)), cuda)
    v = _t(np.tile(vel, (64, 1, 1, 1)), cuda)
    p = _t(np.tile(pres, (64, 1, 1, 1)), cuda)
    v1 = ns_step.update_velocity(v, p, DT, DX)
    p1 = ns_step.update_pressure(p, v1, DT, DX)
    d1 = ns_step.update_density(d, v1, DT, DX)
    for t in (v1, p1, d1):
        assert torch.isfinite(t).all()
        assert torch.equal(t[:4], t[252:256])          # replicas give identical bits
    ref = ns_step_ref.update_velocity(vel[:1], pres[:1], DT, DX)
    err = np.abs(v1[:1].cpu().numpy() - ref).max() / np.abs(ref).max()
    assert err <= 1e-5
    const = torch.full_like(d, 0.7)
    torch.testing.assert_close(ns_step.update_density(const, v, DT, DX), const, rtol=1e-6, atol=0)
