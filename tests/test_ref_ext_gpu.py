"""Pins against the reference's OWN CUDA extensions (oracle/_ref, built by oracle/build_ref_ext.py from the unmodified
sources under /root/reference/op and shipped to the GPU box with the snapshot).

The reference has no CPU code for `ns_step` (op/ns_step.cpp:46-47: CHECK_CUDA everywhere) and none for the second
derivative of `grid_sample` (op/grid_sample.py:68 asserts CUDA), so the CPU restatements in oracle/ could not be pinned
in the authoring container.  Here, on the GPU, three things are compared on the same inputs:

    reference extension  <->  oracle restatement (oracle/ns_step_ref.c, oracle/grid_sample_ref.py)   [pins the oracle]
    reference extension  <->  libb200sde.so kernels                                                    [direct parity]

`ns_step_forward.update_velocity` is correct only at B = 1 (op/ns_step.cpp:70-72 hands non-contiguous views to kernels
that index with a dense stride): B = 1 pins the per-sample semantics, B = 3 pins `bug_compat=True`.
These tests FAIL (not skip) when a GPU is present and oracle/_ref is missing.
"""
import numpy as np
import pytest
import torch

from oracle import build_ref_ext, grid_sample_ref, ns_step_ref, upfirdn2d_ref, fused_act_ref

pytestmark = pytest.mark.gpu

DT, DX = 0.0025, 1 / 200        # pinn_kalman/simulator.py:35-36


def _ext(name):
    m = build_ref_ext.load_ext(name)
    if m is None:
        pytest.fail("oracle/_ref/%s is not built: run `python oracle/build_ref_ext.py %s` in the authoring container "
                    "before the snapshot ships" % (name, name))
    return m


def _t(a, dev):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


def _err(got, ref):
    got, ref = np.asarray(got, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    return np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30)


# ---------------------------------------------------------------- ns_step
@pytest.mark.parametrize("S", [16, 24, 64, 192])
def test_ns_step_oracle_and_kernels_match_reference_extension_b1(cuda, S):
    """B = 1 (the only batch at which the reference's update_velocity is the intended per-sample scheme)."""
    from b_pinn_kalman_filter_b200.op import ns_step
    ext = _ext("ns_step_forward")
    dens, vel, pres = ns_step_ref.synthetic_fields(1, S, seed=300 + S)
    d, v, p = _t(dens, cuda), _t(vel, cuda), _t(pres, cuda)
    ref_v = ext.update_velocity(v, p, DT, DX)
    ref_p = ext.update_pressure(p, v, DT, DX)
    ref_d = ext.update_density(d, v, DT, DX)
    torch.cuda.synchronize()
    for name, ref, ora, ours in [
        ("velocity", ref_v, ns_step_ref.update_velocity(vel, pres, DT, DX), ns_step.update_velocity(v, p, DT, DX)),
        ("pressure", ref_p, ns_step_ref.update_pressure(pres, vel, DT, DX), ns_step.update_pressure(p, v, DT, DX)),
        ("density", ref_d, ns_step_ref.update_density(dens, vel, DT, DX), ns_step.update_density(d, v, DT, DX)),
    ]:
        ref = ref.cpu().numpy()
        assert np.isfinite(ref).all(), name
        assert ref.shape == ora.shape == tuple(ours.shape), name
        # north_star: ns_step fields within 1e-5 relative (of the field scale)
        assert _err(ora, ref) <= 1e-5, ("oracle vs reference extension", name, _err(ora, ref))
        assert _err(ours.cpu().numpy(), ref) <= 1e-5, ("kernel vs reference extension", name, _err(ours.cpu().numpy(), ref))


def test_ns_step_bug_compat_matches_reference_extension_b3(cuda):
    """B = 3: the reference mixes velocity planes between samples; `bug_compat` reproduces exactly that, and
    pressure / density (which take contiguous inputs) stay per-sample."""
    from b_pinn_kalman_filter_b200.op import ns_step
    ext = _ext("ns_step_forward")
    dens, vel, pres = ns_step_ref.synthetic_fields(3, 32, seed=17)
    d, v, p = _t(dens, cuda), _t(vel, cuda), _t(pres, cuda)
    ref_v = ext.update_velocity(v, p, DT, DX).cpu().numpy()
    ora = ns_step_ref.update_velocity(vel, pres, DT, DX, bug_compat=True)
    ours = ns_step.update_velocity(v, p, DT, DX, reference_bug_compat=True).cpu().numpy()
    assert _err(ora, ref_v) <= 1e-5, _err(ora, ref_v)
    assert _err(ours, ref_v) <= 1e-5, _err(ours, ref_v)
    # the defect is real: the per-sample scheme differs from what the reference returns at B > 1
    assert _err(ns_step_ref.update_velocity(vel, pres, DT, DX), ref_v) > 1e-3
    for ref, ora, ours in [
        (ext.update_pressure(p, v, DT, DX), ns_step_ref.update_pressure(pres, vel, DT, DX), ns_step.update_pressure(p, v, DT, DX)),
        (ext.update_density(d, v, DT, DX), ns_step_ref.update_density(dens, vel, DT, DX), ns_step.update_density(d, v, DT, DX)),
    ]:
        ref = ref.cpu().numpy()
        assert _err(ora, ref) <= 1e-5 and _err(ours.cpu().numpy(), ref) <= 1e-5


def test_ns_rollout_matches_reference_extension_b1(cuda):
    """20 steps of simulator.step's loop (pinn_kalman/simulator.py:54-61) driven by the reference extension and by
    `op.ns_step.step` (one C-ABI call per step)."""
    from b_pinn_kalman_filter_b200.op import ns_step
    ext = _ext("ns_step_forward")
    dens, vel, pres = ns_step_ref.synthetic_fields(1, 48, seed=5)
    f, v, p = _t(dens, cuda), _t(vel, cuda), _t(pres, cuda)
    f2, v2, p2 = f.clone(), v.clone(), p.clone()
    for _ in range(20):
        v = ext.update_velocity(v, p, DT, DX)
        p = ext.update_pressure(p, v, DT, DX)
        f = ext.update_density(f, v, DT, DX)
        f2, v2, p2 = ns_step.step(f2, v2, p2, DT, DX)
    for got, ref in [(f2, f), (v2, v), (p2, p)]:
        assert _err(got.cpu().numpy(), ref.cpu().numpy()) <= 1e-4      # 20 compounding steps


# ---------------------------------------------------------------- grid_sample second derivative
@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("padding_mode", ["zeros", "border"])
@pytest.mark.parametrize("align_corners", [True, False])
def test_grid_sample_grad2_matches_reference_extension(cuda, dim, padding_mode, align_corners):
    """gridsample_grad2.grad2_{2d,3d} (op/grid_sample.cpp:26-56) vs the fp64 autograd restatement (pins the oracle) and
    vs b200_grid_sample_grad2 through the op's double-backward."""
    from b_pinn_kalman_filter_b200 import _lib
    ext = _ext("gridsample_grad2")
    inp, grid = grid_sample_ref.synthetic(dim, seed=11 + dim)
    gen = torch.Generator().manual_seed(3)
    out_shape = inp.shape[:2] + tuple(grid.shape[1:-1])
    go = torch.randn(out_shape, generator=gen)
    a = torch.randn(inp.shape, generator=gen)        # grad2_grad_input
    b = torch.randn(grid.shape, generator=gen)       # grad2_grad_grid
    pm = ["zeros", "border"].index(padding_mode)
    fn = ext.grad2_2d if dim == 2 else ext.grad2_3d
    ref = fn(a.to(cuda), b.to(cuda), go.to(cuda), inp.to(cuda), grid.to(cuda), bool(pm), bool(align_corners))
    torch.cuda.synchronize()
    ref = [r.cpu() for r in ref]
    ora = grid_sample_ref.grad2(a, b, go, inp, grid, padding_mode, align_corners)
    # ours, through the C ABI directly (the same call the op's double-backward makes)
    N, C = inp.shape[:2]
    if dim == 2:
        D, H, W, Do, Ho, Wo = 1, inp.shape[2], inp.shape[3], 1, grid.shape[1], grid.shape[2]
    else:
        D, H, W, Do, Ho, Wo = inp.shape[2], inp.shape[3], inp.shape[4], grid.shape[1], grid.shape[2], grid.shape[3]
    x, g, gd, ad, bd = (t.to(cuda).contiguous() for t in (inp, grid, go, a, b))
    gg_out, g_in, g_grid = torch.empty_like(gd), torch.zeros_like(x), torch.empty_like(g)
    _lib.check(_lib.lib().b200_grid_sample_grad2(dim, _lib.ptr(gg_out), _lib.ptr(g_in), _lib.ptr(g_grid), _lib.ptr(ad),
                                                 _lib.ptr(bd), _lib.ptr(gd), _lib.ptr(x), _lib.ptr(g), N, C, D, H, W, Do,
                                                 Ho, Wo, pm, int(align_corners), _lib.stream()), "b200_grid_sample_grad2")
    ours = [gg_out.cpu(), g_in.cpu(), g_grid.cpu()]
    for name, r, o, u in zip(("grad_grad_output", "grad_input", "grad_grid"), ref, ora, ours):
        assert torch.isfinite(r).all(), name
        if o is None:
            o = torch.zeros_like(r)
        # atomics into grad_input: summation order differs run to run -> 1e-5 of the tensor's scale
        assert _err(o.numpy(), r.numpy()) <= 1e-5, ("oracle vs reference extension", name, _err(o.numpy(), r.numpy()))
        assert _err(u.numpy(), r.numpy()) <= 1e-5, ("kernel vs reference extension", name, _err(u.numpy(), r.numpy()))


# ---------------------------------------------------------------- upfirdn2d / fused_bias_act: the reference's CUDA kernels
@pytest.mark.parametrize("up,down,pad", [(1, 2, (1, 1)), (2, 1, (2, 1)), (1, 1, (2, 2))])
def test_upfirdn2d_matches_reference_cuda_kernel(cuda, up, down, pad):
    """op/upfirdn2d_kernel.cu:107-207 (through upfirdn2d.cpp:12-19) at the call shapes of the C3 network."""
    from b_pinn_kalman_filter_b200 import op
    ext = _ext("upfirdn2d")
    k = upfirdn2d_ref.setup_kernel((1, 3, 3, 1)) * (up ** 2)
    x = torch.randn(4, 32, 16, 16, generator=torch.Generator().manual_seed(2)).to(cuda)
    kt = torch.from_numpy(k.astype(np.float32)).to(cuda)
    # the reference's python wrapper views [N, C, H, W] as [N*C, H, W, 1] (op/upfirdn2d.py:99)
    ref = ext.upfirdn2d(x.reshape(-1, 16, 16, 1), kt, up, up, down, down, pad[0], pad[1], pad[0], pad[1])
    ref = ref.view(4, 32, ref.shape[1], ref.shape[2])
    got = op.upfirdn2d(x, kt, up, down, pad)
    assert got.shape == ref.shape
    assert _err(got.cpu().numpy(), ref.cpu().numpy()) <= 1e-6
    ora = upfirdn2d_ref.upfirdn2d(x.cpu().numpy(), k, up, down, pad)
    assert _err(ora, ref.cpu().numpy()) <= 1e-6


def test_fused_bias_act_matches_reference_cuda_kernel(cuda):
    """op/fused_bias_act_kernel.cu:18-49 with a NON-default slope (the reference's CPU branch ignores the argument,
    op/fused_act.py:87-94, so only its CUDA kernel can pin that case)."""
    from b_pinn_kalman_filter_b200 import op
    ext = _ext("fused")
    g = torch.Generator().manual_seed(4)
    x = torch.randn(3, 10, 9, 7, generator=g).to(cuda)
    b = torch.randn(10, generator=g).to(cuda)
    empty = x.new_empty(0)
    for slope, scale in [(0.2, 2 ** 0.5), (0.05, 1.25)]:
        ref = ext.fused_bias_act(x, b, empty, 3, 0, slope, scale)
        got = op.fused_leaky_relu(x, b, slope, scale)
        assert torch.equal(got, ref)
        assert np.array_equal(fused_act_ref.fused_leaky_relu(x.cpu().numpy(), b.cpu().numpy(), slope, scale), ref.cpu().numpy())
