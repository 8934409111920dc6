"""op.grid_sample (forward, backward, second derivative) through the C-ABI against the oracle
(oracle/grid_sample_ref.py: ATen on the CPU for value / first derivative, fp64 double-backward of the restatement
for the second), fp32 tolerance 1e-5 relative to the result's scale (north_star)."""
import pytest
import torch

from oracle import grid_sample_ref as gs

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _rel(a, b):
    return float((a.detach().double().cpu() - b.detach().double()).abs().max() / b.detach().double().abs().max().clamp_min(1e-30))


def _fn(dim):
    from b_pinn_kalman_filter_b200.op import grid_sample
    return grid_sample.grid_sample_2d if dim == 2 else grid_sample.grid_sample_3d


@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("padding_mode", ["zeros", "border"])
@pytest.mark.parametrize("align_corners", [True, False])
def test_forward_backward_grad2_vs_oracle(cuda, dim, padding_mode, align_corners):
    inp, grid = gs.synthetic(dim, seed=dim * 10 + align_corners)
    x = inp.to(cuda).requires_grad_(True)
    g = grid.to(cuda).requires_grad_(True)
    y = _fn(dim)(x, g, padding_mode, align_corners)
    assert _rel(y, gs.forward(inp, grid, padding_mode, align_corners)) <= TOL
    gen = torch.Generator().manual_seed(7)
    go = torch.randn(y.shape, generator=gen)
    go_d = go.to(cuda).requires_grad_(True)
    gi, gg = torch.autograd.grad(y, (x, g), go_d, create_graph=True)
    gi_ref, gg_ref = gs.backward(go, inp, grid, padding_mode, align_corners)
    assert _rel(gi, gi_ref) <= TOL and _rel(gg, gg_ref) <= TOL
    # second derivative: cotangents on (grad_input, grad_grid) -> gradients wrt (grad_output, input, grid)
    a = torch.randn(inp.shape, generator=gen)
    b = torch.randn(grid.shape, generator=gen)
    L = (gi * a.to(cuda)).sum() + (gg * b.to(cuda)).sum()
    ggo, g_in, g_grid = torch.autograd.grad(L, (go_d, x, g))
    r_ggo, r_in, r_grid = gs.grad2(a, b, go, inp, grid, padding_mode, align_corners)
    assert _rel(ggo, r_ggo) <= TOL, _rel(ggo, r_ggo)
    assert _rel(g_in, r_in) <= TOL, _rel(g_in, r_in)
    assert _rel(g_grid, r_grid) <= 5 * TOL, _rel(g_grid, r_grid)


def test_project_shape_like_flownet(cuda):
    """models/flownet.py:8-25 `project`: f [B,1,H,W], grid [B,H,W,2], border padding, align_corners=True, at the PINN
    size 192 x 192; identity grid reproduces the field, and the op needs no grid gradient when grid is constant."""
    from b_pinn_kalman_filter_b200.op import grid_sample
    B, H, W = 4, 192, 192
    f = torch.randn(B, 1, H, W, device=cuda, requires_grad=True)
    ys, xs = torch.meshgrid(torch.linspace(-1, 1, H, device=cuda), torch.linspace(-1, 1, W, device=cuda), indexing="ij")
    ident = torch.stack([xs, ys], -1).expand(B, H, W, 2).contiguous()
    out = grid_sample.grid_sample_2d(f, ident, padding_mode='border', align_corners=True)
    # the source index (x + 1) / 2 * 191 carries ~191 * 2^-24 = 1e-5 px of fp32 rounding in ANY fp32 implementation
    # (ATen included); times neighbour differences of a few units that is a few 1e-5 absolute -> compare in fp64
    torch.testing.assert_close(out, f.detach(), rtol=0, atol=2e-4)
    shifted = ident + 0.013
    out = grid_sample.grid_sample_2d(f, shifted, padding_mode='border', align_corners=True)
    ref = gs.forward(f.detach().cpu().double(), shifted.cpu().double(), 'border', True)
    assert _rel(out, ref) <= 5e-5, _rel(out, ref)
    out.sum().backward()
    ref_gi, _ = gs.backward(torch.ones_like(ref), f.detach().cpu().double(), shifted.cpu().double(), 'border', True)
    assert _rel(f.grad, ref_gi) <= 5e-5, _rel(f.grad, ref_gi)


def test_rejects_cpu_and_bad_padding(cuda):
    from b_pinn_kalman_filter_b200.op import grid_sample
    with pytest.raises(RuntimeError):
        grid_sample.grid_sample_2d(torch.zeros(1, 1, 4, 4), torch.zeros(1, 2, 2, 2))
    with pytest.raises(AssertionError):
        grid_sample.grid_sample_2d(torch.zeros(1, 1, 4, 4, device=cuda), torch.zeros(1, 2, 2, 2, device=cuda), 'reflection')
