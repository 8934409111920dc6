"""TEST INFRASTRUCTURE ONLY -- CPU oracle for op.grid_sample (reference op/grid_sample.py:15-146).

Forward and first derivative of the reference ARE ATen (`F.grid_sample`, `aten::grid_sampler_{2d,3d}_backward`,
op/grid_sample.py:38-39, 54-60), so on the CPU the oracle for those is `torch.nn.functional.grid_sample` + autograd.
The second derivative exists in the reference only as CUDA (op/grid_sample_kernel.cu:27-533; grid_sample.py:68 asserts
CUDA), so it is restated here: `sample` is bilinear / trilinear sampling written with differentiable torch ops, checked
against ATen for value and first derivative (tests/test_oracle.py), and differentiated twice by autograd in fp64.
The per-point formulas the CUDA kernel follows (weights, clip multipliers, corner masks) are the ones at
op/grid_sample_kernel.cu:92-129: floor corner + 1, products of (1 - t) / t, zero outside the input, `border` clamps
the source index (derivative 0 where clamped).
"""
import torch
import torch.nn.functional as F


def forward(input, grid, padding_mode='zeros', align_corners=True):
    """What the reference's forward computes (op/grid_sample.py:38-39, 107-108)."""
    return F.grid_sample(input, grid, mode='bilinear', padding_mode=padding_mode, align_corners=align_corners)


def backward(grad_output, input, grid, padding_mode='zeros', align_corners=True):
    """(grad_input, grad_grid) as aten::grid_sampler_*_backward returns them (op/grid_sample.py:54-60)."""
    x = input.detach().clone().requires_grad_(True)
    g = grid.detach().clone().requires_grad_(True)
    y = forward(x, g, padding_mode, align_corners)
    return torch.autograd.grad(y, (x, g), grad_output)


def sample(input, grid, padding_mode='zeros', align_corners=True):
    """Differentiable restatement (any order) of bilinear (4-D input) / trilinear (5-D input) sampling."""
    dim = grid.shape[-1]
    assert input.ndim == dim + 2 and grid.ndim == dim + 2
    N, C = input.shape[:2]
    sizes = list(input.shape[2:])[::-1]              # x indexes the last (fastest) axis
    lo, t = [], []
    for d in range(dim):
        x, size = grid[..., d], sizes[d]
        v = (x + 1) / 2 * (size - 1) if align_corners else ((x + 1) * size - 1) / 2
        if padding_mode == 'border':
            v = v.clamp(0, size - 1)
        f = torch.floor(v).detach()
        lo.append(f.long())
        t.append(v - f)
    flat = input.reshape(N, C, -1)
    out = 0
    for k in range(1 << dim):
        w, idx, stride, inb = 1, 0, 1, True
        for d in range(dim):
            bit = (k >> d) & 1
            i = lo[d] + bit
            w = w * (t[d] if bit else 1 - t[d])
            inb = inb & (i >= 0) & (i < sizes[d])
            idx = idx + i.clamp(0, sizes[d] - 1) * stride
            stride *= sizes[d]
        val = torch.gather(flat, 2, idx.reshape(N, 1, -1).expand(N, C, -1)).reshape((N, C) + tuple(grid.shape[1:-1]))
        out = out + val * (w * inb).unsqueeze(1)
    return out


def grad2(grad2_grad_input, grad2_grad_grid, grad_output, input, grid, padding_mode='zeros', align_corners=True):
    """[grad_grad_output, grad_input, grad_grid]: what gridsample_grad2.grad2_{2d,3d} returns
    (op/grid_sample.cpp:26-56): the vector-Jacobian product of the backward function, in fp64."""
    go = grad_output.detach().double().requires_grad_(True)
    x = input.detach().double().requires_grad_(True)
    g = grid.detach().double().requires_grad_(True)
    y = sample(x, g, padding_mode, align_corners)
    gi, gg = torch.autograd.grad(y, (x, g), go, create_graph=True)
    L = (gi * grad2_grad_input.double()).sum() + (gg * grad2_grad_grid.double()).sum()
    return torch.autograd.grad(L, (go, x, g), allow_unused=True)


def synthetic(dim, N=2, C=3, size=(7, 9, 5), out=(6, 8, 4), seed=0, spread=1.15):
    """Inputs whose grid reaches a little outside [-1, 1] so both padding modes are exercised."""
    gen = torch.Generator().manual_seed(seed)
    if dim == 2:
        inp = torch.randn(N, C, size[0], size[1], generator=gen)
        grid = (torch.rand(N, out[0], out[1], 2, generator=gen) * 2 - 1) * spread
    else:
        inp = torch.randn(N, C, size[2], size[0], size[1], generator=gen)
        grid = (torch.rand(N, out[2], out[0], out[1], 3, generator=gen) * 2 - 1) * spread
    return inp, grid
