"""grid_sample_2d / grid_sample_3d with the reference's signatures (op/grid_sample.py:15-22), differentiable twice.

Same autograd structure as the reference (Forward -> Backward -> grad2, op/grid_sample.py:28-146) but every stage is a
kernel of libb200sde.so (csrc/grid_sample.cu): the reference calls ATen for forward/backward and its own extension
only for the second derivative.  fp32 CUDA tensors only; no CPU fallback.
"""
import torch

from .. import _lib

_PADDING = ['zeros', 'border']


def grid_sample_2d(input, grid, padding_mode='zeros', align_corners=True):
    assert padding_mode in _PADDING
    return _GridSampleForward.apply(input, grid, _PADDING.index(padding_mode), bool(align_corners), 2)


def grid_sample_3d(input, grid, padding_mode='zeros', align_corners=True):
    assert padding_mode in _PADDING
    return _GridSampleForward.apply(input, grid, _PADDING.index(padding_mode), bool(align_corners), 3)


def _dims(input, grid, dim):
    assert input.ndim == dim + 2
    assert grid.ndim == dim + 2
    assert input.shape[0] == grid.shape[0]
    assert grid.shape[-1] == dim
    N, C = input.shape[:2]
    if dim == 2:
        return N, C, 1, input.shape[2], input.shape[3], 1, grid.shape[1], grid.shape[2]
    return N, C, input.shape[2], input.shape[3], input.shape[4], grid.shape[1], grid.shape[2], grid.shape[3]


def _prep(*ts):
    _lib.require_cuda(*ts)
    out = []
    for t in ts:
        if t is None:
            out.append(None)
            continue
        if t.dtype != torch.float32:
            raise RuntimeError("grid_sample: fp32 tensors only (got %s)" % t.dtype)
        out.append(t.detach().contiguous())
    return out


class _GridSampleForward(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input, grid, padding_mode, align_corners, dim):
        N, C, D, H, W, Do, Ho, Wo = _dims(input, grid, dim)
        x, g = _prep(input, grid)
        out = torch.empty((N, C) + tuple(grid.shape[1:-1]), device=x.device, dtype=torch.float32)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().b200_grid_sample_fwd(dim, _lib.ptr(out), _lib.ptr(x), _lib.ptr(g), N, C, D, H, W, Do,
                                                       Ho, Wo, padding_mode, int(align_corners), _lib.stream()),
                       "b200_grid_sample_fwd")
        ctx.save_for_backward(input, grid)
        ctx.cfg = (padding_mode, align_corners, dim)
        return out

    @staticmethod
    def backward(ctx, grad_output):
        input, grid = ctx.saved_tensors
        grad_input, grad_grid = _GridSampleBackward.apply(grad_output, input, grid, *ctx.cfg)
        return grad_input, grad_grid, None, None, None


class _GridSampleBackward(torch.autograd.Function):
    @staticmethod
    def forward(ctx, grad_output, input, grid, padding_mode, align_corners, dim):
        N, C, D, H, W, Do, Ho, Wo = _dims(input, grid, dim)
        go, x, g = _prep(grad_output, input, grid)
        need_in, need_grid = ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        grad_input = torch.zeros_like(x) if need_in else None
        grad_grid = torch.empty_like(g) if need_grid else None
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().b200_grid_sample_bwd(dim, _lib.ptr(grad_input), _lib.ptr(grad_grid), _lib.ptr(go),
                                                       _lib.ptr(x), _lib.ptr(g), N, C, D, H, W, Do, Ho, Wo, padding_mode,
                                                       int(align_corners), _lib.stream()), "b200_grid_sample_bwd")
        ctx.save_for_backward(grad_output, input, grid)
        ctx.cfg = (padding_mode, align_corners, dim)
        return grad_input, grad_grid

    @staticmethod
    def backward(ctx, grad2_grad_input, grad2_grad_grid):
        grad_output, input, grid = ctx.saved_tensors
        padding_mode, align_corners, dim = ctx.cfg
        N, C, D, H, W, Do, Ho, Wo = _dims(input, grid, dim)
        a, b, go, x, g = _prep(grad2_grad_input, grad2_grad_grid, grad_output, input, grid)
        need_go, need_in, need_grid = ctx.needs_input_grad[:3]
        gg_out = torch.empty_like(go) if need_go else None
        g_in = torch.zeros_like(x) if need_in else None
        g_grid = torch.empty_like(g) if need_grid else None
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().b200_grid_sample_grad2(dim, _lib.ptr(gg_out), _lib.ptr(g_in), _lib.ptr(g_grid),
                                                         _lib.ptr(a), _lib.ptr(b), _lib.ptr(go), _lib.ptr(x), _lib.ptr(g),
                                                         N, C, D, H, W, Do, Ho, Wo, padding_mode, int(align_corners),
                                                         _lib.stream()), "b200_grid_sample_grad2")
        return gg_out, g_in, g_grid, None, None, None
