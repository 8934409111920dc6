"""fused_leaky_relu / FusedLeakyReLU with the reference's signatures (op/fused_act.py:74-97).

Unlike the reference's CPU branch (which hard-codes negative_slope=0.2, op/fused_act.py:87-94) the slope argument
is always honoured, as in the reference's CUDA branch.
"""
import torch
from torch import nn
from torch.autograd import Function

from .. import _lib


def fused_bias_act(x, bias, ref, act, grad, alpha, scale):
    """Mirror of the `fused.fused_bias_act` pybind (op/fused_bias_act.cpp:11-20): empty tensors mean "absent"."""
    _lib.require_cuda(x)
    dtype = x.dtype
    xc = x.detach().to(torch.float32).contiguous()
    b = bias.detach().to(torch.float32).contiguous() if bias is not None and bias.numel() else None
    r = ref.detach().to(torch.float32).contiguous() if ref is not None and ref.numel() else None
    step_b = 1
    for i in range(2, xc.dim()):
        step_b *= xc.size(i)
    out = torch.empty_like(xc)
    with torch.cuda.device(x.device):
        rc = _lib.lib().b200_fused_bias_act_f32(_lib.ptr(out), _lib.ptr(xc), _lib.ptr(b), _lib.ptr(r), xc.numel(),
                                                step_b, b.numel() if b is not None else 0, act, grad, float(alpha),
                                                float(scale), _lib.stream())
    _lib.check(rc, "b200_fused_bias_act_f32")
    return out.to(dtype)


class FusedLeakyReLUFunctionBackward(Function):
    @staticmethod
    def forward(ctx, grad_output, out, negative_slope, scale):
        ctx.save_for_backward(out)
        ctx.negative_slope, ctx.scale = negative_slope, scale
        grad_input = fused_bias_act(grad_output, None, out, 3, 1, negative_slope, scale)
        dim = [0] + list(range(2, grad_input.ndim))
        return grad_input, grad_input.sum(dim).detach()

    @staticmethod
    def backward(ctx, gradgrad_input, gradgrad_bias):
        (out,) = ctx.saved_tensors
        gg = fused_bias_act(gradgrad_input, gradgrad_bias, out, 3, 1, ctx.negative_slope, ctx.scale)
        return gg, None, None, None


class FusedLeakyReLUFunction(Function):
    @staticmethod
    def forward(ctx, input, bias, negative_slope, scale):
        out = fused_bias_act(input, bias, None, 3, 0, negative_slope, scale)
        ctx.save_for_backward(out)
        ctx.negative_slope, ctx.scale = negative_slope, scale
        return out

    @staticmethod
    def backward(ctx, grad_output):
        (out,) = ctx.saved_tensors
        gi, gb = FusedLeakyReLUFunctionBackward.apply(grad_output, out, ctx.negative_slope, ctx.scale)
        return gi, gb, None, None


class FusedLeakyReLU(nn.Module):
    def __init__(self, channel, negative_slope=0.2, scale=2 ** 0.5):
        super().__init__()
        self.bias = nn.Parameter(torch.zeros(channel))
        self.negative_slope = negative_slope
        self.scale = scale

    def forward(self, input):
        return fused_leaky_relu(input, self.bias, self.negative_slope, self.scale)


def fused_leaky_relu(input, bias, negative_slope=0.2, scale=2 ** 0.5):
    if input.device.type != "cuda":
        raise RuntimeError("b200 fused_leaky_relu runs on CUDA tensors only (no CPU fallback); got %s" % input.device)
    return FusedLeakyReLUFunction.apply(input, bias, negative_slope, scale)
