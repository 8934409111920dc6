"""upfirdn2d with the reference's call signature (op/upfirdn2d.py:145-156) on the sm_100a kernel.

Differentiable to any order like the reference's autograd Functions (op/upfirdn2d.py:19-142): the gradient of an
upfirdn2d is another upfirdn2d with the flipped filter and up/down swapped, so one recursive Function suffices.
"""
import torch
from torch.autograd import Function

from .. import _lib


def _launch(inp, kernel, up, down, pad):
    """inp: [planes, H, W] fp32 contiguous CUDA. Returns [planes, out_h, out_w]."""
    up_x, up_y = up
    down_x, down_y = down
    px0, px1, py0, py1 = pad
    planes, in_h, in_w = inp.shape
    kh, kw = kernel.shape
    out_h = (in_h * up_y + py0 + py1 - kh) // down_y + 1
    out_w = (in_w * up_x + px0 + px1 - kw) // down_x + 1
    if out_h <= 0 or out_w <= 0:
        raise RuntimeError("upfirdn2d: empty output (%d x %d)" % (out_h, out_w))
    out = torch.empty((planes, out_h, out_w), device=inp.device, dtype=torch.float32)
    if planes:
        rc = _lib.lib().b200_upfirdn2d_f32(_lib.ptr(inp), _lib.ptr(kernel), _lib.ptr(out), planes, in_h, in_w, kh, kw,
                                           up_x, up_y, down_x, down_y, px0, px1, py0, py1, _lib.stream())
        _lib.check(rc, "b200_upfirdn2d_f32")
    return out


class UpFirDn2d(Function):
    @staticmethod
    def forward(ctx, input, kernel, up, down, pad):
        _lib.require_cuda(input, kernel)
        batch, channel, in_h, in_w = input.shape
        dtype = input.dtype
        k = kernel.detach().to(torch.float32).contiguous()
        x = input.detach().to(torch.float32).contiguous().view(-1, in_h, in_w)
        with torch.cuda.device(input.device):
            out = _launch(x, k, up, down, pad)
        out_h, out_w = out.shape[-2:]
        up_x, up_y = up
        down_x, down_y = down
        px0, px1, py0, py1 = pad
        kh, kw = k.shape
        # padding of the adjoint op (same algebra as op/upfirdn2d.py:111-116)
        ctx.g_pad = (kw - px0 - 1, in_w * up_x - out_w * down_x + px0 - up_x + 1,
                     kh - py0 - 1, in_h * up_y - out_h * down_y + py0 - up_y + 1)
        ctx.up, ctx.down = up, down
        ctx.in_hw = (in_h, in_w)
        ctx.save_for_backward(k)
        return out.view(batch, channel, out_h, out_w).to(dtype)

    @staticmethod
    def backward(ctx, grad_output):
        (k,) = ctx.saved_tensors
        gk = torch.flip(k, [0, 1])
        # adjoint: swap up/down, flipped filter; recursion gives double-backward (op/upfirdn2d.py:62-85) for free
        g = UpFirDn2d.apply(grad_output, gk, ctx.down, ctx.up, ctx.g_pad)
        in_h, in_w = ctx.in_hw
        g = g[..., :in_h, :in_w] if (g.shape[-2] != in_h or g.shape[-1] != in_w) else g
        return g, None, None, None, None


def upfirdn2d(input, kernel, up=1, down=1, pad=(0, 0)):
    if input.device.type != "cuda":
        raise RuntimeError("b200 upfirdn2d runs on CUDA tensors only (no CPU fallback); got %s" % input.device)
    return UpFirDn2d.apply(input, kernel, (up, up), (down, down), (pad[0], pad[1], pad[0], pad[1]))
