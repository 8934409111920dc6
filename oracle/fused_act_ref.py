"""CPU oracle for op.fused_leaky_relu / fused_bias_act (test infrastructure only).

Restates op/fused_bias_act_kernel.cu:18-49 (all act/grad cases) in numpy.  The reference's own CPU branch
(op/fused_act.py:87-94) covers only the forward with slope 0.2; the pin test uses it for exactly that case.
"""
import numpy as np


def fused_bias_act(x, bias=None, ref=None, act=3, grad=0, alpha=0.2, scale=2 ** 0.5):
    x = np.asarray(x, dtype=np.float32)
    y = x.copy()
    if bias is not None and np.size(bias):
        shape = [1, -1] + [1] * (x.ndim - 2)
        y = y + np.asarray(bias, dtype=np.float32).reshape(shape)
    mode = act * 10 + grad
    if mode in (10, 11):
        out = y
    elif mode in (12, 32):
        out = np.zeros_like(y)
    elif mode == 30:
        out = np.where(y > 0, y, y * np.float32(alpha))
    elif mode == 31:
        out = np.where(np.asarray(ref) > 0, y, y * np.float32(alpha))
    else:
        raise ValueError(mode)
    return (out * np.float32(scale)).astype(np.float32)


def fused_leaky_relu(x, bias, negative_slope=0.2, scale=2 ** 0.5):
    return fused_bias_act(x, bias, None, 3, 0, negative_slope, scale)
