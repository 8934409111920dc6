"""CPU oracle for op.upfirdn2d (test infrastructure only).

Restates the algorithm of the reference's `upfirdn2d_native` (op/upfirdn2d.py:159-200): zero-insertion upsampling,
padding (negative pad = crop), correlation with the flipped filter, decimation.  Written with numpy slicing in float64
accumulation so it can arbitrate fp32 implementations; pinned against the reference's own function in
tests/test_oracle_pins.py and by tests/golden/upfirdn2d.npz.
"""
import numpy as np


def upfirdn2d(x, k, up=1, down=1, pad=(0, 0)):
    return upfirdn2d_xy(x, k, up, up, down, down, pad[0], pad[1], pad[0], pad[1])


def upfirdn2d_xy(x, k, up_x, up_y, down_x, down_y, pad_x0, pad_x1, pad_y0, pad_y1):
    x = np.asarray(x)
    k = np.asarray(k, dtype=np.float64)
    n, c, in_h, in_w = x.shape
    kh, kw = k.shape
    # 1) zero insertion
    u = np.zeros((n, c, in_h * up_y, in_w * up_x), dtype=np.float64)
    u[:, :, ::up_y, ::up_x] = x
    # 2) pad / crop
    ph, pw = in_h * up_y + pad_y0 + pad_y1, in_w * up_x + pad_x0 + pad_x1
    p = np.zeros((n, c, max(ph, 0), max(pw, 0)), dtype=np.float64)
    sy0, sx0 = max(-pad_y0, 0), max(-pad_x0, 0)
    sy1, sx1 = u.shape[2] - max(-pad_y1, 0), u.shape[3] - max(-pad_x1, 0)
    dy0, dx0 = max(pad_y0, 0), max(pad_x0, 0)
    src = u[:, :, sy0:sy1, sx0:sx1]
    p[:, :, dy0:dy0 + src.shape[2], dx0:dx0 + src.shape[3]] = src
    # 3) correlate with the flipped filter ("valid")
    oh, ow = ph - kh + 1, pw - kw + 1
    kf = k[::-1, ::-1]
    acc = np.zeros((n, c, oh, ow), dtype=np.float64)
    for a in range(kh):
        for b in range(kw):
            acc += kf[a, b] * p[:, :, a:a + oh, b:b + ow]
    # 4) decimate
    out = acc[:, :, ::down_y, ::down_x]
    return out.astype(np.float32 if x.dtype != np.float64 else np.float64)


def setup_kernel(k):
    """models/up_or_down_sampling.py:181-188"""
    k = np.asarray(k, dtype=np.float32)
    if k.ndim == 1:
        k = np.outer(k, k)
    k = k / np.sum(k)
    return k


def upsample_2d(x, k=(1, 3, 3, 1), factor=2, gain=1):
    """models/up_or_down_sampling.py:195-224"""
    k = setup_kernel(k) * (gain * factor ** 2)
    p = k.shape[0] - factor
    return upfirdn2d(x, k, up=factor, pad=((p + 1) // 2 + factor - 1, p // 2))


def downsample_2d(x, k=(1, 3, 3, 1), factor=2, gain=1):
    """models/up_or_down_sampling.py:227-257"""
    k = setup_kernel(k) * gain
    p = k.shape[0] - factor
    return upfirdn2d(x, k, down=factor, pad=((p + 1) // 2, p // 2))
