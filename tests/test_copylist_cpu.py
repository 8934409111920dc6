"""Host side of b200_copy_batch (models/copylist.py): the descriptor table built from torch views must describe exactly the
copy the views stand for.  No GPU: the packed descriptors are decoded and executed by a few lines of numpy here, and the
result is compared with the torch expression (permute / slice / transpose / flip / scale / second source / bf16)."""
import struct

import numpy as np
import torch

from b_pinn_kalman_filter_b200 import _lib
from b_pinn_kalman_filter_b200.models.copylist import CopyList

FMT = "<QQQ4q4q4iifii"


def _execute(cl, tensors):
    """Runs every descriptor of `cl` on the CPU.  tensors: the storages the views point into (found by address)."""
    def locate(ptr, dt):
        for t in tensors:
            base, nbytes = t.data_ptr(), t.numel() * t.element_size()
            if base <= ptr < base + nbytes and {torch.float32: 0, torch.bfloat16: 1}[t.dtype] == dt:
                return t.view(-1), (ptr - base) // t.element_size()
        raise AssertionError("descriptor points outside the known storages")
    raw = bytes(cl.descs.numpy().tobytes())
    size = struct.calcsize(FMT)
    for i in range(len(raw) // size):
        f = struct.unpack(FMT, raw[i * size:(i + 1) * size])
        src, src2, dst = f[0:3]
        ss, ds, n, flip, scale, sdt, ddt = f[3:7], f[7:11], f[11:15], f[15], f[16], f[17], f[18]
        s, so = locate(src, sdt)
        s2 = locate(src2, sdt) if src2 else None
        d, do = locate(dst, ddt)
        for idx in np.ndindex(*n):
            j = [n[k] - 1 - idx[k] if flip >> k & 1 else idx[k] for k in range(4)]
            v = float(s[so + sum(j[k] * ss[k] for k in range(4))])
            if s2 is not None:
                v += float(s2[0][s2[1] + sum(j[k] * ss[k] for k in range(4))])
            d[do + sum(idx[k] * ds[k] for k in range(4))] = float(np.float32(v) * np.float32(scale))


def test_descriptor_size_matches_the_library():
    lib = _lib.lib()
    assert lib.b200_copy_desc_bytes() == struct.calcsize(FMT)
    assert lib.b200_copy_chunk() > 0


def test_descriptors_describe_the_views():
    g = torch.Generator().manual_seed(0)
    w = torch.randn(6, 5, 3, 3, generator=g)                       # conv weight [Cout, Cin, kh, kw]
    nin = torch.randn(5, 7, generator=g)                           # NIN.W [Cin, Cout]
    b1, b2 = torch.randn(6, generator=g), torch.randn(6, generator=g)
    packed = torch.zeros(8, 3, 3, 8)                               # [Cout_store][kh][kw][Cin_store], zero padded
    packed_nin = torch.zeros(8, 1, 1, 8, dtype=torch.bfloat16)
    flipped = torch.zeros(8, 3, 3, 8)                              # backward-data weight [ci][kh][kw][co]
    bias = torch.zeros(6)
    stacked = torch.zeros(6, 9 * 5 + 4)
    w2 = torch.randn(6, 4, 1, 1, generator=g)
    cl = CopyList("cpu")
    cl.add(packed[:6, :, :, :5], w.permute(0, 2, 3, 1))
    cl.add(packed_nin[:7, 0, 0, :5], nin.t())
    cl.add(flipped[:5, :, :, :6], packed[:6, :, :, :5].permute(3, 1, 2, 0), scale=0.5, flip=(1, 2))
    cl.add(bias, b1, src2=b2)
    cl.add(stacked[:, :45].unflatten(1, (3, 3, 5)), w.permute(0, 2, 3, 1))
    cl.add(stacked[:, 45:], w2.reshape(6, 4))
    cl.finalize()
    assert len(cl) == 6 and cl.nblocks >= 6 and cl.elements == 6 * 45 + 35 + 5 * 54 + 6 + 6 * 45 + 24
    _execute(cl, [w, nin, b1, b2, packed, packed_nin, flipped, bias, stacked, w2])
    assert torch.equal(packed[:6, :, :, :5], w.permute(0, 2, 3, 1)) and float(packed[6:].abs().sum()) == 0
    assert torch.equal(packed_nin[:7, 0, 0, :5], nin.t().to(torch.bfloat16))
    assert torch.equal(flipped[:5, :, :, :6], packed[:6, :, :, :5].flip(1, 2).permute(3, 1, 2, 0) * 0.5)
    assert torch.equal(bias, b1 + b2)
    assert torch.equal(stacked, torch.cat([w.permute(0, 2, 3, 1).reshape(6, 45), w2.reshape(6, 4)], 1))


def test_shape_mismatch_is_refused():
    cl = CopyList("cpu")
    try:
        cl.add(torch.zeros(3, 4), torch.zeros(4, 3))
    except ValueError:
        return
    raise AssertionError("shape mismatch accepted")
