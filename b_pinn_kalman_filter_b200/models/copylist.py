"""Descriptor tables for b200_copy_batch (csrc/train.cu): many small strided copies, one launch.

    cl = CopyList(device)
    cl.add(dst_view, src_view, scale=1.0, flip=(), src2=None)     # torch VIEWS over persistent storage, <= 4-D
    cl.finalize()
    cl.run()                                                      # every time the sources have changed

The views are only read for their address, extents and strides: nothing is copied on the host, and the table stays valid
as long as the storages it points into do (the list keeps them alive)."""
import struct

import torch

from .. import _lib
from .._lib import c_int, c_vp

_lib.register_signature("b200_copy_desc_bytes", [])
_lib.register_signature("b200_copy_chunk", [])
_lib.register_signature("b200_copy_batch", [c_vp, c_vp, c_int, c_vp])

_DT = {torch.float32: 0, torch.bfloat16: 1}


class CopyList:
    def __init__(self, device):
        self.device = device
        self.items = []
        self.keep = []
        self.descs = self.blocks = None
        self.nblocks = 0
        self.elements = 0

    def __len__(self):
        return len(self.items)

    def add(self, dst, src, scale=1.0, flip=(), src2=None):
        """dst[...] = scale * (src[...] + src2[...]); `flip`: dimensions (of the <= 4-D shape) along which the SOURCE is
        read backwards.  src2 must have src's extents and strides."""
        if tuple(dst.shape) != tuple(src.shape):
            raise ValueError("CopyList.add: shapes differ %s vs %s" % (tuple(dst.shape), tuple(src.shape)))
        nd = dst.dim()
        if nd > 4:
            raise ValueError("CopyList.add: at most 4 dimensions")
        if dst.numel() == 0:
            return
        if src2 is not None and (tuple(src2.shape) != tuple(src.shape) or src2.stride() != src.stride() or src2.dtype != src.dtype):
            raise ValueError("CopyList.add: src2 must match src")
        pad = 4 - nd
        n = (1,) * pad + tuple(int(v) for v in dst.shape)
        ss = (0,) * pad + tuple(int(v) for v in src.stride())
        ds = (0,) * pad + tuple(int(v) for v in dst.stride())
        fl = 0
        for d in flip:
            fl |= 1 << (pad + (d % nd))
        self.items.append((src.data_ptr(), 0 if src2 is None else src2.data_ptr(), dst.data_ptr(), ss, ds, n, fl, float(scale),
                           _DT[src.dtype], _DT[dst.dtype]))
        self.keep += [dst, src] + ([src2] if src2 is not None else [])

    def finalize(self):
        lib = _lib.lib()
        size, chunk = lib.b200_copy_desc_bytes(), lib.b200_copy_chunk()
        fmt = "<QQQ4q4q4iifii"
        assert struct.calcsize(fmt) == size, (struct.calcsize(fmt), size)
        raw = bytearray()
        blocks = []
        self.elements = 0
        for i, (s, s2, d, ss, ds, n, fl, sc, sdt, ddt) in enumerate(self.items):
            raw += struct.pack(fmt, s, s2, d, *ss, *ds, *n, fl, sc, sdt, ddt)
            total = n[0] * n[1] * n[2] * n[3]
            self.elements += total
            blocks += [(i, c) for c in range((total + chunk - 1) // chunk)]
        self.nblocks = len(blocks)
        if self.nblocks:
            self.descs = torch.frombuffer(raw, dtype=torch.uint8).clone().to(self.device)
            self.blocks = torch.tensor(blocks, dtype=torch.int32).to(self.device)
        return self

    def run(self):
        if self.nblocks:
            _lib.check(_lib.lib().b200_copy_batch(_lib.ptr(self.descs), _lib.ptr(self.blocks), self.nblocks, _lib.stream()),
                       "b200_copy_batch")
